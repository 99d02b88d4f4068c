"""Host-side utilities: configuration, metrics / timers, result files, statistics, logging."""
from .config import load_config, setup_environment, validate_tp  # noqa: F401
from .io import collect_system_info, load_json, save_json, save_results  # noqa: F401
from .metrics import (DeviceTimer, MetricsCollector, Timer, gather_metrics_from_all_ranks,  # noqa: F401
                      print_summary, rank_statistics)
