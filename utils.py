"""Entry-point module with the reference's name and symbols (utils.py): metrics, timers, config and
result IO, implemented in ``distributed_llm_backend_benchmark_b200.utils``."""
from distributed_llm_backend_benchmark_b200.utils import (  # noqa: F401
    DeviceTimer, MetricsCollector, Timer, collect_system_info, gather_metrics_from_all_ranks, load_config,
    load_json, print_summary, rank_statistics, save_json, save_results, setup_environment, validate_tp)


def run_experiment(model, dataset, config, metrics_collector):
    """Warm-up + timed loop over ``model(batch)`` (ref utils.py:212-244)."""
    ex = config["execution"]
    for phase, count in (("warmup", ex["warmup_iterations"]), ("bench", ex["benchmark_iterations"])):
        for _ in range(count):
            batch = dataset.get_batch()
            with Timer() as t:
                model(batch)
            (metrics_collector.record_warmup_time if phase == "warmup" else metrics_collector.record_forward_time)(t.elapsed)
    return metrics_collector
