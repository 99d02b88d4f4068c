"""distributed_llm_backend_benchmark_b200 - a B200-native (sm_100a) rebuild of the capabilities of
hardik-jinda/distributed-llm-backend-benchmark: collective sweeps, a tensor-/data-parallel LLM
benchmark, a launcher and a torch.distributed backend, all on hand-written CUDA kernels that use
tcgen05/TMEM/TMA on one GPU and peer / multicast memory over NVLink 5 / NVSwitch across GPUs.

Sub-packages
    ops        tensor-level kernels (GEMM with fused epilogues, LayerNorm, GELU, MSE, Adam)
    models     the reference's LLM (models.py) as device-resident forward / backward engines
    parallel   symmetric memory, collectives, process-group backend, TP / DP engines
    utils      config, metrics, timers, result files, statistics
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401


def native_available() -> bool:
    """True when libdlbb.so is built and loadable (it is required on GPU machines)."""
    return _lib.available()
