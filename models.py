"""Entry-point module with the reference's name and symbols (models.py): tensor-parallel LLM for
the forward benchmark, implemented by ``distributed_llm_backend_benchmark_b200.models`` on the
framework's own sm_100a kernels and NVLink collectives."""
from distributed_llm_backend_benchmark_b200.models import (  # noqa: F401
    LLM, MODEL_CONFIGS, ColumnParallelLinear, LLMConfig, LLMEngine, RowParallelLinear, TransformerBlock,
    create_model, create_model_from_config)
