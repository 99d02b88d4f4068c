"""Model family of the reference (models.py): the 1B / 7B / 13B pre-LN transformer with
column/row tensor-parallel linears, as a device-resident engine plus reference-named wrappers."""
from .llm import MODEL_CONFIGS, LLMConfig, LLMEngine  # noqa: F401
from .tp import (LLM, ColumnParallelLinear, RowParallelLinear, TransformerBlock, create_model,  # noqa: F401
                 create_model_from_config)
