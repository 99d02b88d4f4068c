"""Entry-point module with the reference's name and symbols (data_gen.py): the fixed synthetic bf16
``[batch, seq, hidden]`` batch, implemented in ``distributed_llm_backend_benchmark_b200.utils.data``."""
from distributed_llm_backend_benchmark_b200.utils.data import (  # noqa: F401
    DeviceFeeder, SyntheticEmbeddingDataset, create_dataset_from_config)
