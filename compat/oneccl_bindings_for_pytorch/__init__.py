"""Drop-in for Intel's ``oneccl_bindings_for_pytorch``: importing it makes ``backend='ccl'`` a valid
torch.distributed backend (ref test/ccl.py:8,24; test/test_ccl_backend.py:3; test/test_deepseed.py:24).
Here 'ccl' is served by the dlbb backend - hand-written sm_100a collectives over NVLink - instead of
oneCCL; on a CPU-only host it resolves to Gloo."""
from distributed_llm_backend_benchmark_b200.parallel.backend import register as _register

__version__ = "2.0+dlbb"
_register("ccl")
_register("dlbb")
