"""dlbb_layernorm_bwd_noalloc: the LayerNorm backward for callers that are not allowed to allocate
device memory (their process pre-assigned all of it to a private pool, or they are a plugin with no
allocator of their own, or they cannot accept cudaMalloc's device-wide synchronisation).  The entry
point keeps its dgamma / dbeta partial sums in an 8 MiB array inside the library, so it must give
the same answers as dlbb_layernorm_bwd (which takes caller scratch) without any help from the
caller - also on a second call, when the array still holds the first call's partials.

The library is driven through its C ABI with ctypes, the way a non-Python host would call it.
This file sorts last on purpose: a `pytest -x` run has finished every other test before it starts.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "distributed_llm_backend_benchmark_b200", "lib", "libdlbb.so")
VP = C.c_void_p


@pytest.fixture(scope="module")
def cabi():
    lib = C.CDLL(LIB)
    lib.dlbb_last_error.restype = C.c_char_p
    lib.dlbb_layernorm_fwd.restype = C.c_int
    lib.dlbb_layernorm_fwd.argtypes = [VP, VP, VP, VP, VP, VP, C.c_int, C.c_int, C.c_float, VP]
    lib.dlbb_layernorm_bwd_scratch_bytes.restype = C.c_size_t
    lib.dlbb_layernorm_bwd_scratch_bytes.argtypes = [C.c_int, C.c_int]
    lib.dlbb_layernorm_bwd.restype = C.c_int
    lib.dlbb_layernorm_bwd.argtypes = [VP] * 10 + [C.c_int, C.c_int, VP]
    lib.dlbb_layernorm_bwd_noalloc.restype = C.c_int
    lib.dlbb_layernorm_bwd_noalloc.argtypes = [VP] * 9 + [C.c_int, C.c_int, VP]
    return lib


def relerr(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def _case(lib, rows, cols, seed, with_dres=True):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    x = (torch.randn(rows, cols, device=dev, generator=g) * 3 + 0.5).bfloat16()
    gamma = (torch.randn(cols, device=dev, generator=g) * 0.5 + 1).bfloat16()
    beta = torch.randn(cols, device=dev, generator=g).bfloat16()
    dy = torch.randn(rows, cols, device=dev, generator=g).bfloat16()
    dres = torch.randn(rows, cols, device=dev, generator=g).bfloat16() if with_dres else None
    y = torch.empty_like(x)
    mean = torch.empty(rows, device=dev)
    rstd = torch.empty(rows, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.dlbb_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), mean.data_ptr(),
                                rstd.data_ptr(), rows, cols, 1e-5, stream)
    assert rc == 0, lib.dlbb_last_error()
    dres_p = dres.data_ptr() if dres is not None else None

    # existing entry point: the caller owns the scratch
    scratch = torch.empty(lib.dlbb_layernorm_bwd_scratch_bytes(rows, cols), dtype=torch.uint8, device=dev)
    dx_a = torch.empty_like(x)
    dg_a, db_a = torch.empty(cols, device=dev), torch.empty(cols, device=dev)
    rc = lib.dlbb_layernorm_bwd(dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                dres_p, dx_a.data_ptr(), dg_a.data_ptr(), db_a.data_ptr(), scratch.data_ptr(),
                                rows, cols, stream)
    assert rc == 0, lib.dlbb_last_error()

    # new entry point: inputs and outputs only; the outputs are pre-filled so stale data cannot pass
    dx_b = torch.full_like(x, 7.0)
    dg_b, db_b = torch.full((cols,), 7.0, device=dev), torch.full((cols,), 7.0, device=dev)
    rc = lib.dlbb_layernorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                        gamma.data_ptr(), dres_p, dx_b.data_ptr(), dg_b.data_ptr(),
                                        db_b.data_ptr(), rows, cols, stream)
    assert rc == 0, lib.dlbb_last_error()
    torch.cuda.synchronize()
    # same tolerances as tests/test_gpu_kernels.py::test_layernorm
    assert relerr(dx_b, dx_a) < 6e-3
    assert relerr(dg_b, dg_a) < 1e-3
    assert relerr(db_b, db_a) < 1e-3


def test_two_calls_in_a_row_match_the_scratch_version(cabi):
    # first call: large shape fills most of the built-in array with its partials
    _case(cabi, rows=1024, cols=4096, seed=1)
    # second call, straight after, different shape and data: must not be disturbed by what is left
    _case(cabi, rows=300, cols=5120, seed=2)
    # and a third through the non-vector path (cols not a multiple of 8), fewer rows than CTAs
    _case(cabi, rows=17, cols=100, seed=3, with_dres=False)


def test_rejects_inputs_beyond_the_built_in_scratch(cabi):
    dev = torch.device("cuda:0")
    rows, cols = 4, 8200  # > 8192 columns
    z = torch.zeros(rows, cols, device=dev, dtype=torch.bfloat16)
    st = torch.zeros(rows, device=dev)
    out = torch.full((rows, cols), 3.0, device=dev, dtype=torch.bfloat16)
    gb = torch.zeros(cols, device=dev)
    rc = cabi.dlbb_layernorm_bwd_noalloc(z.data_ptr(), z.data_ptr(), st.data_ptr(), st.data_ptr(),
                                         z[0].data_ptr(), None, out.data_ptr(), gb.data_ptr(), gb.data_ptr(),
                                         rows, cols, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = cabi.dlbb_last_error().decode()
    assert "8192" in msg and "cols=8200" in msg
    torch.cuda.synchronize()
    assert (out == 3.0).all(), "a rejected call must not launch anything"
