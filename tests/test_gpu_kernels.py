"""GPU numerics: every native kernel against a plain fp32 PyTorch statement of the same op."""
import pytest
import torch

from distributed_llm_backend_benchmark_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def relerr(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def _mk(shape, pad=0, scale=1.0):
    full = (torch.randn(shape[0], shape[1] + pad, device=DEV) * scale).bfloat16()
    return full[:, :shape[1]]


@pytest.mark.parametrize("ta", [False, True])
@pytest.mark.parametrize("tb", [False, True])
@pytest.mark.parametrize("bn", [128, 256])
def test_gemm_layouts(native_lib, ta, tb, bn):
    native_lib.dlbb_gemm_set_tile_n(bn)
    try:
        M, N, K = 384, 640, 328
        a, b = _mk((K, M) if ta else (M, K)), _mk((N, K) if tb else (K, N), scale=0.5)
        out = ops.matmul(a, b, transpose_a=ta, transpose_b=tb)
        ref = (a.t() if ta else a).float() @ (b.t() if tb else b).float()
        assert relerr(out, ref) < 6e-3
    finally:
        native_lib.dlbb_gemm_set_tile_n(0)


@pytest.mark.parametrize("ta", [False, True])
@pytest.mark.parametrize("tb", [False, True])
@pytest.mark.parametrize("shape", [(512, 768, 320), (300, 520, 200), (4096, 1024, 512)])
def test_gemm_cta_pair_matches_one_cta_bitwise(native_lib, ta, tb, shape):
    """The 256 x 256 CTA-pair kernel (tcgen05.mma.cta_group::2, the default) accumulates in the same k order
    as the one-CTA kernel, so the two must agree bit for bit - also where the pair's second CTA hangs over
    the matrix edge."""
    M, N, K = shape
    a, b = _mk((K, M) if ta else (M, K), pad=8), _mk((N, K) if tb else (K, N), scale=0.5)
    ref = (a.t() if ta else a).float() @ (b.t() if tb else b).float()
    outs = []
    try:
        for mode in (2, 1, 0):     # capped-register pair build (default), uncapped pair build, one CTA
            native_lib.dlbb_gemm_set_pair_mode(mode)
            outs.append(ops.matmul(a, b, transpose_a=ta, transpose_b=tb))
    finally:
        native_lib.dlbb_gemm_set_pair_mode(-1)
    assert relerr(outs[0], ref) < 6e-3
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[2])


@pytest.mark.parametrize("ta", [False, True])
@pytest.mark.parametrize("tb", [False, True])
@pytest.mark.parametrize("shape", [(4096, 4096, 16384), (4096, 16384, 4096), (4096, 12288, 4096)])
def test_gemm_model_shapes(native_lib, ta, tb, shape):
    """The 7B model's shapes (K = 16384: the smem ring wraps 256 k-blocks; N = 4096: 256 pair tiles = 3.46 waves with a
    partial last wave), all four operand majors, against fp32 torch.matmul on a 512-row slice spread over the M range
    (the first rows, rows that straddle a tile edge, the last rows)."""
    M, N, K = shape
    a, b = _mk((K, M) if ta else (M, K), scale=0.25), _mk((N, K) if tb else (K, N), scale=0.25)
    out = ops.matmul(a, b, transpose_a=ta, transpose_b=tb)
    rows = torch.cat([torch.arange(0, 128), torch.arange(1984, 2240), torch.arange(M - 128, M)]).to(DEV)
    A = (a.t() if ta else a)[rows].float()
    ref = A @ (b.t() if tb else b).float()
    assert relerr(out[rows], ref) < 6e-3
    assert torch.isfinite(out.float()).all()


@pytest.mark.parametrize("epi", ["gelu", "add", "mul_dgelu"])
@pytest.mark.parametrize("shape", [(256, 512, 192), (200, 136, 72), (1024, 4096, 1024)])
def test_gemm_epilogues(native_lib, epi, shape):
    M, N, K = shape
    a, b = _mk((M, K), pad=8), _mk((K, N), pad=16, scale=0.3)
    acc = a.float() @ b.float()
    if epi == "gelu":
        aux = torch.zeros(M, N, device=DEV, dtype=torch.bfloat16)
        ref = ops.gelu_ref(acc)
    elif epi == "add":
        aux = _mk((M, N), scale=8.0)
        ref = acc + aux.float()
    else:
        aux = _mk((M, N))
        ref = acc * ops.dgelu_ref(aux.float())
    out = ops.matmul(a, b, epilogue=epi, aux=aux)
    assert relerr(out, ref) < 6e-3
    if epi == "gelu":
        assert relerr(aux, acc) < 6e-3


def test_gemm_strided_output_and_slice_operand(native_lib):
    """The QKV slice (models.py:166-167) as a strided A operand, and a column block of a wider
    gradient matrix as the output."""
    T, H = 256, 128
    qkv = _mk((T, 3 * H))
    w = _mk((H, H), scale=0.2)
    out = ops.matmul(qkv[:, :H], w)
    assert relerr(out, qkv[:, :H].float() @ w.float()) < 6e-3
    big = torch.zeros(H, 3 * H, device=DEV, dtype=torch.bfloat16)
    h1, d = _mk((T, H)), _mk((T, H))
    ops.matmul(h1, d, transpose_a=True, out=big[:, :H])
    assert relerr(big[:, :H], h1.float().t() @ d.float()) < 6e-3
    assert big[:, H:].abs().max().item() == 0


def test_gemm_simt_matches_tcgen05(native_lib):
    a, b = _mk((130, 72)), _mk((72, 88))
    x = ops.matmul(a, b)
    y = ops.matmul(a, b, force_simt=True)
    assert relerr(x, y) < 4e-3
    # unaligned pitch -> automatic fallback
    a2, b2 = _mk((8, 10)), _mk((10, 20))
    assert relerr(ops.matmul(a2, b2), a2.float() @ b2.float()) < 6e-3


@pytest.mark.parametrize("rows,cols", [(512, 4096), (300, 5120), (17, 100), (33, 8200)])
def test_layernorm(native_lib, rows, cols):
    x = (torch.randn(rows, cols, device=DEV) * 3 + 0.5).bfloat16()
    g = (torch.randn(cols, device=DEV) * 0.5 + 1).bfloat16()
    b = torch.randn(cols, device=DEV).bfloat16()
    y, mean, rstd = ops.layernorm_fwd(x, g, b)
    yr, mr, rr = ops.layernorm_fwd(x.cpu(), g.cpu(), b.cpu())
    assert relerr(y, yr) < 4e-3 and relerr(mean, mr) < 1e-4 and relerr(rstd, rr) < 1e-4
    dy, dres = torch.randn(rows, cols, device=DEV).bfloat16(), torch.randn(rows, cols, device=DEV).bfloat16()
    dx, dg, db = ops.layernorm_bwd(dy, x, mean, rstd, g, dres)
    dxr, dgr, dbr = ops.layernorm_bwd(dy.cpu(), x.cpu(), mr, rr, g.cpu(), dres.cpu())
    assert relerr(dx, dxr) < 6e-3 and relerr(dg, dgr) < 1e-3 and relerr(db, dbr) < 1e-3


@pytest.mark.parametrize("n", [1 << 20, 12345])
def test_elementwise_loss_adam(native_lib, n):
    x = (torch.randn(n, device=DEV) * 2).bfloat16()
    d = torch.randn(n, device=DEV).bfloat16()
    assert relerr(ops.gelu(x), ops.gelu(x.cpu())) < 4e-3
    assert relerr(ops.gelu_bwd(d, x), ops.gelu_bwd(d.cpu(), x.cpu())) < 4e-3
    assert relerr(ops.add(x, d), ops.add(x.cpu(), d.cpu())) < 4e-3
    for _ in range(2):
        loss, g = ops.mse_loss(x, d, grad_scale=0.5)
    lr_, gr = ops.mse_loss(x.cpu(), d.cpu(), grad_scale=0.5)
    assert relerr(loss, lr_) < 1e-4 and relerr(g, gr) < 4e-3
    for dt in (torch.float16, torch.float32, torch.float64):
        assert relerr(ops.cast(x, dt), x.cpu().to(dt)) < 1e-6
    for use_master in (True, False):
        p = torch.randn(n, device=DEV).bfloat16()
        pr = p.cpu().clone()
        m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
        mr, vr = torch.zeros(n), torch.zeros(n)
        ma, mar = (p.float(), pr.float()) if use_master else (None, None)
        for step in (1, 2, 3):
            ops.adam_step(p, d, m, v, ma, lr=1e-2, weight_decay=0.01, grad_scale=0.5, step=step)
            ops.adam_step(pr, d.cpu(), mr, vr, mar, lr=1e-2, weight_decay=0.01, grad_scale=0.5, step=step)
        assert relerr(p, pr) < 4e-3 and relerr(m, mr) < 1e-3 and relerr(v, vr) < 1e-3
