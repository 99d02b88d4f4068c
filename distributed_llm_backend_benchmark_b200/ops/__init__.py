"""Tensor-level ops.

CUDA tensors go to the hand-written sm_100a kernels in libdlbb.so (tcgen05/TMEM/TMA GEMM with fused
epilogues, LayerNorm, GELU, MSE, Adam).  CPU tensors take a plain PyTorch path that exists only so
the host-side plumbing (gloo process groups, config, metrics, result files) can be exercised on a
machine without a GPU; it is also the fp32 oracle the GPU tests compare against.  On a CUDA tensor a
missing native library raises - there is no silent fallback.
"""
from __future__ import annotations

import math

import torch

from .. import _lib

EPI_NONE, EPI_GELU, EPI_ADD, EPI_MUL_DGELU = 0, 1, 2, 3
_EPI = {"none": EPI_NONE, None: EPI_NONE, "gelu": EPI_GELU, "add": EPI_ADD, "mul_dgelu": EPI_MUL_DGELU}
_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2, torch.float64: 3}


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _row_major_2d(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise ValueError(f"{name}: expected a 2-D tensor with unit inner stride, got shape {tuple(t.shape)} "
                         f"strides {t.stride()}")


def _ld(t: torch.Tensor) -> int:
    return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)


# ------------------------------------------------------------------------------------------ GEMM
def gelu_ref(x: torch.Tensor) -> torch.Tensor:
    return 0.5 * x * (1.0 + torch.erf(x * 0.7071067811865476))


def dgelu_ref(x: torch.Tensor) -> torch.Tensor:
    cdf = 0.5 * (1.0 + torch.erf(x * 0.7071067811865476))
    pdf = torch.exp(-0.5 * x * x) * 0.3989422804014327
    return cdf + x * pdf


def matmul(a: torch.Tensor, b: torch.Tensor, *, transpose_a: bool = False, transpose_b: bool = False,
           epilogue: str | None = None, aux: torch.Tensor | None = None, out: torch.Tensor | None = None,
           force_simt: bool = False) -> torch.Tensor:
    """C = epi(op(a) @ op(b)) in bf16 with fp32 accumulation.

    ``a`` and ``b`` are 2-D views with unit inner stride (row pitch is free, so column slices of a
    wider matrix are passed without a copy - this is how the reference's ``qkv[:, :, :H/ws]`` slice,
    models.py:166-167, is consumed).  ``transpose_*`` selects the transposed operand without moving
    data: the kernel switches the UMMA descriptor between K-major and MN-major.

    epilogue: None | "gelu" (``aux`` receives the pre-activation if given) | "add" (C = acc + aux) |
    "mul_dgelu" (C = acc * gelu'(aux)).
    """
    _row_major_2d(a, "a")
    _row_major_2d(b, "b")
    M, K = (a.shape[1], a.shape[0]) if transpose_a else (a.shape[0], a.shape[1])
    Kb, N = (b.shape[1], b.shape[0]) if transpose_b else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"matmul: inner dimensions differ ({K} vs {Kb})")
    epi = _EPI[epilogue]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    _row_major_2d(out, "out")
    if tuple(out.shape) != (M, N):
        raise ValueError(f"matmul: out has shape {tuple(out.shape)}, expected {(M, N)}")
    if aux is not None:
        _row_major_2d(aux, "aux")
        if tuple(aux.shape) != (M, N):
            raise ValueError("matmul: aux must have the shape of the output")
    if epi in (EPI_ADD, EPI_MUL_DGELU) and aux is None:
        raise ValueError(f"matmul: epilogue {epilogue!r} needs aux")

    if not a.is_cuda:
        A = (a.t() if transpose_a else a).float()
        B = (b.t() if transpose_b else b).float()
        acc = A @ B
        if epi == EPI_GELU:
            if aux is not None:
                aux.copy_(acc.to(aux.dtype))
            acc = gelu_ref(acc)
        elif epi == EPI_ADD:
            acc = acc + aux.float()
        elif epi == EPI_MUL_DGELU:
            acc = acc * dgelu_ref(aux.float())
        out.copy_(acc.to(out.dtype))
        return out

    for t, n in ((a, "a"), (b, "b"), (out, "out")):
        if t.dtype != torch.bfloat16:
            raise TypeError(f"matmul: {n} must be bfloat16 on CUDA (got {t.dtype})")
    lib = _lib.load()
    fn = lib.dlbb_gemm_bf16_simt if force_simt else lib.dlbb_gemm_bf16
    rc = fn(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, _ld(a), _ld(b), _ld(out),
            1 if transpose_a else 0, 0 if transpose_b else 1, epi, _ptr(aux), _ld(aux) if aux is not None else 0,
            _stream(a))
    _lib.check(rc, "gemm_bf16")
    return out


# ------------------------------------------------------------------------------------ LayerNorm
def layernorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
                  out: torch.Tensor | None = None, save_stats: bool = True):
    """y = LN(x) over the last dim.  Returns (y, mean, rstd); stats are fp32 [rows] (or None)."""
    cols = x.shape[-1]
    rows = x.numel() // cols
    if out is None:
        out = torch.empty_like(x)
    if not x.is_cuda:
        xf = x.float().reshape(rows, cols)
        mean = xf.mean(-1)
        var = xf.var(-1, unbiased=False)
        rstd = torch.rsqrt(var + eps)
        y = (xf - mean[:, None]) * rstd[:, None] * gamma.float() + beta.float()
        out.copy_(y.reshape(x.shape).to(out.dtype))
        return out, (mean if save_stats else None), (rstd if save_stats else None)
    assert x.is_contiguous() and out.is_contiguous() and x.dtype == torch.bfloat16
    mean = torch.empty(rows, dtype=torch.float32, device=x.device) if save_stats else None
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device) if save_stats else None
    rc = _lib.load().dlbb_layernorm_fwd(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), out.data_ptr(),
                                        _ptr(mean), _ptr(rstd), rows, cols, eps, _stream(x))
    _lib.check(rc, "layernorm_fwd")
    return out, mean, rstd


_ln_scratch: dict = {}


def _scratch(device: torch.device, nbytes: int, key: str) -> torch.Tensor:
    """Per-device, per-purpose scratch buffer that only ever grows (no allocation in steady state)."""
    k = (device, key)
    buf = _ln_scratch.get(k)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(nbytes, 1), dtype=torch.uint8, device=device)
        _ln_scratch[k] = buf
    return buf


def layernorm_bwd(dy: torch.Tensor, x: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor,
                  gamma: torch.Tensor, dres: torch.Tensor | None = None, want_param_grads: bool = True,
                  dx: torch.Tensor | None = None, dgamma_out: torch.Tensor | None = None,
                  dbeta_out: torch.Tensor | None = None):
    """Returns (dx, dgamma, dbeta); dx already includes ``dres`` (the residual-branch gradient).

    ``dgamma_out`` / ``dbeta_out`` (bf16 [cols], e.g. slices of a flat gradient bucket) receive the parameter
    gradients directly from the kernel that folds the partial sums; they are then what is returned."""
    cols = x.shape[-1]
    rows = x.numel() // cols
    if dx is None:
        dx = torch.empty_like(x)
    if not x.is_cuda:
        xf, dyf = x.float().reshape(rows, cols), dy.float().reshape(rows, cols)
        xh = (xf - mean[:, None]) * rstd[:, None]
        g = dyf * gamma.float()
        c1 = g.mean(-1, keepdim=True)
        c2 = (g * xh).mean(-1, keepdim=True)
        r = rstd[:, None] * (g - c1 - xh * c2)
        if dres is not None:
            r = r + dres.float().reshape(rows, cols)
        dx.copy_(r.reshape(x.shape).to(dx.dtype))
        if dgamma_out is not None:
            dgamma_out.copy_((dyf * xh).sum(0).to(dgamma_out.dtype))
            dbeta_out.copy_(dyf.sum(0).to(dbeta_out.dtype))
            return dx, dgamma_out, dbeta_out
        if want_param_grads:
            return dx, (dyf * xh).sum(0), dyf.sum(0)
        return dx, None, None
    lib = _lib.load()
    nbytes = lib.dlbb_layernorm_bwd_scratch_bytes(rows, cols)
    scratch = _scratch(x.device, nbytes, "ln_bwd")
    if dgamma_out is not None:
        assert dbeta_out is not None and dgamma_out.dtype == torch.bfloat16 and dbeta_out.dtype == torch.bfloat16
        assert dgamma_out.is_contiguous() and dbeta_out.is_contiguous() and dgamma_out.numel() == cols
        rc = lib.dlbb_layernorm_bwd_ex(dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                       _ptr(dres), dx.data_ptr(), None, None, dgamma_out.data_ptr(), dbeta_out.data_ptr(),
                                       scratch.data_ptr(), rows, cols, _stream(x))
        _lib.check(rc, "layernorm_bwd")
        return dx, dgamma_out, dbeta_out
    dgamma = torch.empty(cols, dtype=torch.float32, device=x.device) if want_param_grads else None
    dbeta = torch.empty(cols, dtype=torch.float32, device=x.device) if want_param_grads else None
    rc = lib.dlbb_layernorm_bwd(dy.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                                _ptr(dres), dx.data_ptr(), _ptr(dgamma), _ptr(dbeta), scratch.data_ptr(), rows,
                                cols, _stream(x))
    _lib.check(rc, "layernorm_bwd")
    return dx, dgamma, dbeta


# ---------------------------------------------------------------------------------- elementwise
def gelu(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(x)
    if not x.is_cuda:
        out.copy_(gelu_ref(x.float()).to(out.dtype))
        return out
    assert x.is_contiguous() and x.dtype == torch.bfloat16
    _lib.check(_lib.load().dlbb_gelu_fwd(x.data_ptr(), out.data_ptr(), x.numel(), _stream(x)), "gelu_fwd")
    return out


def gelu_bwd(dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(x)
    if not x.is_cuda:
        out.copy_((dy.float() * dgelu_ref(x.float())).to(out.dtype))
        return out
    assert x.is_contiguous() and dy.is_contiguous() and x.dtype == torch.bfloat16
    _lib.check(_lib.load().dlbb_gelu_bwd(dy.data_ptr(), x.data_ptr(), out.data_ptr(), x.numel(), _stream(x)),
               "gelu_bwd")
    return out


def add(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(a)
    if not a.is_cuda:
        out.copy_((a.float() + b.float()).to(out.dtype))
        return out
    assert a.is_contiguous() and b.is_contiguous() and a.dtype == torch.bfloat16
    _lib.check(_lib.load().dlbb_add(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _stream(a)), "add")
    return out


def copy2d(src: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """Materialise a strided 2-D bf16 view (unit inner stride) as a contiguous tensor."""
    _row_major_2d(src, "src")
    if out is None:
        out = torch.empty(src.shape, dtype=src.dtype, device=src.device)
    if not src.is_cuda:
        out.copy_(src)
        return out
    _lib.check(_lib.load().dlbb_copy2d(src.data_ptr(), _ld(src), out.data_ptr(), _ld(out), src.shape[0],
                                       src.shape[1], _stream(src)), "copy2d")
    return out


def cast(src: torch.Tensor, dtype: torch.dtype, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty(src.shape, dtype=dtype, device=src.device)
    if not src.is_cuda:
        out.copy_(src.to(dtype))
        return out
    assert src.is_contiguous() and out.is_contiguous()
    _lib.check(_lib.load().dlbb_cast(src.data_ptr(), _DT[src.dtype], out.data_ptr(), _DT[dtype], src.numel(),
                                     _stream(src)), "cast")
    return out


# ----------------------------------------------------------------------------------------- loss
def mse_loss(y: torch.Tensor, target: torch.Tensor, want_grad: bool = True, grad_scale: float = 1.0,
             loss_out: torch.Tensor | None = None, dy: torch.Tensor | None = None):
    """Returns (loss fp32 [1], dy or None) with loss = mean((y-t)^2), dy = 2 (y-t)/n * grad_scale."""
    n = y.numel()
    if loss_out is None:
        loss_out = torch.empty(1, dtype=torch.float32, device=y.device)
    if want_grad and dy is None:
        dy = torch.empty_like(y)
    if not y.is_cuda:
        d = y.float() - target.float()
        loss_out.copy_((d * d).mean().reshape(1))
        if want_grad:
            dy.copy_((d * (2.0 / n * grad_scale)).to(dy.dtype))
        return loss_out, dy
    assert y.is_contiguous() and target.is_contiguous() and y.dtype == torch.bfloat16
    lib = _lib.load()
    scratch = _scratch(y.device, lib.dlbb_mse_scratch_bytes(n), "mse")
    rc = lib.dlbb_mse_loss(y.data_ptr(), target.data_ptr(), _ptr(dy) if want_grad else None, loss_out.data_ptr(),
                           scratch.data_ptr(), n, grad_scale, _stream(y))
    _lib.check(rc, "mse_loss")
    return loss_out, dy


# ---------------------------------------------------------------------------------------- optim
def adam_step(param: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor,
              master: torch.Tensor | None, *, lr: float, beta1: float = 0.9, beta2: float = 0.999,
              eps: float = 1e-8, weight_decay: float = 0.0, grad_scale: float = 1.0, step: int = 1) -> None:
    """In-place Adam on a flat parameter (bf16 with optional fp32 master, or plain fp32)."""
    n = param.numel()
    if not param.is_cuda:
        w = master if master is not None else param.float()
        g = grad.float() * grad_scale
        if weight_decay:
            g = g + weight_decay * w
        m.mul_(beta1).add_(g, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        bc1 = 1 - beta1 ** step
        bc2 = 1 - beta2 ** step
        denom = v.sqrt() / math.sqrt(bc2) + eps
        w = w - lr * (m / bc1) / denom
        if master is not None:
            master.copy_(w)
        param.copy_(w.to(param.dtype))
        return
    lib = _lib.load()
    if param.dtype == torch.bfloat16:
        rc = lib.dlbb_adam_bf16(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), _ptr(master), n, lr,
                                beta1, beta2, eps, weight_decay, grad_scale, step, _stream(param))
    elif param.dtype == torch.float32:
        rc = lib.dlbb_adam_f32(param.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), n, lr, beta1, beta2,
                               eps, weight_decay, grad_scale, step, _stream(param))
    else:
        raise TypeError(f"adam_step: unsupported parameter dtype {param.dtype}")
    _lib.check(rc, "adam")


def flush_l2(buf: torch.Tensor) -> None:
    """Write through a buffer larger than L2 (benchmark hygiene between timed iterations)."""
    if buf.is_cuda:
        _lib.check(_lib.load().dlbb_fill_u32(buf.data_ptr(), 0, buf.numel() * buf.element_size() // 4,
                                             _stream(buf)), "fill_u32")
    else:
        buf.zero_()
