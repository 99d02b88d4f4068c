#!/usr/bin/env python3
"""One-shot GPU self-check of every single-GPU kernel: numerics vs fp32 PyTorch and timing vs the
library GEMM.  Run under gpurun; writes gpurun_out/kernels_check.json and prints a summary."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributed_llm_backend_benchmark_b200 as dlbb
from distributed_llm_backend_benchmark_b200 import ops, _lib

dev = torch.device("cuda:0")
torch.manual_seed(0)
res = {"build": _lib.load().dlbb_build_info().decode(), "gemm": [], "perf": [], "other": {}}
ok_all = True

def relerr(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()

def gemm_case(M, N, K, ta, tb, epi, pad=0, bn=0, simt=False):
    global ok_all
    _lib.load().dlbb_gemm_set_tile_n(bn)
    a_shape = (K, M) if ta else (M, K)
    b_shape = (N, K) if tb else (K, N)
    def mk(shape, scale=1.0):
        full = torch.randn(shape[0], shape[1] + pad, device=dev, dtype=torch.float32) * scale
        return full.bfloat16()[:, :shape[1]]
    a, b = mk(a_shape), mk(b_shape, 0.5)
    A = (a.t() if ta else a).float(); B = (b.t() if tb else b).float()
    acc = A @ B
    aux = None
    if epi == "gelu":
        aux = torch.zeros(M, N + pad, device=dev, dtype=torch.bfloat16)[:, :N]
        ref = ops.gelu_ref(acc)
    elif epi == "add":
        aux = mk((M, N), 10.0); ref = acc + aux.float()
    elif epi == "mul_dgelu":
        aux = mk((M, N)); ref = acc * ops.dgelu_ref(aux.float())
    else:
        ref = acc
    out = torch.full((M, N + pad), 7.0, device=dev, dtype=torch.bfloat16)[:, :N]
    ops.matmul(a, b, transpose_a=ta, transpose_b=tb, epilogue=epi, aux=aux, out=out, force_simt=simt)
    torch.cuda.synchronize()
    e = relerr(out, ref)
    e_aux = relerr(aux, acc) if epi == "gelu" else 0.0
    # bf16 output rounding: ~2^-9 relative per element
    ok = e < 6e-3 and e_aux < 6e-3
    ok_all &= ok
    rec = dict(M=M, N=N, K=K, ta=ta, tb=tb, epi=epi, pad=pad, bn=bn, simt=simt, relerr=e, relerr_aux=e_aux, ok=ok)
    res["gemm"].append(rec)
    print(("PASS " if ok else "FAIL ") + json.dumps(rec), flush=True)

try:
    # smallest first: a failure here is the cheapest to diagnose
    gemm_case(128, 128, 64, False, True, None, bn=128)
    gemm_case(128, 256, 64, False, True, None, bn=256)
    gemm_case(128, 256, 256, False, True, None, bn=256)
    for ta in (False, True):
        for tb in (False, True):
            for bn in (128, 256):
                gemm_case(256, 512, 320, ta, tb, None, bn=bn)
    for epi in ("gelu", "add", "mul_dgelu"):
        gemm_case(384, 768, 256, False, False, epi, bn=256)
        gemm_case(384, 384, 256, True, False, epi, bn=128)
    # ragged shapes + padded row pitch (TMA zero fill / clipping)
    gemm_case(200, 136, 72, False, False, "add", pad=8)
    gemm_case(200, 136, 72, True, True, None, pad=16)
    gemm_case(1000, 1000, 1000, False, True, "gelu")
    gemm_case(130, 40, 24, True, False, None)
    # many tiles per CTA (phase wrap-around of every pipeline)
    gemm_case(4096, 4096, 512, False, False, None)
    gemm_case(2048, 8192, 1024, True, False, "add", bn=256)
    # SIMT fallback, incl. the toy shapes of the DeepSpeed smoke tests
    gemm_case(8, 20, 10, False, True, None, simt=True)
    gemm_case(33, 65, 70, True, False, "gelu", simt=True)
except Exception as e:  # keep going to the report
    ok_all = False
    res["gemm_exception"] = repr(e)
    print("EXCEPTION", repr(e), flush=True)

_lib.load().dlbb_gemm_set_tile_n(0)

def bench(fn, iters=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters

if ok_all:
    try:
        H, F, T = 4096, 16384, 4096
        shapes = [("qkv_fwd", T, 3 * H, H, False, False), ("out_fwd", T, H, H, False, False),
                  ("up_fwd", T, F, H, False, False), ("down_fwd", T, H, F, False, False),
                  ("down_dgrad", T, F, H, False, True), ("up_dgrad", T, H, F, False, True),
                  ("down_wgrad", F, H, T, True, False), ("up_wgrad", H, F, T, True, False)]
        for name, M, N, K, ta, tb in shapes:
            a = torch.randn((K, M) if ta else (M, K), device=dev).bfloat16()
            b = torch.randn((N, K) if tb else (K, N), device=dev).bfloat16()
            out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            A = a.t() if ta else a; B = b.t() if tb else b
            row = dict(name=name, M=M, N=N, K=K, ta=ta, tb=tb)
            for bn in (128, 256):
                _lib.load().dlbb_gemm_set_tile_n(bn)
                ms = bench(lambda: ops.matmul(a, b, transpose_a=ta, transpose_b=tb, out=out))
                row[f"ours_bn{bn}_ms"] = ms; row[f"ours_bn{bn}_tflops"] = 2.0 * M * N * K / ms / 1e9
            _lib.load().dlbb_gemm_set_tile_n(0)
            ms = bench(lambda: ops.matmul(a, b, transpose_a=ta, transpose_b=tb, out=out))
            row["ours_auto_ms"] = ms; row["ours_auto_tflops"] = 2.0 * M * N * K / ms / 1e9
            ms = bench(lambda: torch.matmul(A, B, out=out))
            row["cublas_ms"] = ms; row["cublas_tflops"] = 2.0 * M * N * K / ms / 1e9
            res["perf"].append(row); print("PERF", json.dumps(row), flush=True)
    except Exception as e:
        res["perf_exception"] = repr(e); print("PERF EXCEPTION", repr(e), flush=True)

# ---- LayerNorm / elementwise / loss / Adam
def check(name, got, ref, tol):
    global ok_all
    e = relerr(got, ref); ok = e < tol; ok_all &= ok
    res["other"][name] = dict(relerr=e, ok=ok); print(("PASS " if ok else "FAIL ") + name, e, flush=True)

try:
    for rows, cols in ((512, 4096), (300, 5120), (64, 2048), (17, 100), (33, 8200)):
        x = (torch.randn(rows, cols, device=dev) * 3 + 0.5).bfloat16()
        g = (torch.randn(cols, device=dev) * 0.5 + 1).bfloat16(); b = torch.randn(cols, device=dev).bfloat16()
        y, mean, rstd = ops.layernorm_fwd(x, g, b)
        yr, mr, rr = ops.layernorm_fwd(x.cpu(), g.cpu(), b.cpu())
        check(f"ln_fwd_{rows}x{cols}", y.cpu(), yr, 4e-3); check(f"ln_mean_{rows}x{cols}", mean.cpu(), mr, 1e-4)
        check(f"ln_rstd_{rows}x{cols}", rstd.cpu(), rr, 1e-4)
        dy = torch.randn(rows, cols, device=dev).bfloat16(); dres = torch.randn(rows, cols, device=dev).bfloat16()
        dx, dg, db = ops.layernorm_bwd(dy, x, mean, rstd, g, dres)
        dxr, dgr, dbr = ops.layernorm_bwd(dy.cpu(), x.cpu(), mr, rr, g.cpu(), dres.cpu())
        check(f"ln_bwd_dx_{rows}x{cols}", dx.cpu(), dxr, 6e-3); check(f"ln_bwd_dg_{rows}x{cols}", dg.cpu(), dgr, 1e-3)
        check(f"ln_bwd_db_{rows}x{cols}", db.cpu(), dbr, 1e-3)
    for n in (1 << 20, 12345):
        x = (torch.randn(n, device=dev) * 2).bfloat16(); dy = torch.randn(n, device=dev).bfloat16()
        check(f"gelu_fwd_{n}", ops.gelu(x).cpu(), ops.gelu(x.cpu()), 4e-3)
        check(f"gelu_bwd_{n}", ops.gelu_bwd(dy, x).cpu(), ops.gelu_bwd(dy.cpu(), x.cpu()), 4e-3)
        check(f"add_{n}", ops.add(x, dy).cpu(), ops.add(x.cpu(), dy.cpu()), 4e-3)
        t = torch.randn(n, device=dev).bfloat16()
        for rep in range(2):  # twice: the ticket counter must reset itself
            l, d = ops.mse_loss(x, t, grad_scale=0.5)
        lr_, dr = ops.mse_loss(x.cpu(), t.cpu(), grad_scale=0.5)
        check(f"mse_loss_{n}", l.cpu(), lr_, 1e-4); check(f"mse_grad_{n}", d.cpu(), dr, 4e-3)
        for dt in (torch.float16, torch.float32, torch.float64):
            check(f"cast_bf16_{dt}_{n}", ops.cast(x, dt).cpu(), x.cpu().to(dt), 1e-6)
        p = torch.randn(n, device=dev).bfloat16(); gr = torch.randn(n, device=dev).bfloat16()
        for use_master in (True, False):
            pc = p.clone(); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
            ms_ = pc.float() if use_master else None
            pr = p.cpu().clone(); mr_ = torch.zeros(n); vr = torch.zeros(n); msr = pr.float() if use_master else None
            for step in (1, 2, 3):
                ops.adam_step(pc, gr, m, v, ms_, lr=1e-2, weight_decay=0.01, grad_scale=0.5, step=step)
                ops.adam_step(pr, gr.cpu(), mr_, vr, msr, lr=1e-2, weight_decay=0.01, grad_scale=0.5, step=step)
            check(f"adam_p_{n}_{use_master}", pc.cpu(), pr, 4e-3); check(f"adam_m_{n}_{use_master}", m.cpu(), mr_, 1e-3)
            check(f"adam_v_{n}_{use_master}", v.cpu(), vr, 1e-3)
    src = torch.randn(64, 96, device=dev).bfloat16()
    check("copy2d", ops.copy2d(src[:, :40]).cpu(), src[:, :40].cpu(), 1e-7)
    # bandwidth of the memory-bound kernels at model shape
    rows, cols = 4096, 4096
    x = torch.randn(rows, cols, device=dev).bfloat16(); g = torch.ones(cols, device=dev).bfloat16(); b = torch.zeros(cols, device=dev).bfloat16()
    y = torch.empty_like(x)
    ms = bench(lambda: ops.layernorm_fwd(x, g, b, out=y)); res["other"]["ln_fwd_gbs"] = 2 * x.numel() * 2 / ms / 1e6
    _, mean, rstd = ops.layernorm_fwd(x, g, b, out=y)
    ms = bench(lambda: ops.layernorm_bwd(y, x, mean, rstd, g, x, dx=y)); res["other"]["ln_bwd_gbs"] = 4 * x.numel() * 2 / ms / 1e6
    n = 1 << 28
    p = torch.zeros(n, device=dev, dtype=torch.bfloat16); gr = torch.ones(n, device=dev, dtype=torch.bfloat16)
    m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev); ma = torch.zeros(n, device=dev)
    ms = bench(lambda: ops.adam_step(p, gr, m, v, ma, lr=1e-3, step=1), iters=5, warm=2)
    res["other"]["adam_gbs"] = n * 28 / ms / 1e6
    print("BW", {k: v for k, v in res["other"].items() if k.endswith("gbs")}, flush=True)
except Exception as e:
    ok_all = False; res["other_exception"] = repr(e); print("EXCEPTION", repr(e), flush=True)

res["ok"] = bool(ok_all)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/kernels_check.json", "w"), indent=1)
print("ALL OK" if ok_all else "SOME FAILED")
sys.exit(0 if ok_all else 1)
