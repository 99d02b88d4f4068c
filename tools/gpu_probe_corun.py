#!/usr/bin/env python3
"""Do the optimizer kernel and the persistent GEMM really share SMs?  Times each alone and both on
two streams, for several optimizer grid sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_llm_backend_benchmark_b200 import ops, _lib
lib = _lib.load()
dev = torch.device("cuda:0")
M, N, K = 4096, 16384, 4096
a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16(); out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
n = 200_000_000
p = torch.zeros(n, device=dev, dtype=torch.bfloat16); g = torch.ones(n, device=dev, dtype=torch.bfloat16)
m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev); ma = torch.zeros(n, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
GI, AI = 60, 4
def gemm_loop():
    for _ in range(GI): ops.matmul(a, b, out=out)
def adam_loop():
    for _ in range(AI): ops.adam_step(p, g, m, v, ma, lr=1e-3, step=1)
def timed(fn_a, fn_b):
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); ea = torch.cuda.Event(enable_timing=True); eb = torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_event(e0); s2.wait_event(e0)
    if fn_a:
        with torch.cuda.stream(s1): fn_a(); ea.record(s1)
    if fn_b:
        with torch.cuda.stream(s2): fn_b(); eb.record(s2)
    torch.cuda.synchronize()
    return (e0.elapsed_time(ea) if fn_a else 0.0, e0.elapsed_time(eb) if fn_b else 0.0)
for _ in range(2): gemm_loop(); adam_loop()
torch.cuda.synchronize()
ga, _ = timed(gemm_loop, None)
print(f"GEMM alone: {ga:.2f} ms ({2*M*N*K*GI/ga/1e9:.0f} TFLOP/s)")
for ctas in (0, 74, 148, 296, 592):
    lib.dlbb_adam_set_max_ctas(ctas)
    _, aa = timed(None, adam_loop)
    gb, ab = timed(gemm_loop, adam_loop)
    print(f"adam ctas={ctas:4d}: adam alone {aa:.2f} ms ({n*28*AI/aa/1e6:.0f} GB/s) | together: gemm {gb:.2f} adam {ab:.2f}  (sum alone {ga+aa:.2f}, max {max(ga,aa):.2f})", flush=True)
lib.dlbb_adam_set_max_ctas(0)
