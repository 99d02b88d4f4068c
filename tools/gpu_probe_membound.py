#!/usr/bin/env python3
"""Bandwidth of the memory-bound kernels at model shape (same box, back to back)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_llm_backend_benchmark_b200 import ops
dev = "cuda"
def bench(fn, iters=30, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize(); return s.elapsed_time(e) / iters
rows, cols = 4096, 4096
x = torch.randn(rows, cols, device=dev).bfloat16(); g = torch.ones(cols, device=dev).bfloat16(); b = torch.zeros(cols, device=dev).bfloat16()
y = torch.empty_like(x); dy = torch.randn_like(x); dres = torch.randn_like(x); dx = torch.empty_like(x)
big = torch.empty(1 << 28, dtype=torch.uint8, device=dev)
ms = bench(lambda: ops.layernorm_fwd(x, g, b, out=y)); print(f"ln_fwd  {ms*1e3:7.1f} us  {2*x.numel()*2/ms/1e6:6.0f} GB/s")
_, mean, rstd = ops.layernorm_fwd(x, g, b, out=y)
ms = bench(lambda: ops.layernorm_bwd(dy, x, mean, rstd, g, dres, dx=dx)); print(f"ln_bwd  {ms*1e3:7.1f} us  {4*x.numel()*2/ms/1e6:6.0f} GB/s (incl. reduce kernel)")
n = 201_342_976
p = torch.zeros(n, device=dev, dtype=torch.bfloat16); gr = torch.ones(n, device=dev, dtype=torch.bfloat16)
m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev); ma = torch.zeros(n, device=dev)
ms = bench(lambda: ops.adam_step(p, gr, m, v, ma, lr=1e-3, step=1), iters=8, warm=2); print(f"adam    {ms*1e3:7.1f} us  {n*28/ms/1e6:6.0f} GB/s (one 7B layer bucket)")
a = torch.empty(1 << 30, dtype=torch.uint8, device=dev); c = torch.empty_like(a)
ms = bench(lambda: c.copy_(a), iters=8, warm=2); print(f"copy    {ms*1e3:7.1f} us  {2*a.numel()/ms/1e6:6.0f} GB/s (torch copy, roofline reference)")
