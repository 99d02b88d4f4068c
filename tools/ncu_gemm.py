#!/usr/bin/env python3
"""Small single-GPU program for ncu: a few launches of the flagship GEMM shapes (and the LayerNorm / Adam
kernels) so that one `ncu --set full` capture covers the hot kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_llm_backend_benchmark_b200 import ops
dev = "cuda"
T, H, F = 4096, 4096, 16384
x = torch.randn(T, H, device=dev).bfloat16(); wup = torch.randn(H, F, device=dev).bfloat16()
z = torch.empty(T, F, device=dev, dtype=torch.bfloat16); a = torch.empty_like(z)
dy = torch.randn(T, H, device=dev).bfloat16(); gw = torch.empty(F, H, device=dev, dtype=torch.bfloat16)
g = torch.ones(H, device=dev).bfloat16(); b = torch.zeros(H, device=dev).bfloat16()
for _ in range(3):
    ops.matmul(x, wup, epilogue="gelu", aux=z, out=a)            # FFN-up + GELU epilogue (K-major x MN-major)
    ops.matmul(a, dy, transpose_a=True, out=gw)                  # FFN-down wgrad (MN-major x MN-major)
    y, mean, rstd = ops.layernorm_fwd(x, g, b)
    ops.layernorm_bwd(dy, x, mean, rstd, g, dy)
torch.cuda.synchronize()
print("ok")
