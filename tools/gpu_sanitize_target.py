#!/usr/bin/env python3
"""Small workload for compute-sanitizer: every single-GPU kernel once, tiny shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_llm_backend_benchmark_b200 import ops
d = "cuda"
a = torch.randn(200, 72, device=d).bfloat16(); b = torch.randn(72, 136, device=d).bfloat16(); aux = torch.randn(200, 136, device=d).bfloat16()
for epi in (None, "gelu", "add", "mul_dgelu"):
    ops.matmul(a, b, epilogue=epi, aux=aux.clone() if epi else None)
ops.matmul(a, b, transpose_a=False, transpose_b=False, force_simt=True)
x = torch.randn(33, 512, device=d).bfloat16(); g = torch.ones(512, device=d).bfloat16(); be = torch.zeros(512, device=d).bfloat16()
y, m, r = ops.layernorm_fwd(x, g, be); ops.layernorm_bwd(y, x, m, r, g, x)
ops.gelu(x); ops.gelu_bwd(y, x); ops.add(x, y); ops.mse_loss(x, y)
p = torch.randn(1000, device=d).bfloat16(); gr = torch.randn(1000, device=d).bfloat16()
ops.adam_step(p, gr, torch.zeros(1000, device=d), torch.zeros(1000, device=d), p.float(), lr=1e-3, step=1)
torch.cuda.synchronize(); print("sanitize target ok")
