#!/usr/bin/env python3
"""Single-GPU program for ncu covering every kernel that is not the plain GEMM: LayerNorm forward / backward,
Adam, MSE, cast, and - through a world-size-1 communicator, where every "peer" pointer is the local one and the
flag waits pass at once - the collective kernels, the fused reduce+Adam bucket kernel and the fused
GEMM + all-reduce.  (ncu replays a kernel ~40 times, so kernels that wait for other ranks cannot be profiled in a
multi-rank job; the world-1 instance shows the instruction mix, occupancy and the local memory traffic, not the
NVLink traffic.)  Model shapes: 4096 tokens, H = 4096, one 7B layer bucket for the optimizers.

usage: python tools/ncu_all.py [iters]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_llm_backend_benchmark_b200 import ops
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
T, H = 4096, 4096
bf = torch.bfloat16
x = torch.randn(T, H, device=dev).to(bf)
dy = torch.randn(T, H, device=dev).to(bf)
g = torch.ones(H, device=dev).to(bf)
b = torch.zeros(H, device=dev).to(bf)
n_layer = 201_342_976 // 4            # a quarter of a 7B layer bucket keeps the capture short (50 M parameters)
p = torch.randn(n_layer, device=dev).to(bf)
gr = torch.randn(n_layer, device=dev).to(bf)
m = torch.zeros(n_layer, device=dev)
v = torch.zeros(n_layer, device=dev)
master = p.float()
f32 = torch.randn(T * H, device=dev)

comm = SymmComm(0, 1, dev, job=f"ncu-{os.getpid()}")
big = torch.randn(32 << 20, device=dev).to(bf)            # 64 MiB
gathered = torch.empty_like(big)
params, grads = comm.alloc_dp_buffers(n_layer)
grads.copy_(gr)
params.copy_(p)
part = comm.alloc_matrix(T, H)
outm = comm.alloc_matrix(T, H)
w = torch.randn(H, H, device=dev).to(bf)

for it in range(iters):
    y, mean, rstd = ops.layernorm_fwd(x, g, b)
    ops.layernorm_bwd(dy, x, mean, rstd, g, dy)
    ops.adam_step(p, gr, m, v, master, lr=1e-4, step=it + 1)
    ops.mse_loss(y, dy)
    ops.cast(f32, bf)
    comm.all_reduce(big)                                   # world 1: pack -> reduce out of the staging copy
    comm.all_gather(big, gathered)                         # pack -> pull
    comm.fused_reduce_adam(0, n_layer, m, v, master, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                           grad_scale=1.0, step=it + 1, use_multicast=False)
    comm.gemm_allreduce(x, w, part, outm, residual=dy, use_multicast=False)
torch.cuda.synchronize()
comm.destroy()
print("ok")
