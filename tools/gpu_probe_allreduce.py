#!/usr/bin/env python3
"""Large-message all-reduce tuning probe (run under torchrun): CTA count x sub-chunk size x algorithm."""
import os, sys, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = SymmComm.from_env(dev, staging_mb=int(os.environ.get("STAGE_MB", "64")))
nbytes = int(os.environ.get("NBYTES", str(256 << 20)))
x = torch.ones(nbytes // 2, dtype=torch.bfloat16, device=dev)
def timeit(fn, iters=8, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev, dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
ms = timeit(lambda: dist.all_reduce(x))
if rank == 0: print(f"NCCL: {ms*1e3:.0f} us busbw {nbytes/ms/1e6*2*(world-1)/world:.0f} GB/s", flush=True)
algos = ["twoshot"] + (["nvls"] if comm.has_multicast else [])
for algo in algos:
    for ctas in [int(v) for v in os.environ.get("CTAS", "32,64,96,128,148").split(",")]:
        comm.set_ctas(ctas)
        ms = timeit(lambda: comm.all_reduce(x, "sum", algo))
        if rank == 0: print(f"{algo} chunkKB={os.environ.get('DLBB_PIPE_CHUNK_KB','64')} pipe_min={os.environ.get('DLBB_PIPELINE_MIN_BYTES','8M')} stage={os.environ.get('STAGE_MB','64')} ctas={ctas}: {ms*1e3:.0f} us busbw {nbytes/ms/1e6*2*(world-1)/world:.0f} GB/s", flush=True)
comm.destroy(); dist.destroy_process_group()
