#!/usr/bin/env python3
"""NVLS kernels vs CTA count (torchrun, N GPUs): is multimem throughput limited by requests in flight?"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = SymmComm.from_env(dev)
comm.set_zero_copy_min_bytes(1 << 40)


def timeit(fn, iters=8, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


rows = []
nbytes = int(os.environ.get("SWEEP_BYTES", str(1 << 27)))
x = torch.ones(nbytes // 2, dtype=torch.bfloat16, device=dev)
for ctas in [int(v) for v in os.environ.get("SWEEP_CTAS", "32,64,96,128,160").split(",")]:
    comm.set_ctas(ctas)
    row = {"ctas": ctas, "bytes": nbytes, "chunk_kb": os.environ.get("DLBB_PIPE_CHUNK_KB", "default")}
    for name, fn in (("broadcast", lambda: comm.broadcast(x, root=0)), ("reduce", lambda: comm.reduce(x, root=0)),
                     ("allreduce_nvls", lambda: comm.all_reduce(x, "sum", "nvls"))):
        ms = timeit(fn); row[name + "_us"] = round(ms * 1e3, 1); row[name + "_GBps"] = round(nbytes / ms / 1e6, 1)
    rows.append(row)
    if rank == 0: print("SWEEP", json.dumps(row), flush=True)
comm.set_ctas(0)
for name, fn in (("nccl_broadcast", lambda: dist.broadcast(x, src=0)), ("nccl_reduce", lambda: dist.reduce(x, dst=0)),
                 ("nccl_allreduce", lambda: dist.all_reduce(x))):
    ms = timeit(fn)
    if rank == 0: print("SWEEP", name, round(ms * 1e3, 1), "us", round(nbytes / ms / 1e6, 1), "GB/s (algbw)", flush=True)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True); json.dump(rows, open(f"gpurun_out/nvls_cta_sweep_{world}gpu.json", "w"), indent=1)
comm.destroy(); dist.destroy_process_group()
