#!/usr/bin/env python3
"""Bucket-fusion study (the analogue of the reference's CCL_FUSION runs, collectives/3d/launch_dsccl.sh:49-59): the
all-reduce of K small tensors as one `all_reduce_coalesced` call on the native backend, device-timed back to back, max
over ranks.  Run once with DLBB_FUSION=1 (multi-tensor kernel: one launch per bucket) and once with DLBB_FUSION=0 (one
collective per tensor); appends one JSON line per configuration to gpurun_out/fusion_study.jsonl."""
import json, os, sys
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "compat")]
import torch, torch.distributed as dist, oneccl_bindings_for_pytorch  # noqa: F401
from distributed_llm_backend_benchmark_b200 import _lib

torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dist.init_process_group("ccl")
r, w = dist.get_rank(), dist.get_world_size()
rows = []
for k, nbytes in ((16, 4096), (64, 4096), (64, 65536), (256, 16384), (32, 1 << 20)):
    ts = [torch.ones(nbytes // 2, device="cuda", dtype=torch.bfloat16) for _ in range(k)]
    for _ in range(3): dist.all_reduce_coalesced(ts)
    torch.cuda.synchronize(); dist.barrier()
    n0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): dist.all_reduce_coalesced(ts)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 10 * 1e3], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    rows.append({"ranks": w, "fusion": os.environ.get("DLBB_FUSION", "1"), "tensors": k, "bytes_each": nbytes,
                 "us_per_call": round(t.item(), 1), "launches_per_call": (_lib.launch_count() - n0 - 0) / 10})
if r == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/fusion_study.jsonl", "a") as f:
        for row in rows:
            f.write(json.dumps(row) + "\n"); print(row)
dist.destroy_process_group()
