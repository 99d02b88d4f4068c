#!/usr/bin/env python3
"""Where one data-parallel training step of the 7B model spends its time on N GPUs (run under torchrun / b200run):
CUDA events on the main stream around forward+backward and after the join with the bucket stream, once with the fused
reduce+Adam bucket kernels running beside backward and once with them skipped (DLBB_DEBUG_SKIP_OPT semantics), all as max
over ranks.  Gives: the forward+backward time alone, its slow-down while the bucket kernels share the SMs, and the tail
of the gradient exchange that is still exposed after the last backward kernel.  Writes gpurun_out/step_breakdown_nN.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_llm_backend_benchmark_b200.train import Trainer
from distributed_llm_backend_benchmark_b200.utils.config import load_config

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
cfg = load_config(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config", "baseline_config.yaml"))
tr = Trainer(cfg, device=dev, rank=rank, world_size=world)
dp, eng = tr.dp, tr.engine
x, t = tr.resident_batch()
main = torch.cuda.current_stream(dev)


def one_step():
    dp.step_count += 1
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record(main)
    eng.forward_backward(x, t, on_bucket_ready=dp._on_bucket_ready if dp.overlap else None)
    e[1].record(main)
    if dp.overlap:
        dp.done.record(dp.side)
        main.wait_event(dp.done)
    else:
        for i in range(len(dp.buckets)):
            dp._update_bucket(i)
    e[2].record(main)
    return e


def measure(skip_opt, steps=6, warm=2):
    dp.skip_optimizer = skip_opt
    fb, tail = [], []
    for i in range(warm + steps):
        e = one_step()
        torch.cuda.synchronize(dev)
        if tr.comm is not None:
            tr.comm.host_barrier()
        if i >= warm:
            fb.append(e[0].elapsed_time(e[1])); tail.append(e[1].elapsed_time(e[2]))
    v = torch.tensor([sum(fb) / len(fb), sum(tail) / len(tail)], dtype=torch.float64, device=dev)
    if tr.comm is not None:
        tr.comm.all_reduce(v, op="max")
    return v.tolist()


with_opt = measure(False)
without = measure(True)
if rank == 0:
    out = {"n_gpus": world, "overlap": bool(dp.overlap), "bucket_mode": eng.bucket_mode, "buckets": len(dp.buckets),
           "fwd_bwd_ms_with_bucket_kernels": with_opt[0], "exposed_tail_ms": with_opt[1],
           "step_ms": with_opt[0] + with_opt[1], "fwd_bwd_ms_alone": without[0],
           "fwd_bwd_slowdown_ms": with_opt[0] - without[0]}
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(f"gpurun_out/step_breakdown_n{world}.json", "w"), indent=1)
    print("BREAKDOWN", json.dumps(out))
if tr.comm is not None:
    tr.comm.destroy()
