#!/usr/bin/env python3
"""'ccl' backend + DeepSpeed-style engine integration test (ref test/ccl.py): basic collectives, a
ZeRO-2 + Adam training step on a small MLP, and a tiny all-reduce throughput print.

    mpirun -n 2 python test/ccl.py"""
import os
import sys
import time

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import oneccl_bindings_for_pytorch  # noqa: E402,F401
import deepspeed  # noqa: E402


def init():
    if "RANK" not in os.environ:
        print("launch with: mpirun -n 2 python test/ccl.py")
        sys.exit(1)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group(backend="ccl")
    return dist.get_rank(), dist.get_world_size()


def basic(rank, world):
    t = torch.tensor([42.0]) if rank == 0 else torch.zeros(1)
    dist.broadcast(t, src=0)
    assert t.item() == 42.0
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t)
    assert abs(t.item() - world * (world + 1) / 2) < 1e-3


def zero2_step(rank, world):
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(10, 20), torch.nn.ReLU(), torch.nn.Linear(20, 5))
    cfg = {"train_batch_size": 8, "gradient_accumulation_steps": 1,
           "optimizer": {"type": "Adam", "params": {"lr": 0.001}}, "fp16": {"enabled": False},
           "zero_optimization": {"stage": 2}}
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=cfg)
    assert (engine.local_rank, engine.global_rank, engine.world_size) == (int(os.environ.get("LOCAL_RANK", rank)), rank, world)
    torch.manual_seed(100 + rank)
    x, y = torch.randn(8, 10), torch.randn(8, 5)
    first = None
    for _ in range(10):
        loss = torch.nn.functional.mse_loss(engine(x), y)
        engine.backward(loss)
        engine.step()
        first = first if first is not None else loss.item()
    assert loss.item() < first
    flat = torch.cat([p.data.reshape(-1) for p in model.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    assert torch.equal(flat, ref), "ranks diverged after ZeRO-2 steps"
    if rank == 0:
        print(f"ZeRO-2 + Adam: loss {first:.4f} -> {loss.item():.4f}, parameters identical on all ranks")


def throughput(rank, world):
    for n in (256, 2560, 25600):
        t = torch.randn(n)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(10):
            dist.all_reduce(t)
        dist.barrier()
        dt = (time.perf_counter() - t0) / 10
        if rank == 0:
            print(f"all_reduce {n * 4 / 1024:.0f} KB: {dt * 1e6:.1f} us  ({n * 4 / dt / 1e6:.1f} MB/s)")


def main():
    rank, world = init()
    try:
        basic(rank, world)
        zero2_step(rank, world)
        throughput(rank, world)
        dist.barrier()
        if rank == 0:
            print("ccl + deepspeed integration test passed")
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
