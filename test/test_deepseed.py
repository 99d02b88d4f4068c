#!/usr/bin/env python3
"""torch.distributed smoke test on backend 'ccl' (= the dlbb backend; Gloo on a CPU-only host):
broadcast, all_reduce (SUM, MAX), all_gather, reduce, scatter, send/recv, staggered barrier - the
checks of the reference's test/test_deepseed.py.

    mpirun -n 2 python test/test_deepseed.py      or      torchrun --nproc_per_node=2 test/test_deepseed.py"""
import os
import sys
import time

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import oneccl_bindings_for_pytorch  # noqa: E402,F401  (registers 'ccl')


def setup():
    if "RANK" not in os.environ:
        print("RANK is not set - launch with mpirun -n <N> or torchrun")
        sys.exit(1)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group(backend="ccl")
    return dist.get_rank(), dist.get_world_size()


def run(rank, world):
    expect_sum = world * (world + 1) / 2
    t = torch.arange(4, dtype=torch.float32) if rank == 0 else torch.zeros(4)
    dist.broadcast(t, src=0)
    assert torch.equal(t, torch.arange(4, dtype=torch.float32)), "broadcast"

    t = torch.full((3,), float(rank + 1))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    assert torch.allclose(t, torch.full((3,), expect_sum)), "all_reduce SUM"

    parts = [torch.zeros(2) for _ in range(world)]
    dist.all_gather(parts, torch.full((2,), float(rank)))
    assert all(torch.equal(p, torch.full((2,), float(i))) for i, p in enumerate(parts)), "all_gather"

    t = torch.full((2,), float(rank + 1))
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        assert torch.allclose(t, torch.full((2,), expect_sum)), "reduce"

    out = torch.zeros(2)
    dist.scatter(out, [torch.full((2,), float(i * 10)) for i in range(world)] if rank == 0 else None, src=0)
    assert torch.equal(out, torch.full((2,), float(rank * 10))), "scatter"

    if world >= 2:
        if rank == 0:
            dist.send(torch.tensor([3.14, 2.71]), dst=1)
        elif rank == 1:
            got = torch.zeros(2)
            dist.recv(got, src=0)
            assert torch.allclose(got, torch.tensor([3.14, 2.71])), "send/recv"

    time.sleep(0.05 * rank)
    dist.barrier()

    t = torch.tensor([float(rank)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert t.item() == world - 1, "all_reduce MAX"


def main():
    rank, world = setup()
    try:
        run(rank, world)
        dist.barrier()
        if rank == 0:
            print(f"all torch.distributed checks passed on backend 'ccl' ({dist.get_backend()}), {world} rank(s)")
    except Exception:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        sys.exit(1)
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
