#!/usr/bin/env python3
"""All-to-all + mixed-collective stress test (BASELINE.json config 5).

    deepspeed --master_port 29503 --include localhost:0,1,2,3,4,5,6,7 stress.py      # bin/deepspeed = the dlbb launcher
    mpirun -np 8 python stress.py

Every round issues a randomly ordered (same order on every rank) mix of all-to-all, all-reduce, all-gather,
reduce-scatter, broadcast and a ring send/recv with round-dependent sizes and verifies every result against a closed-form
expectation, so lost flags, stale staging data or epoch mix-ups between back-to-back kernels show up as wrong numbers,
not just as hangs.  It then times each op in isolation (device-timed, max over ranks) and writes one JSON
(`stress_ranks<P>.json`) with the all-to-all bus bandwidth by size and the mixed-round throughput - the "launcher scaling
1 -> 8 ranks" row comes from running it at every rank count (collectives/launch_stress.sh).

The all-to-all / sendrecv micro-benchmarks of the reference are collectives/1d/openmpi.py:154-198 and
collectives/1d/dsccl.py:134-160; the mixed stress has no counterpart there.
"""
import json
import os
import random
import sys
import time

_ROOT = os.path.abspath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
for _p in (_ROOT, os.path.join(_ROOT, "compat")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import deepspeed  # noqa: E402
import deepspeed.comm as dist  # noqa: E402

BACKEND = os.environ.get("DLBB_SWEEP_BACKEND", "ccl")
ROUNDS = int(os.environ.get("DLBB_STRESS_ROUNDS", "200"))
ALLTOALL_SIZES = [1 << 10, 1 << 14, 1 << 18, 1 << 22, 1 << 26, 1 << 28]      # bytes per rank (bf16)
OUTPUT_DIR = os.environ.get("DLBB_OUTPUT_DIR", "results/stress")


def main():
    deepspeed.init_distributed(dist_backend=BACKEND)
    rank, world = dist.get_rank(), dist.get_world_size()
    use_cuda = torch.cuda.is_available() and BACKEND != "gloo" and os.environ.get("DLBB_SWEEP_DEVICE", "cuda") != "cpu"
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count()) if use_cuda else torch.device("cpu")
    helper = dev if BACKEND == "nccl" else torch.device("cpu")
    rng = random.Random(1234)                      # same stream of choices on every rank
    ops = ["alltoall", "allreduce", "allgather", "reducescatter", "broadcast"] + (["sendrecv"] if world > 1 else [])
    if os.environ.get("DLBB_STRESS_OPS"):       # e.g. the NCCL arm: unbatched isend/irecv rings can deadlock there
        ops = [o for o in ops if o in os.environ["DLBB_STRESS_OPS"].split(",")]
    errors = 0
    t0 = time.perf_counter()
    for rnd in range(ROUNDS):
        order = ops[:]
        rng.shuffle(order)
        n = world * rng.choice([8, 64, 1000, 4096, 65536 + 8])       # elements, multiple of world
        for op in order:
            base = float(rnd % 7 + 1)
            if op == "alltoall":
                src = (torch.arange(n, device=dev) // (n // world)).float() + 10.0 * rank        # chunk p holds p + 10*rank
                out = torch.empty_like(src)
                dist.all_to_all_single(out, src)
                want = float(rank) + 10.0 * (torch.arange(n, device=dev) // (n // world)).float()
                errors += int(not torch.equal(out, want))
            elif op == "allreduce":
                t = torch.full((n,), base * (rank + 1), device=dev)
                dist.all_reduce(t)
                errors += int(not torch.allclose(t, torch.full_like(t, base * world * (world + 1) / 2)))
            elif op == "allgather":
                out = torch.empty(world * n, device=dev)
                dist.all_gather_into_tensor(out, torch.full((n,), base + rank, device=dev))
                want = (torch.arange(world * n, device=dev) // n).float() + base
                errors += int(not torch.equal(out, want))
            elif op == "reducescatter":
                out = torch.empty(n // world, device=dev)
                dist.reduce_scatter_tensor(out, torch.full((n,), base + rank, device=dev))
                errors += int(not torch.allclose(out, torch.full_like(out, world * base + world * (world - 1) / 2)))
            elif op == "broadcast":
                root = rnd % world
                t = torch.full((n,), base * 3 if rank == root else -1.0, device=dev)
                dist.broadcast(t, src=root)
                errors += int(not torch.equal(t, torch.full_like(t, base * 3)))
            elif op == "sendrecv":
                nxt, prv = (rank + 1) % world, (rank - 1) % world
                recv = torch.empty(n, device=dev)
                s = dist.isend(torch.full((n,), base + rank, device=dev), dst=nxt)
                r = dist.irecv(recv, src=prv)
                s.wait()
                r.wait()
                if use_cuda:
                    torch.cuda.synchronize()
                errors += int(not torch.equal(recv, torch.full_like(recv, base + prv)))
    if use_cuda:
        torch.cuda.synchronize()
    mixed_s = time.perf_counter() - t0
    bad = torch.tensor([float(errors)], device=helper)
    dist.all_reduce(bad)
    total_errors = int(bad.item())

    # all-to-all bandwidth by size (device-timed, max over ranks)
    rows = []
    for nbytes in ALLTOALL_SIZES:
        n = max(world, nbytes // 2 // world * world)
        if n * 2 * 2 > 3 * (1 << 30):
            continue
        src = torch.ones(n, dtype=torch.bfloat16, device=dev)
        out = torch.empty_like(src)
        iters = 20 if nbytes <= (1 << 22) else 5
        for _ in range(3):
            dist.all_to_all_single(out, src)
        dist.barrier()
        if use_cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                dist.all_to_all_single(out, src)
            e1.record()
            e1.synchronize()
            sec = e0.elapsed_time(e1) / 1e3 / iters
        else:
            t1 = time.perf_counter()
            for _ in range(iters):
                dist.all_to_all_single(out, src)
            sec = (time.perf_counter() - t1) / iters
        tt = torch.tensor([sec], dtype=torch.float64, device=helper)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = tt.item()
        rows.append({"bytes_per_rank": n * 2, "time_us": sec * 1e6,
                     "busbw_gbps": n * 2 * (world - 1) / world / sec / 1e9 if world > 1 else n * 2 / sec / 1e9})
    if rank == 0:
        os.makedirs(OUTPUT_DIR, exist_ok=True)
        res = {"implementation": f"stress_{BACKEND}", "backend": BACKEND, "num_ranks": world, "device": str(dev),
               "rounds": ROUNDS, "ops_per_round": len(ops), "mixed_seconds": mixed_s,
               "mixed_ops_per_second": ROUNDS * len(ops) / mixed_s, "errors": total_errors, "alltoall": rows}
        path = os.path.join(OUTPUT_DIR, f"stress_{BACKEND}_ranks{world}.json")
        with open(path, "w") as fh:
            json.dump(res, fh, indent=2)
        print(f"stress on {world} rank(s), backend {BACKEND}: {ROUNDS} mixed rounds, {total_errors} wrong results, "
              f"{res['mixed_ops_per_second']:.0f} collectives/s; saved {path}")
        for r in rows:
            print(f"  all-to-all {r['bytes_per_rank']:>11d} B/rank: {r['time_us']:9.1f} us  busbw {r['busbw_gbps']:7.1f} GB/s")
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if total_errors else 0)


if __name__ == "__main__":
    main()
