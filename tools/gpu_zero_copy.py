#!/usr/bin/env python3
"""Zero-copy collectives (peer_registry.cpp) on N GPUs, run under torchrun: numerics of every op with the
zero-copy path forced at all sizes (plain torch tensors, views at offsets inside an allocation, changing
buffers), then bandwidth of zero-copy vs the staged kernels vs NCCL on large messages.
Writes gpurun_out/zero_copy_rank0.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = SymmComm.from_env(dev)
P = lambda *a: print(f"[r{rank}]", *a, flush=True)
res = {"world": world, "multicast": comm.has_multicast, "zero_copy_default_min": comm.zero_copy_min_bytes, "checks": {},
       "perf": []}
ok_all = True
if rank == 0: P("comm up; multicast:", comm.has_multicast, "zero-copy min bytes:", comm.zero_copy_min_bytes)
if comm.zero_copy_min_bytes == 0:
    P("zero-copy unavailable on this box"); res["ok"] = False
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True); json.dump(res, open("gpurun_out/zero_copy_rank0.json", "w"))
    dist.destroy_process_group(); sys.exit(1)


def gen(r, n, dtype, seed=0):
    g = torch.Generator(device=dev); g.manual_seed(1000 * seed + r)
    return (torch.randn(n, device=dev, generator=g, dtype=torch.float32) * 2).to(dtype)


def check(name, got, ref, tol=0.0):
    global ok_all
    a, b = got.double(), ref.double()
    err = ((a - b).abs().max() / (b.abs().max() + 1e-30)).item() if a.numel() else 0.0
    ok = err <= tol
    ok_all &= ok
    res["checks"][name] = dict(err=err, ok=ok)
    if not ok or rank == 0: P(("PASS " if ok else "FAIL ") + name, err)


TOL = {torch.bfloat16: 2e-2, torch.float16: 4e-3, torch.float32: 1e-5, torch.float64: 1e-12}
ROUNDS = 0 if os.environ.get("ZC_SKIP_NUMERICS") == "1" else int(os.environ.get("ZC_ROUNDS", "2"))   # 0: perf only
PERF_OPS = os.environ.get("ZC_PERF_OPS", "allreduce,allgather,broadcast,reduce,reduce_scatter,alltoall").split(",")
try:
    comm.set_zero_copy_min_bytes(1)          # every size goes zero-copy
    for rnd in range(ROUNDS):                     # second round: new tensors -> other allocations / cached mappings
        for dtype in (torch.bfloat16, torch.float32, torch.float64):
            tol = TOL[dtype]
            for n in (1000, (1 << 18) + 8, (5 << 20) + 3):
                dn = f"{str(dtype)[6:]}_{n}_r{rnd}"
                xs = [gen(r, n, dtype, seed=n % 97 + rnd) for r in range(world)]
                x = xs[rank].clone(); comm.all_reduce(x)
                check(f"allreduce_{dn}", x, sum(v.double() for v in xs).to(dtype), tol)
                if n == 1000:
                    for op, f in (("max", torch.maximum), ("min", torch.minimum)):
                        x = xs[rank].clone(); comm.all_reduce(x, op)
                        ref = xs[0].double()
                        for v in xs[1:]: ref = f(ref, v.double())
                        check(f"allreduce_{op}_{dn}", x, ref.to(dtype), tol)
                # a view at an odd (but 16 B aligned) offset inside a bigger allocation
                big = torch.zeros(n + 4096, dtype=dtype, device=dev)
                v = big[2048:2048 + n]; v.copy_(xs[rank]); comm.all_reduce(v)
                check(f"allreduce_view_{dn}", v, sum(t.double() for t in xs).to(dtype), tol)
                check(f"allreduce_view_guard_{dn}", torch.cat([big[:2048], big[2048 + n:]]).abs().max().reshape(1),
                      torch.zeros(1, device=dev))
                out = torch.empty(world * n, dtype=dtype, device=dev); comm.all_gather(xs[rank], out)
                check(f"allgather_{dn}", out, torch.cat(xs))
                out = torch.cat([torch.zeros_like(x) for x in xs]); out[rank * n:(rank + 1) * n].copy_(xs[rank])
                comm.all_gather(out[rank * n:(rank + 1) * n], out)            # in place
                check(f"allgather_inplace_{dn}", out, torch.cat(xs))
                b = xs[rank].clone(); comm.broadcast(b, root=world - 1); check(f"broadcast_{dn}", b, xs[world - 1])
                root = 1 % world
                out = torch.zeros(world * n, dtype=dtype, device=dev); comm.gather(xs[rank], out, root=root)
                if rank == root: check(f"gather_{dn}", out, torch.cat(xs))
                bigs = [torch.cat([gen(r + 10 * q, n, dtype, seed=3 + rnd) for q in range(world)]) for r in range(world)]
                out = torch.empty(n, dtype=dtype, device=dev); comm.scatter(bigs[root] if rank == root else None, out, root=root)
                check(f"scatter_{dn}", out, bigs[root][rank * n:(rank + 1) * n])
                out = torch.empty(world * n, dtype=dtype, device=dev); comm.all_to_all(bigs[rank], out)
                check(f"alltoall_{dn}", out, torch.cat([bigs[p][rank * n:(rank + 1) * n] for p in range(world)]))
                out = torch.empty(n, dtype=dtype, device=dev); comm.reduce_scatter(bigs[rank], out)
                check(f"reduce_scatter_{dn}", out, sum(b[rank * n:(rank + 1) * n].double() for b in bigs).to(dtype), tol)
                o = xs[rank].clone(); comm.reduce(o, root=root)
                if rank == root: check(f"reduce_{dn}", o, sum(v.double() for v in xs).to(dtype), tol)
                del xs, bigs, out, big
    if world > 2:
        # large broadcast / reduce on unaligned views (the chain kernels take any alignment)
        n = 9 * (1 << 20) + 5
        xs = [gen(r, n, torch.bfloat16, seed=31) for r in range(world)]
        big = torch.zeros(n + 64, dtype=torch.bfloat16, device=dev); v = big[1:1 + n]; v.copy_(xs[rank])
        comm.broadcast(v, root=1); check("broadcast_unaligned_18MiB", v, xs[1])
        check("broadcast_unaligned_guard", torch.stack([big[0], big[1 + n]]).float(), torch.zeros(2, device=dev))
        v.copy_(xs[rank]); comm.reduce(v, root=world - 1)
        if rank == world - 1: check("reduce_unaligned_18MiB", v, sum(t.double() for t in xs).to(torch.bfloat16), 3e-2)
        del xs, big
    # results identical on every rank
    x = gen(rank, 100003 * 8, torch.bfloat16, seed=9); comm.all_reduce(x)
    g = [torch.empty_like(x) for _ in range(world)]; dist.all_gather(g, x)
    check("allreduce_bitwise_equal_across_ranks", g[0].float(), g[-1].float())
    # symmetric (VMM) memory cannot be exported through legacy IPC: must fall back to staging, on every rank
    t, seg = comm.alloc(1 << 16, torch.float32); t.fill_(float(rank + 1)); comm.all_reduce(t)
    check("allreduce_symmetric_tensor_falls_back", t[:8], torch.full((8,), world * (world + 1) / 2.0, device=dev))
    # alternate zero-copy and staged calls (epochs and staging parity stay consistent)
    for i in range(6):
        comm.set_zero_copy_min_bytes(1 if i % 2 == 0 else 1 << 40)
        x = gen(rank, 70001, torch.float32, seed=20 + i); xs = [gen(r, 70001, torch.float32, seed=20 + i) for r in range(world)]
        comm.all_reduce(x); check(f"allreduce_alternating_{i}", x, sum(v.double() for v in xs).float(), 1e-5)
    torch.cuda.synchronize()
except Exception as e:
    import traceback; traceback.print_exc()
    ok_all = False; res["exception"] = repr(e)


def timeit(fn, iters, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


flag = torch.tensor([1 if ok_all else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if flag.item() == 1 and os.environ.get("DLBB_SKIP_PERF") != "1":
    try:
        ZC, STAGED = 1, 1 << 40
        f = 2.0 * (world - 1) / world
        for nbytes in ((1 << 22, 1 << 25, 1 << 27, 1 << 30) if "allreduce" in PERF_OPS else ()):
            n = nbytes // 2
            x = torch.ones(n, dtype=torch.bfloat16, device=dev)
            iters = 20 if nbytes <= (1 << 25) else (8 if nbytes <= (1 << 27) else 4)
            row = {"op": "allreduce", "bytes": nbytes}
            for name, zc in (("zero_copy", ZC), ("staged", STAGED)):
                comm.set_zero_copy_min_bytes(zc)
                ms = timeit(lambda: comm.all_reduce(x), iters)
                row[f"{name}_us"] = ms * 1e3; row[f"{name}_busbw"] = nbytes / ms / 1e6 * f
            ms = timeit(lambda: dist.all_reduce(x), iters)
            row["nccl_us"] = ms * 1e3; row["nccl_busbw"] = nbytes / ms / 1e6 * f
            res["perf"].append(row)
            if rank == 0: P("PERF", json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}))
        # per-rank message sizes for the data-movement ops
        for nbytes in (1 << 25, 1 << 27, 1 << 29):
            n = nbytes // 2
            x = torch.ones(n, dtype=torch.bfloat16, device=dev)
            wide = torch.ones(world * (n // world), dtype=torch.bfloat16, device=dev)      # nbytes in total, split per peer
            out_w = torch.empty(world * n, dtype=torch.bfloat16, device=dev)
            out_1 = torch.empty(n // world, dtype=torch.bfloat16, device=dev)
            out_a = torch.empty_like(wide)
            iters = 10 if nbytes <= (1 << 27) else 4
            ops_ = {
                "allgather": (lambda: comm.all_gather(x, out_w), lambda: dist.all_gather_into_tensor(out_w, x), nbytes * (world - 1)),
                "broadcast": (lambda: comm.broadcast(x, root=0), lambda: dist.broadcast(x, src=0), nbytes),
                "reduce": (lambda: comm.reduce(x, root=0), lambda: dist.reduce(x, dst=0), nbytes),
                "reduce_scatter": (lambda: comm.reduce_scatter(wide, out_1), lambda: dist.reduce_scatter_tensor(out_1, wide),
                                   wide.numel() * 2 * (world - 1) / world),
                "alltoall": (lambda: comm.all_to_all(wide, out_a), lambda: dist.all_to_all_single(out_a, wide),
                             wide.numel() * 2 * (world - 1) / world),
            }
            for op, (ours, nccl, wire) in ops_.items():
                if op not in PERF_OPS: continue
                row = {"op": op, "bytes": nbytes}
                for name, zc in (("zero_copy", ZC), ("staged", STAGED)):
                    comm.set_zero_copy_min_bytes(zc)
                    ms = timeit(ours, iters); row[f"{name}_us"] = ms * 1e3; row[f"{name}_busbw"] = wire / ms / 1e6
                ms = timeit(nccl, iters); row["nccl_us"] = ms * 1e3; row["nccl_busbw"] = wire / ms / 1e6
                res["perf"].append(row)
                if rank == 0: P("PERF", json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}))
            del x, wide, out_w, out_1, out_a
    except Exception as e:
        import traceback; traceback.print_exc(); res["perf_exception"] = repr(e); ok_all = False

res["ok"] = bool(ok_all)
flag = torch.tensor([1 if ok_all else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/zero_copy_{world}gpu.json", "w"), indent=1)
    P("ZERO-COPY CHECK", "OK" if flag.item() == 1 else "FAILED", f"({len(res['checks'])} checks)")
comm.destroy()
dist.destroy_process_group()
sys.exit(0 if flag.item() == 1 else 1)
