#!/usr/bin/env python3
"""Collectives on symmetric tensors (SymmComm.alloc: in-place NVLS all-reduce / reduce-scatter / reduce, relay
broadcast) and the direct forms on plain tensors, on N GPUs under torchrun: numerics against a gathered oracle,
then device-timed bandwidth (tight loops, max over ranks) against NCCL on the same sizes in the same process.
Variants are switched with the DLBB_MC_* / DLBB_BCAST_* environment knobs read by the library at call time.
Writes gpurun_out/symm_{N}gpu.json."""
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
res = {"world": world, "multicast": comm.has_multicast, "checks": {}, "perf": []}
ok_all = True
MAXB = int(os.environ.get("SYMM_MAX_BYTES", str(1 << 30)))
bf = torch.bfloat16


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


def setenv(**kw):
    for k, v in kw.items():
        os.environ[k] = str(v)


# one symmetric arena per role; tests and timings take views of it (allocation is collective and 2 MiB granular)
arena_a, _ = comm.alloc(MAXB // 2 + 4096, bf)      # in-place tensor
arena_b, _ = comm.alloc(MAXB // 2 + 4096, bf)      # second tensor (reduce-scatter output etc.)
TOL = {torch.bfloat16: 2e-2, torch.float32: 1e-5}
os.environ.setdefault("DLBB_SYMM_MIN_BYTES", "1")   # this tool exercises the direct kernels at every size
VARIANTS = [("mc", dict(DLBB_MC_ALLREDUCE=1, DLBB_MC_REDUCE=1, DLBB_MC_REDUCE_SCATTER=1, DLBB_BCAST_MC=1, DLBB_MC_WORLD2=1, DLBB_BCAST_MC_MAX_BYTES=1 << 40)),
            ("p2p", dict(DLBB_MC_ALLREDUCE=0, DLBB_MC_REDUCE=0, DLBB_MC_REDUCE_SCATTER=0, DLBB_BCAST_MC=0, DLBB_MC_WORLD2=0))]
try:
    for vname, env in (VARIANTS if os.environ.get("SYMM_SKIP_NUMERICS") != "1" else []):
        setenv(**env)
        for dtype in (bf, torch.float32):
            es = torch.empty((), dtype=dtype).element_size()
            for n in (8 * world, 4096 * world + 8 * world, (3 << 20) + 64 * world):
                n = n // (world * 16) * (world * 16) or world * 16
                tag = f"{vname}_{str(dtype)[6:]}_{n}"
                xs = [gen(r, n, dtype, seed=n % 89) for r in range(world)]
                ref = sum(v.double() for v in xs).to(dtype)
                a = arena_a.view(torch.uint8)[256:256 + n * es].view(dtype)      # offset view inside the segment
                a.copy_(xs[rank]); comm.all_reduce(a); check(f"symm_allreduce_{tag}", a, ref, TOL[dtype])
                a.copy_(xs[rank]); comm.all_reduce(a, "max")
                mref = xs[0].double()
                for v in xs[1:]: mref = torch.maximum(mref, v.double())
                check(f"symm_allreduce_max_{tag}", a, mref.to(dtype), TOL[dtype])
                for root in (0, world - 1):
                    a.copy_(xs[rank]); comm.reduce(a, root=root)
                    if rank == root: check(f"symm_reduce_r{root}_{tag}", a, ref, TOL[dtype])
                    a.copy_(xs[rank]); comm.broadcast(a, root=root); check(f"symm_broadcast_r{root}_{tag}", a, xs[root])
                a.copy_(xs[rank])
                o = arena_b.view(torch.uint8)[:n // world * es].view(dtype)
                comm.reduce_scatter(a, o)
                check(f"symm_reduce_scatter_{tag}", o, ref[rank * (n // world):(rank + 1) * (n // world)], TOL[dtype])
                # plain (caching-allocator) tensors: CUDA-IPC direct forms, every size forced zero-copy
                comm.set_zero_copy_min_bytes(1)
                t = xs[rank].clone(); comm.reduce(t, root=1 % world)
                if rank == 1 % world: check(f"ipc_reduce_{tag}", t, ref, TOL[dtype])
                t = xs[rank].clone(); comm.broadcast(t, root=world - 1); check(f"ipc_broadcast_{tag}", t, xs[world - 1])
                o2 = torch.empty(n // world, dtype=dtype, device=dev); comm.reduce_scatter(xs[rank], o2)
                check(f"ipc_reduce_scatter_{tag}", o2, ref[rank * (n // world):(rank + 1) * (n // world)], TOL[dtype])
                comm.set_zero_copy_min_bytes(16 << 20)
                del xs
    # same result on every rank, bit for bit
    a = arena_a[:1 << 20]; a.copy_(gen(rank, 1 << 20, bf, seed=5)); comm.all_reduce(a)
    g = [torch.empty_like(a) for _ in range(world)]; dist.all_gather(g, a.clone())
    check("symm_allreduce_bitwise_equal_across_ranks", g[0].float(), g[-1].float())
    torch.cuda.synchronize()
except Exception as e:
    import traceback; traceback.print_exc()
    ok_all = False; res["exception"] = repr(e)


def timeit(fn, iters, warm=3):
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
        sizes = [s for s in (1 << 10, 1 << 16, 1 << 20, 1 << 22, 1 << 25, 1 << 27, 1 << 29, 1 << 30) if int(os.environ.get('SYMM_MIN_BYTES', '0')) <= s <= MAXB]
        ctas_list = [int(x) for x in os.environ.get("SYMM_CTAS", "0").split(",")]
        for nbytes in sizes:
            n = nbytes // 2
            a = arena_a[:n]; a.fill_(1.0)
            plain = torch.ones(n, dtype=bf, device=dev)
            o_sym = arena_b[:n // world]
            o_plain = torch.empty(n // world, dtype=bf, device=dev)
            iters = 30 if nbytes <= (1 << 22) else (12 if nbytes <= (1 << 27) else 5)
            wire = {"allreduce": nbytes * 2.0 * (world - 1) / world, "reduce_scatter": nbytes * (world - 1) / world,
                    "reduce": nbytes, "broadcast": nbytes}
            ours = {"allreduce": lambda t, o: comm.all_reduce(t), "reduce": lambda t, o: comm.reduce(t, root=0),
                    "broadcast": lambda t, o: comm.broadcast(t, root=0), "reduce_scatter": lambda t, o: comm.reduce_scatter(t, o)}
            nccl = {"allreduce": lambda: dist.all_reduce(plain), "reduce": lambda: dist.reduce(plain, dst=0),
                    "broadcast": lambda: dist.broadcast(plain, src=0),
                    "reduce_scatter": lambda: dist.reduce_scatter_tensor(o_plain, plain)}
            for op in os.environ.get("SYMM_OPS", "allreduce,reduce_scatter,reduce,broadcast").split(","):
                row = {"op": op, "bytes": nbytes}
                for ctas in ctas_list:
                    comm.set_ctas(ctas)
                    sfx = "" if ctas == 0 else f"_c{ctas}"
                    for vname, env in VARIANTS:
                        setenv(**env)
                        ms = timeit(lambda: ours[op](a, o_sym), iters)
                        row[f"symm_{vname}{sfx}_us"] = ms * 1e3; row[f"symm_{vname}{sfx}_busbw"] = wire[op] / ms / 1e6
                    if nbytes >= (1 << 22):
                        setenv(**VARIANTS[1][1])
                        comm.set_zero_copy_min_bytes(1)
                        ms = timeit(lambda: ours[op](plain, o_plain), iters)
                        row[f"ipc{sfx}_us"] = ms * 1e3; row[f"ipc{sfx}_busbw"] = wire[op] / ms / 1e6
                        comm.set_zero_copy_min_bytes(16 << 20)
                comm.set_ctas(0)
                ms = timeit(nccl[op], iters)
                row["nccl_us"] = ms * 1e3; row["nccl_busbw"] = wire[op] / ms / 1e6
                res["perf"].append(row)
                if rank == 0: P("PERF", json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}))
            del plain, o_plain
    except Exception as e:
        import traceback; traceback.print_exc(); res["perf_exception"] = repr(e); ok_all = False

res["ok"] = bool(ok_all)
flag = torch.tensor([1 if ok_all else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/symm_{world}gpu.json", "w"), indent=1)
    P("SYMM CHECK", "OK" if flag.item() == 1 else "FAILED", f"({len(res['checks'])} checks)")
comm.destroy()
dist.destroy_process_group()
sys.exit(0 if flag.item() == 1 else 1)
