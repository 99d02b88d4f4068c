#!/usr/bin/env python3
"""Multi-GPU self-check of the native communicator (run under torchrun, one rank per GPU):
numerics of every collective against locally computed expectations, the fused DP step against an
explicit average+Adam, and all-reduce bus bandwidth next to NCCL.  Writes gpurun_out/comm_check_rank0.json."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_llm_backend_benchmark_b200 import ops
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
comm = SymmComm.from_env(dev, staging_mb=64)     # small staging: the 80 MB cases below exercise the chunked paths
res = {"world": world, "multicast": comm.has_multicast, "checks": {}, "perf": []}
ok_all = True
P = lambda *a: print(f"[r{rank}]", *a, flush=True)
if rank == 0: P("comm up; multicast:", comm.has_multicast)

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

try:
    # symmetric memory + peer visibility
    t, seg = comm.alloc(1 << 16, torch.float32)
    t.fill_(float(rank + 1)); torch.cuda.synchronize(); comm.host_barrier()
    for p in range(world):
        check(f"peer_view_{p}", comm.peer_view(seg, p, 1 << 16, torch.float32)[:8], torch.full((8,), float(p + 1), device=dev))
    comm.barrier(); torch.cuda.synchronize()

    sizes = [1, 7, 256, 4099, 1 << 16, (1 << 20) + 24, 5 << 20]
    for dtype, tol in ((torch.bfloat16, 2e-2), (torch.float16, 4e-3), (torch.float32, 1e-5), (torch.float64, 1e-12)):
        for n in sizes:
            xs = [gen(r, n, dtype, seed=n % 97) for r in range(world)]
            for algo in ("oneshot", "twoshot", "auto") + (("nvls",) if comm.has_multicast and dtype != torch.float64 else ()):
                x = xs[rank].clone()
                comm.all_reduce(x, "sum", algo)
                ref = sum(v.double() for v in xs)
                check(f"allreduce_{algo}_{str(dtype)[6:]}_{n}", x, ref.to(dtype), tol)
        n = 4099
        xs = [gen(r, n, dtype, seed=5).abs() * 0.5 + 0.5 for r in range(world)]
        for op, f in (("max", torch.maximum), ("min", torch.minimum), ("prod", torch.mul)):
            x = xs[rank].clone(); comm.all_reduce(x, op)
            ref = xs[0].double()
            for v in xs[1:]: ref = f(ref, v.double())
            check(f"allreduce_{op}_{str(dtype)[6:]}", x, ref.to(dtype), tol)
    # identical results on every rank (deterministic rank-order reduction)
    x = gen(rank, 100003, torch.bfloat16, seed=9); comm.all_reduce(x)
    g = [torch.empty_like(x) for _ in range(world)]; dist.all_gather(g, x)
    check("allreduce_bitwise_equal_across_ranks", g[0].float(), g[-1].float())

    # strided 3-D view packed in-kernel
    base = gen(rank, 4 * 6 * 64, torch.bfloat16, seed=3).view(4, 6, 64)
    bases = [gen(r, 4 * 6 * 64, torch.bfloat16, seed=3).view(4, 6, 64) for r in range(world)]
    for name, sl in (("inner", (slice(None), slice(None), slice(0, 32))), ("outer", (slice(1, 3), slice(None), slice(None))),
                     ("mixed", (slice(None), slice(0, 6, 2), slice(8, 40))), ("transposed", None)):
        v = base.transpose(0, 1) if sl is None else base[sl]
        out = torch.empty(v.shape, dtype=v.dtype, device=dev)
        comm.all_reduce(v, out=out)
        ref = sum((b.transpose(0, 1) if sl is None else b[sl]).double() for b in bases)
        check(f"allreduce_strided_{name}", out, ref.to(torch.bfloat16), 2e-2)

    for n in (1, 1000, (1 << 18) + 8):
        for dtype in (torch.float16, torch.bfloat16, torch.float32, torch.float64, torch.uint8):
            mk = (lambda r: gen(r, n, dtype, seed=2)) if dtype != torch.uint8 else (lambda r: (torch.arange(n, device=dev) * (r + 3) % 251).to(torch.uint8))
            xs = [mk(r) for r in range(world)]
            dn = str(dtype)[6:]
            out = torch.empty(world * n, dtype=dtype, device=dev); comm.all_gather(xs[rank], out)
            check(f"allgather_{dn}_{n}", out, torch.cat(xs))
            b = xs[rank].clone(); comm.broadcast(b, root=world - 1); check(f"broadcast_{dn}_{n}", b, xs[world - 1])
            root = 1 % world
            out = torch.zeros(world * n, dtype=dtype, device=dev); comm.gather(xs[rank], out, root=root)
            if rank == root: check(f"gather_{dn}_{n}", out, torch.cat(xs))
            big = [torch.cat([mk(r + 10 * q) for q in range(world)]) for r in range(world)]
            out = torch.empty(n, dtype=dtype, device=dev); comm.scatter(big[root] if rank == root else None, out, root=root)
            check(f"scatter_{dn}_{n}", out, big[root][rank * n:(rank + 1) * n])
            out = torch.empty(world * n, dtype=dtype, device=dev); comm.all_to_all(big[rank], out)
            check(f"alltoall_{dn}_{n}", out, torch.cat([big[p][rank * n:(rank + 1) * n] for p in range(world)]))
            if dtype != torch.uint8:
                tol = {torch.bfloat16: 2e-2, torch.float16: 4e-3, torch.float32: 1e-5, torch.float64: 1e-12}[dtype]
                out = torch.empty(n, dtype=dtype, device=dev); comm.reduce_scatter(big[rank], out)
                check(f"reduce_scatter_{dn}_{n}", out, sum(b[rank * n:(rank + 1) * n].double() for b in big).to(dtype), tol)
                o = xs[rank].clone(); comm.reduce(o, root=root)
                if rank == root: check(f"reduce_{dn}_{n}", o, sum(v.double() for v in xs).to(dtype), tol)
            # ring sendrecv + plain send/recv
            nxt, prv = (rank + 1) % world, (rank - 1) % world
            if world > 1:
                out = torch.empty(n, dtype=dtype, device=dev); comm.sendrecv(xs[rank], nxt, out, prv)
                check(f"sendrecv_{dn}_{n}", out, xs[prv])
                if rank == 0: comm.send(xs[0], 1)
                elif rank == 1:
                    out = torch.empty(n, dtype=dtype, device=dev); comm.recv(out, 0); check(f"send_recv_{dn}_{n}", out, xs[0])
    # a message larger than the staging region (chunked inside the call)
    n = (80 << 20) // 2 + 1032
    xs = [gen(r, n, torch.bfloat16, seed=11) for r in range(world)]
    x = xs[rank].clone(); comm.all_reduce(x); check("allreduce_big_bf16", x, sum(v.double() for v in xs).to(torch.bfloat16), 2e-2)
    out = torch.empty(world * n, dtype=torch.bfloat16, device=dev); comm.all_gather(xs[rank], out); check("allgather_big", out, torch.cat(xs))
    if world > 1:
        out = torch.empty(n, dtype=torch.bfloat16, device=dev); comm.sendrecv(xs[rank], (rank + 1) % world, out, (rank - 1) % world)
        check("sendrecv_big", out, xs[(rank - 1) % world])
    del xs, x, out
    torch.cuda.synchronize()

    # fused GEMM + all-reduce + residual (row-parallel linear in one kernel)
    for (M, N, Kt) in ((512, 1024, 256 * world), (200, 136, 64 * world), (4096, 4096, 512 * world)):
        Kr = Kt // world
        def mk(r, rows, cols, seed, scale=1.0):
            g = torch.Generator(device=dev); g.manual_seed(7000 + 31 * seed + r)
            return (torch.randn(rows, cols, device=dev, generator=g) * scale).bfloat16()
        As = [mk(r, M, Kr, 1) for r in range(world)]; Bs = [mk(r, Kr, N, 2, 0.2) for r in range(world)]
        resid_t = mk(0, M, N, 3, 4.0)
        part, out = comm.alloc_matrix(M, N), comm.alloc_matrix(M, N)
        ref = sum(ops.matmul(a, b).float() for a, b in zip(As, Bs)) + resid_t.float()   # bf16 partials, like the kernel
        for use_mc in ([False, True] if comm.has_multicast else [False]):
            for rep in range(2):
                out.tensor.fill_(7.0); torch.cuda.synchronize(); comm.host_barrier()
                comm.gemm_allreduce(As[rank], Bs[rank], part, out, residual=resid_t, use_multicast=use_mc)
                torch.cuda.synchronize(); comm.host_barrier()
            check(f"gemm_allreduce_{M}x{N}x{Kt}_mc{int(use_mc)}", out.tensor, ref.bfloat16(), 3e-2)
        # transposed B (weights stored [out, in]) and no residual
        Bt = [b.t().contiguous() for b in Bs]
        comm.gemm_allreduce(As[rank], Bt[rank], part, out, transpose_b=True, use_multicast=False)
        torch.cuda.synchronize(); comm.host_barrier()
        check(f"gemm_allreduce_tb_{M}x{N}x{Kt}", out.tensor, (ref - resid_t.float()).bfloat16(), 3e-2)

    # fused reduce + Adam + broadcast
    for use_mc in ([False, True] if comm.has_multicast else [False]):
        n = 3 * 8 * world * 1000 + 8 * 5 + 3   # ragged on purpose
        numel = n + 64
        params, grads = comm.alloc_dp_buffers(numel)
        off = 32
        p0 = gen(0, n, torch.bfloat16, seed=21)
        params[off:off + n].copy_(p0)
        gs = [gen(r, n, torch.bfloat16, seed=22) for r in range(world)]
        grads[off:off + n].copy_(gs[rank])
        per = -(-n // (world * 8)) * 8
        lo, hi = min(rank * per, n), min(rank * per + per, n)
        m = torch.zeros(hi - lo, device=dev); v = torch.zeros(hi - lo, device=dev); master = p0[lo:hi].float()
        pr = p0.clone(); mr = torch.zeros(n, device=dev); vr = torch.zeros(n, device=dev); mar = p0.float()
        for step in (1, 2):
            torch.cuda.synchronize(); comm.host_barrier()
            comm.fused_reduce_adam(off, n, m, v, master, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01,
                                   grad_scale=1.0 / world, step=step, use_multicast=use_mc)
            gsum = sum(g.float() for g in gs)
            if use_mc: gsum = gsum.bfloat16().float()   # the switch returns a bf16-rounded sum
            # reference Adam in torch (fp32 gradient)
            g = gsum / world + 0.01 * mar
            mr.mul_(0.9).add_(g, alpha=0.1); vr.mul_(0.999).addcmul_(g, g, value=0.001)
            mar -= 1e-2 * (mr / (1 - 0.9 ** step)) / (vr.sqrt() / (1 - 0.999 ** step) ** 0.5 + 1e-8)
            torch.cuda.synchronize(); comm.host_barrier()
        check(f"fused_adam_params_mc{int(use_mc)}", params[off:off + n], mar.bfloat16(), 1e-2)
        check(f"fused_adam_m_mc{int(use_mc)}", m, mr[lo:hi], 2e-2 if use_mc else 1e-4)
        check(f"fused_adam_untouched_mc{int(use_mc)}", params[:off].float().abs().max().reshape(1), torch.zeros(1, device=dev))
        allp = [torch.empty(n, dtype=torch.bfloat16, device=dev) for _ in range(world)]
        dist.all_gather(allp, params[off:off + n].contiguous())
        check(f"fused_adam_params_identical_mc{int(use_mc)}", allp[0].float(), allp[-1].float())
except Exception as e:
    import traceback; traceback.print_exc()
    ok_all = False; res["exception"] = repr(e)

# ---------------------------------------------------------------- bandwidth next to NCCL
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

if ok_all and os.environ.get("DLBB_SKIP_PERF") != "1":
    comm.destroy()
    comm = SymmComm.from_env(dev)                    # default staging for the bandwidth tables
if ok_all and os.environ.get("DLBB_SKIP_PERF") != "1":
    try:
        for nbytes in (1 << 10, 1 << 13, 1 << 16, 1 << 19, 1 << 22, 1 << 25, 1 << 27, 1 << 30):
            n = nbytes // 2
            x = torch.ones(n, dtype=torch.bfloat16, device=dev)
            iters = 50 if nbytes <= (1 << 22) else (10 if nbytes <= (1 << 27) else 4)
            row = {"bytes": nbytes}
            algos = ["oneshot", "twoshot"] + (["nvls"] if comm.has_multicast else [])
            for algo in algos:
                if algo == "oneshot" and nbytes > (1 << 25): continue
                ms = timeit(lambda: comm.all_reduce(x, "sum", algo), iters)
                row[f"{algo}_us"] = ms * 1e3; row[f"{algo}_busbw"] = nbytes / ms / 1e6 * 2 * (world - 1) / world
            ms = timeit(lambda: dist.all_reduce(x), iters)
            row["nccl_us"] = ms * 1e3; row["nccl_busbw"] = nbytes / ms / 1e6 * 2 * (world - 1) / world
            out = torch.empty(world * n, dtype=torch.bfloat16, device=dev) if nbytes * world <= (4 << 30) else None
            if out is not None:
                ms = timeit(lambda: comm.all_gather(x, out), iters); row["allgather_busbw"] = nbytes * (world - 1) / ms / 1e6
                ms = timeit(lambda: dist.all_gather_into_tensor(out, x), iters); row["nccl_allgather_busbw"] = nbytes * (world - 1) / ms / 1e6
            res["perf"].append(row)
            if rank == 0: P("PERF", json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}))
    except Exception as e:
        import traceback; traceback.print_exc(); res["perf_exception"] = repr(e)

if ok_all and os.environ.get("DLBB_SKIP_PERF") != "1":
    try:
        # 7B out-projection / FFN-down shapes under tensor parallelism: fused kernel vs GEMM followed by all-reduce
        for name, M, N, Kt in (("out_proj", 4096, 4096, 4096), ("ffn_down", 4096, 4096, 16384)):
            Kr = Kt // world
            A = torch.randn(M, Kr, device=dev).bfloat16(); B = (torch.randn(Kr, N, device=dev) * 0.1).bfloat16()
            resid = torch.randn(M, N, device=dev).bfloat16()
            part, out = comm.alloc_matrix(M, N), comm.alloc_matrix(M, N)
            tmp = torch.empty(M, N, device=dev, dtype=torch.bfloat16); o2 = torch.empty_like(tmp)
            row = {"shape": name, "M": M, "N": N, "K_total": Kt}
            row["gemm_only_us"] = timeit(lambda: ops.matmul(A, B, out=tmp), 20) * 1e3
            row["unfused_ours_us"] = timeit(lambda: (ops.matmul(A, B, out=tmp), comm.all_reduce_residual(tmp, resid, o2)), 20) * 1e3
            def nccl_path():
                torch.matmul(A, B, out=tmp); dist.all_reduce(tmp); torch.add(tmp, resid, out=o2)
            row["cublas_nccl_us"] = timeit(nccl_path, 20) * 1e3
            row["fused_p2p_us"] = timeit(lambda: comm.gemm_allreduce(A, B, part, out, residual=resid, use_multicast=False), 20) * 1e3
            if comm.has_multicast:
                row["fused_nvls_us"] = timeit(lambda: comm.gemm_allreduce(A, B, part, out, residual=resid, use_multicast=True), 20) * 1e3
            flops = 2.0 * M * N * Kr; wire = 2.0 * (world - 1) / world * M * N * 2
            row["roofline_us"] = max(flops / 1.4118e15, wire / 770e9) * 1e6
            res["perf"].append(row)
            if rank == 0: P("FUSED", json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in row.items()}))
    except Exception as e:
        import traceback; traceback.print_exc(); res["fused_perf_exception"] = repr(e)

res["ok"] = bool(ok_all)
flag = torch.tensor([1 if ok_all else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/comm_check_rank0.json", "w"), indent=1)
    print("COMM ALL OK" if flag.item() else "COMM SOME FAILED", flush=True)
comm.destroy()
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
