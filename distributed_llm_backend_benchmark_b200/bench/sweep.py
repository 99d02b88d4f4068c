"""Collective micro-benchmark engine behind collectives/{1d,3d}/*.py.

The reference ships the same sweep eight times (once per backend and tensor rank);
here one engine runs all of them.  A *flavour* fixes what differed between those files:

  ``mpi``        mpi4py communicator; numpy fp16 vectors (1-D) or fp32 up-cast of bf16 tensors (3-D,
                 ref collectives/3d/openmpi.py:43); object API for every op except Bcast / Isend /
                 Irecv (ref collectives/1d/openmpi.py:55-198); MPI result schema (§2.9)
  ``deepspeed``  deepspeed.comm (= torch.distributed); torch tensors; DeepSpeed result schema

Per iteration the engine keeps the reference's measurement (barrier, host clock around the call) and,
on CUDA, adds a device measurement with CUDA events; both go into the result file (``timings`` keeps
its meaning so stats.py works unchanged, ``timings_device`` is what this project reports, reduced as
max over ranks).  Tensors live on the GPU by default (``DLBB_SWEEP_DEVICE=cpu`` restores the
reference's host tensors, which the backend then stages through the GPU).

Environment knobs: DLBB_SWEEP_DEVICE, DLBB_SWEEP_ITERS, DLBB_SWEEP_WARMUP,
DLBB_SWEEP_OPS (comma list), DLBB_SWEEP_SIZES (1-D: comma list of labels or element counts),
DLBB_SWEEP_MAX_MB (3-D: skip tensors above this size), DLBB_SWEEP_STRIDED=1 (3-D: communicate a
non-contiguous [b,s,h] view, packed inside the kernel), DLBB_OUTPUT_DIR, DLBB_SKIP_EXISTING=1,
DLBB_SWEEP_SYMMETRIC (default 1 on the native backend: the communicated tensors live in symmetric memory, which the
NVLS / direct kernels work on in place; 0 = ordinary caching-allocator tensors), DLBB_SWEEP_TIGHT (default 1 on CUDA: after
the reference-style loop the same call is also timed back to back between two CUDA events - ``device_tight_us`` - because
the barrier-then-event bracket of the per-call measurement includes the host's time to enqueue the call on an idle GPU).
"""
from __future__ import annotations

import json
import os
import sys
import time
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np
import torch

DATA_SIZES_1D = {"1KB": 256, "64KB": 16384, "1MB": 262144, "16MB": 4194304}
EXTENDED_SIZES_1D = {"1KB": 512, "8KB": 4096, "64KB": 32768, "512KB": 262144, "4MB": 2097152, "32MB": 16777216,
                     "128MB": 67108864, "1GB": 536870912}      # true payload bytes / 2 (fp16 elements)
OPS_1D = ["allreduce", "allgather", "broadcast", "gather", "scatter", "reduce", "alltoall", "sendrecv"]
OPS_3D = ["allreduce", "allgather", "broadcast", "gather", "reduce"]


def _env_list(name, default):
    v = os.environ.get(name)
    return [s.strip() for s in v.split(",") if s.strip()] if v else default


def sweep_device() -> torch.device:
    want = os.environ.get("DLBB_SWEEP_DEVICE", "cuda" if torch.cuda.is_available() else "cpu")
    if want == "cuda" and torch.cuda.is_available():
        return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())
    return torch.device("cpu")


class _Clock:
    """Host + device time of one call.  ``host`` matches the reference's perf_counter pair."""

    def __init__(self, device: torch.device):
        self.cuda = device.type == "cuda"
        self.device = device
        if self.cuda:
            self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def measure(self, fn):
        if self.cuda:
            self.e0.record()
            t0 = time.perf_counter()
            fn()
            self.e1.record()
            self.e1.synchronize()
            t1 = time.perf_counter()
            return t1 - t0, self.e0.elapsed_time(self.e1) / 1e3
        t0 = time.perf_counter()
        fn()
        t1 = time.perf_counter()
        return t1 - t0, None


# ------------------------------------------------------------------------------------------------
# transports
# ------------------------------------------------------------------------------------------------
class TorchTransport:
    """deepspeed.comm / torch.distributed call sites (ref collectives/1d/dsccl.py:59-163)."""

    def __init__(self, dist, rank, world, device, symmetric_alloc=None):
        self.d, self.rank, self.world, self.device = dist, rank, world, device
        self.clock = _Clock(device)
        self.symm_alloc = symmetric_alloc          # callable(nbytes) -> uint8 tensor in symmetric memory, or None
        self._arenas: dict[str, torch.Tensor] = {}
        self.tight = device.type == "cuda" and os.environ.get("DLBB_SWEEP_TIGHT", "1") == "1"

    def barrier(self):
        self.d.barrier()

    def buffer(self, role: str, like: torch.Tensor) -> torch.Tensor:
        """A tensor shaped like `like`: in symmetric memory when the backend offers it (one growing arena per role,
        allocated collectively - every rank asks for the same sizes in the same order), else an ordinary tensor."""
        if self.symm_alloc is None:
            return torch.empty_like(like, memory_format=torch.contiguous_format)
        nbytes = like.numel() * like.element_size()
        arena = self._arenas.get(role)
        if arena is None or arena.numel() < nbytes:
            arena = self.symm_alloc(max(nbytes, 2 * arena.numel() if arena is not None else 0, 1 << 21))
            self._arenas[role] = arena
        return arena[:nbytes].view(like.dtype).view(like.shape)

    def _loop(self, iterations, prepare, call):
        host, dev = [], []
        for _ in range(iterations):
            args = prepare()
            self.barrier()
            h, dv = self.clock.measure(lambda: call(*args))
            host.append(h)
            dev.append(dv)
        tight = None
        if self.tight:
            args = prepare()
            for _ in range(2):
                call(*args)
            torch.cuda.synchronize(self.device)
            self.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = max(4, min(iterations, 20))
            e0.record()
            for _ in range(n):
                call(*args)
            e1.record()
            e1.synchronize()
            tight = e0.elapsed_time(e1) / 1e3 / n
        return host, (dev if dev and dev[0] is not None else None), tight

    def _work_copy(self, data):
        """The reference clones its input before every in-place call (outside the timed region); here the copy goes
        into a persistent work buffer so that it can live in symmetric memory."""
        if data.is_contiguous():
            work = self.buffer("work", data)
        else:
            # keep the view's geometry: the work tensor is a view with the same sizes and strides over its own storage,
            # so the backend sees the strided tensor (and packs / unpacks it inside the kernel), not a dense copy
            span = 1 + sum((sz - 1) * st for sz, st in zip(data.shape, data.stride()))
            flat = self.buffer("work", torch.empty(0, dtype=data.dtype, device=data.device).new_empty(span))
            work = torch.as_strided(flat, data.shape, data.stride())

        def prepare():
            work.copy_(data)
            return (work,)

        return prepare

    def allreduce(self, data, iterations):
        return self._loop(iterations, self._work_copy(data), lambda t: self.d.all_reduce(t, op=self.d.ReduceOp.SUM))

    def allgather(self, data, iterations):
        out = [torch.zeros_like(data, memory_format=torch.contiguous_format) for _ in range(self.world)]
        return self._loop(iterations, lambda: (), lambda: self.d.all_gather(out, data))

    def broadcast(self, data, iterations):
        return self._loop(iterations, self._work_copy(data), lambda t: self.d.broadcast(t, src=0))

    def gather(self, data, iterations):
        out = [torch.zeros_like(data, memory_format=torch.contiguous_format) for _ in range(self.world)] if self.rank == 0 else None
        return self._loop(iterations, lambda: (), lambda: self.d.gather(data, gather_list=out, dst=0))

    def scatter(self, data, iterations):
        src = [data.clone() for _ in range(self.world)] if self.rank == 0 else None
        out = torch.zeros_like(data)
        return self._loop(iterations, lambda: (), lambda: self.d.scatter(out, scatter_list=src, src=0))

    def reduce(self, data, iterations):
        return self._loop(iterations, self._work_copy(data), lambda t: self.d.reduce(t, dst=0, op=self.d.ReduceOp.SUM))

    def reducescatter(self, data, iterations):
        n = data.numel() // self.world
        flat_src = data.reshape(-1)[:n * self.world]
        flat = self.buffer("work", flat_src)
        flat.copy_(flat_src)
        out = self.buffer("out", flat[:n])
        return self._loop(iterations, lambda: (), lambda: self.d.reduce_scatter_tensor(out, flat))

    def alltoall(self, data, iterations):
        chunk = data.numel() // self.world
        ins = [data.reshape(-1)[i * chunk:(i + 1) * chunk].clone() for i in range(self.world)]
        outs = [torch.zeros_like(c) for c in ins]
        return self._loop(iterations, lambda: (), lambda: self.d.all_to_all(outs, ins))

    def sendrecv(self, data, iterations):
        nxt, prv = (self.rank + 1) % self.world, (self.rank - 1) % self.world
        recv = torch.zeros_like(data)

        def ring():
            s = self.d.isend(data, dst=nxt)
            r = self.d.irecv(recv, src=prv)
            s.wait()
            r.wait()

        return self._loop(iterations, lambda: (), ring)


class MpiTransport:
    """mpi4py call sites: object (pickle) API everywhere except Bcast and the Isend/Irecv ring
    (ref collectives/1d/openmpi.py:55-198, collectives/3d/openmpi.py:37-123)."""

    def __init__(self, MPI, comm):
        self.MPI, self.c = MPI, comm
        self.rank, self.world = comm.Get_rank(), comm.Get_size()

    def barrier(self):
        self.c.Barrier()

    def _loop(self, iterations, call):
        host = []
        for _ in range(iterations):
            self.c.Barrier()
            t0 = self.MPI.Wtime()
            call()
            host.append(self.MPI.Wtime() - t0)
        return host, None, None

    def allreduce(self, data, iterations):
        return self._loop(iterations, lambda: self.c.allreduce(data, op=self.MPI.SUM))

    def allgather(self, data, iterations):
        return self._loop(iterations, lambda: self.c.allgather(data))

    def broadcast(self, data, iterations):
        buf = data.copy() if self.rank == 0 else np.empty_like(data)
        return self._loop(iterations, lambda: self.c.Bcast(buf, root=0))

    def gather(self, data, iterations):
        return self._loop(iterations, lambda: self.c.gather(data, root=0))

    def scatter(self, data, iterations):
        src = [data.copy() for _ in range(self.world)] if self.rank == 0 else None
        return self._loop(iterations, lambda: self.c.scatter(src, root=0))

    def reduce(self, data, iterations):
        return self._loop(iterations, lambda: self.c.reduce(data, op=self.MPI.SUM, root=0))

    def reducescatter(self, data, iterations):
        n = data.size // self.world
        out = np.empty(n, dtype=data.dtype)
        flat = np.ascontiguousarray(data.reshape(-1)[:n * self.world])

        def call():
            full = np.empty_like(flat)
            self.c.Allreduce(flat, full, op=self.MPI.SUM)
            out[:] = full[self.rank * n:(self.rank + 1) * n]

        return self._loop(iterations, call)

    def alltoall(self, data, iterations):
        chunk = data.size // self.world
        src = [data.reshape(-1)[i * chunk:(i + 1) * chunk].copy() for i in range(self.world)]
        return self._loop(iterations, lambda: self.c.alltoall(src))

    def sendrecv(self, data, iterations):
        nxt, prv = (self.rank + 1) % self.world, (self.rank - 1) % self.world
        recv = np.empty_like(data)

        def ring():
            s = self.c.Isend(data, dest=nxt)
            r = self.c.Irecv(recv, source=prv)
            s.Wait()
            r.Wait()

        return self._loop(iterations, ring)


# ------------------------------------------------------------------------------------------------
# sweep specification and driver
# ------------------------------------------------------------------------------------------------
@dataclass
class SweepSpec:
    flavour: str                      # "mpi" | "deepspeed"
    implementation: str               # name written into the result files and file names
    backend: str | None = None        # deepspeed: dist backend ("ccl", "gloo", "dlbb", "nccl")
    rank_counts: list = field(default_factory=lambda: [1, 2, 4, 8, 16, 32, 56])
    operations: list = field(default_factory=lambda: list(OPS_1D))
    data_sizes: dict = field(default_factory=lambda: dict(DATA_SIZES_1D))
    batch_sizes: list = field(default_factory=lambda: [1, 8, 16, 32])
    seq_lengths: list = field(default_factory=lambda: [1, 2048, 4096, 8192])
    hidden_dims: list = field(default_factory=lambda: [2048, 4096])
    warmup: int = 10
    iterations: int = 100
    output_dir: str = "results"
    exit_on_bad_world: bool = True


def _apply_env(spec: SweepSpec, three_d: bool) -> SweepSpec:
    spec.iterations = int(os.environ.get("DLBB_SWEEP_ITERS", spec.iterations))
    spec.warmup = int(os.environ.get("DLBB_SWEEP_WARMUP", spec.warmup))
    spec.operations = _env_list("DLBB_SWEEP_OPS", spec.operations)
    spec.output_dir = os.environ.get("DLBB_OUTPUT_DIR", spec.output_dir)
    if os.environ.get("DLBB_SWEEP_BACKEND") and spec.flavour == "deepspeed":
        # baseline arm: the same loops on another torch.distributed backend (e.g. nccl), SURVEY.md 6.4
        spec.backend = os.environ["DLBB_SWEEP_BACKEND"]
        spec.implementation = f"deepspeed_{spec.backend}"
    if not three_d and os.environ.get("DLBB_SWEEP_SIZES"):
        chosen = {}
        for tok in _env_list("DLBB_SWEEP_SIZES", []):
            if tok == "extended":
                chosen.update(EXTENDED_SIZES_1D)
            elif tok in spec.data_sizes:
                chosen[tok] = spec.data_sizes[tok]
            elif tok in EXTENDED_SIZES_1D:
                chosen[tok] = EXTENDED_SIZES_1D[tok]
            else:
                chosen[f"{int(tok)}el"] = int(tok)
        spec.data_sizes = chosen
    for env, attr in (("DLBB_SWEEP_BATCH", "batch_sizes"), ("DLBB_SWEEP_SEQ", "seq_lengths"),
                      ("DLBB_SWEEP_HIDDEN", "hidden_dims")):
        if os.environ.get(env):
            setattr(spec, attr, [int(x) for x in _env_list(env, [])])
    return spec


class _World:
    """Initialised transport + helpers common to both flavours."""

    def __init__(self, spec: SweepSpec):
        self.spec = spec
        self.device = sweep_device()
        if spec.flavour == "mpi":
            from mpi4py import MPI
            self.MPI = MPI
            self.comm = MPI.COMM_WORLD
            self.rank, self.world = self.comm.Get_rank(), self.comm.Get_size()
            self.transport = MpiTransport(MPI, self.comm)
            self.device = torch.device("cpu")   # numpy buffers; the facade stages them through the GPU
        else:
            import deepspeed
            import deepspeed.comm as dist
            deepspeed.init_distributed(dist_backend=spec.backend)
            self.deepspeed, self.dist = deepspeed, dist
            self.rank, self.world = dist.get_rank(), dist.get_world_size()
            if spec.backend == "gloo":
                self.device = torch.device("cpu")
            # bookkeeping collectives (timing gather, skip votes) use host tensors, except on backends
            # that only serve CUDA tensors (nccl)
            self.helper_device = self.device if spec.backend == "nccl" else torch.device("cpu")
            self.transport = TorchTransport(dist, self.rank, self.world, self.device, self._symmetric_allocator())

    def _symmetric_allocator(self):
        """`nbytes -> uint8 tensor` in the native backend's symmetric memory, or None (other backends, CPU tensors,
        one rank, DLBB_SWEEP_SYMMETRIC=0)."""
        if (self.device.type != "cuda" or self.world == 1 or self.spec.backend not in ("ccl", "dlbb")
                or os.environ.get("DLBB_SWEEP_SYMMETRIC", "1") != "1"):
            return None
        try:
            from ..parallel.backend import native_backend
            be = native_backend()
            if not hasattr(be, "symm_empty"):
                return None
            return lambda nbytes: be.symm_empty([int(nbytes)], torch.uint8)
        except Exception:  # noqa: BLE001 - a Gloo fallback has no symmetric memory
            return None

    def gather_timings(self, values):
        """[world][iters] on rank 0 (ref: comm.gather of lists / dist.gather of a float64 tensor)."""
        if values is None:
            return None
        if self.spec.flavour == "mpi":
            return self.comm.gather(list(values), root=0)
        t = torch.tensor(values, dtype=torch.float64, device=self.helper_device)
        if self.world == 1:
            return [t.tolist()]
        bucket = [torch.zeros_like(t) for _ in range(self.world)] if self.rank == 0 else None
        self.dist.gather(t, gather_list=bucket, dst=0)
        return [b.tolist() for b in bucket] if self.rank == 0 else None

    def any_failed(self, mine: bool) -> bool:
        """True on every rank if any rank failed - so all ranks skip the same configurations."""
        if self.world == 1:
            return mine
        if self.spec.flavour == "mpi":
            return bool(self.comm.allreduce(1 if mine else 0, op=self.MPI.MAX))
        flag = torch.tensor([1.0 if mine else 0.0], device=self.helper_device)
        self.dist.all_reduce(flag, op=self.dist.ReduceOp.MAX)
        return bool(flag.item() > 0)

    def finish(self):
        if self.spec.flavour == "deepspeed" and self.dist.is_initialized():
            self.dist.destroy_process_group()


def _check_world(w: _World) -> bool:
    spec = w.spec
    if w.world in spec.rank_counts:
        return True
    if w.rank == 0:
        print(f"Error: Current world size ({w.world}) not in RANK_COUNTS {spec.rank_counts}")
    if spec.exit_on_bad_world:
        sys.exit(1)
    return False


def _run_config(w: _World, op: str, data, label: str):
    """warm-up + measurement with the reference's per-configuration try/except (a failing op is
    reported and skipped, collectives/1d/dsccl.py:207-223) - but decided collectively, so ranks
    cannot desynchronise when only some of them fail."""
    fn = getattr(w.transport, op, None)
    err = None
    host = dev = tight = None
    try:
        if fn is None:
            raise NotImplementedError(f"operation {op!r} is not available for this flavour")
        fn(data, w.spec.warmup)
        host, dev, tight = fn(data, w.spec.iterations)
    except Exception as e:  # noqa: BLE001
        err = e
    if w.any_failed(err is not None):
        if err is not None and w.rank == 0:
            print(f"  ERROR in {op} {label}: {err}  (configuration skipped on every rank)")
        return None, None, None
    return host, dev, tight


def _tight_fields(w: "_World", op: str, nbytes: int, tight) -> dict:
    """Back-to-back device time of one call (max over ranks) and the bus bandwidth it gives."""
    per_rank = w.gather_timings([tight]) if tight is not None else None
    if w.rank != 0 or per_rank is None:
        return {}
    from ..utils.stats import bus_bandwidth
    worst = max(v[0] for v in per_rank)
    alg, bus = bus_bandwidth(op, w.world, nbytes, worst)
    return {"device_tight_us": worst * 1e6, "algbw_tight_gbps": alg, "busbw_tight_gbps": bus}


def _busbw(op, world, nbytes, dev_or_host):
    from ..utils.stats import bus_bandwidth
    arr = np.array(dev_or_host)
    med = float(np.median(arr.max(axis=0)))
    alg, bus = bus_bandwidth(op, world, nbytes, med)
    return med, alg, bus


def run_sweep_1d(spec: SweepSpec) -> None:
    spec = _apply_env(spec, three_d=False)
    w = _World(spec)
    if not _check_world(w):
        return
    out_dir = Path(spec.output_dir)
    if w.rank == 0:
        out_dir.mkdir(parents=True, exist_ok=True)
        print(f"Starting benchmarks for {spec.implementation}")
        print(f"World size: {w.world}   device: {w.device}   torch {torch.__version__}")
        print(f"Operations: {spec.operations}")
        print(f"Data sizes: {list(spec.data_sizes.keys())}")
        print(f"Warmup iterations: {spec.warmup}   Measurement iterations: {spec.iterations}")
        print("-" * 80)
    for op in spec.operations:
        for size_name, n in spec.data_sizes.items():
            fname = f"{spec.implementation}_{op}_ranks{w.world}_{size_name}.json"
            if os.environ.get("DLBB_SKIP_EXISTING") == "1" and (out_dir / fname).exists():
                continue
            if w.rank == 0:
                print(f"Benchmarking {op} with {size_name} ({n} elements)...")
            if spec.flavour == "mpi":
                np.random.seed(42 + w.rank)
                data = np.random.randn(n).astype(np.float16)
            else:
                g = torch.Generator(device="cpu")
                g.manual_seed(42 + w.rank)
                data = torch.randn(n, dtype=torch.float16, generator=g).to(w.device)
            host, dev, tight = _run_config(w, op, data, size_name)
            if host is None:
                continue
            all_host, all_dev = w.gather_timings(host), w.gather_timings(dev)
            tight_fields = _tight_fields(w, op, n * 2, tight)
            if w.rank != 0:
                continue
            med, alg, bus = _busbw(op, w.world, n * 2, all_dev or all_host)
            if spec.flavour == "mpi":
                result = {"mpi_implementation": spec.implementation, "operation": op, "num_ranks": w.world,
                          "data_size_name": size_name, "num_elements": n, "dtype": str(np.float16),
                          "warmup_iterations": spec.warmup, "measurement_iterations": spec.iterations,
                          "timings": all_host}
            else:
                result = {"implementation": spec.implementation, "backend": spec.backend, "operation": op,
                          "num_ranks": w.world, "data_size_name": size_name, "num_elements": n, "dtype": "float16",
                          "warmup_iterations": spec.warmup, "measurement_iterations": spec.iterations,
                          "timing_method": "time.perf_counter()", "timings": all_host}
            result.update({"device": str(w.device), "element_size": 2, "timings_device": all_dev,
                           "median_max_over_ranks_us": med * 1e6, "algbw_gbps": alg, "busbw_gbps": bus,
                           "symmetric_memory": getattr(w.transport, "symm_alloc", None) is not None, **tight_fields})
            (out_dir / fname).write_text(json.dumps(result, indent=2))
            print(f"  Saved: {fname}   median {med * 1e6:.1f} us   busbw {bus:.2f} GB/s"
                  + (f"   back-to-back {tight_fields['device_tight_us']:.1f} us, {tight_fields['busbw_tight_gbps']:.2f} GB/s"
                     if tight_fields else ""))
    if w.rank == 0:
        print("-" * 80)
        print(f"Benchmarking complete!  Results saved in '{spec.output_dir}/' directory")
    w.finish()


def run_sweep_3d(spec: SweepSpec) -> None:
    spec = _apply_env(spec, three_d=True)
    w = _World(spec)
    if not _check_world(w):
        return
    out_dir = Path(spec.output_dir)
    max_mb = float(os.environ.get("DLBB_SWEEP_MAX_MB", "4096"))
    strided = os.environ.get("DLBB_SWEEP_STRIDED", "0") == "1" and spec.flavour != "mpi"
    if w.rank == 0:
        out_dir.mkdir(parents=True, exist_ok=True)
        print(f"Starting 3-D tensor benchmarks for {spec.implementation}")
        print(f"World size: {w.world}   device: {w.device}   strided views: {strided}")
        print(f"Operations: {spec.operations}")
        print(f"batch {spec.batch_sizes} x seq {spec.seq_lengths} x hidden {spec.hidden_dims}")
        print("-" * 80)
    for op in spec.operations:
        for b in spec.batch_sizes:
            for s in spec.seq_lengths:
                for h in spec.hidden_dims:
                    n = b * s * h
                    nbytes = n * 2                              # reported as bf16 bytes, as in the reference
                    mb = nbytes / (1024 ** 2)
                    if mb > max_mb:
                        continue
                    fname = f"{spec.implementation}_{op}_ranks{w.world}_b{b}_s{s}_h{h}.json"
                    if os.environ.get("DLBB_SKIP_EXISTING") == "1" and (out_dir / fname).exists():
                        continue
                    if w.rank == 0:
                        print(f"Benchmarking {op} with shape ({b}, {s}, {h}) = {mb:.2f} MB...")
                    g = torch.Generator(device="cpu")
                    g.manual_seed(42 + w.rank)
                    if strided:
                        base = torch.randn(b, s, 2 * h, dtype=torch.bfloat16, generator=g).to(w.device)
                        tensor = base[:, :, :h]                 # non-contiguous [b, s, h] view
                    else:
                        tensor = torch.randn(b, s, h, dtype=torch.bfloat16, generator=g).to(w.device)
                    data = tensor.float().cpu().numpy() if spec.flavour == "mpi" else tensor
                    host, dev, tight = _run_config(w, op, data, f"({b},{s},{h})")
                    if host is None:
                        continue
                    all_host, all_dev = w.gather_timings(host), w.gather_timings(dev)
                    wire = nbytes * (2 if spec.flavour == "mpi" else 1)
                    tight_fields = _tight_fields(w, op, wire, tight)
                    del tensor, data
                    if w.rank != 0:
                        continue
                    med, alg, bus = _busbw(op, w.world, wire, all_dev or all_host)
                    key = "mpi_implementation" if spec.flavour == "mpi" else "implementation"
                    result = {key: spec.implementation, "operation": op, "num_ranks": w.world,
                              "tensor_shape": {"batch": b, "seq_len": s, "hidden_dim": h}, "num_elements": n,
                              "tensor_size_bytes": nbytes, "tensor_size_mb": mb, "dtype": "bfloat16",
                              "warmup_iterations": spec.warmup, "measurement_iterations": spec.iterations,
                              "timing_method": "MPI.Wtime()" if spec.flavour == "mpi" else "time.perf_counter()",
                              "timings": all_host}
                    if spec.flavour == "mpi":
                        result["dtype_note"] = "BF16 converted to FP32 for MPI (numpy limitation)"
                    else:
                        result["backend"] = spec.backend
                    result.update({"device": str(w.device), "strided_view": strided, "timings_device": all_dev,
                                   "median_max_over_ranks_us": med * 1e6, "algbw_gbps": alg, "busbw_gbps": bus,
                                   "symmetric_memory": getattr(w.transport, "symm_alloc", None) is not None and not strided,
                                   **tight_fields})
                    (out_dir / fname).write_text(json.dumps(result, indent=2))
                    print(f"  Saved: {fname}   median {med * 1e3:.3f} ms   busbw {bus:.2f} GB/s"
                          + (f"   back-to-back {tight_fields['device_tight_us'] / 1e3:.3f} ms, "
                             f"{tight_fields['busbw_tight_gbps']:.2f} GB/s" if tight_fields else ""))
    if w.rank == 0:
        print("-" * 80)
        print(f"Benchmarking complete!  Results saved in '{spec.output_dir}/' directory")
    w.finish()
