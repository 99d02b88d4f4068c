"""The native process-group backend, the shims and the entry points on real GPUs.  The collective self-checks run at
every world size of {2, 4, 8} the box offers; everything else runs at the largest of them (world size 1 on a one-GPU
box still runs the symmetric-memory kernels against itself)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
BIN = ROOT / "bin"


def nranks():
    """The largest supported world size this box can host: 8, 4, 2 or 1."""
    have = torch.cuda.device_count()
    for w in (8, 4, 2):
        if have >= w:
            return w
    return 1


WORLDS = [w for w in (2, 4, 8) if w <= torch.cuda.device_count()] or [1]


def run(cmd, cwd=ROOT, env=None, timeout=600):
    e = dict(os.environ)
    e.pop("RANK", None), e.pop("WORLD_SIZE", None)
    e.update(env or {})
    return subprocess.run([str(c) for c in cmd], cwd=cwd, env=e, capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("n", WORLDS)
def test_comm_selfcheck(native_lib, n):
    r = run([BIN / "mpirun", "-np", n, sys.executable, ROOT / "tools" / "gpu_check_comm.py"],
            env={"DLBB_SKIP_PERF": "1", "MASTER_PORT": "29631"})
    assert r.returncode == 0 and "COMM ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_comm_selfcheck_with_poisoned_staging(native_lib):
    """DLBB_DEBUG_POISON=1: every staging region is NaN-filled right before the kernel that rewrites it, so
    a protocol bug that reads staging data early shows up as NaN in the self-check instead of as stale but plausible
    numbers."""
    n = nranks()
    r = run([BIN / "mpirun", "-np", n, sys.executable, ROOT / "tools" / "gpu_check_comm.py"],
            env={"DLBB_SKIP_PERF": "1", "DLBB_DEBUG_POISON": "1", "MASTER_PORT": "29632"})
    assert r.returncode == 0 and "COMM ALL OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("n", WORLDS)
def test_zero_copy_collectives(native_lib, n):
    """Every collective directly on the ranks' own tensors (CUDA IPC mappings, csrc/bootstrap/peer_registry.cpp),
    forced at all sizes, against locally computed expectations."""
    if n < 2:
        pytest.skip("zero-copy collectives need two GPUs")
    r = run([BIN / "mpirun", "-np", n, sys.executable, ROOT / "tools" / "gpu_zero_copy.py"],
            env={"DLBB_SKIP_PERF": "1", "MASTER_PORT": "29633"})
    assert r.returncode == 0 and "ZERO-COPY CHECK OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("n", WORLDS)
def test_symmetric_tensor_collectives(native_lib, n):
    """In-place collectives on tensors in symmetric memory (multimem.ld_reduce / multimem.st where the box has NVLS,
    peer loads / stores otherwise) and the direct reduce / reduce-scatter / relay broadcast on plain tensors."""
    if n < 2:
        pytest.skip("needs two GPUs")
    r = run([BIN / "mpirun", "-np", n, sys.executable, ROOT / "tools" / "gpu_symm.py"],
            env={"DLBB_SKIP_PERF": "1", "SYMM_MAX_BYTES": str(64 << 20), "MASTER_PORT": "29634"})
    assert r.returncode == 0 and "SYMM CHECK OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_device_watchdog_names_the_missing_rank(native_lib, tmp_path):
    """a rank that never issues a collective must not hang the job.  With a 2 s deadline the waiting ranks'
    kernels trap with a message that names the absent peer, their processes fail and the launcher tears the job down."""
    n = nranks()
    if n < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "wd.py"
    script.write_text(f"""
import sys, os, time; sys.path[:0] = [{str(ROOT)!r}]
import torch
from distributed_llm_backend_benchmark_b200.parallel.symm import SymmComm
r = int(os.environ['RANK']); torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
comm = SymmComm.from_env()
x = torch.ones(1024, device='cuda'); comm.all_reduce(x); torch.cuda.synchronize(); comm.host_barrier()
if r == 1:
    time.sleep(30)          # never issues the second collective
    sys.exit(0)
comm.all_reduce(x)
torch.cuda.synchronize()    # the trap surfaces here
print('UNEXPECTED: all_reduce completed without rank 1')
""")
    r = run([BIN / "b200run", "-np", n, sys.executable, script], env={"DLBB_WATCHDOG_SECONDS": "2", "MASTER_PORT": "29635"},
            timeout=120)
    assert r.returncode != 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "UNEXPECTED" not in r.stdout
    assert "device barrier timeout" in r.stdout + r.stderr and "for rank 1" in r.stdout + r.stderr, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("script", ["test_open.py", "test_deepseed.py", "ccl.py", "test_ccl_backend.py"])
def test_smoke_scripts_on_native_backend(native_lib, script):
    r = run([BIN / "mpirun", "-n", nranks(), sys.executable, ROOT / "test" / script], env={"DLBB_LOG_LEVEL": "info"})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    if nranks() > 1:   # the native backend came up (not a Gloo fallback): rank 0 of the communicator says so
        assert "dlbb: comm up" in r.stderr, r.stderr[-2000:]


def test_backend_is_native_not_gloo(native_lib, tmp_path):
    script = tmp_path / "b.py"
    script.write_text(f"""
import sys; sys.path[:0] = [{str(ROOT)!r}, {str(ROOT / 'compat')!r}]
import torch, torch.distributed as dist, oneccl_bindings_for_pytorch
from distributed_llm_backend_benchmark_b200 import _lib
from distributed_llm_backend_benchmark_b200.parallel.backend import native_backend
dist.init_process_group('ccl')
r, w = dist.get_rank(), dist.get_world_size()
be = native_backend()
assert type(be).__name__ == 'BackendDlbb', type(be)
n0 = _lib.launch_count()
for dev in ('cuda', 'cpu'):
    for dt in (torch.float16, torch.bfloat16, torch.float32, torch.float64, torch.int64):
        t = torch.full((1000,), r + 1, dtype=dt, device=dev); dist.all_reduce(t)
        assert (t == w * (w + 1) // 2).all(), (dev, dt)
    t = torch.arange(6, dtype=torch.float32, device=dev).view(2, 3).t()   # non-contiguous
    ref = t.clone() * w; dist.all_reduce(t); assert torch.equal(t, ref)
    out = torch.empty(w * 4, device=dev); dist.all_gather_into_tensor(out, torch.full((4,), float(r), device=dev))
    assert torch.equal(out, torch.arange(w, device=dev).repeat_interleave(4).float())
    o = torch.empty(4, device=dev); dist.reduce_scatter_tensor(o, torch.ones(4 * w, device=dev)); assert (o == w).all()
    a = torch.arange(w * 2, dtype=torch.float32, device=dev) + 100 * r; b = torch.empty_like(a); dist.all_to_all_single(b, a)
    assert torch.equal(b, torch.cat([torch.arange(2 * r, 2 * r + 2, device=dev).float() + 100 * p for p in range(w)]))
    objs = [None] * w; dist.all_gather_object(objs, {{'r': r}}); assert objs == [{{'r': i}} for i in range(w)]
    if w > 1:
        nxt, prv = (r + 1) % w, (r - 1) % w
        rb = torch.empty(5, device=dev); rq = dist.irecv(rb, src=prv); sq = dist.isend(torch.full((5,), float(r), device=dev), dst=nxt)
        sq.wait(); rq.wait(); torch.cuda.synchronize(); assert (rb == prv).all()
dist.barrier()
assert _lib.launch_count() > n0, 'no native kernels were launched'
print('NATIVE_OK', r)
dist.destroy_process_group()
""")
    n = nranks()
    r = run([BIN / "mpirun", "-np", n, sys.executable, script])
    assert r.returncode == 0 and r.stdout.count("NATIVE_OK") == n, r.stdout[-3000:] + r.stderr[-3000:]


def test_strided_views_are_packed_and_unpacked_in_kernel(native_lib, tmp_path):
    """K12: non-contiguous tensors through all-reduce / broadcast / reduce / all-gather / gather of the process group -
    3-D [b, s, h] slices of a wider tensor (vector path) and a transposed fp32 matrix (element path); the bytes of the
    base tensor outside the view must stay untouched."""
    script = tmp_path / "v.py"
    script.write_text(f"""
import sys; sys.path[:0] = [{str(ROOT)!r}, {str(ROOT / 'compat')!r}]
import torch, torch.distributed as dist, oneccl_bindings_for_pytorch
dist.init_process_group('ccl')
r, w = dist.get_rank(), dist.get_world_size()
dev = 'cuda'
def mk(rank, kind):
    g = torch.Generator().manual_seed(100 + rank)
    if kind == '3d':
        base = torch.randn(3, 40, 256, generator=g).bfloat16().to(dev); return base, base[:, :, :128]
    if kind == 'big':
        base = torch.randn(8, 512, 2048, generator=g).bfloat16().to(dev); return base, base[:, :, 1024:]
    base = torch.randn(37, 53, generator=g).to(dev); return base, base.t()
for kind in ('3d', 'big', 't'):
    views = [mk(q, kind)[1] for q in range(w)]
    tot = sum(v.double() for v in views)
    tol = 2e-2 if kind != 't' else 1e-5
    def fresh():
        base, v = mk(r, kind); return base, v, base.clone()
    def outside_untouched(base, v, before):
        b2, bb = base.clone(), before.clone()
        (b2[:, :, :128] if kind == '3d' else b2[:, :, 1024:] if kind == 'big' else b2.t()).zero_()
        (bb[:, :, :128] if kind == '3d' else bb[:, :, 1024:] if kind == 'big' else bb.t()).zero_()
        return torch.equal(b2, bb)
    base, v, before = fresh(); assert not v.is_contiguous()
    dist.all_reduce(v)
    assert ((v.double() - tot).abs().max() / tot.abs().max()).item() < tol, ('allreduce', kind)
    assert outside_untouched(base, v, before), ('allreduce outside', kind)
    base, v, before = fresh(); dist.broadcast(v, src=w - 1)
    assert torch.equal(v, views[w - 1]), ('broadcast', kind)
    assert outside_untouched(base, v, before), ('broadcast outside', kind)
    base, v, before = fresh(); dist.reduce(v, dst=0)
    if r == 0: assert ((v.double() - tot).abs().max() / tot.abs().max()).item() < tol, ('reduce', kind)
    else: assert torch.equal(v, views[r]), ('reduce left a non-root tensor changed', kind)
    base, v, before = fresh()
    outs = [torch.empty(v.shape, dtype=v.dtype, device=dev) for _ in range(w)]; dist.all_gather(outs, v)
    assert all(torch.equal(o, views[q]) for q, o in enumerate(outs)), ('allgather', kind)
    outs = [torch.empty(v.shape, dtype=v.dtype, device=dev) for _ in range(w)] if r == 0 else None
    dist.gather(v, gather_list=outs, dst=0)
    if r == 0: assert all(torch.equal(o, views[q]) for q, o in enumerate(outs)), ('gather', kind)
torch.cuda.synchronize(); dist.barrier()
print(f'VIEWS_OK {{r}}', flush=True)
dist.destroy_process_group()
""")
    n = nranks()
    r = run([BIN / "mpirun", "-np", n, sys.executable, script])
    assert r.returncode == 0 and r.stdout.count("VIEWS_OK") == n, r.stdout[-3000:] + r.stderr[-3000:]


def test_coalesced_allreduce_is_one_launch(native_lib, tmp_path):
    """K13: all_reduce_coalesced of many small tensors goes through the multi-tensor kernel (one launch per bucket),
    DLBB_FUSION=0 through one collective per tensor; both give the same sums."""
    script = tmp_path / "c.py"
    script.write_text(f"""
import sys, os; sys.path[:0] = [{str(ROOT)!r}, {str(ROOT / 'compat')!r}]
import torch, torch.distributed as dist, oneccl_bindings_for_pytorch
from distributed_llm_backend_benchmark_b200 import _lib
dist.init_process_group('ccl')
r, w = dist.get_rank(), dist.get_world_size()
ts = [torch.full((n,), float(r + 1 + i), device='cuda', dtype=torch.float32) for i, n in enumerate([5, 1000, 77, 4096, 3] * 8)]
want = [sum(q + 1 + i for q in range(w)) for i in range(len(ts))]
torch.cuda.synchronize(); n0 = _lib.launch_count()
dist.all_reduce_coalesced(ts)
torch.cuda.synchronize(); launches = _lib.launch_count() - n0
for t, x in zip(ts, want): assert (t.double() == x).all(), (t.dtype, t[0].item(), x)
fused = os.environ.get('DLBB_FUSION', '1') == '1'
assert (launches <= 4) if (fused and w > 1) else (launches >= len(ts) or w == 1), launches
print(f'COALESCED_OK {{r}} {{launches}}', flush=True)
dist.destroy_process_group()
""")
    n = nranks()
    for fusion in ("1", "0"):
        r = run([BIN / "mpirun", "-np", n, sys.executable, script], env={"DLBB_FUSION": fusion})
        assert r.returncode == 0 and r.stdout.count("COALESCED_OK") == n, r.stdout[-3000:] + r.stderr[-3000:]


def test_run_mpi_on_gpu(native_lib, tmp_path):
    n = nranks()
    cfg = tmp_path / "small.yaml"
    cfg.write_text(f"""
experiment: {{name: small_world{n}, output_dir: {tmp_path}/results}}
model: {{size: small, hidden_size: 1024, num_layers: 2, num_heads: 8, ffn_intermediate: 4096}}
parallelism: {{world_size: {n}, cores_per_rank: 1}}
input: {{batch_size: 4, sequence_length: 256, seed: 42}}
execution: {{warmup_iterations: 2, benchmark_iterations: 3}}
system: {{omp_num_threads: 1, mkl_num_threads: 1}}
""")
    r = run([BIN / "mpirun", "-np", n, sys.executable, ROOT / "run_mpi.py", "--config", cfg, "--backend", "openmpi"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = json.loads((tmp_path / "results" / f"openmpi_small_world{n}.json").read_text())
    assert out["rank_0_summary"]["forward_device_mean"] > 0 and "gpu_name" in out["system_info"]


def test_sweep_on_gpu(native_lib, tmp_path):
    n = nranks()
    env = {"DLBB_SWEEP_ITERS": "5", "DLBB_SWEEP_WARMUP": "2", "DLBB_SWEEP_SIZES": "1KB,1MB",
           "DLBB_SWEEP_OPS": "allreduce,allgather,broadcast,reducescatter,alltoall,sendrecv" if n > 1 else "allreduce,allgather",
           "DLBB_OUTPUT_DIR": str(tmp_path / "r")}
    inc = ",".join(str(i) for i in range(n))
    r = run([BIN / "deepspeed", "--master_port", 29641, "--include", f"localhost:{inc}", "dsccl.py"],
            cwd=ROOT / "collectives" / "1d", env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads((tmp_path / "r" / f"deepspeed_oneccl_allreduce_ranks{n}_1MB.json").read_text())
    assert res["device"].startswith("cuda") and res["timings_device"] is not None and res["busbw_gbps"] > 0
    env3 = {"DLBB_SWEEP_ITERS": "3", "DLBB_SWEEP_WARMUP": "1", "DLBB_SWEEP_BATCH": "8", "DLBB_SWEEP_SEQ": "1,512",
            "DLBB_SWEEP_HIDDEN": "2048", "DLBB_SWEEP_OPS": "allreduce", "DLBB_SWEEP_STRIDED": "1",
            "DLBB_OUTPUT_DIR": str(tmp_path / "r3")}
    r = run([BIN / "deepspeed", "--num_gpus", n, "--master_port", 29642, "dsccl.py"], cwd=ROOT / "collectives" / "3d", env=env3)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads((tmp_path / "r3" / f"deepspeed_oneccl_allreduce_ranks{n}_b8_s512_h2048.json").read_text())
    assert res["strided_view"] is True


def test_dp_training_two_ranks_matches_single_process(native_lib, tmp_path):
    """N-rank data-parallel step (fused reduce+Adam over NVLink) == one process that averages the
    per-rank gradients itself."""
    n = nranks()
    script = tmp_path / "dp.py"
    script.write_text(f"""
import sys, os; sys.path[:0] = [{str(ROOT)!r}]
import torch
from distributed_llm_backend_benchmark_b200.train import Trainer
from distributed_llm_backend_benchmark_b200.models.llm import LLMEngine, LLMConfig
from distributed_llm_backend_benchmark_b200 import ops
cfg = {{"experiment": {{"name": "t", "output_dir": "x"}}, "model": {{"size": "t", "hidden_size": 256, "num_layers": 2, "num_heads": 4, "ffn_intermediate": 512}},
       "parallelism": {{"world_size": 1, "cores_per_rank": 1}}, "input": {{"batch_size": 2, "sequence_length": 64, "seed": 5}},
       "execution": {{"warmup_iterations": 1, "benchmark_iterations": 1}}, "system": {{}}}}
r, w = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
tr = Trainer(cfg, lr=1e-3)
for lv in tr.engine.layers:
    for t_ in (lv.wqkv, lv.wo, lv.wup, lv.wdown): t_.mul_(0.03)
tr.dp.master.copy_(torch.cat([tr.engine.params[o:o+n_] for o, n_ in tr.dp.shards]).float())
if tr.comm: tr.comm.barrier()
torch.cuda.synchronize()
p0 = tr.engine.params.clone()
stats = tr.fit(steps=3)
# reference: same weights, every rank's data, explicit average, plain Adam
ref = LLMEngine(LLMConfig(256, 2, 4, 512), tr.device, seed=0, training=True); ref.params.copy_(p0)
m = torch.zeros_like(p0, dtype=torch.float32); v = torch.zeros_like(m); master = p0.float()
from distributed_llm_backend_benchmark_b200.utils.data import create_dataset_from_config
sets = [create_dataset_from_config(cfg, rank_offset=q, with_target=True) for q in range(w)]
for step in (1, 2, 3):
    g = torch.zeros_like(m)
    for ds in sets:
        ref.forward_backward(ds.get_batch().to(tr.device), ds.get_target().to(tr.device)); g += ref.grads.float()
    g = g / w
    m.mul_(0.9).add_(g, alpha=0.1); v.mul_(0.999).addcmul_(g, g, value=0.001)
    master -= 1e-3 * (m / (1 - 0.9 ** step)) / (v.sqrt() / (1 - 0.999 ** step) ** 0.5 + 1e-8)
    ref.params.copy_(master.bfloat16())
mine = tr.dp.master
want = torch.cat([master[o:o + n_] for o, n_ in tr.dp.shards])
start = torch.cat([p0[o:o + n_] for o, n_ in tr.dp.shards]).float()
err = ((mine - want).norm() / (want - start).norm()).item()
same = torch.equal(tr.engine.params, ref.params) or ((tr.engine.params.float() - ref.params.float()).abs().max().item() < 0.02)
print("DP_OK" if (err < 0.05 and same) else "DP_BAD", r, err, stats["losses"])
""")
    r = run([BIN / "b200run", "-np", n, sys.executable, script], env={"MASTER_PORT": "29651"})
    assert r.returncode == 0 and r.stdout.count("DP_OK") == n, r.stdout[-3000:] + r.stderr[-3000:]


def test_tp_forward_fused_gemm_allreduce_matches_unsharded(native_lib, tmp_path):
    """Tensor-parallel forward (fused tcgen05 GEMM + NVLink all-reduce + residual in one kernel) against
    the same model on a single rank with the full weights."""
    n = nranks()
    script = tmp_path / "tp.py"
    script.write_text(f"""
import sys, os; sys.path[:0] = [{str(ROOT)!r}, {str(ROOT / 'compat')!r}]
import torch
from mpi4py import MPI
from distributed_llm_backend_benchmark_b200 import _lib
from distributed_llm_backend_benchmark_b200.models import LLM, LLMConfig, LLMEngine
comm = MPI.COMM_WORLD; r, w = comm.Get_rank(), comm.Get_size()
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
H, F, L = 1024, 2048, 3
full = LLMEngine(LLMConfig(H, L, 8, F), torch.device('cuda', torch.cuda.current_device()), seed=7)
for lv in full.layers:
    for t in (lv.wqkv, lv.wo, lv.wup, lv.wdown): t.mul_(0.02)
m = LLM(H, L, 8, F, r, w, comm)
for lf, ls in zip(full.layers, m.engine.layers):
    ls.wqkv.copy_(torch.cat([lf.wqkv[:, :H][:, r*H//w:(r+1)*H//w], lf.wqkv[:, H:][:, :3*H//w - H//w]], 1))
    ls.wo.copy_(lf.wo[r*H//w:(r+1)*H//w]); ls.wup.copy_(lf.wup[:, r*F//w:(r+1)*F//w]); ls.wdown.copy_(lf.wdown[r*F//w:(r+1)*F//w])
x = torch.randn(4, 200, H, generator=torch.Generator().manual_seed(1)).bfloat16().cuda()
for rep in range(3):
    y = m(x)
yf = full(x)
torch.cuda.synchronize()
err = ((y.float() - yf.float()).norm() / yf.float().norm()).item()
fused = w > 1 and m.engine._act['xa'].data_ptr() in m._reduce._by_ptr
print('TP_OK' if err < 3e-2 and (fused or w == 1) else 'TP_BAD', r, err, fused)
""")
    r = run([BIN / "mpirun", "-np", n, sys.executable, script])
    assert r.returncode == 0 and r.stdout.count("TP_OK") == n, r.stdout[-3000:] + r.stderr[-3000:]


def test_mixed_collective_stress_native(native_lib, tmp_path):
    n = nranks()
    inc = ",".join(str(i) for i in range(n))
    r = run([BIN / "deepspeed", "--master_port", 29661, "--include", f"localhost:{inc}", "stress.py"], cwd=ROOT / "collectives",
            env={"DLBB_STRESS_ROUNDS": "60", "DLBB_OUTPUT_DIR": str(tmp_path)})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads((tmp_path / f"stress_ccl_ranks{n}.json").read_text())
    assert res["errors"] == 0 and res["device"].startswith("cuda")
