"""Host-side plumbing of every entry point on a machine without a GPU: the launcher's personalities,
the compat shims (mpi4py / deepspeed / oneccl_bindings_for_pytorch) over Gloo with 2 ranks, the six
smoke scripts, the TP forward driver and the sweep + stats scripts (BASELINE.json config 1)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
BIN = ROOT / "bin"
LAUNCHER = BIN / "b200run"

pytestmark = pytest.mark.skipif(not LAUNCHER.exists(), reason="launcher not built (python build.py)")


def run(cmd, cwd=ROOT, env=None, timeout=300):
    e = dict(os.environ)
    e["DLBB_FORCE_GLOO"] = "1"            # these tests exercise the host-side plumbing only
    e["DLBB_DEVICE"] = "cpu"
    e["DLBB_SWEEP_DEVICE"] = "cpu"
    e.pop("RANK", None), e.pop("WORLD_SIZE", None)
    e.update(env or {})
    return subprocess.run([str(c) for c in cmd], cwd=cwd, env=e, capture_output=True, text=True, timeout=timeout)


def test_launcher_mpirun_flags_and_failure_propagation(tmp_path):
    script = tmp_path / "p.py"
    script.write_text("import os,sys\nprint('R', os.environ['RANK'], os.environ['WORLD_SIZE'], os.environ.get('FOO'),"
                      " os.environ['OMPI_COMM_WORLD_RANK'], os.environ['PMI_SIZE'], len(os.sched_getaffinity(0)))\n"
                      "sys.exit(7 if os.environ['RANK']=='1' and len(sys.argv)>1 else 0)\n")
    r = run([BIN / "mpirun", "-np", 2, "--report-bindings", "--map-by", "slot:PE=1", "-x", "FOO=bar", sys.executable, script])
    assert r.returncode == 0, r.stderr
    lines = sorted(l for l in r.stdout.splitlines() if l.startswith("R "))
    assert lines == ["R 0 2 bar 0 2 1", "R 1 2 bar 1 2 1"]
    assert "rank 0 -> gpu" in r.stderr
    r = run([BIN / "mpirun", "-n", 2, "-ppn", 2, "-genv", "FOO=1", sys.executable, script, "fail"])
    assert r.returncode == 7


def test_launcher_under_address_sanitizer(tmp_path):
    """SURVEY 5.2 for the launcher: the same binary built with ASan + UBSan spawns, binds, forwards the environment
    and propagates a failing rank's exit code without a sanitizer report."""
    exe = tmp_path / "b200run_asan"
    srcs = sorted((ROOT / "csrc" / "launcher").glob("*.cpp"))
    b = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
                        "-o", str(exe), *map(str, srcs), "-lpthread", "-ldl"], capture_output=True, text=True)
    if b.returncode != 0:
        pytest.skip(f"sanitizer runtime not available: {b.stderr[-300:]}")
    script = tmp_path / "p.py"
    script.write_text("import os,sys\nprint('R', os.environ['RANK'])\nsys.exit(5 if os.environ['RANK']=='1' and len(sys.argv)>1 else 0)\n")
    env = {"ASAN_OPTIONS": "detect_leaks=0:halt_on_error=1"}
    r = run([exe, "--as", "mpirun", "-np", 2, "--map-by", "slot:PE=1", "-x", "FOO=bar", sys.executable, script], env=env)
    assert r.returncode == 0 and "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-3000:]
    assert sorted(l for l in r.stdout.splitlines() if l.startswith("R ")) == ["R 0", "R 1"]
    r = run([exe, "--as", "mpirun", "-np", 2, sys.executable, script, "fail"], env=env)
    assert r.returncode == 5 and "ERROR: AddressSanitizer" not in r.stderr, r.stderr[-3000:]


def test_launcher_deepspeed_flags(tmp_path):
    script = tmp_path / "d.py"
    script.write_text("import os\nprint('D', os.environ['RANK'], os.environ['MASTER_PORT'], os.environ['CUDA_VISIBLE_DEVICES'],"
                      " sorted(os.sched_getaffinity(0)))\n")
    r = run([BIN / "deepspeed", "--master_port", 29577, "--include", "localhost:0,1", "--bind_cores_to_rank",
             "--bind_core_list", "0,1", script.name], cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert sorted(l for l in r.stdout.splitlines() if l.startswith("D ")) == ["D 0 29577 0,1 [0]", "D 1 29577 0,1 [1]"]


@pytest.mark.parametrize("script", ["test_open.py", "test_deepseed.py", "ccl.py", "test_ccl_backend.py", "ds_mpi_test.py"])
def test_smoke_scripts_two_ranks(script):
    r = run([BIN / "mpirun", "-n", 2, sys.executable, ROOT / "test" / script])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_smoke_script_single_process():
    r = run([sys.executable, ROOT / "test" / "test.py"])
    assert r.returncode == 0 and "SUCCESS" in r.stdout, r.stderr[-2000:]


def test_run_mpi_tp2_writes_reference_schema(tmp_path):
    cfg = tmp_path / "tiny.yaml"
    cfg.write_text(f"""
experiment: {{name: tiny_world2, output_dir: {tmp_path}/results}}
model: {{size: tiny, hidden_size: 64, num_layers: 2, num_heads: 4, ffn_intermediate: 128}}
parallelism: {{world_size: 2, cores_per_rank: 1}}
input: {{batch_size: 2, sequence_length: 8, seed: 42}}
execution: {{warmup_iterations: 1, benchmark_iterations: 3}}
system: {{omp_num_threads: 1, mkl_num_threads: 1}}
""")
    r = run([BIN / "mpirun", "-np", 2, sys.executable, ROOT / "run_mpi.py", "--config", cfg, "--backend", "openmpi"])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = json.loads((tmp_path / "results" / "openmpi_tiny_world2.json").read_text())
    assert set(out) == {"experiment", "backend", "config", "system_info", "rank_0_summary", "rank_statistics",
                        "raw_metrics_rank_0"}
    for k in ("forward_mean", "forward_std", "forward_min", "forward_max", "forward_median", "forward_p95",
              "forward_p99", "init_time", "num_iterations", "rank", "world_size"):
        assert k in out["rank_0_summary"]
    assert len(out["rank_statistics"]["forward_mean_per_rank"]) == 2
    assert len(out["raw_metrics_rank_0"]["forward_times"]) == 3
    # world-size mismatch is a hard error, as in the reference
    r = run([BIN / "mpirun", "-np", 1, sys.executable, ROOT / "run_mpi.py", "--config", cfg, "--backend", "openmpi"])
    assert r.returncode == 1 and "World size mismatch" in r.stdout


def test_tp2_forward_equals_single_rank(tmp_path):
    """Column/row sharding + all-reduce reproduces the unsharded model (same full weights)."""
    script = tmp_path / "tp.py"
    script.write_text(f"""
import sys; sys.path[:0] = [{str(ROOT)!r}, {str(ROOT / 'compat')!r}]
import torch
from mpi4py import MPI
from distributed_llm_backend_benchmark_b200.models import LLM, LLMConfig, LLMEngine
comm = MPI.COMM_WORLD; r, w = comm.Get_rank(), comm.Get_size()
H, F, L = 64, 128, 2
full = LLMEngine(LLMConfig(H, L, 4, F), 'cpu', seed=7)
for lv in full.layers:
    for t in (lv.wqkv, lv.wo, lv.wup, lv.wdown): t.mul_(0.05)
m = LLM(H, L, 4, F, r, w, comm)
for lf, ls in zip(full.layers, m.engine.layers):
    # the reference's fake attention consumes the first H/ws columns of each rank's QKV shard
    ls.wqkv.copy_(torch.cat([lf.wqkv[:, :H][:, r*H//w:(r+1)*H//w], lf.wqkv[:, H:][:, :3*H//w - H//w]], 1))
    ls.wo.copy_(lf.wo[r*H//w:(r+1)*H//w]); ls.wup.copy_(lf.wup[:, r*F//w:(r+1)*F//w]); ls.wdown.copy_(lf.wdown[r*F//w:(r+1)*F//w])
x = torch.randn(2, 8, H, generator=torch.Generator().manual_seed(1)).bfloat16()
y, yf = m(x), full(x)
err = ((y.float() - yf.float()).norm() / yf.float().norm()).item()
assert err < 3e-2, err
print('TP_OK', r, err)
""")
    r = run([BIN / "mpirun", "-np", 2, sys.executable, script])
    assert r.returncode == 0 and r.stdout.count("TP_OK") == 2, r.stdout[-1500:] + r.stderr[-1500:]


def test_sweeps_and_stats(tmp_path):
    env = {"DLBB_SWEEP_ITERS": "3", "DLBB_SWEEP_WARMUP": "1", "DLBB_SWEEP_SIZES": "1KB",
           "DLBB_SWEEP_OPS": "allreduce,broadcast,sendrecv"}
    r = run([BIN / "mpirun", "-np", 2, sys.executable, "openmpi.py"], cwd=ROOT / "collectives" / "1d",
            env={**env, "DLBB_OUTPUT_DIR": str(tmp_path / "r1")})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    res = json.loads((tmp_path / "r1" / "openmpi_allreduce_ranks2_1KB.json").read_text())
    assert res["mpi_implementation"] == "openmpi" and res["num_elements"] == 256
    assert res["dtype"] == "<class 'numpy.float16'>" and len(res["timings"]) == 2 and len(res["timings"][0]) == 3
    r = run([BIN / "deepspeed", "--master_port", 29571, "--include", "localhost:0,1", "dsgloo.py"],
            cwd=ROOT / "collectives", env={**env, "DLBB_OUTPUT_DIR": str(tmp_path / "r2"), "DLBB_SWEEP_OPS": "allreduce,gather"})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    res = json.loads((tmp_path / "r2" / "deepspeed_gloo_gather_ranks2_1KB.json").read_text())
    assert res["implementation"] == "deepspeed_gloo" and res["backend"] == "gloo" and res["dtype"] == "float16"
    assert res["timing_method"] == "time.perf_counter()"
    r = run([sys.executable, "stats.py", tmp_path / "r2", tmp_path / "s2"], cwd=ROOT / "collectives" / "1d")
    assert r.returncode == 0 and (tmp_path / "s2" / "benchmark_statistics.csv").exists()
    env3 = {"DLBB_SWEEP_ITERS": "2", "DLBB_SWEEP_WARMUP": "1", "DLBB_SWEEP_BATCH": "1", "DLBB_SWEEP_SEQ": "1,16",
            "DLBB_SWEEP_HIDDEN": "64", "DLBB_SWEEP_OPS": "allreduce,allgather", "DLBB_OUTPUT_DIR": str(tmp_path / "r3")}
    r = run([BIN / "deepspeed", "--num_gpus", 2, "--master_port", 29572, "dsgloo.py"], cwd=ROOT / "collectives" / "3d", env=env3)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    res = json.loads((tmp_path / "r3" / "deepspeed_gloo_allreduce_ranks2_b1_s16_h64.json").read_text())
    assert res["tensor_shape"] == {"batch": 1, "seq_len": 16, "hidden_dim": 64} and res["tensor_size_bytes"] == 2048
    r = run([sys.executable, "stats.py", "dsgloo", tmp_path / "r3", tmp_path / "s3"], cwd=ROOT / "collectives" / "3d")
    assert r.returncode == 0 and (tmp_path / "s3" / "benchmark_statistics_3d_dsgloo_transpose.csv").exists()


def test_mixed_collective_stress_two_ranks(tmp_path):
    r = run([BIN / "deepspeed", "--master_port", 29573, "--include", "localhost:0,1", "stress.py"], cwd=ROOT / "collectives",
            env={"DLBB_STRESS_ROUNDS": "4", "DLBB_OUTPUT_DIR": str(tmp_path)})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    res = json.loads((tmp_path / "stress_ccl_ranks2.json").read_text())
    assert res["errors"] == 0 and res["num_ranks"] == 2 and len(res["alltoall"]) >= 4
