"""The reference's own entry points, byte for byte, on this framework's launcher and shims (Gloo, 2 ranks).

The north star says "the same entrypoints must work unchanged".  The files are copied out of /root/reference into a
temporary directory (the reference tree is read-only and the scripts write ``results/`` next to themselves) and run
through ``bin/mpirun`` / ``bin/deepspeed``; nothing in them is edited.  The result files the sweeps write are then fed
to the reference's own ``stats.py``.  Skipped when the reference tree is not present (e.g. on a GPU box).

ref: test/*.py, run_mpi.py, collectives/dsgloo.py, collectives/1d/{openmpi,dsccl,stats}.py.
"""
import json
import os
import shutil
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
REF = Path(os.environ.get("DLBB_REFERENCE_DIR", "/root/reference"))
BIN = ROOT / "bin"

pytestmark = [
    pytest.mark.skipif(not (REF / "run_mpi.py").exists(), reason="reference tree not present"),
    pytest.mark.skipif(not (BIN / "b200run").exists(), reason="launcher not built (python build.py)"),
]


def run(cmd, cwd, env=None, timeout=600):
    e = dict(os.environ)
    e["DLBB_FORCE_GLOO"] = "1"      # host-side plumbing; the GPU tier runs the repo's entry points natively
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    e.update(env or {})
    return subprocess.run([str(c) for c in cmd], cwd=cwd, env=e, capture_output=True, text=True, timeout=timeout)


def same_bytes(a: Path, b: Path) -> bool:
    return a.read_bytes() == b.read_bytes()


@pytest.mark.parametrize("script", ["test_open.py", "test_deepseed.py", "ccl.py", "test_ccl_backend.py", "ds_mpi_test.py"])
def test_reference_smoke_script_two_ranks(tmp_path, script):
    shutil.copy(REF / "test" / script, tmp_path / script)
    r = run([BIN / "mpirun", "-n", 2, sys.executable, script], cwd=tmp_path)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert same_bytes(REF / "test" / script, tmp_path / script)


def test_reference_single_process_deepspeed(tmp_path):
    shutil.copy(REF / "test" / "test.py", tmp_path / "test.py")
    r = run([sys.executable, "test.py"], cwd=tmp_path, env={"PYTHONPATH": f"{ROOT / 'compat'}:{ROOT}"})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_reference_ccl_backend_without_the_bindings_import(tmp_path):
    """test/test_deepseed.py:24 asks for backend='ccl' and never imports oneccl_bindings_for_pytorch: the launcher's
    start-up hook (compat/sitecustomize.py) has to make the name known in every rank."""
    src = (REF / "test" / "test_deepseed.py").read_text()
    assert "oneccl_bindings_for_pytorch" not in src and "backend='ccl'" in src
    p = tmp_path / "probe.py"
    p.write_text("import torch.distributed as dist\n"
                 "dist.init_process_group(backend='ccl')\n"
                 "import sys; sys.stdout.write(f'BACKEND={dist.get_backend()} rank={dist.get_rank()}\\n')\n"
                 "dist.destroy_process_group()\n")
    r = run([BIN / "mpirun", "-n", 2, sys.executable, p.name], cwd=tmp_path)
    assert r.returncode == 0 and r.stdout.count("BACKEND=ccl") == 2, r.stdout[-1500:] + r.stderr[-1500:]


def test_reference_run_mpi_tiny_yaml(tmp_path):
    for f in ("run_mpi.py", "models.py", "data_gen.py", "utils.py"):
        shutil.copy(REF / f, tmp_path / f)
    cfg = tmp_path / "tiny.yaml"
    cfg.write_text(f"""
experiment: {{name: tiny_ref, output_dir: {tmp_path}/results}}
model: {{size: tiny, hidden_size: 64, num_layers: 2, num_heads: 4, ffn_intermediate: 128}}
parallelism: {{world_size: 2, cores_per_rank: 1}}
input: {{batch_size: 2, sequence_length: 8, seed: 42}}
execution: {{warmup_iterations: 1, benchmark_iterations: 3}}
system: {{omp_num_threads: 1, mkl_num_threads: 1}}
""")
    r = run([BIN / "mpirun", "-np", 2, sys.executable, "run_mpi.py", "--config", cfg, "--backend", "openmpi"], cwd=tmp_path)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = json.loads((tmp_path / "results" / "openmpi_tiny_ref.json").read_text())
    assert out["rank_0_summary"]["world_size"] == 2 and len(out["raw_metrics_rank_0"]["forward_times"]) == 3


def test_reference_1d_sweeps_and_its_own_stats(tmp_path):
    d = tmp_path / "collectives" / "1d"
    d.mkdir(parents=True)
    for f in ("dsccl.py", "openmpi.py", "stats.py"):
        shutil.copy(REF / "collectives" / "1d" / f, d / f)
    shutil.copy(REF / "collectives" / "dsgloo.py", tmp_path / "collectives" / "dsgloo.py")
    # DeepSpeed + 'ccl' flavour (8 ops x 4 sizes; Gloo has no all-to-all, which the script skips by itself)
    r = run([BIN / "deepspeed", "--master_port", 29583, "--include", "localhost:0,1", "dsccl.py"], cwd=d)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    files = sorted((d / "results" / "dsccl").glob("deepspeed_oneccl_*_ranks2_*.json"))
    assert len(files) >= 28, [f.name for f in files]
    res = json.loads((d / "results" / "dsccl" / "deepspeed_oneccl_allreduce_ranks2_16MB.json").read_text())
    assert res["num_elements"] == 4194304 and len(res["timings"]) == 2 and len(res["timings"][0]) == 100
    # the reference's own post-processing, unmodified (INPUT_DIR = results/dsccl, collectives/1d/stats.py:13-17)
    r = run([sys.executable, "stats.py"], cwd=d)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert (d / "stats" / "dsccl" / "benchmark_statistics.csv").exists()
    st = json.loads((d / "stats" / "dsccl" / "deepspeed_oneccl_allreduce_ranks2_16MB_stats.json").read_text())
    assert st["mean_time_us"] > 0 and st["num_ranks"] == 2 and len(st["per_rank_means_us"]) == 2
    # mpi4py flavour (pickle path for the lowercase methods, buffer path for Bcast / Isend / Irecv)
    r = run([BIN / "mpirun", "-np", 2, sys.executable, "openmpi.py"], cwd=d)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert len(list((d / "results" / "openmpi").glob("openmpi_*_ranks2_*.json"))) == 32
    # the DeepSpeed + Gloo copy at collectives/dsgloo.py
    r = run([BIN / "deepspeed", "--master_port", 29584, "--include", "localhost:0,1", "dsgloo.py"], cwd=tmp_path / "collectives")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert len(list((tmp_path / "collectives" / "results" / "dsgloo").glob("deepspeed_gloo_*_ranks2_*.json"))) >= 24
