"""The parts of the bench.py / __graft_entry__ contract that can be checked without a GPU."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_reference_arm_prints_one_json_line_and_exits_zero():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and d["unavailable"]


def test_graft_entry_exposes_build_and_smoke():
    sys.path.insert(0, str(ROOT))
    import __graft_entry__ as g
    assert callable(g.build) and callable(g.smoke)


def test_published_sweep_results_have_statistics():
    """collectives/{1d,3d}/results/<impl>/*.json (reference C17) each have their stats.py output (C18)."""
    n = 0
    for dim in ("1d", "3d"):
        for impl_dir in sorted((ROOT / "collectives" / dim / "results").iterdir()):
            stats_dir = ROOT / "collectives" / dim / "stats" / impl_dir.name
            for f in impl_dir.glob("*.json"):
                d = json.loads(f.read_text())
                assert "timings" in d and len(d["timings"]) == d["num_ranks"], f
                assert (stats_dir / (f.stem + "_stats.json")).exists(), f
                n += 1
            assert any(stats_dir.glob("*.csv")), stats_dir
    assert n > 100


def test_published_sweeps_cover_2_4_8_ranks_for_both_arms():
    """BASELINE.json configs 2 and 3: the 1-D sweep (all-reduce / all-gather / broadcast / reduce-scatter, 1 KB - 1 GB) and the
    3-D sweep incl. strided views at 2 / 4 / 8 B200, native backend and the NCCL arm, plus the tuning study at 8 ranks."""
    d1 = ROOT / "collectives" / "1d" / "results"
    for ranks in (2, 4, 8):
        for arm, impl in (("dsccl", "deepspeed_oneccl"), ("nccl", "deepspeed_nccl")):
            for op in ("allreduce", "allgather", "broadcast", "reducescatter", "alltoall", "reduce"):
                for size in ("1KB", "4MB", "128MB", "1GB"):
                    f = d1 / arm / f"{impl}_{op}_ranks{ranks}_{size}.json"
                    assert f.exists(), f
                    d = json.loads(f.read_text())
                    assert d["num_ranks"] == ranks and d["busbw_gbps"] > 0 and d.get("device_tight_us", 0) > 0, f
        d3 = ROOT / "collectives" / "3d" / "results"
        for op in ("allreduce", "allgather", "broadcast", "reduce", "gather"):
            assert (d3 / "dsccl" / f"deepspeed_oneccl_{op}_ranks{ranks}_b8_s2048_h4096.json").exists(), (op, ranks)
            assert (d3 / "nccl" / f"deepspeed_nccl_{op}_ranks{ranks}_b8_s2048_h4096.json").exists(), (op, ranks)
            s = json.loads((d3 / "dsccl_strided" / f"deepspeed_oneccl_{op}_ranks{ranks}_b8_s2048_h4096.json").read_text())
            assert s["strided_view"] is True, (op, ranks)
    for study in ("dlbb_directnvls_allreduce", "dlbb_direct_allreduce", "dlbb_nvls_allreduce", "dlbb_twoshot_allreduce", "dlbb_ctas24"):
        assert (ROOT / "collectives" / "3d" / "results" / study / "deepspeed_oneccl_allreduce_ranks8_b16_s4096_h4096.json").exists(), study
