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
