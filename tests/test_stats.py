"""Statistics post-processing against the reference's own published outputs (golden files copied
from collectives/{1d,3d}/{results,stats}/dsccl)."""
import csv
import json
from pathlib import Path

import pytest

from distributed_llm_backend_benchmark_b200.utils import stats

GOLD = Path(__file__).parent / "golden"


def _inputs(sub, tmp_path):
    src = tmp_path / "in"
    src.mkdir()
    for f in (GOLD / sub).glob("*.json"):
        if not f.name.endswith("_stats.json"):
            (src / f.name).write_text(f.read_text())
    return src


def test_1d_matches_published_stats(tmp_path):
    out = tmp_path / "out"
    res = stats.process_1d(_inputs("1d", tmp_path), out, verbose=False)
    assert len(res) == 2
    for gold in (GOLD / "1d").glob("*_stats.json"):
        want = json.loads(gold.read_text())
        got = json.loads((out / gold.name).read_text())
        for k, v in want.items():
            if isinstance(v, float):
                assert got[k] == pytest.approx(v, rel=1e-12), k
            elif isinstance(v, list):
                assert got[k] == pytest.approx(v, rel=1e-12), k
            else:
                assert got[k] == v, k
    rows = list(csv.DictReader(open(out / "benchmark_statistics.csv")))
    assert list(rows[0].keys()) == stats.CSV_1D
    assert {r["operation"] for r in rows} == {"allreduce", "broadcast"}


def test_3d_matches_published_stats(tmp_path):
    out = tmp_path / "out"
    res = stats.process_3d(_inputs("3d", tmp_path), out, "dsccl", verbose=False)
    assert len(res) == 2
    for gold in (GOLD / "3d").glob("*_stats.json"):
        want = json.loads(gold.read_text())
        got = json.loads((out / gold.name).read_text())
        for k, v in want.items():
            assert got[k] == (pytest.approx(v, rel=1e-12) if isinstance(v, float) else v), k
    std = list(csv.reader(open(out / "benchmark_statistics_3d_dsccl_standard.csv")))
    assert std[0] == stats.CSV_3D
    assert [r[1] for r in std[1:]] == ["allgather", "allreduce"]          # sorted by operation first
    tr = list(csv.reader(open(out / "benchmark_statistics_3d_dsccl_transpose.csv")))
    assert tr[0][0] == "Metric" and tr[0][1] == "allgather_r16_h2048_s1_b16"
    assert [r[0] for r in tr[1:5]] == ["mean_time_ms", "median_time_ms", "min_time_ms", "max_time_ms"]
    assert tr[5] == [] and tr[6] == ["--- Metadata ---"]
    assert [r[0] for r in tr[7:]] == ["operation", "num_ranks", "hidden_dim", "seq_len", "batch", "tensor_size_mb"]


def test_bus_bandwidth_conventions():
    alg, bus = stats.bus_bandwidth("allreduce", 8, 1e9, 1e-3)
    assert alg == pytest.approx(1000.0) and bus == pytest.approx(1750.0)
    alg, bus = stats.bus_bandwidth("allgather", 4, 1e9, 1e-3)       # per-rank contribution 1 GB -> 4 GB gathered
    assert alg == pytest.approx(4000.0) and bus == pytest.approx(3000.0)
    assert stats.bus_bandwidth("broadcast", 8, 1e9, 1e-3) == (pytest.approx(1000.0), pytest.approx(1000.0))
    assert stats.reference_bandwidth_1d(4194304, 1.0, "allreduce", 8) == pytest.approx(4194304 * 2 * 8 / 2 ** 30)
    assert stats.reference_bandwidth_1d(1, 1.0, "nonsense", 8) is None
