"""Index arithmetic of strided views (csrc/collectives/view_math.h - the code the collective kernels compile) on the host:
random 1-4-dimensional views against nested-loop walks, the vector-path conditions, and a gather + scatter written like
the kernels' loops against a plain copy; built with AddressSanitizer + UBSan (K12, SURVEY.md section 4: non-contiguous
3-D views)."""
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_view_math_matches_nested_loops(tmp_path):
    exe = tmp_path / "view_math_test"
    src = ROOT / "csrc" / "collectives" / "tests" / "view_math_test.cpp"
    b = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-Wall", "-Wno-unknown-pragmas", "-fsanitize=address,undefined",
                        "-fno-sanitize-recover=all", "-o", str(exe), str(src)], capture_output=True, text=True)
    if b.returncode != 0 and "sanitize" in b.stderr:
        b = subprocess.run(["g++", "-O1", "-std=c++17", "-Wno-unknown-pragmas", "-o", str(exe), str(src)], capture_output=True, text=True)
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([str(exe), "5000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "VIEW MATH OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def _build(tmp_path, name):
    exe = tmp_path / name
    src = ROOT / "csrc" / "collectives" / "tests" / f"{name}.cpp"
    b = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-Wall", "-Wno-unknown-pragmas", "-fsanitize=address,undefined",
                        "-fno-sanitize-recover=all", "-o", str(exe), str(src)], capture_output=True, text=True)
    if b.returncode != 0 and "sanitize" in b.stderr:
        b = subprocess.run(["g++", "-O1", "-std=c++17", "-Wno-unknown-pragmas", "-o", str(exe), str(src)], capture_output=True, text=True)
    assert b.returncode == 0, b.stderr[-2000:]
    return exe


def test_two_hop_partition_covers_every_vector_once(tmp_path):
    """direct_math.h (compiled into bcast_relay_kernel / reduce_pull_kernel): the schedules of both kernels replayed on index
    sets for worlds 2-8 (3, 5, 6, 7 never run on a box), messages smaller than the grid, every root: each vector arrives /
    is reduced exactly once, every waited flag has its signaller with the same epoch, relay numbering is a bijection."""
    exe = _build(tmp_path, "direct_math_test")
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DIRECT MATH OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
