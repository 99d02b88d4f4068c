#!/usr/bin/env python3
"""Build the host emulation of the collectives (csrc/collectives/emu/): the kernel + launcher source files are compiled
with g++ -DDLBB_HOST_EMU after ONE textual change - `kernel<<<grid, block, smem, stream>>>(args)` becomes
`EMU_LAUNCH(grid, block, kernel)(args)` - so the code under test is the code nvcc compiles for sm_100a.

    python tools/emu_build.py --out /tmp/emu_test [--sanitize thread|address|none]
"""
import argparse
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
COLL = ROOT / "csrc" / "collectives"
# translation units of the emulation build (everything else is included from them)
UNITS = ["collectives.cu", "fused_adam.cu"]
INCLUDED = ["direct_kernels.cuh", "coll_device.cuh", "view_math.h", "direct_math.h", "gemm_allreduce_fusion.cuh"]

LAUNCH = re.compile(r"([A-Za-z_]\w*(?:\s*<[^<>;(){}]*>)?)\s*<<<(.*?)>>>", re.S)


def split_args(s: str) -> list[str]:
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "(<[{":
            depth += 1
        elif ch in ")>]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    out.append(cur.strip())
    return out


def rewrite(text: str) -> tuple[str, int]:
    n = 0

    def sub(m: re.Match) -> str:
        nonlocal n
        n += 1
        a = split_args(m.group(2))
        assert len(a) in (2, 3, 4), m.group(0)
        return f"EMU_LAUNCH({a[0]}, {a[1]}, {m.group(1)})"

    return LAUNCH.sub(sub, text).replace("__noinline__", "__attribute__((noinline))"), n


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--sanitize", default="none", choices=["none", "thread", "address"])
    ap.add_argument("--test", default=str(COLL / "tests" / "emu_collectives_test.cpp"))
    ap.add_argument("--opt", default="-O1")
    ap.add_argument("--define", action="append", default=[], help="extra -D (e.g. EMU_WEAK_FLAGS: negative control)")
    a = ap.parse_args()
    out = Path(a.out)
    gen = out.parent / (out.name + "_src")
    (gen / "csrc" / "collectives" / "emu").mkdir(parents=True, exist_ok=True)
    total = 0
    # mirror the directory layout so that the sources' relative includes keep working
    for name in UNITS + INCLUDED:
        text, n = rewrite((COLL / name).read_text())
        total += n
        dst = gen / "csrc" / "collectives" / (name.replace(".cu", ".cpp") if name in UNITS else name)
        dst.write_text(text)
    for f in (COLL / "emu").glob("*.h"):
        (gen / "csrc" / "collectives" / "emu" / f.name).write_text(f.read_text())
    for sub in ("csrc/bootstrap", "csrc/kernels", "include"):
        d = gen / sub
        if not d.exists():
            d.symlink_to(ROOT / sub)
    assert total >= 20, f"only {total} kernel launches rewritten"
    flags = ["g++", "-std=c++17", a.opt, "-g", "-pthread", "-DDLBB_HOST_EMU", "-Wno-unknown-pragmas", "-Wno-unused-function",
             "-I/usr/local/cuda/include", f"-I{gen / 'csrc' / 'collectives'}", "-include", str(gen / "csrc" / "collectives" / "emu" / "cuda_emu.h")]
    flags += [f"-D{d}" for d in a.define]
    if a.sanitize != "none":
        flags += [f"-fsanitize={a.sanitize}", "-fno-omit-frame-pointer"]
    srcs = [str(gen / "csrc" / "collectives" / u.replace(".cu", ".cpp")) for u in UNITS] + [a.test]
    r = subprocess.run(flags + ["-o", str(out)] + srcs, capture_output=True, text=True)
    Path(str(out) + ".log").write_text(r.stderr)
    if r.returncode != 0:
        errs = [l for l in r.stderr.splitlines() if " error" in l]
        sys.stderr.write("\n".join(errs[:40]) + f"\n(full compiler output: {out}.log)\n")
        return r.returncode
    print(f"built {out} ({total} launches rewritten, sanitize={a.sanitize})")
    return 0


if __name__ == "__main__":
    sys.exit(main())
