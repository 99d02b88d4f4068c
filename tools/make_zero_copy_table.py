#!/usr/bin/env python3
"""Render profiles/raw/zero_copy_{N}gpu.json (written by tools/gpu_zero_copy.py) as markdown tables."""
import json, sys
from pathlib import Path

root = Path(__file__).resolve().parent.parent / "profiles" / "raw"


def human(n):
    return f"{n >> 30} GiB" if n >= 1 << 30 else f"{n >> 20} MiB"


for world in (2, 4, 8):
    f = root / f"zero_copy_{world}gpu.json"
    if not f.exists():
        continue
    d = json.load(open(f))
    print(f"### {world} x B200 ({len(d['checks'])} numerics checks passed: {all(c['ok'] for c in d['checks'].values())})\n")
    print("| op | message per rank | zero-copy us | staged us | NCCL us | zero-copy busbw GB/s | staged busbw | NCCL busbw | "
          "best ours / NCCL |")
    print("|---|---|---|---|---|---|---|---|---|")
    for r in d["perf"]:
        best = max(r["zero_copy_busbw"], r["staged_busbw"])
        print(f"| {r['op']} | {human(r['bytes'])} | {r['zero_copy_us']:.0f} | {r['staged_us']:.0f} | {r['nccl_us']:.0f} | "
              f"{r['zero_copy_busbw']:.0f} | {r['staged_busbw']:.0f} | {r['nccl_busbw']:.0f} | {best / r['nccl_busbw']:.2f} |")
    print()
