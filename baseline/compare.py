#!/usr/bin/env python3
"""Side-by-side busbw table (dlbb vs reference-loops-on-NCCL) from two sweep result trees."""
import json, sys
from pathlib import Path

root = Path(sys.argv[1])


def load(d):
    out = {}
    for f in sorted(Path(d).glob("*.json")):
        r = json.loads(f.read_text())
        key = (r["operation"], r.get("data_size_name") or "b{batch}_s{seq_len}_h{hidden_dim}".format(**r["tensor_shape"]))
        nbytes = r.get("tensor_size_bytes") or r["num_elements"] * 2
        out[key] = (nbytes, r["median_max_over_ranks_us"], r["busbw_gbps"], r["num_ranks"])
    return out


for dim in ("1d", "3d"):
    a, b = load(root / f"{dim}_dlbb"), load(root / f"{dim}_nccl")
    if not a:
        continue
    ranks = next(iter(a.values()))[3]
    print(f"\n### {dim} sweep, {ranks} x B200: dlbb backend vs the same loops on NCCL (device-timed median of the slowest rank)\n")
    print("| op | message | bytes | dlbb us | NCCL us | dlbb busbw GB/s | NCCL busbw GB/s | ratio |")
    print("|---|---|---|---|---|---|---|---|")
    for key in sorted(a, key=lambda k: (k[0], a[k][0])):
        nb, us, bw, _ = a[key]
        if key in b:
            _, us2, bw2, _ = b[key]
            print(f"| {key[0]} | {key[1]} | {nb} | {us:.1f} | {us2:.1f} | {bw:.1f} | {bw2:.1f} | {bw / bw2 if bw2 else 0:.2f} |")
        else:
            print(f"| {key[0]} | {key[1]} | {nb} | {us:.1f} | - | {bw:.1f} | - | - |")
s = load(root / "3d_dlbb_strided")
if s:
    print("\n### strided [b,s,h] views packed inside the kernel (dlbb only)\n")
    print("| op | shape | bytes | us | busbw GB/s |\n|---|---|---|---|---|")
    for key in sorted(s, key=lambda k: (k[0], s[k][0])):
        print(f"| {key[0]} | {key[1]} | {s[key][0]} | {s[key][1]:.1f} | {s[key][2]:.1f} |")
