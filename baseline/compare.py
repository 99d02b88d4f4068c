#!/usr/bin/env python3
"""Side-by-side bus-bandwidth table (dlbb vs the same sweep loops on NCCL) from two result trees written by
baseline/run_sweeps.sh.  Two timings per cell: "call" = the reference-style bracket (barrier, then CUDA events around ONE
call on an idle GPU - it includes the host's time to enqueue the call, which is what makes NCCL look flat at ~200 us
for small messages) and "b2b" = the same call issued back to back between two events (device_tight_us), which is the
kernel-level comparison.  The ratio column uses b2b when both arms have it."""
import json
import sys
from pathlib import Path

root = Path(sys.argv[1])


def load(d):
    out = {}
    for f in sorted(Path(d).glob("*.json")):
        r = json.loads(f.read_text())
        key = (r["operation"], r.get("data_size_name") or "b{batch}_s{seq_len}_h{hidden_dim}".format(**r["tensor_shape"]))
        nbytes = r.get("tensor_size_bytes") or r["num_elements"] * 2
        out[key] = dict(bytes=nbytes, us=r["median_max_over_ranks_us"], bw=r["busbw_gbps"], ranks=r["num_ranks"],
                        tus=r.get("device_tight_us"), tbw=r.get("busbw_tight_gbps"))
    return out


def fmt(v, nd=1):
    return "-" if v is None else f"{v:.{nd}f}"


for dim in ("1d", "3d"):
    a, b = load(root / f"{dim}_dlbb"), load(root / f"{dim}_nccl")
    if not a:
        continue
    ranks = next(iter(a.values()))["ranks"]
    print(f"\n### {dim} sweep, {ranks} x B200: dlbb backend vs the same loops on NCCL (device-timed, slowest rank)\n")
    print("| op | message | bytes | dlbb call us | NCCL call us | dlbb b2b us | NCCL b2b us | dlbb busbw GB/s (b2b) | NCCL busbw GB/s (b2b) | ratio |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for key in sorted(a, key=lambda k: (k[0], a[k]["bytes"])):
        x, y = a[key], b.get(key)
        if y:
            xb, yb = (x["tbw"], y["tbw"]) if x["tbw"] and y["tbw"] else (x["bw"], y["bw"])
            print(f"| {key[0]} | {key[1]} | {x['bytes']} | {x['us']:.1f} | {y['us']:.1f} | {fmt(x['tus'])} | {fmt(y['tus'])} | "
                  f"{xb:.1f} | {yb:.1f} | {xb / yb if yb else 0:.2f} |")
        else:
            print(f"| {key[0]} | {key[1]} | {x['bytes']} | {x['us']:.1f} | - | {fmt(x['tus'])} | - | {fmt(x['tbw'] or x['bw'])} | - | - |")
s = load(root / "3d_dlbb_strided")
if s:
    print("\n### strided [b,s,h] views packed and unpacked inside the kernel (dlbb only)\n")
    print("| op | shape | bytes | call us | b2b us | busbw GB/s (b2b) |\n|---|---|---|---|---|---|")
    for key in sorted(s, key=lambda k: (k[0], s[k]["bytes"])):
        x = s[key]
        print(f"| {key[0]} | {key[1]} | {x['bytes']} | {x['us']:.1f} | {fmt(x['tus'])} | {fmt(x['tbw'] or x['bw'])} |")
