#!/usr/bin/env python3
"""profiles/tuning_study.md: the counterpart of the reference's oneCCL tuning study (collectives/3d/launch_dsccl.sh:37-59,
results dsccl_{ring,direct,...}_allreduce, dsccl_worker{1..4}, dsccl_fusion{,0}) from collectives/3d/results/dlbb_*:
all-reduce algorithm x CTA count x fusion on the reference's shapes, plus the bucket-fusion study of small tensors."""
import json, re
from pathlib import Path
root = Path(__file__).resolve().parent.parent
res = root / "collectives" / "3d" / "results"
rows = {}
for d in sorted(res.glob("dlbb_*")):
    for f in sorted(d.glob("*.json")):
        r = json.loads(f.read_text())
        sh = r["tensor_shape"]
        key = (r["num_ranks"], f"({sh['batch']},{sh['seq_len']},{sh['hidden_dim']})", r["tensor_size_bytes"])
        rows.setdefault(key, {})[d.name] = (r.get("device_tight_us") or r["median_max_over_ranks_us"], r.get("busbw_tight_gbps") or r["busbw_gbps"])
names = sorted({n for v in rows.values() for n in v})
LABEL = {"dlbb_oneshot_allreduce": "one-shot (staged)", "dlbb_twoshot_allreduce": "two-shot peer (staged)", "dlbb_nvls_allreduce": "two-shot NVLS (staged, pipelined)",
         "dlbb_direct_allreduce": "in place, peer loads/stores", "dlbb_directnvls_allreduce": "in place, multimem (default)",
         "dlbb_fusion": "DLBB_FUSION=1", "dlbb_fusion0": "DLBB_FUSION=0"}
out = ["# Tuning study: all-reduce algorithm x CTA count x fusion (3-D sweep entry point, bf16, symmetric tensors)\n",
       "Counterpart of the reference's oneCCL study (`CCL_ALLREDUCE` x `CCL_WORKER_COUNT` x `CCL_FUSION`, 15 result directories at 8 ranks). "
       "`collectives/3d/dsccl.py` under `bin/deepspeed` with `DLBB_ALGO` / `DLBB_CTAS` / `DLBB_FUSION` set (`tools/gpu_publish.sh`, `tools/gpu_final8.sh`); "
       "raw files `collectives/3d/results/dlbb_*`, statistics `collectives/3d/stats/dlbb_*`.  Cells: back-to-back device time in us (bus bandwidth GB/s), max over ranks.  "
       "A single large tensor cannot be fused with anything, so the two fusion columns agree by construction - the fusion knob is studied on small tensors below.\n"]
for ranks in sorted({k[0] for k in rows}):
    out.append(f"\n### {ranks} ranks\n")
    order = ["dlbb_directnvls_allreduce", "dlbb_direct_allreduce", "dlbb_nvls_allreduce", "dlbb_twoshot_allreduce", "dlbb_oneshot_allreduce",
             "dlbb_ctas24", "dlbb_ctas48", "dlbb_ctas96", "dlbb_ctas128", "dlbb_fusion", "dlbb_fusion0"]
    cols = [n for n in sorted(names, key=lambda n: order.index(n) if n in order else 99) if any(n in v for k, v in rows.items() if k[0] == ranks)]
    out.append("| shape | bytes | " + " | ".join(LABEL.get(n, n.replace("dlbb_", "DLBB_").replace("ctas", "CTAS=")) for n in cols) + " |")
    out.append("|---|---|" + "---|" * len(cols))
    for k in sorted((k for k in rows if k[0] == ranks), key=lambda k: k[2]):
        cells = [f"{rows[k][n][0]:.0f} ({rows[k][n][1]:.0f})" if n in rows[k] else "-" for n in cols]
        out.append(f"| {k[1]} | {k[2]} | " + " | ".join(cells) + " |")
fs = sorted((root / "profiles" / "raw").glob("fusion_study_np*.jsonl"))
for f in fs:
    recs = [json.loads(l) for l in f.read_text().splitlines() if l.strip()]
    if not recs:
        continue
    out.append(f"\n### Bucket fusion of small tensors, {recs[0]['ranks']} ranks (`tools/gpu_fusion_study.py`: one `all_reduce_coalesced` call of K tensors)\n")
    out.append("| tensors x bytes | fused: us per call (launches) | unfused: us per call (launches) | speed-up |")
    out.append("|---|---|---|---|")
    by = {}
    for r in recs:
        by.setdefault((r["tensors"], r["bytes_each"]), {})[r["fusion"]] = r
    for (k, b), v in sorted(by.items()):
        if "1" in v and "0" in v:
            out.append(f"| {k} x {b} | {v['1']['us_per_call']:.1f} ({v['1']['launches_per_call']:.0f}) | {v['0']['us_per_call']:.1f} ({v['0']['launches_per_call']:.0f}) | "
                       f"{v['0']['us_per_call'] / v['1']['us_per_call']:.1f}x |")
(root / "profiles" / "tuning_study.md").write_text("\n".join(out) + "\n")
print("wrote profiles/tuning_study.md")
