#!/usr/bin/env python3
"""Copy the sweep results of tools/gpu_publish.sh (gpurun_out/publish_npN) into the published tree
(collectives/{1d,3d}/results/...), regenerate the statistics with the repo's stats.py (the reference's post-processing,
collectives/{1d,3d}/stats.py) and write the comparison table to profiles/sweeps_npN_vs_nccl.md.

    python tools/install_published.py gpurun_out/publish_np8
"""
import json, re, shutil, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
src = Path(sys.argv[1]).resolve()
np_ = re.search(r"np(\d+)", src.name).group(1)
MAP = {"1d_dlbb": ("1d", "dsccl"), "1d_nccl": ("1d", "nccl"), "3d_dlbb": ("3d", "dsccl"), "3d_nccl": ("3d", "nccl"),
       "3d_dlbb_strided": ("3d", "dsccl_strided")}
touched = set()


def slim(path: Path, dst: Path):
    """Result files keep the reference's schema; the per-iteration device timings are kept, rounded to 0.1 us."""
    r = json.loads(path.read_text())
    for k in ("timings", "timings_device"):
        if r.get(k):
            r[k] = [[round(v, 7) for v in row] for row in r[k]]
    dst.write_text(json.dumps(r, indent=1))


for sub, (dim, name) in MAP.items():
    d = src / "sweeps" / sub
    if not d.is_dir():
        continue
    out = ROOT / "collectives" / dim / "results" / name
    out.mkdir(parents=True, exist_ok=True)
    for f in sorted(d.glob("*.json")):
        slim(f, out / f.name)
    touched.add((dim, name))
tun = src / "tuning"
if tun.is_dir():
    for d in sorted(tun.iterdir()):
        out = ROOT / "collectives" / "3d" / "results" / d.name
        out.mkdir(parents=True, exist_ok=True)
        for f in sorted(d.glob("*.json")):
            slim(f, out / f.name)
        touched.add(("3d", d.name))
for dim, name in sorted(touched):
    cwd = ROOT / "collectives" / dim
    cmd = [sys.executable, "stats.py", f"results/{name}", f"stats/{name}"] if dim == "1d" else \
          [sys.executable, "stats.py", name, f"results/{name}", f"stats/{name}"]
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True)
    print(dim, name, "stats rc", r.returncode, (r.stderr or "")[-200:])
cmp_md = src / "sweeps" / "compare.md"
if cmp_md.exists():
    head = (f"# Collective sweeps on {np_} x B200 through the entry points: native backend vs the same loops on NCCL\n\n"
            f"`tools/gpu_publish.sh {np_}` = `baseline/run_sweeps.sh {np_}`: `collectives/1d/dsccl.py` (extended sizes 1 KB - 1 GB, plus "
            "reduce-scatter and all-to-all) and `collectives/3d/dsccl.py` under `bin/deepspeed`, once on backend `ccl` (ours; the "
            "communicated tensors live in symmetric memory, `DLBB_SWEEP_SYMMETRIC=1`) and once with `DLBB_SWEEP_BACKEND=nccl`.  "
            "\"call\" = the reference's loop (barrier, then events around one call: includes the host's enqueue time on an idle GPU - "
            "NCCL's flat ~230 us); \"b2b\" = the same call back to back between two events.  Raw files: "
            "`collectives/{1d,3d}/results/{dsccl,nccl,dsccl_strided}/*ranks" + np_ + "*`.\n")
    (ROOT / "profiles" / f"sweeps_np{np_}_vs_nccl.md").write_text(head + cmp_md.read_text())
    print("wrote", f"profiles/sweeps_np{np_}_vs_nccl.md")
