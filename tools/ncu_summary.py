#!/usr/bin/env python3
"""Markdown table of an `ncu --set full` report: one line per captured kernel launch with the numbers the profiling
recipe asks for (duration, DRAM bytes, DRAM / tensor-pipe / SM utilisation, achieved occupancy, registers, grid).

    python tools/ncu_summary.py gpurun_out/prof_nongemm.ncu-rep [> profiles/ncu_nongemm.md]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {n: i for i, n in enumerate(hdr)}
want = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "DRAM read"),
        ("dram__bytes_write.sum", "DRAM write"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("lts__t_bytes.sum", "L2 bytes")]
have = [(k, t) for k, t in want if k in col]
print("| " + " | ".join(f"{t} [{units[col[k]]}]" if units[col[k]] else t for k, t in have) + " |")
print("|" + "---|" * len(have))
for r in rows[2:]:
    cells = []
    for k, _ in have:
        v = r[col[k]]
        if k == "Kernel Name":
            v = v.split("(")[0].replace("void ", "").replace("dlbb::", "")[:70]
        else:
            try:
                v = f"{float(v.replace(',', '')):.1f}"
            except ValueError:
                pass
        cells.append(v)
    print("| " + " | ".join(cells) + " |")
