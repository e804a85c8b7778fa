#!/usr/bin/env python
"""Turn an `ncu --set full` report into the small JSON kept under profiles/: one row per captured launch with duration, DRAM bytes,
DRAM / L2 / tensor-pipe / issue utilisation and registers.   usage: ncu_summarize.py gpurun_out/X.ncu-rep|X_raw.csv profiles/X_summary.json"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed.sum", "sm__cycles_elapsed.max"]
if sys.argv[1].endswith(".csv"):      # already exported with `ncu -i X.ncu-rep --page raw --csv`
    txt = open(sys.argv[1]).read()
else:
    txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
out = []
for r in data:
    e = {"kernel": r[idx["Kernel Name"]].split("(")[0], "grid": r[idx["Grid Size"]].replace(" ", "")}
    for k in KEEP:
        if k in idx:
            try:
                v = float(r[idx[k]].replace(",", ""))
            except ValueError:
                continue
            u = units[idx[k]]
            if k.startswith("dram__bytes"):      # normalise to MB
                v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
            if k == "gpu__time_duration.sum":    # normalise to us
                v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
            e[k] = round(v, 3)
    out.append(e)
json.dump(out, open(sys.argv[2], "w"), indent=0)
print(f"{len(out)} launches -> {sys.argv[2]}")
for e in out:
    print(f"{e['kernel'][-34:]:<34} {e.get('gpu__time_duration.sum', 0):9.1f} us  dram r/w {e.get('dram__bytes_read.sum', 0):8.1f}/{e.get('dram__bytes_write.sum', 0):8.1f} MB"
          f"  dram {e.get('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 0):5.1f}%  tensor {e.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0):5.1f}%"
          f"  issue {e.get('smsp__issue_active.avg.pct_of_peak_sustained_active', 0):5.1f}%")
