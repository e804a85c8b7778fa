#!/bin/bash
# per-launch CUDA-event times of one bench step set -> gpurun_out/per_launch.csv (aggregated by label)
rm -f gpurun_out/per_launch_raw.csv
KDL_PROFILE_DUMP=gpurun_out/per_launch_raw.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/pl_bench.json 2>gpurun_out/pl_bench.err
python - <<'PY'
import csv, collections
agg=collections.OrderedDict()
for r in csv.reader(open('gpurun_out/per_launch_raw.csv'), delimiter=';'):
    k=(r[0],r[1]); a=agg.setdefault(k,[0,0.0,0.0,0.0]); a[0]+=1; a[1]+=float(r[2]); a[2]+=float(r[3]); a[3]+=float(r[4])
steps=5
print(f"{'kernel':<32} {'label':<28} {'n/step':>6} {'us/launch':>10} {'TF/s':>8} {'GB/s':>8}")
for (k,l),(n,ms,fl,by) in agg.items():
    print(f"{k:<32} {l:<28} {n//steps:>6} {1e3*ms/n:>10.1f} {fl/ms/1e9:>8.1f} {by/ms/1e6:>8.1f}")
PY
