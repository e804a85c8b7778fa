#!/bin/bash
# per-launch CUDA-event times at batch 1 (what batch-1 latency is made of)
rm -f gpurun_out/pl_b1_raw.csv
KDL_PROFILE_DUMP=gpurun_out/pl_b1_raw.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --batch 1 > gpurun_out/pl_b1.json 2>gpurun_out/pl_b1.err
python - <<'PY'
import collections
agg=collections.OrderedDict()
for line in open('gpurun_out/pl_b1_raw.csv'):
    f=line.rstrip('\n').split(';')
    a=agg.setdefault((f[0],f[1]),[0,0.0]); a[0]+=1; a[1]+=float(f[2])
steps=20; tot=0
for (k,l),(n,ms) in agg.items():
    print(f"{(l or k)[:34]:<34} n/step={n//steps:3d} us/launch={1e3*ms/n:7.1f}"); tot+=ms/steps
print("sum of kernel times per step: %.3f ms"%tot)
PY
