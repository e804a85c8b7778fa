#!/bin/bash
for mb in 0 16 32 48 64; do
  KDL_MICROBATCH_MB=$mb python bench.py --steps 10 --warmup 4 --no-cpu-baseline > gpurun_out/mb_$mb.json 2>gpurun_out/mb_$mb.err || tail -3 gpurun_out/mb_$mb.err
  python - <<PY
import json
d=json.load(open("gpurun_out/mb_$mb.json"))
k={x["kernel"]:x["ms_per_step"] for x in d["kernels"]}
print("mb=$mb value",round(d["value"],1),"launches/step",d["gpu_launches"]//d["steps"],"dw",round(k["k_dw3x3"],3),"gemm_hbm",round(k["k_gemm_tc[K<=256,hbm-bound]"],3))
PY
done
