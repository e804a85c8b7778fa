#!/bin/bash
# inner loop for kernel work: bit-identity of the kernel variants, then the bench's per-kernel table (no CPU baseline, no configs)
set -u
mkdir -p gpurun_out
TAG=${1:-q}
timeout 600 python -m pytest tests/test_gpu_variants.py -x -q -k "variants_are_bit or batch_256" > gpurun_out/${TAG}_variants.log 2>&1
echo "variants rc=$?"; tail -5 gpurun_out/${TAG}_variants.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-configs --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    j = json.load(open("gpurun_out/${TAG}_bench.json"))
    print("value", round(j["value"]), "ms", round(j["ms_per_step"], 3), "e2e", round(j["e2e"]["value"]), "b1 p50", j["latency_b1"]["p50_ms"], j["clocks"])
    for k in j["kernels"]:
        print("%-44s %3d %7.3f ms  frac %.3f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["frac"]))
except Exception as e:
    print("bench json unreadable:", e)
PY
