#!/bin/bash
# Round-2 staged GPU run: (1) the new fused middle-flow kernel against the unfused path in its own process (a protocol bug traps
# there without poisoning the rest), (2) the whole GPU suite, (3) bench.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
TAG=${1:-r02a}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/${TAG}_gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_variants.py -x -q -k "variants_are_bit or graph_replay" > gpurun_out/${TAG}_variants.log 2>&1
echo "variants rc=$?" | tee -a gpurun_out/${TAG}_variants.log
tail -15 gpurun_out/${TAG}_variants.log
timeout 2400 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -40 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"
tail -5 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    j = json.load(open("gpurun_out/${TAG}_bench.json"))
    print("value", j["value"], "ms", j["ms_per_step"], "e2e", j["e2e"]["value"])
    for k in j["kernels"]:
        print("%-44s %3d %7.3f ms  frac %.3f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["frac"]))
    print(json.dumps(j.get("configs"))[:1500])
except Exception as e:
    print("bench json unreadable:", e)
PY
