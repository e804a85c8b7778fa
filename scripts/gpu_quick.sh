#!/bin/bash
# quick loop: tc gemm tests, bf16 network parity, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm_tc.py -q -m gpu -x > gpurun_out/q_gemm.log 2>&1; echo "gemm rc=$?"; tail -4 gpurun_out/q_gemm.log
timeout 900 python -m pytest tests/test_gpu_forward.py -q -m gpu -k "bf16" -x -s > gpurun_out/q_fwd.log 2>&1; echo "fwd rc=$?"; grep -E "bf16|passed|failed|Error|error" gpurun_out/q_fwd.log | tail -12
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/q_bench.err
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/q_bench.json'))
    print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],3),'lat',d.get('latency_b1'))
    for k in d['kernels']: print(f"  {k['kernel']:<34} n={k['launches_per_step']:3d} {k['ms_per_step']:8.3f} ms {100*k['share']:5.1f}%  {k['achieved_tflops']:8.1f} TF {k['achieved_gbs']:8.1f} GB/s frac {k['frac']:.3f}")
except Exception as e: print('no bench json',e)
PY
