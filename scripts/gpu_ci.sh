#!/bin/bash
# Staged GPU check: safe kernels first, the tcgen05 GEMM in its own process under a timeout.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== stage 1: preprocess + fp32 + bf16(simple gemm)" 
timeout 900 python -m pytest tests/test_gpu_preprocess.py tests/test_gpu_forward.py -q -m gpu -k "not auto" -x -s > gpurun_out/stage1.log 2>&1; echo "stage1 rc=$?"
tail -15 gpurun_out/stage1.log
echo "== stage 2: tcgen05 gemm kernel"
timeout 600 python -m pytest tests/test_gpu_gemm_tc.py -q -m gpu -x > gpurun_out/stage2.log 2>&1; echo "stage2 rc=$?"
tail -25 gpurun_out/stage2.log
echo "== stage 3: bf16 auto (tcgen05 in the network) + gateway"
timeout 900 python -m pytest tests/test_gpu_forward.py tests/test_gpu_gateway.py -q -m gpu -k "auto or gateway or fused or apply or predict" -x -s > gpurun_out/stage3.log 2>&1; echo "stage3 rc=$?"
tail -15 gpurun_out/stage3.log
echo "== stage 4: smoke + bench"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err; cat gpurun_out/bench.json
