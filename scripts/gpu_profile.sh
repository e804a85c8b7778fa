#!/bin/bash
# ncu evidence: (1) launch list with per-launch device time, (2) --set full capture of the kernels named in $1 (regex).
# Usage: bash scripts/gpu_profile.sh <tag> <kernel-regex> <skip> <count>
TAG=${1:-r01}; KRE=${2:-k_gemm_tc}; SKIP=${3:-12}; CNT=${4:-2}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --batch ${BATCH:-256}"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s ${SKIP} -c ${CNT} -f -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/
