#!/bin/bash
# ncu evidence: (1) launch list with per-launch device time, (2) --set full capture of the kernels named in $2 (regex), exported as the
# raw-page CSV (the .ncu-rep itself is kept only when it is small enough to travel back: gpurun_out/ is capped at 64 MiB).
# Usage: bash scripts/gpu_profile.sh <tag> <kernel-regex> <skip> <count> [--import-source]
TAG=${1:-r01}; KRE=${2:-k_gemm_tc}; SKIP=${3:-12}; CNT=${4:-2}; SRC=${5:-}
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs --batch ${BATCH:-256}"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
REP=/tmp/${TAG}_full
SRCFLAG=""; [ "$SRC" = "--import-source" ] && SRCFLAG="--import-source on"
ncu --set full --clock-control none $SRCFLAG -k regex:${KRE} -s ${SKIP} -c ${CNT} -f -o $REP $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "full capture rc=$?"
ncu -i $REP.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2> gpurun_out/${TAG}_raw.err; echo "raw export rc=$?"
SZ=$(stat -c %s $REP.ncu-rep); echo "report size $SZ"
[ "$SZ" -lt 30000000 ] && cp $REP.ncu-rep gpurun_out/${TAG}_full.ncu-rep
ls -la gpurun_out/ | grep ${TAG}
