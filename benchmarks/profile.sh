#!/bin/bash
# ncu evidence for one round (run under gpurun, 1 GPU, ONE ncu pass per call).  $1 = tag (e.g. r2k), $2 = launches | gemm.
# The launch list covers ONE steady-state step of a depth-2 model (the per-depth kernel mix is that of depth 32).
TAG=${1:-r1}
WHAT=${2:-launches}
CMD="python bench.py --steps 1 --warmup 1 --depth 2 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
if [ "$WHAT" = launches ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -s 800 -c 800 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
  echo "launch list rc=$?"
else
  ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 60 -c 16 -o gpurun_out/${TAG}_gemm -f $CMD > gpurun_out/${TAG}_ncu_gemm.log 2>&1
  echo "gemm capture rc=$?"
fi
ls -la gpurun_out | tail -4
