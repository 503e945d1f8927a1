#!/bin/bash
# ncu evidence for one round (run under gpurun, 1 GPU).  $1 = tag (e.g. r1b).  Outputs land in gpurun_out/.
# The launch list covers ONE steady-state step of a depth-2 model (the per-depth kernel mix is that of depth 32).
TAG=${1:-r1}
CMD="python bench.py --steps 1 --warmup 1 --depth 2 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 800 -c 760 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 60 -c 16 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu_gemm.log 2>&1
echo "gemm capture rc=$?"
ls -la gpurun_out | tail -8
