#!/bin/bash
# ncu evidence for one round (run under gpurun, 1 GPU).  $1 = tag (e.g. r1).  Outputs land in gpurun_out/.
TAG=${1:-r1}
CMD="python bench.py --steps 1 --warmup 1 --depth 2 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 40 -c 14 -o gpurun_out/${TAG}_gemm $CMD > gpurun_out/${TAG}_ncu_gemm.log 2>&1
echo "gemm capture rc=$?"
ls -la gpurun_out | tail -12
