#!/bin/bash
# Same-box alternating A/B of one environment switch on the default bench: ab_env.sh VAR A_VALUE B_VALUE [out] [extra bench args]
var=$1; a=$2; b=$3; out=${4:-gpurun_out/ab_env.txt}; shift 4
: > "$out"
for v in $a $b $a $b; do
  env $var=$v timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('$var=$v', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['ms_per_step'],2), 'sm_mhz', d['clocks']['sm_mhz'], 'gemm_ms', round(d['roofline']['kernel_ms_per_step'],1), 'frac', round(d['roofline']['frac'],4))" >> "$out"
done
cat "$out"
