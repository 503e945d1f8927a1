#!/bin/bash
# Same-box alternating A/B of one bench.py flag: ab_flag.sh "<flag or empty>" out.txt [extra bench args]
flag=$1; out=${2:-gpurun_out/ab_flag.txt}; shift 2
: > "$out"
for f in "$flag" "" "$flag" ""; do
  timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline $f "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('flags=[$f]', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['ms_per_step'],2), 'sm_mhz', d['clocks']['sm_mhz'], 'gemm_ms', round(d['roofline']['kernel_ms_per_step'],1), 'frac', round(d['roofline']['frac'],4), 'launches', d['gpu_launches'])" >> "$out"
done
cat "$out"
