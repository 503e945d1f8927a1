#!/bin/bash
# 8-GPU A/B of environment switches on the headline bench (one box, back to back).
# usage: scale8_ab.sh out.txt "VAR=val VAR2=val" "VAR=val" ...   (each quoted argument is one arm; "-" = no switch)
out=$1; shift
: > "$out"
port=29500
for arm in "$@" "$1"; do
  port=$((port+1))
  envs=""; [ "$arm" != "-" ] && envs="$arm"
  env $envs timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('[$arm]', round(d['value']), 'e2e', round(d['e2e']['value']), 'ms', round(d['ms_per_step'],2), 'sm_mhz', d['clocks']['sm_mhz'], 'gemm_ms', round(d['roofline']['kernel_ms_per_step'],1), 'frac', round(d['roofline']['frac'],4))" >> "$out"
done
cat "$out"
