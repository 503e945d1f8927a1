#!/bin/bash
# Sweep of the update kernel's work-list grading (OBST_OPT_DRAIN = tail fraction : divisor 1 : divisor 2).
# Prints the update kernel's duration and the whole update per setting; results under gpurun_out/.
out=${1:-gpurun_out/opt_drain_sweep.txt}
: > "$out"
for d in 0:1:1 0.03:2:4 0.05:2:2 0.05:2:4 0.05:3:6 0.08:2:4 0.1:2:8 0:1:1 0.05:2:4; do
  OBST_OPT_DRAIN=$d timeout 300 python bench.py --workload optimizer --steps 30 --warmup 3 --no-cpu-baseline 2>/dev/null \
    | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); c=d['config']
print('$d', 'update_ms', round(c['ms_update_kernel'],4), 'no_t_ms', round(c['ms_without_transposed_refresh'],4), 'whole_ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))" >> "$out"
done
cat "$out"
