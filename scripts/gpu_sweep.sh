#!/bin/bash
# Throughput-mode sweep: bench step time under environment knobs / lane counts.
# Usage: gpurun -- 'bash scripts/gpu_sweep.sh "ENV=1 ENV2=2|--lanes 2" "…|…"'
run() {
  echo "== $*"
  env $1 python bench.py --steps 300 --warmup 10 --no-cpu-baseline $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']
print(round(d['value']), round(d['ms_per_step']*1e3,2), round(d['config']['ms_per_step_one_in_flight']*1e3,2), {n[:12]:round(v['ms_per_launch']*1e3,1) for n,v in k.items()})"
}
for spec in "$@"; do run "${spec%%|*}" "${spec#*|}"; done
