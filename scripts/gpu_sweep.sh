#!/bin/bash
# Throughput-mode sweep: bench step time under environment knobs / ring shapes.
# Usage: gpurun -- 'bash scripts/gpu_sweep.sh "ENV=1 ENV2=2|--lanes 2" "…|…"'   (output: one line per spec)
run() {
  echo "== $*"
  env $1 python bench.py --steps ${SWEEP_STEPS:-200} --warmup 10 --no-cpu-baseline --quick --no-e2e --sets ${SWEEP_SETS:-4} $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernels']
print(round(d['value']), 'us/step', round(d['ms_per_step']*1e3,2), 'one-in-flight', round(d['schedule']['ms_per_step_one_in_flight']*1e3,2), {n.replace('_kernel','')[:14]:round(v['ms_per_launch']*1e3,1) for n,v in k.items()})"
}
for spec in "$@"; do run "${spec%%|*}" "${spec#*|}"; done
