#!/bin/bash
# One gpurun call of the build->measure loop: GPU tests, the bench line, a knob sweep.  Usage: gpu_round.sh TAG [sweep specs...]
TAG=$1; shift
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_$TAG.log
tail -25 gpurun_out/pytest_$TAG.log
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit $?"
tail -c 2000 gpurun_out/bench_$TAG.err
if [ $# -gt 0 ]; then
  SWEEP_STEPS=200 timeout 500 bash scripts/gpu_sweep.sh "$@" > gpurun_out/sweep_$TAG.log 2>&1
  cat gpurun_out/sweep_$TAG.log
fi
