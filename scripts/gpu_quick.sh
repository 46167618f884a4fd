#!/bin/bash
# Quick GPU iteration: parity tests + short bench.  Usage: gpurun --timeout 900 -- 'bash scripts/gpu_quick.sh tag'
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest exit $?"
tail -25 $OUT/pytest_$TAG.log
timeout 300 python bench.py --steps 200 --warmup 10 --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"; cat $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
