#!/bin/bash
# One call: GPU tests, the driver-style bench line (and the reference arm), then — after the same command line has exited
# 0 without a profiler — the ncu launch list and one ncu --set full capture of the streaming / NMS kernels.
TAG=${1:-r02}
OUT=gpurun_out; mkdir -p $OUT
timeout 500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu_$TAG.log; tail -3 $OUT/pytest_gpu_$TAG.log
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 200 > $OUT/clocks_$TAG.csv &
SMI=$!
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2>/dev/null; echo "ref exit $?"
timeout 400 python bench.py --steps 20 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench exit $?"
timeout 400 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --quick --no-e2e > $OUT/bench_${TAG}_200.json 2>/dev/null; echo "bench200 exit $?"
kill $SMI
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --quick --no-e2e --sets 4"
$CMD > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_list_$TAG.log 2>&1
echo "ncu list exit $?"
$CMD > $OUT/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"decode_score_kernel|nms_image_kernel|encode_rows|assign_greedy" -s 12 -c 8 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_$TAG.log 2>&1
echo "ncu full exit $?"
