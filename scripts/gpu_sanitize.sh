#!/bin/bash
# compute-sanitizer over a reduced `-m gpu` subset (small shapes: tiny128 / mid80 / mid100 / odd7 / voc, strict NMS with
# spill, the uncached-key NMS path, fallback assignment, TF-style mode, losses, AP kernels; no CUDA-graph replay).
# ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun -- 'bash scripts/gpu_sanitize.sh memcheck'     (then racecheck, synccheck, initcheck in later calls)
# Output: gpurun_out/sanitizer_<tool>.log (pytest output + the sanitizer's report and "ERROR SUMMARY" line).
TOOL=${1:-memcheck}
K='other_class_counts or decode_score or nms_kat or nms_golden or fused_decode_nms or class_limit_and_spill or many_rounds or global_sort_path or assign_boxes_golden or encode_truth_golden or compact_expands or empty_and_ragged or (loss_terms and (tiny128 or odd7)) or (loss_gradient and (tiny128 or odd7)) or voc_ap_micro or coco_ap_micro or randomised_small or tf_style or guard_band or channel_argmax or decode_locs or (ignore_mask_scatter and not stress) or box_augmenter_vs_oracle'
mkdir -p gpurun_out
LOG=gpurun_out/sanitizer_${TOOL}.log
EXTRA=""
[ "$TOOL" = "memcheck" ] && EXTRA="--leak-check no"
[ "$TOOL" = "initcheck" ] && EXTRA="--track-unused-memory no"
echo "# compute-sanitizer --tool $TOOL $EXTRA ; pytest -m gpu -k '$K'" > $LOG
ODK_SINGLE_STREAM=${ODK_SINGLE_STREAM-} timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $TOOL $EXTRA --print-limit 40 --error-exitcode 9 \
  python -m pytest tests -m gpu -q -p no:cacheprovider -k "$K" >> $LOG 2>&1
echo "exit $?" >> $LOG
grep -E "ERROR SUMMARY|passed|failed|exit " $LOG | tail -5
