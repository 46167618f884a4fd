// ml.compute_scores / ml.search_conf_threshold (ml.py:328-340, 352-394): per-class TP / FP / FN counts of a
// detection set against the ground truth, for MANY confidence thresholds in one launch
// (search_conf_threshold re-runs compute_scores 50 times on the CPU; here a threshold is one more grid row).
//
// One warp per (image, threshold).  The reference's loop over the image's GT boxes is sequential (a matched
// detection is disabled for the following GTs), lanes share the detections:
//   for each GT (in annotation order): among the still enabled detections of its class with conf >= thr,
//   arg-max of the fp32 IoU (first maximum wins); IoU >= iou_thr -> TP (unless difficult) and the detection is
//   disabled, else FN (unless difficult).  Afterwards every still enabled detection is a FP of its own class.
// Precision / recall / F are formed on the host from the exact integer counts with the reference's expressions.
#include "odk_common.cuh"

namespace {

constexpr int kScoreWarps = 4;

struct ScoreArgs {
  int n_img, top_k, num_classes, n_thr;
  const float* dets;           // [n_img, K, 6] rows [class, conf, x1, y1, x2, y2]; class < 0 = padding
  const float4* gt_boxes;      // [n_gt]
  const int32_t* gt_classes;   // [n_gt]
  const uint8_t* gt_difficult; // [n_gt]
  const int32_t* gt_offsets;   // [n_img + 1]
  const float* conf_thr;       // [n_thr] (device)
  float iou_thr;
  int32_t* counts;             // [n_thr, 3, C] = tp, fp, fn (pre-zeroed)
};

__global__ void __launch_bounds__(kScoreWarps * 32) match_counts_kernel(const ScoreArgs p) {
  extern __shared__ __align__(16) unsigned char score_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int img = blockIdx.x * kScoreWarps + warp;
  const int t = blockIdx.y;
  if (img >= p.n_img) return;
  const int K = p.top_k;
  unsigned char* base = score_smem + (size_t)warp * odk_align_up((size_t)K * 21, 16);
  float4* dbox = reinterpret_cast<float4*>(base);                 // [K]
  int* dcls = reinterpret_cast<int*>(base + (size_t)K * 16);      // [K] class, or -1 when disabled / not eligible
  unsigned char* dlive = base + (size_t)K * 20;                   // [K] still enabled
  const float thr = p.conf_thr[t];
  const float* rows = p.dets + (size_t)img * K * 6;
  for (int d = lane; d < K; d += 32) {
    const float* r = rows + d * 6;
    const int c = (int)r[0];
    const bool ok = r[0] >= 0.0f && c < p.num_classes && r[1] >= thr;  // pred_enabled = confs >= conf_threshold (ml.py:367)
    dbox[d] = make_float4(r[2], r[3], r[4], r[5]);
    dcls[d] = ok ? c : -1;
    dlive[d] = ok ? 1 : 0;
  }
  __syncwarp();
  int32_t* tp = p.counts + ((size_t)t * 3 + 0) * p.num_classes;
  int32_t* fp = p.counts + ((size_t)t * 3 + 1) * p.num_classes;
  int32_t* fn = p.counts + ((size_t)t * 3 + 2) * p.num_classes;
  const int g0 = p.gt_offsets[img], g1 = p.gt_offsets[img + 1];
  for (int g = g0; g < g1; ++g) {
    const int gc = p.gt_classes[g];
    const float4 gb = p.gt_boxes[g];
    const float ga = odk_area(gb);
    float best = -1.0f;  // "no detection of this class": pred_iou = -1 (ml.py:377)
    int best_d = 0x7fffffff;
    for (int d = lane; d < K; d += 32) {
      if (!dlive[d] || dcls[d] != gc) continue;
      const float4 db = dbox[d];
      const float v = odk_iou_value(gb, ga, db, odk_area(db));  // ml.compute_iou (ml.py:502-519), fp32
      if (v > best) { best = v; best_d = d; }                  // ascending d per lane: the first maximum stays
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, s);
      const int od = __shfl_xor_sync(0xffffffffu, best_d, s);
      if (ov > best || (ov == best && od < best_d)) { best = ov; best_d = od; }
    }
    const bool difficult = p.gt_difficult[g] != 0;
    const bool in_range = gc >= 0 && gc < p.num_classes;
    if (best_d != 0x7fffffff && best >= p.iou_thr) {  // detected (ml.py:378-382)
      if (lane == 0) {
        if (!difficult && in_range) atomicAdd(tp + gc, 1);
        dlive[best_d] = 0;
      }
    } else if (lane == 0 && !difficult && in_range) {
      atomicAdd(fn + gc, 1);
    }
    __syncwarp();
  }
  for (int d = lane; d < K; d += 32)
    if (dlive[d]) atomicAdd(fp + dcls[d], 1);  // ml.py:388-389
}

}  // namespace

extern "C" int odk_match_counts(int n_img, int top_k, int num_classes, const float* dets, const float* gt_boxes, const int32_t* gt_classes,
                                const uint8_t* gt_difficult, const int32_t* gt_offsets, int n_thr, const float* conf_thr, float iou_thr,
                                int32_t* counts, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(n_img >= 1 && top_k >= 1 && num_classes >= 1 && n_thr >= 1, ODK_ERR_INVALID_ARGUMENT,
              "match_counts: bad sizes n_img=%d top_k=%d classes=%d n_thr=%d", n_img, top_k, num_classes, n_thr);
  ODK_REQUIRE(dets && gt_offsets && conf_thr && counts, ODK_ERR_INVALID_ARGUMENT, "match_counts: null pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "match_counts: gt_boxes must be 16-byte aligned");
  ODK_REQUIRE(n_thr <= 65535, ODK_ERR_UNSUPPORTED, "match_counts: n_thr=%d > 65535", n_thr);
  const size_t smem = odk_align_up((size_t)top_k * 21, 16) * kScoreWarps;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "match_counts: top_k=%d needs too much shared memory", top_k);
  ODK_CUDA(cudaFuncSetAttribute(match_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  ODK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * (size_t)n_thr * 3 * num_classes, stream));
  ScoreArgs a;
  a.n_img = n_img; a.top_k = top_k; a.num_classes = num_classes; a.n_thr = n_thr;
  a.dets = dets; a.gt_boxes = reinterpret_cast<const float4*>(gt_boxes); a.gt_classes = gt_classes;
  a.gt_difficult = gt_difficult; a.gt_offsets = gt_offsets; a.conf_thr = conf_thr; a.iou_thr = iou_thr; a.counts = counts;
  dim3 grid((n_img + kScoreWarps - 1) / kScoreWarps, n_thr);
  {
    OdkProfileScope prof_("match_counts_kernel", stream);
    match_counts_kernel<<<grid, kScoreWarps * 32, smem, stream>>>(a);
  }
  ODK_LAUNCH_CHECK("match_counts_kernel");
  return ODK_OK;
}
