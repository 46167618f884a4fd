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

// ---- ml.od_confusion_matrix (ml.py:397-451) ---------------------------------------------------------------------------
// One warp per image; detections live in shared memory.  Per image: (1) descending-confidence order of the
// detections (rank by counting; equal confidences: higher index first); (2) per detection the GT of its largest
// fp32 IoU (first maximum) and whether that IoU reaches the threshold; (3) the reference's sequential loop over the
// GTs.  Step 3 reproduces ml.py:421 literally: the per-detection mask m is applied to the POSITIONS of the sorted index
// array, so position j contributes detection order[j] when detection j passes the mask.  For one GT the visited set T
// splits into A = {class == gt class} and B = {still enabled, other class}; pass 1 (:424-431) books the first of A as
// the hit and the rest of A as duplicates, pass 2 (:433-442) books the first of B as a class confusion if A was empty
// and everything else as duplicates; all of T is disabled afterwards.  Leftover enabled detections are false
// positives (:446-447).
namespace {

struct CmArgs {
  int n_img, top_k, num_classes;
  const float* dets;           // [n_img, K, 6]; the real rows (class >= 0) come first
  const float4* gt_boxes;      // [n_gt]
  const int32_t* gt_classes;   // [n_gt]
  const int32_t* gt_offsets;   // [n_img + 1]
  float conf_thr, iou_thr;
  int32_t* cm;                 // [(C+1), (C+1)] pre-zeroed
};

__global__ void __launch_bounds__(kScoreWarps * 32) confusion_kernel(const CmArgs p) {
  extern __shared__ __align__(16) unsigned char cm_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int img = blockIdx.x * kScoreWarps + warp;
  if (img >= p.n_img) return;
  const int K = p.top_k, C = p.num_classes, C1 = p.num_classes + 1;
  unsigned char* base = cm_smem + (size_t)warp * odk_align_up((size_t)K * 32, 16);
  float4* dbox = reinterpret_cast<float4*>(base);                        // [K]
  float* dconf = reinterpret_cast<float*>(base + (size_t)K * 16);        // [K]
  short* dcls = reinterpret_cast<short*>(base + (size_t)K * 20);         // [K]
  short* order = reinterpret_cast<short*>(base + (size_t)K * 22);        // [K] detection at sorted position j
  short* dgt = reinterpret_cast<short*>(base + (size_t)K * 24);          // [K] GT of the largest IoU, -1 if below threshold
  short* tlist = reinterpret_cast<short*>(base + (size_t)K * 26);        // [K] detections visited for the current GT
  unsigned char* dlive = base + (size_t)K * 28;                          // [K] pred_enabled
  const float* rows = p.dets + (size_t)img * K * 6;
  int n_local = 0;
  for (int d = lane; d < K; d += 32) {
    const float* r = rows + d * 6;
    const bool real = r[0] >= 0.0f;
    dbox[d] = make_float4(r[2], r[3], r[4], r[5]);
    dconf[d] = r[1];
    dcls[d] = real ? (short)min((int)r[0], C) : (short)-1;
    dlive[d] = (real && r[1] >= p.conf_thr) ? 1 : 0;  // ml.py:410
    n_local += real ? 1 : 0;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) n_local += __shfl_xor_sync(0xffffffffu, n_local, s);
  const int n = n_local;  // number of detections of this image
  __syncwarp();
  const int g0 = p.gt_offsets[img], G = p.gt_offsets[img + 1] - g0;
  if (G > 0) {
    for (int d = lane; d < n; d += 32) {
      // (1) sorted position of detection d
      const float cd = dconf[d];
      int rk = 0;
      for (int q = 0; q < n; ++q) {
        const float cq = dconf[q];
        rk += (cq > cd) || (cq == cd && q > d);
      }
      order[rk] = (short)d;
      // (2) nearest GT (ml.py:413-415)
      const float4 db = dbox[d];
      const float da = odk_area(db);
      float best = -1.0f;
      int bg = 0;
      for (int g = 0; g < G; ++g) {
        const float4 gb = p.gt_boxes[g0 + g];
        const float v = odk_iou_value(gb, odk_area(gb), db, da);
        if (v > best) { best = v; bg = g; }
      }
      dgt[d] = (best >= p.iou_thr) ? (short)bg : (short)-1;
    }
    __syncwarp();
    // (3) the loop over the GTs (ml.py:418-444)
    for (int g = 0; g < G; ++g) {
      const int gc = p.gt_classes[g0 + g];
      int nA = 0, nT = 0, firstA = 0x7fffffff, firstB = 0x7fffffff;
      for (int j0 = 0; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        const bool mj = j < n && dlive[j] && dgt[j] == g;
        const int t = mj ? order[j] : 0;
        const bool inA = mj && dcls[t] == gc;
        const bool inB = mj && dlive[t] && dcls[t] != gc;
        const unsigned bm = __ballot_sync(0xffffffffu, mj), ba = __ballot_sync(0xffffffffu, inA), bb = __ballot_sync(0xffffffffu, inB);
        if (mj) tlist[nT + __popc(bm & ((1u << lane) - 1u))] = (short)t;
        nT += __popc(bm);
        nA += __popc(ba);
        if (ba && firstA == 0x7fffffff) firstA = j0 + __ffs(ba) - 1;
        if (bb && firstB == 0x7fffffff) firstB = j0 + __ffs(bb) - 1;
        if (inA) atomicAdd(p.cm + (size_t)((j == firstA) ? gc : C) * C1 + gc, 1);
        // pass 2 needs to know whether A is empty: decided after the scan (below) for the first of B only
        if (inB && j != firstB) atomicAdd(p.cm + (size_t)C * C1 + dcls[t], 1);
      }
      if (lane == 0) {
        if (firstB != 0x7fffffff) {
          const int t = order[firstB];
          atomicAdd(p.cm + (size_t)((nA == 0) ? gc : C) * C1 + dcls[t], 1);
        }
        if (nA == 0 && firstB == 0x7fffffff) atomicAdd(p.cm + (size_t)gc * C1 + C, 1);  // missed (:443-444)
      }
      __syncwarp();
      for (int i = lane; i < nT; i += 32) dlive[tlist[i]] = 0;
      __syncwarp();
    }
  }
  for (int d = lane; d < n; d += 32)
    if (dlive[d]) atomicAdd(p.cm + (size_t)C * C1 + dcls[d], 1);
}

}  // namespace

extern "C" int odk_confusion_matrix(int n_img, int top_k, int num_classes, const float* dets, const float* gt_boxes, const int32_t* gt_classes,
                                    const int32_t* gt_offsets, float conf_thr, float iou_thr, int32_t* cm, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(n_img >= 1 && top_k >= 1 && num_classes >= 1, ODK_ERR_INVALID_ARGUMENT, "confusion_matrix: bad sizes n_img=%d top_k=%d classes=%d",
              n_img, top_k, num_classes);
  ODK_REQUIRE(dets && gt_offsets && cm, ODK_ERR_INVALID_ARGUMENT, "confusion_matrix: null pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "confusion_matrix: gt_boxes must be 16-byte aligned");
  ODK_REQUIRE(top_k <= 32767 && num_classes <= 32766, ODK_ERR_UNSUPPORTED, "confusion_matrix: top_k / num_classes exceed 16-bit indices");
  const size_t smem = odk_align_up((size_t)top_k * 32, 16) * kScoreWarps;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "confusion_matrix: top_k=%d needs too much shared memory", top_k);
  ODK_CUDA(cudaFuncSetAttribute(confusion_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  ODK_CUDA(cudaMemsetAsync(cm, 0, sizeof(int32_t) * (size_t)(num_classes + 1) * (num_classes + 1), stream));
  CmArgs a;
  a.n_img = n_img; a.top_k = top_k; a.num_classes = num_classes;
  a.dets = dets; a.gt_boxes = reinterpret_cast<const float4*>(gt_boxes); a.gt_classes = gt_classes; a.gt_offsets = gt_offsets;
  a.conf_thr = conf_thr; a.iou_thr = iou_thr; a.cm = cm;
  {
    OdkProfileScope prof_("confusion_kernel", stream);
    confusion_kernel<<<(n_img + kScoreWarps - 1) / kScoreWarps, kScoreWarps * 32, smem, stream>>>(a);
  }
  ODK_LAUNCH_CHECK("confusion_kernel");
  return ODK_OK;
}

// ---- ObjectsPrediction.is_match / ml.od_accuracy (ml.py:159-182, 343-349) -----------------------------------------------
// One warp per image: the detections with conf >= conf_thr must pair off one-to-one with the ground truth — as many as
// there are GTs, each attached (largest fp32 IoU, first maximum) to a different GT with IoU >= iou_thr and of that GT's
// class.
namespace {

__global__ void __launch_bounds__(kScoreWarps * 32) exact_match_kernel(const CmArgs p, uint8_t* __restrict__ match) {
  extern __shared__ __align__(16) unsigned char em_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int img = blockIdx.x * kScoreWarps + warp;
  if (img >= p.n_img) return;
  const int K = p.top_k;
  short* dgt = reinterpret_cast<short*>(em_smem + (size_t)warp * odk_align_up((size_t)K * 4, 16));  // [K] GT of an enabled detection
  short* dcls = dgt + K;                                                                              // [K]
  const float* rows = p.dets + (size_t)img * K * 6;
  const int g0 = p.gt_offsets[img], G = p.gt_offsets[img + 1] - g0;
  int cnt = 0;
  bool ok = true;
  for (int d = lane; d < K; d += 32) {
    const float* r = rows + d * 6;
    const bool on = r[0] >= 0.0f && r[1] >= p.conf_thr;  // mask = confs >= conf_threshold (ml.py:163)
    short tgt = -1;
    if (on) {
      ++cnt;
      const float4 db = make_float4(r[2], r[3], r[4], r[5]);
      const float da = odk_area(db);
      float best = -1.0f;
      int bg = 0;
      for (int g = 0; g < G; ++g) {
        const float4 gb = p.gt_boxes[g0 + g];
        const float v = odk_iou_value(gb, odk_area(gb), db, da);
        if (v > best) { best = v; bg = g; }
      }
      if (G > 0 && best < p.iou_thr) ok = false;  // ml.py:171-172
      tgt = (short)bg;
    }
    dgt[d] = tgt;
    dcls[d] = (short)(int)r[0];
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
  __syncwarp();
  if (cnt != G) ok = false;  // ml.py:164-165
  if (cnt == G && G > 0) {
    for (int g = lane; g < G; g += 32) {  // ml.py:175-181: one detection per GT, of its class
      int seen = 0, cls = -1;
      for (int d = 0; d < K; ++d)
        if (dgt[d] == g) {
          if (seen == 0) cls = dcls[d];
          ++seen;
        }
      if (seen != 1 || cls != p.gt_classes[g0 + g]) ok = false;
    }
  }
  const bool all_ok = __all_sync(0xffffffffu, ok);
  if (lane == 0) match[img] = all_ok ? 1 : 0;
}

}  // namespace

extern "C" int odk_exact_match(int n_img, int top_k, const float* dets, const float* gt_boxes, const int32_t* gt_classes,
                               const int32_t* gt_offsets, float conf_thr, float iou_thr, uint8_t* match, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(n_img >= 1 && top_k >= 1, ODK_ERR_INVALID_ARGUMENT, "exact_match: bad sizes n_img=%d top_k=%d", n_img, top_k);
  ODK_REQUIRE(dets && gt_offsets && match, ODK_ERR_INVALID_ARGUMENT, "exact_match: null pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "exact_match: gt_boxes must be 16-byte aligned");
  ODK_REQUIRE(top_k <= 32767, ODK_ERR_UNSUPPORTED, "exact_match: top_k=%d exceeds 16-bit indices", top_k);
  const size_t smem = odk_align_up((size_t)top_k * 4, 16) * kScoreWarps;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "exact_match: top_k=%d needs too much shared memory", top_k);
  ODK_CUDA(cudaFuncSetAttribute(exact_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CmArgs a;
  a.n_img = n_img; a.top_k = top_k; a.num_classes = 0;
  a.dets = dets; a.gt_boxes = reinterpret_cast<const float4*>(gt_boxes); a.gt_classes = gt_classes; a.gt_offsets = gt_offsets;
  a.conf_thr = conf_thr; a.iou_thr = iou_thr; a.cm = nullptr;
  {
    OdkProfileScope prof_("exact_match_kernel", stream);
    exact_match_kernel<<<(n_img + kScoreWarps - 1) / kScoreWarps, kScoreWarps * 32, smem, stream>>>(a, match);
  }
  ODK_LAUNCH_CHECK("exact_match_kernel");
  return ODK_OK;
}
