// K6 — MS COCO detection metrics (data/coco.py:75-99 -> chainercv.evaluations.eval_detection_coco ->
// pycocotools COCOeval, iouType 'bbox').
//
// Neither chainercv nor pycocotools is vendored or pinned by the reference, so this stage follows the published
// algorithm (COCOeval.computeIoU / evaluateImg / accumulate) as restated in oracle/restated.py (coco_evaluate) —
// PARITY UNPINNED against the real packages.
//
//   coco_count_kernel   per detection row: conf >= conf_thr and class in range -> class histogram (segment sizes)
//   coco_layout_kernel  class segment offsets (each padded to a power of two for the sort)
//   coco_match_kernel   one CTA per image: boxes -> integer pixels (ml.py:62-65,184-186); rank of every detection
//                       within its (image, class) by score (stable, mergesort order), the first 100 kept
//                       (maxDets[-1]); then one thread per (class present, area range, IoU threshold) walks that
//                       class's detections in score order through evaluateImg's greedy matching (non-ignored GT
//                       first, crowd GT re-matchable, ties to the later GT) and records per detection a 40-bit
//                       matched mask and a 40-bit ignore mask (bit = area*10 + threshold); counts the non-ignored
//                       GT per (class, area range); emits one sort key per kept detection into its class segment.
//   coco_class_kernel   one CTA per class: bitonic sort (score descending, ties in image / rank order = mergesort on
//                       the concatenated list), then one warp per (area range, maxDets, IoU threshold): ballot
//                       prefix counts of TP / FP, precision tp/(fp+tp+eps) at every TP, bucketed by the last recall
//                       threshold it reaches, suffix maximum = the right-to-left precision envelope sampled by
//                       searchsorted(recall, recThrs, 'left'); writes precision[T,R,C,A,M] and recall[T,C,A,M]
//                       (-1 where the class has no non-ignored GT in that area range).
// The 12 summary numbers (chainercv _summarize: nanmean over thresholds / recall points, then over classes) are a few
// hundred flops on these arrays and are taken on the host by object_detector_b200/coco.py.
#include "odk_common.cuh"
#include "odk_sort.cuh"

namespace {

constexpr int kT = 10, kR = 101, kA = 4, kM = 3, kMaxDet = 100;
constexpr int kMatchThreads = 128;
constexpr int kClassThreads = 1024;
constexpr int kClassChunk = 16384;
constexpr int kGtWords = 4;  // matched-GT bitmask: up to 256 GT boxes per image

struct CocoWorkspace {
  int* cls_count;   // [C]
  int* cls_off;     // [C+1]
  int* cls_cursor;  // [C]
  int* npig;        // [C*4]
  unsigned long long* keys;   // [cap]
  unsigned long long* mbits;  // [n_img*K]
  unsigned long long* igbits; // [n_img*K]
  uint8_t* rank;              // [n_img*K]
  unsigned long long* s_m;    // [cap]   payloads in sorted order
  unsigned long long* s_ig;   // [cap]
  uint8_t* s_rank;            // [cap]
  size_t head_bytes;          // the int arrays (cleared per call)
  size_t bytes;
};

CocoWorkspace coco_layout(void* base, int n_img, int top_k, int num_classes) {
  CocoWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = odk_align_up(off + bytes, 256); return o; };
  const size_t rows = (size_t)n_img * top_k, cap = 2 * rows + 2 * (size_t)num_classes;
  const size_t o0 = take(sizeof(int) * (size_t)num_classes);
  const size_t o1 = take(sizeof(int) * ((size_t)num_classes + 1));
  const size_t o2 = take(sizeof(int) * (size_t)num_classes);
  const size_t o3 = take(sizeof(int) * (size_t)num_classes * kA);
  w.head_bytes = off;
  const size_t o4 = take(8 * cap), o5 = take(8 * rows), o6 = take(8 * rows), o7 = take(rows);
  const size_t o8 = take(8 * cap), o9 = take(8 * cap), o10 = take(cap);
  char* p = static_cast<char*>(base);
  w.cls_count = reinterpret_cast<int*>(p + o0);
  w.cls_off = reinterpret_cast<int*>(p + o1);
  w.cls_cursor = reinterpret_cast<int*>(p + o2);
  w.npig = reinterpret_cast<int*>(p + o3);
  w.keys = reinterpret_cast<unsigned long long*>(p + o4);
  w.mbits = reinterpret_cast<unsigned long long*>(p + o5);
  w.igbits = reinterpret_cast<unsigned long long*>(p + o6);
  w.rank = reinterpret_cast<uint8_t*>(p + o7);
  w.s_m = reinterpret_cast<unsigned long long*>(p + o8);
  w.s_ig = reinterpret_cast<unsigned long long*>(p + o9);
  w.s_rank = reinterpret_cast<uint8_t*>(p + o10);
  w.bytes = off;
  return w;
}

__device__ __forceinline__ bool coco_det_valid(const float* row, float conf_thr, int num_classes, int& cls) {
  cls = (int)row[0];  // p[:, 0].astype(np.int32), od.py:261
  return row[1] >= conf_thr && row[0] >= 0.0f && cls < num_classes;
}

__global__ void __launch_bounds__(256) coco_count_kernel(const float* __restrict__ dets, long long rows, float conf_thr, int num_classes,
                                                         int* __restrict__ cls_count) {
  extern __shared__ int s_hist[];
  for (int c = threadIdx.x; c < num_classes; c += blockDim.x) s_hist[c] = 0;
  __syncthreads();
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int cls;
  if (r < rows && coco_det_valid(dets + r * 6, conf_thr, num_classes, cls)) atomicAdd(s_hist + cls, 1);
  __syncthreads();
  for (int c = threadIdx.x; c < num_classes; c += blockDim.x)
    if (s_hist[c]) atomicAdd(cls_count + c, s_hist[c]);
}

__global__ void coco_layout_kernel(int num_classes, const int* __restrict__ cls_count, int* __restrict__ cls_off) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int off = 0;
    for (int c = 0; c < num_classes; ++c) {
      cls_off[c] = off;
      int pad = 2;
      while (pad < cls_count[c]) pad <<= 1;
      off += pad;
    }
    cls_off[num_classes] = off;
  }
}

__device__ __forceinline__ int coco_to_pixel(float v, int scale) { return (int)rint((double)v * (double)scale); }

struct CocoMatchArgs {
  int n_img, top_k, num_classes, gt_cap;
  const float* dets;
  float conf_thr;
  const int32_t* img_wh;
  const float4* gt_boxes;
  const int32_t* gt_classes;
  const double* gt_areas;
  const uint8_t* gt_crowd;
  const int32_t* gt_offsets;
  const int* cls_off;
  int* cls_cursor;
  int* npig;
  unsigned long long* keys;
  unsigned long long* mbits;
  unsigned long long* igbits;
  uint8_t* rank;
  double iou_thr[kT];
  double area_lo[kA], area_hi[kA];
};

// maskApi bbIou on [x, y, w, h] given as integer pixel corners (exact in fp64)
__device__ __forceinline__ double coco_iou(const int4 d, const int4 g, bool crowd) {
  const double w = (double)(min(d.z, g.z) - max(d.x, g.x));
  const double h = (double)(min(d.w, g.w) - max(d.y, g.y));
  if (w <= 0.0 || h <= 0.0) return 0.0;
  const double inter = w * h;
  const double da = (double)(d.z - d.x) * (double)(d.w - d.y);
  const double ga = (double)(g.z - g.x) * (double)(g.w - g.y);
  const double u = crowd ? da : da + ga - inter;
  return inter / u;
}

__global__ void __launch_bounds__(kMatchThreads) coco_match_kernel(const CocoMatchArgs p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int K = p.top_k, img = blockIdx.x, tid = threadIdx.x;
  const int g0 = p.gt_offsets[img], G = p.gt_offsets[img + 1] - g0;
  // shared layout
  int4* s_box = reinterpret_cast<int4*>(smem);                               // [K]
  int4* s_gbox = s_box + K;                                                  // [gt_cap]
  unsigned long long* s_m = reinterpret_cast<unsigned long long*>(s_gbox + p.gt_cap);  // [K]
  unsigned long long* s_ig = s_m + K;                                        // [K]
  double* s_garea = reinterpret_cast<double*>(s_ig + K);                     // [gt_cap]
  float* s_score = reinterpret_cast<float*>(s_garea + p.gt_cap);             // [K]
  int* s_cls = reinterpret_cast<int*>(s_score + K);                          // [K]
  int* s_rank = s_cls + K;                                                   // [K]
  int* s_pos = s_rank + K;                                                   // [K]
  int* s_perm = s_pos + K;                                                   // [K]
  int* s_seg = s_perm + K;                                                   // [K+1] segment starts in perm
  int* s_gcls = s_seg + K + 1;                                               // [gt_cap]
  unsigned char* s_gcrowd = reinterpret_cast<unsigned char*>(s_gcls + p.gt_cap);  // [gt_cap]
  __shared__ int s_nseg, s_nvalid;

  const int W = p.img_wh[2 * img], H = p.img_wh[2 * img + 1];
  for (int i = tid; i < K; i += kMatchThreads) {
    const float* row = p.dets + ((size_t)img * K + i) * 6;
    int cls;
    const bool ok = coco_det_valid(row, p.conf_thr, p.num_classes, cls);
    s_cls[i] = ok ? cls : 0x7fffffff;
    s_score[i] = row[1];
    s_box[i] = make_int4(coco_to_pixel(row[2], W), coco_to_pixel(row[3], H), coco_to_pixel(row[4], W), coco_to_pixel(row[5], H));
    s_m[i] = 0ull;
    s_ig[i] = 0ull;
  }
  for (int g = tid; g < G; g += kMatchThreads) {
    const float4 b = p.gt_boxes[g0 + g];
    s_gbox[g] = make_int4(coco_to_pixel(b.x, W), coco_to_pixel(b.y, H), coco_to_pixel(b.z, W), coco_to_pixel(b.w, H));
    s_gcls[g] = p.gt_classes[g0 + g];
    s_garea[g] = p.gt_areas[g0 + g];
    s_gcrowd[g] = p.gt_crowd[g0 + g] ? 1 : 0;
  }
  __syncthreads();
  // rank within (image, class): argsort(-score, kind='mergesort') -> equal scores keep their input order
  for (int i = tid; i < K; i += kMatchThreads) {
    const int ci = s_cls[i];
    const float si = s_score[i];
    int rank = 0, below = 0;
    for (int j = 0; j < K; ++j) {
      const int cj = s_cls[j];
      const float sj = s_score[j];
      below += cj < ci;
      rank += (cj == ci) && (sj > si || (sj == si && j < i));
    }
    s_rank[i] = rank;
    s_pos[i] = below + rank;
  }
  __syncthreads();
  for (int i = tid; i < K; i += kMatchThreads) s_perm[s_pos[i]] = i;  // a permutation: (class, rank) is a total order
  __syncthreads();
  if (tid == 0) {
    int nseg = 0, nvalid = 0;
    for (int q = 0; q < K; ++q) {
      const int c = s_cls[s_perm[q]];
      if (c == 0x7fffffff) break;
      ++nvalid;
      if (q == 0 || s_cls[s_perm[q - 1]] != c) s_seg[nseg++] = q;
    }
    s_seg[nseg] = nvalid;
    s_nseg = nseg;
    s_nvalid = nvalid;
  }
  __syncthreads();
  // evaluateImg: one thread per (class segment, area range, IoU threshold)
  const int items = s_nseg * (kA * kT);
  for (int item = tid; item < items; item += kMatchThreads) {
    const int seg = item / (kA * kT), a = (item / kT) % kA, t = item % kT;
    const int q0 = s_seg[seg], D = min(kMaxDet, s_seg[seg + 1] - q0);
    const int c = s_cls[s_perm[q0]];
    const double lo = p.area_lo[a], hi = p.area_hi[a];
    const double thr = fmin(p.iou_thr[t], 1.0 - 1e-10);
    const unsigned long long bit = 1ull << (a * kT + t);
    unsigned long long used[kGtWords] = {0ull, 0ull, 0ull, 0ull};
    for (int d = 0; d < D; ++d) {
      const int i = s_perm[q0 + d];
      const int4 db = s_box[i];
      double best = thr;
      int m = -1;
      // non-ignored GT first (gtind = argsort(_ignore, mergesort)), in input order; ties go to the later GT
      for (int g = 0; g < G; ++g) {
        if (s_gcls[g] != c) continue;
        const double ar = s_garea[g];
        if (s_gcrowd[g] || ar < lo || ar > hi) continue;
        if ((used[g >> 6] >> (g & 63)) & 1ull) continue;
        const double iou = coco_iou(db, s_gbox[g], false);
        if (iou < best) continue;
        best = iou;
        m = g;
      }
      bool m_ign = false;
      if (m < 0) {  // then the ignored ones; a crowd GT can be matched again
        for (int g = 0; g < G; ++g) {
          if (s_gcls[g] != c) continue;
          const double ar = s_garea[g];
          const bool crowd = s_gcrowd[g] != 0;
          if (!(crowd || ar < lo || ar > hi)) continue;
          if (((used[g >> 6] >> (g & 63)) & 1ull) && !crowd) continue;
          const double iou = coco_iou(db, s_gbox[g], crowd);
          if (iou < best) continue;
          best = iou;
          m = g;
          m_ign = true;
        }
      }
      if (m >= 0) {
        used[m >> 6] |= 1ull << (m & 63);
        atomicOr(&s_m[i], bit);
        if (m_ign) atomicOr(&s_ig[i], bit);
      } else {
        const double da = (double)(db.z - db.x) * (double)(db.w - db.y);
        if (da < lo || da > hi) atomicOr(&s_ig[i], bit);
      }
    }
  }
  // non-ignored GT per (class, area range): npig of accumulate()
  for (int g = tid; g < G; g += kMatchThreads) {
    const int c = s_gcls[g];
    if (c < 0 || c >= p.num_classes) continue;
    const double ar = s_garea[g];
#pragma unroll
    for (int a = 0; a < kA; ++a)
      if (!(s_gcrowd[g] || ar < p.area_lo[a] || ar > p.area_hi[a])) atomicAdd(p.npig + c * kA + a, 1);
  }
  __syncthreads();
  // one key per kept detection: [flip(score):32][~(image*K + position in (class, rank) order):32]
  for (int i = tid; i < K; i += kMatchThreads) {
    const int c = s_cls[i];
    if (c == 0x7fffffff || s_rank[i] >= kMaxDet) continue;
    const unsigned ordinal = (unsigned)img * (unsigned)K + (unsigned)s_pos[i];
    p.mbits[ordinal] = s_m[i];
    p.igbits[ordinal] = s_ig[i];
    p.rank[ordinal] = (uint8_t)s_rank[i];
    const int slot = atomicAdd(p.cls_cursor + c, 1);
    p.keys[p.cls_off[c] + slot] = ((unsigned long long)odk_flip_score(s_score[i]) << 32) | (unsigned long long)(0xffffffffu - ordinal);
  }
}

struct CocoClassArgs {
  int num_classes;
  const int* cls_cursor;
  const int* cls_off;
  const int* npig;
  unsigned long long* keys;
  const unsigned long long* mbits;
  const unsigned long long* igbits;
  const uint8_t* rank;
  unsigned long long* s_m;
  unsigned long long* s_ig;
  uint8_t* s_rank;
  double* precision;  // [T,R,C,A,M]
  double* recall;     // [T,C,A,M]
  double rec_thr[kR];
};

__global__ void __launch_bounds__(kClassThreads) coco_class_kernel(const CocoClassArgs p) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ double s_thr[kR];
  __shared__ unsigned long long s_bmax[kClassThreads / 32][kR + 3];
  const int c = blockIdx.x, C = p.num_classes;
  const int cnt = p.cls_cursor[c];
  const int off = p.cls_off[c];
  unsigned long long* g = p.keys + off;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int r = tid; r < kR; r += kClassThreads) s_thr[r] = p.rec_thr[r];
  odk_block_sort_desc(g, cnt, s_keys, kClassChunk);
  // payloads in sorted order
  for (int e = tid; e < cnt; e += kClassThreads) {
    const unsigned ordinal = 0xffffffffu - (unsigned)(g[e] & 0xffffffffull);
    p.s_m[off + e] = p.mbits[ordinal];
    p.s_ig[off + e] = p.igbits[ordinal];
    p.s_rank[off + e] = p.rank[ordinal];
  }
  __syncthreads();
  const unsigned long long* sm = p.s_m + off;
  const unsigned long long* sig = p.s_ig + off;
  const uint8_t* srank = p.s_rank + off;
  constexpr int kCombos = kA * kM * kT;
  for (int combo = warp; combo < kCombos; combo += kClassThreads / 32) {
    const int a = combo / (kM * kT), mi = (combo / kT) % kM, t = combo % kT;
    const int maxdet = mi == 0 ? 1 : (mi == 1 ? 10 : 100);
    const int npig = p.npig[c * kA + a];
    const size_t rec_ix = (((size_t)t * C + c) * kA + a) * kM + mi;
    auto prec_ix = [&](int r) { return ((((size_t)t * kR + r) * C + c) * kA + a) * kM + mi; };
    if (npig == 0) {  // accumulate(): `if npig == 0: continue` -> the -1 initialisation stays
      for (int r = lane; r < kR; r += 32) p.precision[prec_ix(r)] = -1.0;
      if (lane == 0) p.recall[rec_ix] = -1.0;
      continue;
    }
    unsigned long long* bmax = s_bmax[warp];
    for (int r = lane; r < kR; r += 32) bmax[r] = 0ull;
    __syncwarp();
    const int bit = a * kT + t;
    int base_tp = 0, base_fp = 0;
    for (int e0 = 0; e0 < cnt; e0 += 32) {
      const int e = e0 + lane;
      const bool in = e < cnt && (int)srank[e] < maxdet;
      const bool mm = in && ((sm[e] >> bit) & 1ull);
      const bool ig = in && ((sig[e] >> bit) & 1ull);
      const unsigned tpm = __ballot_sync(0xffffffffu, in && mm && !ig);
      const unsigned fpm = __ballot_sync(0xffffffffu, in && !mm && !ig);
      if ((tpm >> lane) & 1u) {
        const unsigned le = 0xffffffffu >> (31 - lane);
        const int tp = base_tp + __popc(tpm & le);
        const int fp = base_fp + __popc(fpm & le);
        const double pr = (double)tp / (((double)fp + (double)tp) + 2.220446049250313e-16);  // tp / (fp + tp + np.spacing(1))
        const double rc = (double)tp / (double)npig;
        int lo = 0, hi = kR;  // number of recall thresholds <= rc (>= 1: recThrs[0] = 0)
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (s_thr[mid] <= rc) lo = mid + 1; else hi = mid;
        }
        atomicMax(&bmax[lo - 1], (unsigned long long)__double_as_longlong(pr));  // pr >= 0: bit order == value order
      }
      base_tp += __popc(tpm);
      base_fp += __popc(fpm);
    }
    __syncwarp();
    if (lane == 0) {
      // q[r] = max precision over positions whose recall reaches recThrs[r] = suffix maximum over the buckets
      unsigned long long run = 0ull;
      for (int r = kR - 1; r >= 0; --r) {
        run = bmax[r] > run ? bmax[r] : run;
        bmax[r] = run;
      }
      p.recall[rec_ix] = (double)base_tp / (double)npig;  // rc[-1]; 0 when there are no detections
    }
    __syncwarp();
    for (int r = lane; r < kR; r += 32) p.precision[prec_ix(r)] = __longlong_as_double((long long)bmax[r]);
    __syncwarp();
  }
}

}  // namespace

extern "C" size_t odk_coco_ap_workspace_bytes(int n_img, int top_k, int num_classes) {
  if (n_img < 1 || top_k < 1 || num_classes < 1) return 0;
  return coco_layout(nullptr, n_img, top_k, num_classes).bytes;
}

extern "C" int odk_coco_ap(int n_img, int top_k, int num_classes, const float* dets, float conf_thr, const int32_t* img_wh,
                           const float* gt_boxes, const int32_t* gt_classes, const double* gt_areas, const uint8_t* gt_crowd,
                           const int32_t* gt_offsets, int gt_max_per_image, const double* iou_thrs_host, const double* rec_thrs_host,
                           const double* area_ranges_host, double* precision, double* recall, void* workspace, size_t workspace_bytes,
                           void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(n_img >= 1 && top_k >= 1 && num_classes >= 1 && gt_max_per_image >= 0, ODK_ERR_INVALID_ARGUMENT,
              "coco_ap: bad sizes n_img=%d top_k=%d classes=%d", n_img, top_k, num_classes);
  ODK_REQUIRE(dets && img_wh && gt_offsets && precision && recall && iou_thrs_host && rec_thrs_host && area_ranges_host,
              ODK_ERR_INVALID_ARGUMENT, "coco_ap: null pointer");
  ODK_REQUIRE(gt_max_per_image == 0 || (gt_boxes && gt_classes && gt_areas && gt_crowd), ODK_ERR_INVALID_ARGUMENT, "coco_ap: null GT pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "coco_ap: gt_boxes must be 16-byte aligned");
  ODK_REQUIRE(gt_max_per_image <= 64 * kGtWords, ODK_ERR_UNSUPPORTED, "coco_ap: %d GT boxes in one image > %d", gt_max_per_image, 64 * kGtWords);
  ODK_REQUIRE(top_k <= 2048, ODK_ERR_UNSUPPORTED, "coco_ap: top_k=%d > 2048", top_k);
  ODK_REQUIRE((long long)n_img * top_k < (1ll << 31), ODK_ERR_UNSUPPORTED, "coco_ap: n_img*top_k exceeds 2^31");
  const CocoWorkspace w = coco_layout(workspace, n_img, top_k, num_classes);
  ODK_REQUIRE(workspace != nullptr && workspace_bytes >= w.bytes, ODK_ERR_WORKSPACE, "coco_ap: workspace of %zu bytes < required %zu",
              workspace_bytes, w.bytes);
  ODK_REQUIRE(((uintptr_t)workspace & 255) == 0, ODK_ERR_INVALID_ARGUMENT, "coco_ap: workspace must be 256-byte aligned");
  const int gcap = gt_max_per_image < 1 ? 1 : gt_max_per_image;
  const size_t match_smem = (size_t)top_k * (16 + 8 + 8 + 4 + 4 + 4 + 4 + 4 + 4) + 4 + (size_t)gcap * (16 + 8 + 4 + 1) + 64;
  ODK_REQUIRE(match_smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "coco_ap: top_k=%d / gt_max=%d need too much shared memory", top_k,
              gt_max_per_image);
  ODK_CUDA(cudaMemsetAsync(workspace, 0, w.head_bytes, stream));
  const long long rows = (long long)n_img * top_k;
  {
    OdkProfileScope prof_("coco_count_kernel", stream);
    coco_count_kernel<<<(unsigned)((rows + 255) / 256), 256, sizeof(int) * (size_t)num_classes, stream>>>(dets, rows, conf_thr, num_classes,
                                                                                                        w.cls_count);
  }
  ODK_LAUNCH_CHECK("coco_count_kernel");
  coco_layout_kernel<<<1, 32, 0, stream>>>(num_classes, w.cls_count, w.cls_off);
  ODK_LAUNCH_CHECK("coco_layout_kernel");
  CocoMatchArgs m;
  m.n_img = n_img; m.top_k = top_k; m.num_classes = num_classes; m.gt_cap = gcap;
  m.dets = dets; m.conf_thr = conf_thr; m.img_wh = img_wh;
  m.gt_boxes = reinterpret_cast<const float4*>(gt_boxes); m.gt_classes = gt_classes; m.gt_areas = gt_areas; m.gt_crowd = gt_crowd;
  m.gt_offsets = gt_offsets; m.cls_off = w.cls_off; m.cls_cursor = w.cls_cursor; m.npig = w.npig; m.keys = w.keys;
  m.mbits = w.mbits; m.igbits = w.igbits; m.rank = w.rank;
  for (int t = 0; t < kT; ++t) m.iou_thr[t] = iou_thrs_host[t];
  for (int a = 0; a < kA; ++a) { m.area_lo[a] = area_ranges_host[2 * a]; m.area_hi[a] = area_ranges_host[2 * a + 1]; }
  ODK_CUDA(cudaFuncSetAttribute(coco_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {
    OdkProfileScope prof_("coco_match_kernel", stream);
    coco_match_kernel<<<n_img, kMatchThreads, match_smem, stream>>>(m);
  }
  ODK_LAUNCH_CHECK("coco_match_kernel");
  CocoClassArgs k;
  k.num_classes = num_classes; k.cls_cursor = w.cls_cursor; k.cls_off = w.cls_off; k.npig = w.npig; k.keys = w.keys;
  k.mbits = w.mbits; k.igbits = w.igbits; k.rank = w.rank; k.s_m = w.s_m; k.s_ig = w.s_ig; k.s_rank = w.s_rank;
  k.precision = precision; k.recall = recall;
  for (int r = 0; r < kR; ++r) k.rec_thr[r] = rec_thrs_host[r];
  ODK_CUDA(cudaFuncSetAttribute(coco_class_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kClassChunk * 8));
  {
    OdkProfileScope prof_("coco_class_kernel", stream);
    coco_class_kernel<<<num_classes, kClassThreads, kClassChunk * 8, stream>>>(k);
  }
  ODK_LAUNCH_CHECK("coco_class_kernel");
  return ODK_OK;
}
