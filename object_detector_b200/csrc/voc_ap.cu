// K5 — PASCAL VOC AP (data/voc.py:165-197 -> chainercv.evaluations.eval_detection_voc).
//
// chainercv is not vendored by the reference and carries no pinned version, so this stage follows the
// published algorithm (calc_detection_voc_prec_rec + calc_detection_voc_ap) as restated in
// oracle/restated.py — PARITY UNPINNED against the real package.
//
//   ap_count_kernel   per detection row: conf >= conf_thr (od.py:265) and class in range -> class histogram
//   ap_layout_kernel  one CTA: class segment offsets (each padded to a power of two for the sort)
//   ap_match_kernel   one WARP per image: boxes -> integer pixels (round-half-even of the fp64 product,
//                     ml.py:62-65,184-186), +1 on the max corners, per-class IoU arg-max against the
//                     image's GT in int64/fp64, greedy matching in score order (difficult -> ignored,
//                     first hit -> TP, repeat -> FP); emits one 64-bit key per detection
//                         [flip(score):32][image*K + rank in (image,class):30][match+1:2]
//                     into its class segment and accumulates n_pos.
//   ap_class_kernel   one CTA per class: bitonic sort (score desc; ties = argsort(stable)[::-1] on the
//                     concatenated list), chunked block scans for cumsum(TP)/cumsum(FP), 11-point maxima,
//                     right-to-left precision envelope and the area sum.
#include "odk_common.cuh"
#include "odk_sort.cuh"

namespace {

constexpr int kMatchWarps = 4;
constexpr int kClassThreads = 1024;
constexpr int kClassChunk = 16384;

struct ApWorkspace {
  int* cls_count;   // [C]
  int* cls_off;     // [C+1]  (padded segments)
  int* cls_cursor;  // [C]
  int* n_pos;       // [C]
  unsigned long long* keys;  // [2*n_img*K + 2*C]
  size_t bytes;
};

ApWorkspace ap_layout(void* base, int n_img, int top_k, int num_classes) {
  ApWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = odk_align_up(off + bytes, 256); return o; };
  const size_t o0 = take(sizeof(int) * (size_t)num_classes);
  const size_t o1 = take(sizeof(int) * ((size_t)num_classes + 1));
  const size_t o2 = take(sizeof(int) * (size_t)num_classes);
  const size_t o3 = take(sizeof(int) * (size_t)num_classes);
  const size_t o4 = take(8 * (2 * (size_t)n_img * top_k + 2 * (size_t)num_classes));
  char* p = static_cast<char*>(base);
  w.cls_count = reinterpret_cast<int*>(p + o0);
  w.cls_off = reinterpret_cast<int*>(p + o1);
  w.cls_cursor = reinterpret_cast<int*>(p + o2);
  w.n_pos = reinterpret_cast<int*>(p + o3);
  w.keys = reinterpret_cast<unsigned long long*>(p + o4);
  w.bytes = off;
  return w;
}

__device__ __forceinline__ bool det_valid(const float* row, float conf_thr, int num_classes, int& cls) {
  cls = (int)row[0];  // p[:, 0].astype(np.int32), od.py:261
  return row[1] >= conf_thr && row[0] >= 0.0f && cls < num_classes;
}

__global__ void __launch_bounds__(256) ap_count_kernel(const float* __restrict__ dets, long long rows, float conf_thr, int num_classes,
                                                       int* __restrict__ cls_count) {
  extern __shared__ int s_hist[];  // block-private class histogram, flushed with one atomic per class
  for (int c = threadIdx.x; c < num_classes; c += blockDim.x) s_hist[c] = 0;
  __syncthreads();
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int cls;
  if (r < rows && det_valid(dets + r * 6, conf_thr, num_classes, cls)) atomicAdd(s_hist + cls, 1);
  __syncthreads();
  for (int c = threadIdx.x; c < num_classes; c += blockDim.x)
    if (s_hist[c]) atomicAdd(cls_count + c, s_hist[c]);
}

__global__ void ap_layout_kernel(int num_classes, const int* __restrict__ cls_count, int* __restrict__ cls_off) {
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

__device__ __forceinline__ int to_pixel(float v, int scale) { return (int)rint((double)v * (double)scale); }

struct MatchArgs {
  int n_img, top_k, num_classes, gt_cap;
  const float* dets;
  float conf_thr;
  const int32_t* img_wh;
  const float4* gt_boxes;
  const int32_t* gt_classes;
  const uint8_t* gt_difficult;
  const int32_t* gt_offsets;
  const int* cls_off;
  int* cls_cursor;
  int* n_pos;
  unsigned long long* keys;
  int warp_smem;
};

__global__ void __launch_bounds__(kMatchWarps * 32) ap_match_kernel(const MatchArgs p) {
  extern __shared__ __align__(16) unsigned char match_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int img = blockIdx.x * kMatchWarps + warp;
  if (img >= p.n_img) return;
  const int K = p.top_k;
  unsigned char* base = match_smem + (size_t)warp * p.warp_smem;
  int4* dbox = reinterpret_cast<int4*>(base);                                  // [K] pixel boxes (+1 applied)
  float* dconf = reinterpret_cast<float*>(base + (size_t)K * 16);             // [K]
  int* dcls = reinterpret_cast<int*>(base + (size_t)K * 20);                  // [K] class or -1
  int* dgt = reinterpret_cast<int*>(base + (size_t)K * 24);                   // [K] matched GT (image-local) or -1
  int* dslot = reinterpret_cast<int*>(base + (size_t)K * 28);                 // [K] position in (class, rank) order
  int* sorted = reinterpret_cast<int*>(base + (size_t)K * 32);                // [K]
  int* drank = reinterpret_cast<int*>(base + (size_t)K * 36);                 // [K]
  int* dsame = reinterpret_cast<int*>(base + (size_t)K * 40);                 // [K] detections of the same class in this image
  signed char* dmatch = reinterpret_cast<signed char*>(base + (size_t)K * 44);  // [K]
  unsigned char* used = base + (size_t)K * 45;                                // [gt_cap]

  const int w = p.img_wh[2 * img], h = p.img_wh[2 * img + 1];
  const int g0 = p.gt_offsets[img], g1 = p.gt_offsets[img + 1];
  const int ng = min(g1 - g0, p.gt_cap);
  const float* rows = p.dets + (size_t)img * K * 6;

  for (int d = lane; d < K; d += 32) {
    const float* r = rows + d * 6;
    int cls;
    const bool ok = det_valid(r, p.conf_thr, p.num_classes, cls);
    dcls[d] = ok ? cls : -1;
    dconf[d] = r[1];
    dbox[d] = make_int4(to_pixel(r[2], w), to_pixel(r[3], h), to_pixel(r[4], w) + 1, to_pixel(r[5], h) + 1);
  }
  for (int g = lane; g < ng; g += 32) {
    used[g] = 0;
    if (!p.gt_difficult[g0 + g]) {
      const int c = p.gt_classes[g0 + g];
      if (c >= 0 && c < p.num_classes) atomicAdd(p.n_pos + c, 1);
    }
  }
  __syncwarp();
  int n_valid = 0;
  for (int d = lane; d < K; d += 32) {
    const int cls = dcls[d];
    int gt_best = -1, rank = 0, before = 0, same = 0;
    if (cls >= 0) {
      const float conf = dconf[d];
      for (int e = 0; e < K; ++e) {
        const int ce = dcls[e];
        if (ce < 0) continue;
        before += ce < cls;
        if (ce == cls) {
          ++same;
          const float se = dconf[e];
          rank += (se > conf) || (se == conf && e > d);  // argsort(stable)[::-1] within the image
        }
      }
      // chainercv bbox_iou on integer boxes: int64 areas, fp64 quotient; arg-max = first maximum
      const int4 b = dbox[d];
      const long long area_d = (long long)(b.z - b.x) * (long long)(b.w - b.y);
      double best = -1.0;
      for (int g = 0; g < ng; ++g) {
        if (p.gt_classes[g0 + g] != cls) continue;
        const float4 gb = p.gt_boxes[g0 + g];
        const int gx1 = to_pixel(gb.x, w), gy1 = to_pixel(gb.y, h), gx2 = to_pixel(gb.z, w) + 1, gy2 = to_pixel(gb.w, h) + 1;
        const int lx = max(b.x, gx1), ly = max(b.y, gy1), rx = min(b.z, gx2), ry = min(b.w, gy2);
        const long long inter = (lx < rx && ly < ry) ? (long long)(rx - lx) * (long long)(ry - ly) : 0ll;
        const long long area_g = (long long)(gx2 - gx1) * (long long)(gy2 - gy1);
        const double iou = (double)inter / (double)(area_d + area_g - inter);
        if (iou > best) { best = iou; gt_best = g; }
      }
      if (best < 0.5) gt_best = -1;  // gt_index[iou.max(axis=1) < iou_thresh] = -1
      n_valid += 1;
    }
    dgt[d] = gt_best;
    drank[d] = rank;
    dslot[d] = (cls >= 0) ? before + rank : -1;
    dsame[d] = same;
  }
  n_valid = __reduce_add_sync(0xffffffffu, n_valid);
  __syncwarp();
  for (int d = lane; d < K; d += 32)
    if (dslot[d] >= 0) sorted[dslot[d]] = d;
  __syncwarp();
  // greedy matching in (class, score) order; GT indices are distinct across classes, one `used` array serves all
  if (lane == 0) {
    for (int i = 0; i < n_valid; ++i) {
      const int d = sorted[i];
      const int g = dgt[d];
      signed char m = 0;
      if (g >= 0) {
        if (p.gt_difficult[g0 + g]) m = -1;
        else m = used[g] ? 0 : 1;
        used[g] = 1;
      }
      dmatch[d] = m;
    }
  }
  __syncwarp();
  // one reservation per (image, class): the rank-0 detection claims `same` slots of the class segment
  int* cbase = sorted;  // reuse: indexed by the class' first slot in (class, rank) order
  for (int d = lane; d < K; d += 32)
    if (dcls[d] >= 0 && drank[d] == 0) cbase[dslot[d]] = atomicAdd(p.cls_cursor + dcls[d], dsame[d]);
  __syncwarp();
  for (int d = lane; d < K; d += 32) {
    const int cls = dcls[d];
    if (cls < 0) continue;
    const int pos = cbase[dslot[d] - drank[d]] + drank[d];
    const unsigned long long key = ((unsigned long long)odk_flip_score(dconf[d]) << 32) |
                                   ((unsigned long long)((unsigned)(img * K + drank[d]) & 0x3fffffffu) << 2) |
                                   (unsigned long long)(dmatch[d] + 1);
    p.keys[(size_t)p.cls_off[cls] + pos] = key;
  }
}

// ---- per-class curve ---------------------------------------------------------------------------------------
__device__ __forceinline__ int2 block_exclusive_scan_int2(int2 v, int2* s_warp, int2& total) {
  const int lane = (int)odk_lane(), warp = threadIdx.x >> 5;
  int2 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int ox = __shfl_up_sync(0xffffffffu, inc.x, d), oy = __shfl_up_sync(0xffffffffu, inc.y, d);
    if (lane >= d) { inc.x += ox; inc.y += oy; }
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int2 wv = (lane < (int)(blockDim.x >> 5)) ? s_warp[lane] : make_int2(0, 0);
    int2 winc = wv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int ox = __shfl_up_sync(0xffffffffu, winc.x, d), oy = __shfl_up_sync(0xffffffffu, winc.y, d);
      if (lane >= d) { winc.x += ox; winc.y += oy; }
    }
    s_warp[lane] = make_int2(winc.x - wv.x, winc.y - wv.y);
    if (lane == 31) s_warp[32] = winc;
  }
  __syncthreads();
  const int2 wbase = s_warp[warp];
  total = s_warp[32];
  __syncthreads();
  return make_int2(wbase.x + inc.x - v.x, wbase.y + inc.y - v.y);
}

__global__ void __launch_bounds__(kClassThreads) ap_class_kernel(int num_classes, const int* __restrict__ cls_count,
                                                                const int* __restrict__ cls_off, const int* __restrict__ n_pos,
                                                                unsigned long long* __restrict__ keys, double* __restrict__ ap_area,
                                                                double* __restrict__ ap07) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ int2 s_scan[33];
  __shared__ double s_max[11][32];
  __shared__ double s_wmax[33];
  __shared__ double s_sum[32];
  const int c = blockIdx.x;
  const int cnt = cls_count[c];
  const int npos = n_pos[c];
  unsigned long long* g = keys + cls_off[c];
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (npos == 0) {  // no non-difficult GT of this class: recall undefined -> AP = nan (chainercv rec[l] = None)
    if (threadIdx.x == 0) { ap_area[c] = nan; ap07[c] = nan; }
    return;
  }
  odk_block_sort_desc(g, cnt, s_keys, kClassChunk);

  const int tid = threadIdx.x, lane = (int)odk_lane(), warp = tid >> 5;
  const int per = (cnt + kClassThreads - 1) / kClassThreads;
  const int e0 = min(cnt, tid * per), e1 = min(cnt, e0 + per);
  int2 loc = make_int2(0, 0);
  for (int e = e0; e < e1; ++e) {
    const int m = (int)(g[e] & 3ull) - 1;
    loc.x += m == 1;
    loc.y += m == 0;
  }
  int2 total;
  const int2 basev = block_exclusive_scan_int2(loc, s_scan, total);
  // forward walk: precision / recall, 11-point maxima and this thread's chunk maximum
  double pmax[11];
#pragma unroll
  for (int t = 0; t < 11; ++t) pmax[t] = -1.0;  // -1 = "no element with rec >= t"
  double chunk_max = 0.0;
  {
    int tp = basev.x, fp = basev.y;
    for (int e = e0; e < e1; ++e) {
      const int m = (int)(g[e] & 3ull) - 1;
      tp += m == 1;
      fp += m == 0;
      const double prec = (tp + fp) > 0 ? (double)tp / (double)(fp + tp) : 0.0;  // nan_to_num
      const double rec = (double)tp / (double)npos;
      chunk_max = fmax(chunk_max, prec);
#pragma unroll
      for (int t = 0; t < 11; ++t)
        if (rec >= (double)t * 0.1) pmax[t] = fmax(pmax[t], prec);  // np.arange(0., 1.1, 0.1)[t] == t*0.1
    }
  }
#pragma unroll
  for (int t = 0; t < 11; ++t) {
    double v = pmax[t];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, d));
    if (lane == 0) s_max[t][warp] = v;
  }
  // suffix maximum of the chunk maxima (exclusive, from the right)
  double incl = chunk_max;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const double o = __shfl_down_sync(0xffffffffu, incl, d);
    if (lane + d < 32) incl = fmax(incl, o);
  }
  if (lane == 0) s_wmax[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    double wv = s_wmax[lane];
    double winc = wv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const double o = __shfl_down_sync(0xffffffffu, winc, d);
      if (lane + d < 32) winc = fmax(winc, o);
    }
    const double excl = __shfl_down_sync(0xffffffffu, winc, 1);
    s_wmax[lane] = (lane == 31) ? 0.0 : excl;  // trailing sentinel mpre[-1] = 0
  }
  __syncthreads();
  double after = __shfl_down_sync(0xffffffffu, incl, 1);  // max over later lanes of this warp
  after = (lane == 31) ? 0.0 : after;
  double run = fmax(after, s_wmax[warp]);
  // backward walk: envelope * recall steps (recall changes exactly at the true positives)
  double acc = 0.0;
  {
    int tp = basev.x + loc.x, fp = basev.y + loc.y;
    for (int e = e1 - 1; e >= e0; --e) {
      const int m = (int)(g[e] & 3ull) - 1;
      const double prec = (tp + fp) > 0 ? (double)tp / (double)(fp + tp) : 0.0;
      run = fmax(run, prec);
      if (m == 1) acc += ((double)tp / (double)npos - (double)(tp - 1) / (double)npos) * run;
      tp -= m == 1;
      fp -= m == 0;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
  if (lane == 0) s_sum[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    double a = 0.0;
    for (int wv = 0; wv < kClassThreads / 32; ++wv) a += s_sum[wv];
    ap_area[c] = a;
    double a07 = 0.0;
    for (int t = 0; t < 11; ++t) {
      double v = -1.0;
      for (int wv = 0; wv < kClassThreads / 32; ++wv) v = fmax(v, s_max[t][wv]);
      a07 += (v < 0.0 ? 0.0 : v) / 11;  // ap[l] += p / 11
    }
    ap07[c] = a07;
  }
}

}  // namespace

extern "C" size_t odk_voc_ap_workspace_bytes(int n_img, int top_k, int num_classes, int gt_max_per_image) {
  (void)gt_max_per_image;
  if (n_img < 1 || top_k < 1 || num_classes < 1) return 0;
  return ap_layout(nullptr, n_img, top_k, num_classes).bytes;
}

extern "C" int odk_voc_ap(int n_img, int top_k, int num_classes, const float* dets, float conf_thr, const int32_t* img_wh,
                          const float* gt_boxes, const int32_t* gt_classes, const uint8_t* gt_difficult, const int32_t* gt_offsets,
                          int gt_max_per_image, double* ap_area, double* ap07, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(n_img >= 1 && top_k >= 1 && num_classes >= 1 && gt_max_per_image >= 0, ODK_ERR_INVALID_ARGUMENT,
              "voc_ap: bad sizes n_img=%d top_k=%d classes=%d", n_img, top_k, num_classes);
  ODK_REQUIRE(dets && img_wh && gt_offsets && ap_area && ap07, ODK_ERR_INVALID_ARGUMENT, "voc_ap: null pointer");
  ODK_REQUIRE(gt_max_per_image == 0 || (gt_boxes && gt_classes && gt_difficult), ODK_ERR_INVALID_ARGUMENT, "voc_ap: null GT pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "voc_ap: gt_boxes must be 16-byte aligned");
  ODK_REQUIRE((long long)n_img * top_k < (1ll << 30), ODK_ERR_UNSUPPORTED, "voc_ap: n_img*top_k exceeds 2^30");
  const ApWorkspace w = ap_layout(workspace, n_img, top_k, num_classes);
  ODK_REQUIRE(workspace != nullptr && workspace_bytes >= w.bytes, ODK_ERR_WORKSPACE, "voc_ap: workspace of %zu bytes < required %zu",
              workspace_bytes, w.bytes);
  ODK_REQUIRE(((uintptr_t)workspace & 255) == 0, ODK_ERR_INVALID_ARGUMENT, "voc_ap: workspace must be 256-byte aligned");
  const size_t warp_smem = odk_align_up((size_t)top_k * 45 + (size_t)gt_max_per_image, 16);
  ODK_REQUIRE(warp_smem * kMatchWarps <= 200 * 1024, ODK_ERR_UNSUPPORTED, "voc_ap: top_k=%d / gt_max=%d need too much shared memory", top_k,
              gt_max_per_image);
  // the four int arrays are contiguous-ish; clear counts, cursors and n_pos
  ODK_CUDA(cudaMemsetAsync(w.cls_count, 0, (size_t)((char*)w.keys - (char*)w.cls_count), stream));
  const long long rows = (long long)n_img * top_k;
  {
    OdkProfileScope prof_("ap_count_kernel", stream);
    ap_count_kernel<<<(unsigned)((rows + 255) / 256), 256, sizeof(int) * (size_t)num_classes, stream>>>(dets, rows, conf_thr, num_classes, w.cls_count);
  }
  ODK_LAUNCH_CHECK("ap_count_kernel");
  ap_layout_kernel<<<1, 32, 0, stream>>>(num_classes, w.cls_count, w.cls_off);
  ODK_LAUNCH_CHECK("ap_layout_kernel");
  MatchArgs m;
  m.n_img = n_img; m.top_k = top_k; m.num_classes = num_classes; m.gt_cap = gt_max_per_image;
  m.dets = dets; m.conf_thr = conf_thr; m.img_wh = img_wh;
  m.gt_boxes = reinterpret_cast<const float4*>(gt_boxes); m.gt_classes = gt_classes; m.gt_difficult = gt_difficult; m.gt_offsets = gt_offsets;
  m.cls_off = w.cls_off; m.cls_cursor = w.cls_cursor; m.n_pos = w.n_pos; m.keys = w.keys; m.warp_smem = (int)warp_smem;
  ODK_CUDA(cudaFuncSetAttribute(ap_match_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {
    OdkProfileScope prof_("ap_match_kernel", stream);
    ap_match_kernel<<<(n_img + kMatchWarps - 1) / kMatchWarps, kMatchWarps * 32, warp_smem * kMatchWarps, stream>>>(m);
  }
  ODK_LAUNCH_CHECK("ap_match_kernel");
  ODK_CUDA(cudaFuncSetAttribute(ap_class_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kClassChunk * 8));
  {
    OdkProfileScope prof_("ap_class_kernel", stream);
    ap_class_kernel<<<num_classes, kClassThreads, kClassChunk * 8, stream>>>(num_classes, w.cls_count, w.cls_off, w.n_pos, w.keys, ap_area, ap07);
  }
  ODK_LAUNCH_CHECK("ap_class_kernel");
  return ODK_OK;
}
