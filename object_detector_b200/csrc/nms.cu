// K2 — per-class NMS + cross-class top-k (layers.py:925-970; inner greedy NMS ml.py:589-625).
//
// Pipeline on the candidate list produced by K1 (or by collect_candidates_kernel for the unfused
// layer entry):
//   1. sort_candidates_kernel  one CTA per image: bitonic sort of the 64-bit keys
//        key1 = [4095-class : 12][flip(score) : 32][anchor : 20]   (descending)
//      => class ascending, score descending, index descending (= argsort(stable)[::-1], SURVEY A-6);
//      in shared memory up to 16384 keys, hybrid shared/global above that.  Emits class segment
//      offsets by binary search.
//   2. nms_sweep_kernel        one WARP per (image, class) segment: candidates are taken 32 at a
//      time in score order; each lane tests its box against the kept list (shared memory), then the
//      32x32 intra-chunk suppression bitmask is built with warp shuffles and resolved by a
//      warp-uniform sweep (ballot-style), stopping at the reference's length limit.
//   3. topk_merge_kernel       one CTA per image: the <= C*L kept keys, re-keyed as
//        key2 = [flip(score) : 32][4095-class : 12][anchor : 20]
//      are sorted descending (score desc; lower class, then per-class NMS order on ties — the order
//      tf.nn.top_k(sorted=True) yields on the class-concatenated list) and the first top_k rows are
//      written as [class, conf, x1, y1, x2, y2], zero padded (layers.py:963-970).
// All IoU tests are the reference's fp32 expression, evaluated exactly (odk_iou_passes).
#include "odk_common.cuh"
#include "odk_sort.cuh"

namespace {

constexpr int kSortThreads = 1024;
constexpr int kSortChunk = 16384;  // keys held in shared memory (128 KB)
constexpr int kSweepWarps = 4;

// ---- candidate collection for the unfused layer entry (layers.py:935) ------------------------------------
__global__ void __launch_bounds__(256) collect_candidates_kernel(int n, int num_classes, const int32_t* __restrict__ classes,
                                                                 const float* __restrict__ confs, float conf_thr, int take_all,
                                                                 unsigned long long* __restrict__ cand_keys,
                                                                 int* __restrict__ cand_count, int cand_cap) {
  const int b = blockIdx.y;
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  bool is_cand = false;
  int cls = 0;
  float conf = 0.f;
  if (a < n) {
    const size_t r = (size_t)b * n + a;
    cls = classes ? classes[r] : 0;
    conf = confs[r];
    is_cand = (take_all || conf > conf_thr) && cls >= 0 && cls < num_classes;
  }
  const unsigned m = __ballot_sync(0xffffffffu, is_cand);
  if (m) {
    const unsigned lane = odk_lane();
    const int leader = __ffs(m) - 1;
    int base = 0;
    if ((int)lane == leader) base = atomicAdd(cand_count + b, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (is_cand) {
      const int pos = base + __popc(m & ((1u << lane) - 1u));
      if (pos < cand_cap) cand_keys[(size_t)b * cand_cap + pos] = odk_make_key1(cls, conf, a);
    }
  }
}

__global__ void __launch_bounds__(kSortThreads) sort_candidates_kernel(unsigned long long* __restrict__ cand_keys,
                                                                      const int* __restrict__ cand_count, int cand_cap,
                                                                      int num_classes, int* __restrict__ seg_off) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  const int b = blockIdx.x;
  unsigned long long* g = cand_keys + (size_t)b * cand_cap;
  const int cnt = min(cand_count[b], cand_cap);

  odk_block_sort_desc(g, cnt, s_keys, kSortChunk);  // in shared memory up to kSortChunk keys, hybrid above
  __syncthreads();
  // class segment offsets: seg_off[b, c] = first sorted position whose class >= c
  for (int c = threadIdx.x; c <= num_classes; c += blockDim.x) {
    int lo = 0, hi = cnt;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (odk_key1_class(g[mid]) >= c) hi = mid; else lo = mid + 1;
    }
    seg_off[(size_t)b * (num_classes + 1) + c] = lo;
  }
}

// ---- greedy sweep, one CTA (4 cooperating warps) per (image, class) ------------------------------------
// Candidates are consumed 32 at a time in score order.  All four warps hold the same 32 candidates
// (lane = candidate); the kept list and the intra-chunk pair tests are split 4 ways across the warps
// and the partial suppression masks are OR-ed through shared memory.
//
// Pair test.  IoU >= thr  <=>  inter >= k*(area_a+area_b) with k = thr/(1+thr).  The hot loops use
// that division-free form with a +-3e-5 relative guard band (two FFMAs against per-lane constants);
// a pair that lands inside the band (or any pair when the fast form does not apply) is re-evaluated
// with the reference's exact fp32 expression (odk_iou_passes), so the decision is always the
// reference's.
struct FastThr {
  float k_hi, k_lo;
  int enabled;
};

__device__ __forceinline__ void fast_pair(const float4 a, float a_hi, float a_lo, const float4 b, float area_b, const FastThr ft,
                                          bool& yes, bool& maybe) {
  const float w = fmaxf(fminf(a.z, b.z) - fmaxf(a.x, b.x), 0.0f);
  const float h = fmaxf(fminf(a.w, b.w) - fmaxf(a.y, b.y), 0.0f);
  const float inter = w * h;
  yes = inter > fmaf(area_b, ft.k_hi, a_hi);
  maybe = (inter >= fmaf(area_b, ft.k_lo, a_lo)) && !yes;
}

template <bool GE>
__global__ void __launch_bounds__(kSweepWarps * 32) nms_sweep_kernel(const unsigned long long* __restrict__ cand_keys, int cand_cap,
                                                                    const int* __restrict__ seg_off, int num_classes, int n,
                                                                    const float4* __restrict__ boxes, float iou_thr, FastThr ft, int limit,
                                                                    int stride, int skip_empty, unsigned long long* __restrict__ kept_keys,
                                                                    int* __restrict__ kept_count) {
  extern __shared__ __align__(16) unsigned char sweep_smem[];
  __shared__ unsigned s_dead[kSweepWarps];
  __shared__ unsigned s_sup[32];
  __shared__ int s_cnt;
  const int warp = threadIdx.x >> 5;
  const unsigned lane = odk_lane();
  const int c = blockIdx.x;
  const int b = blockIdx.y;
  float4* kbox = reinterpret_cast<float4*>(sweep_smem);
  float* karea = reinterpret_cast<float*>(sweep_smem + (size_t)stride * 16);
  const bool fast = GE && ft.enabled;  // the fast form needs positive-extent boxes: guaranteed by the zero-area skip

  const int s = seg_off[(size_t)b * (num_classes + 1) + c];
  const int e = seg_off[(size_t)b * (num_classes + 1) + c + 1];
  const unsigned long long* keys = cand_keys + (size_t)b * cand_cap;
  const float4* bx = boxes + (size_t)b * n;
  unsigned long long* out = kept_keys + ((size_t)b * num_classes + c) * stride;
  int nkept = 0;  // identical in every thread of the CTA
  if (threadIdx.x < 32) s_sup[threadIdx.x] = 0u;
  __syncthreads();

  for (int base = s; base < e && nkept < limit; base += 32) {
    const int i = base + (int)lane;
    const bool have = i < e;
    const unsigned long long key = have ? keys[i] : 0ull;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    if (have) box = bx[(int)(key & ODK_IDX_MASK)];
    const float w = __fsub_rn(box.z, box.x), h = __fsub_rn(box.w, box.y);
    const float area = __fmul_rn(w, h);
    const bool alive0 = have && !(skip_empty && (w <= 0.0f || h <= 0.0f));  // ml.py:610-611
    const float a_hi = area * ft.k_hi, a_lo = area * ft.k_lo;
    // (1) this warp's quarter of the kept list (uniform trip count, no divergent exits)
    bool hit_kept = false;
    if (fast) {
      bool unsure = false;
#pragma unroll 4
      for (int k = warp; k < nkept; k += kSweepWarps) {
        bool yes, maybe;
        fast_pair(box, a_hi, a_lo, kbox[k], karea[k], ft, yes, maybe);
        hit_kept |= yes;
        unsure |= maybe;
      }
      if (__any_sync(0xffffffffu, unsure && !hit_kept && alive0)) {  // rare: some pair sits inside the guard band
        for (int k = warp; k < nkept; k += kSweepWarps) hit_kept |= odk_iou_passes<GE>(box, area, kbox[k], karea[k], iou_thr);
      }
    } else {
#pragma unroll 2
      for (int k = warp; k < nkept; k += kSweepWarps) hit_kept |= odk_iou_passes<GE>(box, area, kbox[k], karea[k], iou_thr);
    }
    const unsigned dead = __ballot_sync(0xffffffffu, hit_kept);
    if (lane == 0) s_dead[warp] = dead;
    // (2) this warp's quarter of the intra-chunk pairs: sup bit j = "this lane's box suppresses lane j"
    const unsigned am0 = __ballot_sync(0xffffffffu, alive0);
    unsigned sup = 0;
    for (unsigned rem = (am0 & (am0 - 1)) & (0x11111111u << warp); rem; rem &= rem - 1) {  // j = warp (mod 4), j != lowest alive
      const int j = __ffs(rem) - 1;
      float4 bj;
      bj.x = __shfl_sync(0xffffffffu, box.x, j);
      bj.y = __shfl_sync(0xffffffffu, box.y, j);
      bj.z = __shfl_sync(0xffffffffu, box.z, j);
      bj.w = __shfl_sync(0xffffffffu, box.w, j);
      const float aj = __shfl_sync(0xffffffffu, area, j);
      bool sj;
      if (fast) {
        bool maybe;
        fast_pair(box, a_hi, a_lo, bj, aj, ft, sj, maybe);
        if (maybe) sj = odk_iou_passes<GE>(bj, aj, box, area, iou_thr);
      } else {
        sj = odk_iou_passes<GE>(bj, aj, box, area, iou_thr);
      }
      if (alive0 && (int)lane < j && sj) sup |= 1u << j;
    }
    if (sup) atomicOr(&s_sup[lane], sup);
    __syncthreads();
    // (3) combine and resolve: warp 0 walks the alive lanes in score order
    if (warp == 0) {
      unsigned am = am0;
#pragma unroll
      for (int q = 0; q < kSweepWarps; ++q) am &= ~s_dead[q];
      const unsigned sup_all = s_sup[lane];
      unsigned removed = 0, keep = 0;
      int cnt = nkept;
      for (unsigned rem = am; rem; rem &= rem - 1) {
        const int l = __ffs(rem) - 1;
        const unsigned sup_l = __shfl_sync(0xffffffffu, sup_all, l);
        if ((removed >> l) & 1u) continue;
        if (cnt >= limit) break;  // ml.py:622-623
        keep |= 1u << l;
        ++cnt;
        removed |= sup_l;
      }
      if ((keep >> lane) & 1u) {
        const int pos = nkept + __popc(keep & ((1u << lane) - 1u));
        kbox[pos] = box;
        karea[pos] = area;
        out[pos] = odk_key1_to_key2(key);
      }
      s_sup[lane] = 0u;
      if (lane == 0) s_cnt = cnt;
    }
    __syncthreads();  // kept list, count and cleared masks visible
    nkept = s_cnt;
  }
  if (threadIdx.x == 0) kept_count[(size_t)b * num_classes + c] = nkept;
}

// ---- cross-class top-k + output rows -------------------------------------------------------------------
__global__ void __launch_bounds__(256) topk_merge_kernel(const unsigned long long* __restrict__ kept_keys,
                                                        const int* __restrict__ kept_count, int num_classes, int limit, int n,
                                                        const float4* __restrict__ boxes, int top_k, float* __restrict__ out,
                                                        int32_t* __restrict__ kept_index) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ int s_total;
  __shared__ int s_scan[256];
  __shared__ int s_off[ODK_MAX_CLASSES + 1];
  const int b = blockIdx.x;
  const int tid = threadIdx.x;
  // exclusive scan of kept_count over classes, 256 classes per round
  int carry = 0;
  for (int c0 = 0; c0 < num_classes; c0 += 256) {
    const int c = c0 + tid;
    const int v = (c < num_classes) ? kept_count[(size_t)b * num_classes + c] : 0;
    s_scan[tid] = v;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
      const int add = (tid >= d) ? s_scan[tid - d] : 0;
      __syncthreads();
      s_scan[tid] += add;
      __syncthreads();
    }
    if (c < num_classes) s_off[c] = carry + s_scan[tid] - v;
    carry += s_scan[255];
    __syncthreads();
  }
  if (tid == 0) { s_total = carry; s_off[num_classes] = carry; }
  __syncthreads();
  // coalesced gather of the per-class kept lists into one dense array
  for (int i = tid; i < num_classes * limit; i += blockDim.x) {
    const int c = i / limit, k = i - c * limit;
    if (k < s_off[c + 1] - s_off[c]) s_keys[s_off[c] + k] = kept_keys[(size_t)b * num_classes * limit + i];
  }
  const int total = s_total;
  int n_pad = 2;
  while (n_pad < total) n_pad <<= 1;
  for (int i = total + tid; i < n_pad; i += blockDim.x) s_keys[i] = 0ull;
  __syncthreads();
  if (total > 1) bitonic_sort_shared(s_keys, n_pad, 0, 2, n_pad);
  const int rows = min(top_k, total);
  for (int r = tid; r < top_k; r += blockDim.x) {
    float* o = out + ((size_t)b * top_k + r) * 6;
    if (r < rows) {
      const unsigned long long k = s_keys[r];
      const int idx = (int)(k & ODK_IDX_MASK);
      const int cls = ODK_MAX_CLASSES - 1 - (int)((k >> ODK_IDX_BITS) & (ODK_MAX_CLASSES - 1));
      const float4 bx = boxes[(size_t)b * n + idx];
      o[0] = (float)cls; o[1] = odk_unflip_score((uint32_t)(k >> 32));
      o[2] = bx.x; o[3] = bx.y; o[4] = bx.z; o[5] = bx.w;
      kept_index[(size_t)b * top_k + r] = idx;
    } else {
      o[0] = o[1] = o[2] = o[3] = o[4] = o[5] = 0.0f;
      kept_index[(size_t)b * top_k + r] = -1;
    }
  }
}

// ml.non_maximum_suppression result export: kept indices in NMS order.
__global__ void export_kept_kernel(const unsigned long long* __restrict__ kept_keys, const int* __restrict__ kept_count,
                                   int32_t* __restrict__ out_index, int32_t* __restrict__ out_count) {
  const int cnt = kept_count[0];
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) out_index[i] = (int32_t)(kept_keys[i] & ODK_IDX_MASK);
  if (threadIdx.x == 0) *out_count = cnt;
}

}  // namespace

// ---- workspace layout ------------------------------------------------------------------------------------
struct NmsWorkspace {
  int* cand_count;              // [B]
  int* seg_off;                 // [B, C+1]
  int* kept_count;              // [B, C]
  unsigned long long* cand_keys;  // [B, cap]
  unsigned long long* kept_keys;  // [B, C, L]
  float4* boxes;                // [B, N]  (fused mode)
  int cap, limit;
  size_t bytes;
};

static NmsWorkspace nms_layout(void* base, int batch, int n, int num_classes, int top_k) {
  NmsWorkspace w;
  w.cap = odk_next_pow2(n < 2 ? 2 : n);
  w.limit = top_k < 2 ? 2 : top_k;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = odk_align_up(off + bytes, 256); return o; };
  const size_t o_cnt = take(sizeof(int) * (size_t)batch);
  const size_t o_seg = take(sizeof(int) * (size_t)batch * (num_classes + 1));
  const size_t o_kc = take(sizeof(int) * (size_t)batch * num_classes);
  const size_t o_ck = take(8 * (size_t)batch * w.cap);
  const size_t o_kk = take(8 * (size_t)batch * num_classes * w.limit);
  const size_t o_bx = take(16 * (size_t)batch * n);
  char* p = static_cast<char*>(base);
  w.cand_count = reinterpret_cast<int*>(p + o_cnt);
  w.seg_off = reinterpret_cast<int*>(p + o_seg);
  w.kept_count = reinterpret_cast<int*>(p + o_kc);
  w.cand_keys = reinterpret_cast<unsigned long long*>(p + o_ck);
  w.kept_keys = reinterpret_cast<unsigned long long*>(p + o_kk);
  w.boxes = reinterpret_cast<float4*>(p + o_bx);
  w.bytes = off;
  return w;
}

size_t odk_nms_workspace_bytes_impl(int batch, int n, int num_classes, int top_k) {
  return nms_layout(nullptr, batch, n, num_classes, top_k).bytes;
}

static int nms_check(int batch, int n, int num_classes, int top_k, const void* ws, size_t ws_bytes) {
  ODK_REQUIRE(batch >= 1 && n >= 1 && num_classes >= 1 && top_k >= 0, ODK_ERR_INVALID_ARGUMENT,
              "nms: bad sizes batch=%d n=%d classes=%d top_k=%d", batch, n, num_classes, top_k);
  ODK_REQUIRE(n <= (int)ODK_IDX_MASK, ODK_ERR_UNSUPPORTED, "nms: n=%d exceeds the 20-bit anchor field", n);
  ODK_REQUIRE(num_classes <= ODK_MAX_CLASSES, ODK_ERR_UNSUPPORTED, "nms: num_classes=%d > %d", num_classes, ODK_MAX_CLASSES);
  const long long kept_cap = (long long)num_classes * (top_k < 2 ? 2 : top_k);
  ODK_REQUIRE(kept_cap <= kSortChunk, ODK_ERR_UNSUPPORTED,
              "nms: num_classes*max(top_k,2)=%lld exceeds the %d-key shared-memory merge", kept_cap, kSortChunk);
  ODK_REQUIRE((size_t)(top_k < 2 ? 2 : top_k) * 20 <= 200 * 1024, ODK_ERR_UNSUPPORTED, "nms: top_k=%d too large", top_k);
  ODK_REQUIRE(ws != nullptr && ws_bytes >= odk_nms_workspace_bytes_impl(batch, n, num_classes, top_k), ODK_ERR_WORKSPACE,
              "nms: workspace of %zu bytes is smaller than the required %zu", ws_bytes,
              odk_nms_workspace_bytes_impl(batch, n, num_classes, top_k));
  ODK_REQUIRE(((uintptr_t)ws & 255) == 0, ODK_ERR_INVALID_ARGUMENT, "nms: workspace must be 256-byte aligned");
  return ODK_OK;
}

// sort -> sweep -> merge on an already filled candidate list
static int nms_run(const NmsWorkspace& w, int batch, int n, int num_classes, const float4* boxes, float iou_thr, int top_k,
                   int iou_ge, float* out, int32_t* kept_index, cudaStream_t stream) {
  ODK_CUDA(cudaFuncSetAttribute(sort_candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortChunk * 8));
  ODK_CUDA(cudaFuncSetAttribute(topk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortChunk * 8));
  ODK_CUDA(cudaFuncSetAttribute(nms_sweep_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  ODK_CUDA(cudaFuncSetAttribute(nms_sweep_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {
    OdkProfileScope prof_("sort_candidates_kernel", stream);
    sort_candidates_kernel<<<batch, kSortThreads, kSortChunk * 8, stream>>>(w.cand_keys, w.cand_count, w.cap, num_classes, w.seg_off);
  }
  ODK_LAUNCH_CHECK("sort_candidates_kernel");
  const int limit = iou_ge ? w.limit : top_k;
  FastThr ft;
  ft.enabled = (iou_thr > 1e-3f && iou_thr < 0.999f) ? 1 : 0;
  const double kk = ft.enabled ? (double)iou_thr / (1.0 + (double)iou_thr) : 0.0;
  ft.k_hi = (float)(kk * (1.0 + 3e-5));
  ft.k_lo = (float)(kk * (1.0 - 3e-5));
  dim3 grid(num_classes, batch);
  const size_t sweep_smem = (size_t)w.limit * 20;
  if (iou_ge)
    {
      OdkProfileScope prof_("nms_sweep_kernel", stream);
      nms_sweep_kernel<true><<<grid, kSweepWarps * 32, sweep_smem, stream>>>(w.cand_keys, w.cap, w.seg_off, num_classes, n, boxes, iou_thr,
                                                                             ft, limit, w.limit, 1, w.kept_keys, w.kept_count);
    }
  else
    {
      OdkProfileScope prof_("nms_sweep_kernel", stream);
      nms_sweep_kernel<false><<<grid, kSweepWarps * 32, sweep_smem, stream>>>(w.cand_keys, w.cap, w.seg_off, num_classes, n, boxes, iou_thr,
                                                                              ft, limit, w.limit, 0, w.kept_keys, w.kept_count);
    }
  ODK_LAUNCH_CHECK("nms_sweep_kernel");
  if (out != nullptr) {
    int pad = odk_next_pow2(num_classes * w.limit);
    if (pad < 2) pad = 2;
    {
      OdkProfileScope prof_("topk_merge_kernel", stream);
      topk_merge_kernel<<<batch, 256, (size_t)pad * 8, stream>>>(w.kept_keys, w.kept_count, num_classes, w.limit, n, boxes, top_k, out,
                                                                 kept_index);
    }
    ODK_LAUNCH_CHECK("topk_merge_kernel");
  }
  return ODK_OK;
}

int odk_launch_decode_score(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs,
                            int32_t* classes, float* confs, float* boxes, bool dense, float conf_thr,
                            unsigned long long* cand_keys, int* cand_count, int cand_cap, cudaStream_t stream);

extern "C" size_t odk_nms_workspace_bytes(int batch, int n, int num_classes, int top_k) {
  if (batch < 1 || n < 1 || num_classes < 1 || top_k < 0) return 0;
  return odk_nms_workspace_bytes_impl(batch, n, num_classes, top_k);
}

extern "C" int odk_nms_layer(int batch, int n, int num_classes, const int32_t* classes, const float* confs, const float* boxes,
                             float conf_thr, float iou_thr, int top_k, int iou_ge, float* out, int32_t* kept_index, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = nms_check(batch, n, num_classes, top_k, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(classes && confs && boxes && out && kept_index, ODK_ERR_INVALID_ARGUMENT, "nms_layer: null pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "nms_layer: boxes must be 16-byte aligned");
  NmsWorkspace w = nms_layout(workspace, batch, n, num_classes, top_k);
  ODK_CUDA(cudaMemsetAsync(w.cand_count, 0, sizeof(int) * (size_t)batch, stream));
  dim3 grid((n + 255) / 256, batch);
  {
    OdkProfileScope prof_("collect_candidates_kernel", stream);
    collect_candidates_kernel<<<grid, 256, 0, stream>>>(n, num_classes, classes, confs, conf_thr, 0, w.cand_keys, w.cand_count, w.cap);
  }
  ODK_LAUNCH_CHECK("collect_candidates_kernel");
  return nms_run(w, batch, n, num_classes, reinterpret_cast<const float4*>(boxes), iou_thr, top_k, iou_ge, out, kept_index, stream);
}

extern "C" int odk_decode_nms(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs, float conf_thr,
                              float iou_thr, int top_k, int iou_ge, float* out, int32_t* kept_index, void* workspace,
                              size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(a != nullptr, ODK_ERR_INVALID_ARGUMENT, "decode_nms: null anchors handle");
  if (int rc = nms_check(batch, a->n, a->num_classes, top_k, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(objs && clfs && locs && out && kept_index, ODK_ERR_INVALID_ARGUMENT, "decode_nms: null pointer");
  ODK_REQUIRE((((uintptr_t)objs | (uintptr_t)clfs | (uintptr_t)locs) & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "decode_nms: inputs must be 4-byte aligned");
  NmsWorkspace w = nms_layout(workspace, batch, a->n, a->num_classes, top_k);
  ODK_CUDA(cudaMemsetAsync(w.cand_count, 0, sizeof(int) * (size_t)batch, stream));
  if (int rc = odk_launch_decode_score(a, batch, objs, clfs, locs, nullptr, nullptr, reinterpret_cast<float*>(w.boxes), false, conf_thr,
                                       w.cand_keys, w.cand_count, w.cap, stream))
    return rc;
  return nms_run(w, batch, a->n, a->num_classes, w.boxes, iou_thr, top_k, iou_ge, out, kept_index, stream);
}

extern "C" int odk_nms(const float* boxes, const float* scores, int n, int top_k, float iou_thr, int iou_ge, int32_t* out_index,
                       int32_t* out_count, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(out_index && out_count, ODK_ERR_INVALID_ARGUMENT, "nms: null output pointer");
  if (n == 0) {  // ml.py:589-625 on an empty set returns an empty index array
    ODK_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t), stream));
    return ODK_OK;
  }
  if (int rc = nms_check(1, n, 1, top_k, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(boxes && scores, ODK_ERR_INVALID_ARGUMENT, "nms: null input pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "nms: boxes must be 16-byte aligned");
  NmsWorkspace w = nms_layout(workspace, 1, n, 1, top_k);
  ODK_CUDA(cudaMemsetAsync(w.cand_count, 0, sizeof(int), stream));
  dim3 grid((n + 255) / 256, 1);
  {
    OdkProfileScope prof_("collect_candidates_kernel", stream);
    collect_candidates_kernel<<<grid, 256, 0, stream>>>(n, 1, nullptr, scores, 0.f, 1, w.cand_keys, w.cand_count, w.cap);
  }
  ODK_LAUNCH_CHECK("collect_candidates_kernel");
  if (int rc = nms_run(w, 1, n, 1, reinterpret_cast<const float4*>(boxes), iou_thr, top_k, iou_ge, nullptr, nullptr, stream)) return rc;
  {
    OdkProfileScope prof_("export_kept_kernel", stream);
    export_kept_kernel<<<1, 256, 0, stream>>>(w.kept_keys, w.kept_count, out_index, out_count);
  }
  ODK_LAUNCH_CHECK("export_kept_kernel");
  return ODK_OK;
}
