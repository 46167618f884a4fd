// K2 — per-class NMS + cross-class top-k (layers.py:925-970; inner greedy NMS ml.py:589-625), one
// kernel, one CTA per image.
//
// Formulation.  The reference runs one greedy NMS per class (each over that class's candidates in
// descending score, limited to min(top_k, n_c) results), concatenates the results in class order and
// takes the global top_k by score (tf.nn.top_k, sorted, lower position first on ties).  Greedy NMS of
// different classes never interact, so visiting the candidates of ALL classes in one global order
//     key2 = [flip(score):32][4095-class:12][anchor:20]  descending
// with a class-aware suppression test yields every class's kept list simultaneously, already merged in
// the output order — and the sweep can stop as soon as top_k boxes are kept: a box cut by its class's
// own limit min(top_k, n_c) has at least top_k better boxes of its own class in front of it, so it can
// never be among the global top_k.  Typically only a few hundred of the ~4k candidates of an image are
// ever touched.
//
// Per image:
//   1. candidates arrive as per-tile compacted pieces (score_emit_kernel / collect_pieces_kernel);
//      a block scan of the piece counts gives a flat index space.
//   2. rounds, until top_k boxes are kept or the candidates run out:
//        a. MSB radix select (11-bit digits, shared-memory histograms) finds a key threshold such that
//           the next band holds <= 512 of the best remaining keys;
//        b. the band is gathered and bitonic-sorted in shared memory by (class, score desc, anchor desc);
//        c. ONLY the band's boxes are fetched — in the fused path they are decoded here from `locs`
//           (od_pb.py:216-222), so the dense decoded-box tensor never exists;
//        d. band candidates suppressed by boxes kept in earlier rounds are dropped (all threads);
//        e. one warp per class segment: 32 candidates at a time, 32x32 suppression bitmask built with
//           warp shuffles, resolved by a warp-uniform ballot-style sweep;
//        f. the round's survivors are sorted into global order and appended to the kept list / written
//           as output rows [class, conf, x1, y1, x2, y2] (layers.py:963-970), cut at top_k.
// All IoU tests evaluate the reference's fp32 expression exactly (odk_iou_passes); the division-free
// fast form is used only outside a guard band.
#include "odk_common.cuh"
#include "odk_sort.cuh"
#include <mutex>

namespace {

constexpr int kNmsThreads = 512;
constexpr int kNmsWarps = kNmsThreads / 32;
constexpr int kBand = 512;
constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kCollectPiece = 256;
constexpr int kMaxPieces = 8192;
// per-warp member lists [kNmsWarps][kBand] u16; in strict mode the same area is re-typed as the suppression bit matrix
// M [kBand][kBand / 32] u32 of the all-class stage, which is larger (found by the checks build: the matrix used to overrun
// the lists' 16 KB into the key cache once a band had more than ~360 class-stage survivors)
constexpr size_t kMemberBytes = (size_t)kNmsWarps * kBand * 2;
constexpr size_t kMatrixBytes = (size_t)kBand * (kBand / 32) * 4;
__host__ __device__ constexpr size_t member_area_bytes(bool strict) { return strict && kMatrixBytes > kMemberBytes ? kMatrixBytes : kMemberBytes; }

// ---- candidate collection for the unfused entries (layers.py:935) ----------------------------------------
// One CTA = one piece of 256 anchors, same piece format as score_emit_kernel.
__global__ void __launch_bounds__(kCollectPiece) collect_pieces_kernel(int n, int num_classes, const int32_t* __restrict__ classes,
                                                                       const float* __restrict__ confs, float conf_thr, int take_all,
                                                                       unsigned long long* __restrict__ cand_keys,
                                                                       int* __restrict__ piece_count) {
  __shared__ int s_wcnt[kCollectPiece / 32];
  const int b = blockIdx.y;
  const int a = blockIdx.x * kCollectPiece + threadIdx.x;
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
  const unsigned lane = odk_lane();
  const int warp = threadIdx.x >> 5;
  if (lane == 0) s_wcnt[warp] = __popc(m);
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kCollectPiece / 32; ++w) {
    const int v = s_wcnt[w];
    base += (w < warp) ? v : 0;
    total += v;
  }
  const size_t piece = (size_t)b * gridDim.x + blockIdx.x;
  if (is_cand) cand_keys[piece * kCollectPiece + base + __popc(m & ((1u << lane) - 1u))] = odk_make_key2(cls, conf, a);
  if (threadIdx.x == 0) piece_count[piece] = total;
}

// Pair test.  IoU >= thr  <=>  inter >= k*(area_a+area_b) with k = thr/(1+thr).  The hot loops use
// that division-free form with a +-3e-5 relative guard band (two FFMAs against per-lane constants);
// a pair that lands inside the band (or any pair when the fast form does not apply) is re-evaluated
// with the reference's exact fp32 expression (odk_iou_passes), so the decision is always the
// reference's.
struct FastThr {
  float k_hi, k_lo;
  int enabled;
};

template <bool GE>
__device__ __forceinline__ bool pair_hit(const float4 a, float area_a, float a_hi, float a_lo, const float4 b, float area_b, float thr,
                                         const FastThr ft, bool fast) {
  if (fast) {
    const float w = fmaxf(fminf(a.z, b.z) - fmaxf(a.x, b.x), 0.0f);
    const float h = fmaxf(fminf(a.w, b.w) - fmaxf(a.y, b.y), 0.0f);
    const float inter = w * h;
    const bool yes = inter > fmaf(area_b, ft.k_hi, a_hi);
    const bool maybe = (inter >= fmaf(area_b, ft.k_lo, a_lo)) && !yes;
    return maybe ? odk_iou_passes<GE>(a, area_a, b, area_b, thr) : yes;
  }
  return odk_iou_passes<GE>(a, area_a, b, area_b, thr);
}

struct NmsArgs {
  AnchorView av;
  int n, piece, n_pieces;
  const unsigned long long* cand_keys;  // [B, n_pieces*piece]
  const int* piece_count;               // [B, n_pieces]
  const int* score_hist;                // [B, ODK_HIST_BINS] per-image score histogram from score_emit, or null
  unsigned long long* tl;               // diagnostics timeline or null
  const float4* boxes;                  // [B, N] decoded boxes, or null ->
  const float4* locs;                   // [B, N] head output, decoded lazily
  float iou_thr;
  FastThr ft;
  int cache_cap;   // candidate keys that fit the shared-memory cache
  int limit;       // sweep stops once this many boxes are kept
  int top_k;       // rows of `out` per image
  int skip_empty;  // ml.py:610-611
  float* out;             // [B, top_k, 6] or null
  int32_t* kept_index;    // [B, top_k] or null
  int32_t* out_index;     // standalone: [limit] kept input indices ...
  int32_t* out_count;     // ... and their number
  // strict mode (layers.py:940,953-957): `limit` is the shared-memory capacity of the class-stage kept list
  float iou_thr2;         // nms_all_threshold
  FastThr ft2;
  int class_cap;          // per-class limit of the class stage (2*top_k)
  float4* spill_box;      // [B, n] class-stage kept entries beyond `limit` (L2-resident workspace)
  int2* spill_meta;       // [B, n] {area bits, class}
};

__device__ __forceinline__ int block_exclusive_scan(int v, int* s_warp /*[kNmsWarps + 1]*/, int& total) {
  const int lane = (int)odk_lane(), warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int wv = (lane < kNmsWarps) ? s_warp[lane] : 0;
    int winc = wv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    if (lane < kNmsWarps) s_warp[lane] = winc - wv;
    if (lane == kNmsWarps - 1) s_warp[kNmsWarps] = winc;
  }
  __syncthreads();
  const int res = s_warp[warp] + inc - v;
  total = s_warp[kNmsWarps];
  __syncthreads();
  return res;
}

// Bitonic sort (descending) of kBand keys, one per thread of the first kBand threads (the others run the
// same network among themselves on dummy keys so that the barriers stay uniform): strides below 32 are
// exchanged with warp shuffles, the rest through a double-buffered shared-memory array (one barrier per
// such stage).
__device__ __forceinline__ unsigned long long block_sort_desc_1(unsigned long long key, unsigned long long* s_x /*[2][kNmsThreads]*/,
                                                                int n_sort /* power of two >= the number of real keys */) {
  const int tid = threadIdx.x;
  int pb = 0;
  for (int k = 2; k <= n_sort; k <<= 1) {
    const bool desc = (tid & k) == 0;
    for (int j = k >> 1; j > 0; j >>= 1) {
      unsigned long long other;
      if (j >= 32) {
        s_x[pb * kNmsThreads + tid] = key;
        __syncthreads();
        other = s_x[pb * kNmsThreads + (tid ^ j)];
        pb ^= 1;
      } else {
        other = __shfl_xor_sync(0xffffffffu, key, j);
      }
      const bool lower = (tid & j) == 0;
      key = (lower == desc) ? (key > other ? key : other) : (key < other ? key : other);
    }
  }
  return key;
}

// BIG: instantiation for images with so many anchors that the candidate keys rarely fit the shared-memory cache (the
// dense-anchor stress shape): no cache, and the passes over the keys keep 16 pieces in flight per warp instead of 4 (more
// registers: kept out of the instantiation that shares its SM with the streaming kernels of other batches).
template <bool GE, bool STRICT, bool BIG>
__global__ void __launch_bounds__(kNmsThreads, BIG ? 1 : 2) nms_image_kernel(const NmsArgs p) {
  OdkTimeline tl_(ODK_TL_NMS, p.tl);
  static_assert(kBand <= kNmsThreads && ODK_MAX_CLASSES / 32 <= kNmsThreads, "one band slot / class word per thread");
  extern __shared__ __align__(16) unsigned char nms_smem[];
  __shared__ int s_hist[kBins];
  __shared__ __align__(16) unsigned long long s_band[kBand];            // key2 of the band, sorted descending (global order)
  __shared__ __align__(16) unsigned long long s_x[2 * kNmsThreads];     // sort exchange buffers
  __shared__ float4 s_box[kBand];
  __shared__ float s_area[kBand];
  __shared__ unsigned short s_cls[kBand];
  __shared__ unsigned char s_alive[kBand];
  __shared__ unsigned char s_keep[kBand];
  __shared__ unsigned s_clsbits[ODK_MAX_CLASSES / 32];
  __shared__ unsigned short s_clslist[kBand];
  __shared__ int s_scan[kNmsWarps + 1];
  __shared__ unsigned long long s_mm[2 * kNmsWarps];
  __shared__ int s_nb, s_sel[4];

  const int tid = threadIdx.x;
  const int lane = (int)odk_lane();
  const int warp = tid >> 5;
  const int b = blockIdx.x;
  const int limit = p.limit;
  unsigned char* dyn = nms_smem;
  float4* kbox = reinterpret_cast<float4*>(dyn);                          // [limit]
  float* karea = reinterpret_cast<float*>(dyn + (size_t)limit * 16);      // [limit]
  int* kcls = reinterpret_cast<int*>(dyn + (size_t)limit * 20);           // [limit]
  dyn += odk_align_up((size_t)(limit < 1 ? 1 : limit) * 24, 16);
  // strict mode: the final (all-class stage) kept list
  float4* k2box = reinterpret_cast<float4*>(dyn);                                              // [top_k]
  float* k2area = reinterpret_cast<float*>(dyn + (size_t)(STRICT ? p.top_k : 0) * 16);         // [top_k]
  if (STRICT) dyn += odk_align_up((size_t)(p.top_k < 1 ? 1 : p.top_k) * 20, 16);
  float4* spill_box = STRICT ? p.spill_box + (size_t)blockIdx.x * p.n : nullptr;
  int2* spill_meta = STRICT ? p.spill_meta + (size_t)blockIdx.x * p.n : nullptr;
  const bool fast2 = GE && p.ft2.enabled;
  int nkept2 = 0;
  int* s_pref = reinterpret_cast<int*>(dyn);                              // [n_pieces + 1]
  dyn += odk_align_up(sizeof(int) * ((size_t)p.n_pieces + 1), 16);
  unsigned short* s_member = reinterpret_cast<unsigned short*>(dyn);      // [kNmsWarps][kBand]
  dyn += member_area_bytes(STRICT);
  unsigned long long* s_cache = reinterpret_cast<unsigned long long*>(dyn);  // [cache_cap] all candidate keys of the image
  const bool fast = GE && p.ft.enabled;  // the fast form needs positive-extent boxes: guaranteed by the zero-area skip

  const unsigned long long* keys = p.cand_keys + (size_t)b * p.n_pieces * p.piece;
  const int* pcount = p.piece_count + (size_t)b * p.n_pieces;

  // ---- flat index space over the pieces ---------------------------------------------------------------
  int total;
  {
    const int per = (p.n_pieces + kNmsThreads - 1) / kNmsThreads;
    const int q0 = min(p.n_pieces, tid * per), q1 = min(p.n_pieces, q0 + per);
    int local = 0;
    for (int q = q0; q < q1; ++q) local += pcount[q];
    int run = block_exclusive_scan(local, s_scan, total);
    for (int q = q0; q < q1; ++q) {
      s_pref[q] = run;
      run += pcount[q];
    }
    if (tid == 0) s_pref[p.n_pieces] = total;
    __syncthreads();
  }
  // The image's keys are pulled into shared memory once when they fit (every later pass is on-chip);
  // otherwise each pass reads them from L2 through the piece prefix sums.
  const bool cached = !BIG && total <= p.cache_cap;
  if (cached) {
    // fire-and-forget 8-byte async copies: every key of the image is in flight at once.  A warp takes a piece, its
    // lanes the piece's entries (pieces are mostly sparse: the loop runs over what is there, not over the slots)
    for (int q = warp; q < p.n_pieces; q += kNmsWarps) {
      const int q_lo = s_pref[q], cnt = s_pref[q + 1] - q_lo;
      const unsigned long long* src = keys + (size_t)q * p.piece;
      ODK_ASSERT(cnt >= 0 && cnt <= p.piece && q_lo + cnt <= p.cache_cap);
      for (int sl = lane; sl < cnt; sl += 32) odk_cp_async<8>(s_cache + q_lo + sl, src + sl);
    }
    odk_cp_async_commit();  // waited for just before the first pass that reads the cache (cache_ready below)
  }
  bool cache_pending = cached;
  auto cache_ready = [&]() {
    if (cache_pending) {
      odk_cp_async_wait<0>();
      __syncthreads();
      cache_pending = false;
    }
  };
  // Visits every candidate key of the image once (any thread, any order).  Cached: a strided walk over shared memory.
  // Not cached (more candidates than the cache holds, e.g. 57 k per image in the dense-anchor stress shape): a warp takes
  // sixteen pieces at a time and its lanes the pieces' entries — coalesced 8-byte loads from L2, no index search.
  auto for_each_key = [&](auto&& f) {
    if (cached) {
      for (int i = tid; i < total; i += kNmsThreads) f(s_cache[i]);
    } else {
      constexpr int kU = BIG ? 16 : 4;  // pieces in flight per warp: their loads are independent, the L2 round trips overlap (a pass over the
                              // 3 216 pieces of a stress-shape image is 4 such round trips per warp instead of 13 at kU = 4)
      for (int q0 = warp * kU; q0 < p.n_pieces; q0 += kNmsWarps * kU) {
        int cnt[kU], maxc = 0;
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          cnt[u] = (q0 + u < p.n_pieces) ? s_pref[q0 + u + 1] - s_pref[q0 + u] : 0;
          maxc = max(maxc, cnt[u]);
        }
        const unsigned long long* src = keys + (size_t)q0 * p.piece;
        for (int sl = lane; sl < maxc; sl += 32) {
          unsigned long long k[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u) k[u] = (sl < cnt[u]) ? src[(size_t)u * p.piece + sl] : 0ull;
#pragma unroll
          for (int u = 0; u < kU; ++u)
            if (sl < cnt[u]) f(k[u]);
        }
      }
    }
  };

  int nkept = 0, processed = 0;
  int hist_top = ODK_HIST_BINS;  // bins below it hold exactly the keys not yet processed (0: histogram no longer usable)
  unsigned long long hi_key = ~0ull;  // keys of earlier bands are >= hi_key (every real key is < ~0)

  // Walks the nb_bins bins of s_hist from the top: n_take = how many top bins to take whole, cnt = keys they hold
  // (incl. `above`).  Prefers the shortest run of top bins that reaches `want` keys (the sweep rarely needs more); if
  // the bin that crosses `want` would overflow the band, the bins before it.  All threads call it; uniform result.
  auto walk_bins = [&](int nb_bins, int above, int want, int& n_take, int& cnt) {
    constexpr int per = (kBins + kNmsThreads - 1) / kNmsThreads;
    int loc[per];
    int lsum = 0;
#pragma unroll
    for (int j = 0; j < per; ++j) {
      const int r = tid * per + j;
      loc[j] = (r < nb_bins) ? s_hist[nb_bins - 1 - r] : 0;
      lsum += loc[j];
    }
    int tot_in_bucket;
    int run = above + block_exclusive_scan(lsum, s_scan, tot_in_bucket);
    // S is monotone over r, so "S < want" and "S <= kBand" each hold on a prefix of the bins
    int n_ok = 0, last_ok_sum = -1, n_lt = 0, first_ge_sum = INT_MAX;
#pragma unroll
    for (int j = 0; j < per; ++j) {
      run += loc[j];
      const int r = tid * per + j;
      if (r < nb_bins) {
        if (run <= kBand) { ++n_ok; last_ok_sum = run; }
        if (run < want) ++n_lt;
        else first_ge_sum = min(first_ge_sum, run);
      }
    }
    if (tid == 0) { s_sel[0] = 0; s_sel[1] = above; s_sel[2] = 0; s_sel[3] = INT_MAX; }
    __syncthreads();
    if (n_ok) { atomicAdd(&s_sel[0], n_ok); atomicMax(&s_sel[1], last_ok_sum); }
    if (n_lt) atomicAdd(&s_sel[2], n_lt);
    if (first_ge_sum != INT_MAX) atomicMin(&s_sel[3], first_ge_sum);
    __syncthreads();
    if (s_sel[2] < s_sel[0]) { n_take = s_sel[2] + 1; cnt = s_sel[3]; }
    else { n_take = s_sel[0]; cnt = s_sel[1]; }
    __syncthreads();
  };

  while (processed < total && (STRICT ? nkept2 < p.top_k : nkept < limit)) {
    const int remaining = total - processed;
    // ---- a. threshold of the next band: MSB radix select starting at the first bit in which the
    //         remaining keys differ (fp32 scores share their leading exponent bits) -------------------------
    unsigned long long lo_key = 0ull;
    const int need = STRICT ? p.top_k - nkept2 : limit - nkept;
    // band size aimed at: a little more than what is still needed, scaled by the survival ratio seen so far (first round:
    // 1.125x + 16, which for top_k = 200 keeps the band within a 256-key sort network when little is suppressed)
    int want = STRICT ? need + (need >> 1) + 16   // the all-class stage removes cross-class duplicates on top: one larger band beats a second round
                      : need + (need >> 3) + 16;
    const int kept_so_far = STRICT ? nkept2 : nkept;
    if (kept_so_far > 0) want = (int)min((long long)kBand, (long long)need * processed / kept_so_far * 9 / 8 + 32);
    want = min(kBand, want);
    bool have_lo = remaining <= kBand;
    if (!have_lo && hist_top > 0 && p.score_hist != nullptr) {
      // The score histogram score_emit built while streaming replaces the passes over the keys (min/max + digit
      // histograms) — in the first round and in every later one, as long as all earlier thresholds were bin boundaries:
      // the keys not yet processed are then exactly the bins below hist_top.  A bin's lower bound is a key threshold; if
      // the bin that crosses `want` overflows the band and too little lies above it, the generic select below refines
      // among the keys themselves (and the histogram is not used again for this image).
      const int* hist = p.score_hist + (size_t)b * ODK_HIST_BINS;
      for (int i = tid; i < hist_top; i += kNmsThreads) s_hist[i] = hist[i];
      __syncthreads();
      int n_take_bins, cnt_take;
      walk_bins(hist_top, 0, want, n_take_bins, cnt_take);
      if (cnt_take >= max(32, want / 2)) {
        const int d_star = hist_top - n_take_bins;  // bins [d_star, hist_top) are taken whole (n_take_bins >= 1 here)
        lo_key = d_star > 0 ? odk_bin_lo_key(d_star) : 0ull;
        have_lo = true;
        hist_top = d_star;
      }
    }
    if (!have_lo) hist_top = 0;  // the generic select cuts inside a bin
    cache_ready();
    if (!have_lo) {
      unsigned long long kmin = ~0ull, kmax = 0ull;
      for_each_key([&](unsigned long long k) {
        if (k < hi_key) {
          kmin = k < kmin ? k : kmin;
          kmax = k > kmax ? k : kmax;
        }
      });
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        const unsigned long long omin = __shfl_xor_sync(0xffffffffu, kmin, d), omax = __shfl_xor_sync(0xffffffffu, kmax, d);
        kmin = omin < kmin ? omin : kmin;
        kmax = omax > kmax ? omax : kmax;
      }
      if (lane == 0) { s_mm[warp] = kmin; s_mm[kNmsWarps + warp] = kmax; }
      __syncthreads();
      for (int w = 0; w < kNmsWarps; ++w) {
        kmin = s_mm[w] < kmin ? s_mm[w] : kmin;
        kmax = s_mm[kNmsWarps + w] > kmax ? s_mm[kNmsWarps + w] : kmax;
      }
      int top = 64 - __clzll((long long)(kmin ^ kmax));  // bits [top-1:0] vary (keys are distinct, so top >= 1)
      unsigned long long prefix = (top >= 64) ? 0ull : (kmax >> top);  // common bits above them
      int above = 0;                                                   // keys < hi_key above the current bucket (all taken)
      while (true) {
        const int shift = max(0, top - kDigitBits);
        const int width = top - shift;
        for (int i = tid; i < kBins; i += kNmsThreads) s_hist[i] = 0;
        __syncthreads();
        for_each_key([&](unsigned long long k) {
          if (k < hi_key && (top >= 64 || (k >> top) == prefix)) atomicAdd(&s_hist[(int)((k >> shift) & (unsigned long long)((1 << width) - 1))], 1);
        });
        __syncthreads();
        // walk the digits from the top: largest r such that above + S(r) <= kBand, S = inclusive sum over reversed bins
        const int nb_bins = 1 << width;
        int n_take_bins, cnt_take;
        walk_bins(nb_bins, above, want, n_take_bins, cnt_take);
        // when the bins taken whole hold too few keys, refine inside the crossing bin
        const int d_star = nb_bins - n_take_bins;  // digits >= d_star are taken whole
        if (cnt_take >= max(32, want / 2) || shift == 0) {
          if (d_star < nb_bins) lo_key = ((prefix << width) | (unsigned long long)d_star) << shift;
          else lo_key = (prefix + 1ull) << top;  // nothing of this bucket: only the keys above it (top < 64 here)
          break;
        }
        prefix = (prefix << width) | (unsigned long long)(d_star - 1);  // refine inside the overflowing bin
        top = shift;
        above = cnt_take;
      }
    }
    // ---- b. gather the band, sort it into the global order (one key per thread) --------------------------
    if (tid == 0) s_nb = 0;
    for (int i = tid; i < ODK_MAX_CLASSES / 32; i += kNmsThreads) s_clsbits[i] = 0u;
    __syncthreads();
    for_each_key([&](unsigned long long k) {
      if (k < hi_key && k >= lo_key) {
        const int pos = atomicAdd(&s_nb, 1);
        if (pos < kBand) s_band[pos] = k;
      }
    });
    __syncthreads();
    const int nb = min(s_nb, kBand);
    int n_sort = 32;
    while (n_sort < nb) n_sort <<= 1;
    const unsigned long long my_key = block_sort_desc_1(tid < nb ? s_band[tid] : 0ull, s_x, n_sort);
    hi_key = lo_key;
    processed += nb;

    // ---- c. boxes of the band (lazy decode), classes present in it -------------------------------------------
    __syncthreads();  // every thread has read its unsorted s_band slot
    if (tid < nb) {
      const int a = (int)(my_key & ODK_IDX_MASK);
      const int cls = odk_key2_class(my_key);
      float4 box;
      if (p.boxes != nullptr) box = p.boxes[(size_t)b * p.n + a];
      else box = odk_decode_box(p.locs[(size_t)b * p.n + a], __ldg(p.av.locs + a), __ldg(p.av.scale + a));
      const float w = __fsub_rn(box.z, box.x), h = __fsub_rn(box.w, box.y);
      s_band[tid] = my_key;
      s_box[tid] = box;
      s_area[tid] = __fmul_rn(w, h);
      s_cls[tid] = (unsigned short)cls;
      s_alive[tid] = !(p.skip_empty && (w <= 0.0f || h <= 0.0f));  // ml.py:610-611
      s_keep[tid] = 0;
      atomicOr(&s_clsbits[cls >> 5], 1u << (cls & 31));
    }
    __syncthreads();
    int n_present;
    {
      const unsigned wbits = (tid < ODK_MAX_CLASSES / 32) ? s_clsbits[tid] : 0u;
      int pos = block_exclusive_scan(__popc(wbits), s_scan, n_present);
      ODK_ASSERT(pos + __popc(wbits) <= kBand);
      for (unsigned rem = wbits; rem; rem &= rem - 1) s_clslist[pos++] = (unsigned short)(tid * 32 + __ffs(rem) - 1);
    }
    __syncthreads();
    // ---- d/e. one warp per class present: its members in score order, 32 at a time ----------------------------
    for (int ci = warp; ci < n_present; ci += kNmsWarps) {
      const int c = s_clslist[ci];
      unsigned short* mem = s_member + (size_t)warp * kBand;
      int nm = 0;
      for (int base = 0; base < nb; base += 32) {
        const int i = base + lane;
        const bool is = i < nb && s_cls[i] == c;
        const unsigned m = __ballot_sync(0xffffffffu, is);
        ODK_ASSERT(nm + __popc(m) <= kBand);
        if (is) mem[nm + __popc(m & ((1u << lane) - 1u))] = (unsigned short)i;
        nm += __popc(m);
      }
      __syncwarp();
      int seg_kept = 0;
      int cls_before = 0;  // STRICT: boxes of this class kept by the class stage in earlier rounds
      for (int mb = 0; mb < nm; mb += 32) {
        const bool have = mb + lane < nm;
        const int i = have ? mem[mb + lane] : 0;
        const float4 box = have ? s_box[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float area = have ? s_area[i] : 0.f;
        const float a_hi = area * p.ft.k_hi, a_lo = area * p.ft.k_lo;
        bool dead = false;
        // boxes of this class kept in earlier rounds (warp-uniform class test)
        const int nk_s = STRICT ? min(nkept, limit) : nkept;
        int cnt = 0;
        for (int k = 0; k < nk_s; ++k)
          if (kcls[k] == c) {
            ++cnt;
            dead |= pair_hit<GE>(box, area, a_hi, a_lo, kbox[k], karea[k], p.iou_thr, p.ft, fast);
          }
        if (STRICT) {
          for (int k = limit; k < nkept; ++k) {  // spilled part of the class-stage list
            const int2 m = spill_meta[k];
            if (m.y == c) {
              ++cnt;
              dead |= pair_hit<GE>(box, area, a_hi, a_lo, spill_box[k], __int_as_float(m.x), p.iou_thr, p.ft, fast);
            }
          }
          cls_before = cnt;
        }
        // ... and in earlier chunks of this round
        if (seg_kept > 0) {
          for (int e = 0; e < mb; ++e) {
            const int ie = mem[e];
            if (s_keep[ie]) dead |= pair_hit<GE>(box, area, a_hi, a_lo, s_box[ie], s_area[ie], p.iou_thr, p.ft, fast);
          }
        }
        const bool alive = have && s_alive[i] && !dead;
        // intra-chunk greedy in score order: the first remaining lane is kept and knocks out what it suppresses
        unsigned keep = 0;
        for (unsigned rem = __ballot_sync(0xffffffffu, alive); rem;) {
          if (STRICT && cls_before + seg_kept + __popc(keep) >= p.class_cap) break;  // layers.py:940-941 (warp-uniform)
          const int l = __ffs(rem) - 1;
          keep |= 1u << l;
          rem &= rem - 1;
          float4 bl;
          bl.x = __shfl_sync(0xffffffffu, box.x, l);
          bl.y = __shfl_sync(0xffffffffu, box.y, l);
          bl.z = __shfl_sync(0xffffffffu, box.z, l);
          bl.w = __shfl_sync(0xffffffffu, box.w, l);
          const float al = __shfl_sync(0xffffffffu, area, l);
          const bool hit = ((rem >> lane) & 1u) && pair_hit<GE>(box, area, a_hi, a_lo, bl, al, p.iou_thr, p.ft, fast);
          rem &= ~__ballot_sync(0xffffffffu, hit);
        }
        if ((keep >> lane) & 1u) s_keep[i] = 1;
        seg_kept += __popc(keep);
        __syncwarp();
      }
    }
    __syncthreads();
    // ---- f. survivors are already in global order: compact, cut at the limit, emit ------------------------------
    if (!STRICT) {
      const bool kept = tid < nb && s_keep[tid];
      int nk;
      const int pos = block_exclusive_scan(kept ? 1 : 0, s_scan, nk);
      const int take = min(nk, limit - nkept);
      if (kept && pos < take) {
        const int a = (int)(my_key & ODK_IDX_MASK);
        const int cls = odk_key2_class(my_key);
        const float4 bx = s_box[tid];
        const int r = nkept + pos;
        ODK_ASSERT(r >= 0 && r < limit && a >= 0 && a < p.n);
        kbox[r] = bx;
        karea[r] = s_area[tid];
        kcls[r] = cls;
        if (p.out != nullptr && r < p.top_k) {
          float* o = p.out + ((size_t)b * p.top_k + r) * 6;
          o[0] = (float)cls;
          o[1] = odk_key2_score(my_key);
          o[2] = bx.x; o[3] = bx.y; o[4] = bx.z; o[5] = bx.w;
          p.kept_index[(size_t)b * p.top_k + r] = a;
        }
        if (p.out_index != nullptr) p.out_index[r] = a;
      }
      nkept += take;
    } else {
      // ---- f'. strict mode.  The class-stage survivors (global order) all join the class-stage kept list;
      //      then the all-class greedy pass (layers.py:953-957) visits them in the same order against the
      //      final kept list at nms_all_threshold, stopping at top_k. ---------------------------------------
      const bool kept1 = tid < nb && s_keep[tid];
      int nk1;
      const int pos = block_exclusive_scan(kept1 ? 1 : 0, s_scan, nk1);
      bool dead2 = false;
      if (kept1) {
        const float4 bx = s_box[tid];
        const float ar = s_area[tid];
        const int r1 = nkept + pos;
        if (r1 < limit) {
          kbox[r1] = bx;
          karea[r1] = ar;
          kcls[r1] = (int)s_cls[tid];
        } else {
          spill_box[r1] = bx;
          spill_meta[r1] = make_int2(__float_as_int(ar), (int)s_cls[tid]);
        }
        s_clslist[pos] = (unsigned short)tid;  // survivors in global order (the class list is no longer needed)
        const float a_hi = ar * p.ft2.k_hi, a_lo = ar * p.ft2.k_lo;
        for (int k = 0; k < nkept2; ++k) dead2 |= pair_hit<GE>(bx, ar, a_hi, a_lo, k2box[k], k2area[k], p.iou_thr2, p.ft2, fast2);
        s_alive[tid] = dead2 ? 0 : 1;
      }
      __syncthreads();
      // in-band greedy order, bitmask form: every thread fills words of the suppression matrix M[i][w] (bit j of word
      // w: survivor 32w+j, later than i, is suppressed by survivor i), then one warp resolves the order serially with
      // a 512-bit "removed" set spread over its lanes — one shared-memory row OR per kept box.
      const int nw = (nk1 + 31) >> 5;
      unsigned* M = reinterpret_cast<unsigned*>(s_member);                  // [nk1][nw] <= 512 x 16 words = 32 KB
      unsigned short* rlist = reinterpret_cast<unsigned short*>(s_x);       // this round's final keeps (band slots)
      // a warp owns row i; its lanes test the 32 survivors of one word at a time and the ballot IS the word
      for (int i = warp; i < nk1; i += kNmsWarps) {
        const int si = s_clslist[i];
        const bool row_alive = s_alive[si] != 0;
        const float4 bi = s_box[si];
        const float ai = s_area[si];
        for (int w = i >> 5; w < nw; ++w) {
          const int j = w * 32 + lane;
          bool hit = false;
          if (row_alive && j > i && j < nk1) {
            const int sj = s_clslist[j];
            const float aj = s_area[sj];
            hit = pair_hit<GE>(s_box[sj], aj, aj * p.ft2.k_hi, aj * p.ft2.k_lo, bi, ai, p.iou_thr2, p.ft2, fast2);
          }
          const unsigned bits = __ballot_sync(0xffffffffu, hit);
          ODK_ASSERT(nk1 <= kBand && (size_t)(i * nw + w) * 4 < member_area_bytes(STRICT));
          if (lane == 0) M[i * nw + w] = bits;
        }
      }
      __syncthreads();
      if (warp == 0) {
        // serial resolve, 32 survivors at a time: the block's own order needs only the diagonal words (registers +
        // shuffles); the rows of the boxes it keeps are then OR-ed into the later words of `removed` in one go
        const int room = p.top_k - nkept2;
        unsigned removed = 0u;  // lane l: removed flags of survivors 32l .. 32l+31
        int rk = 0;
        for (int w = 0; w < nw && rk < room; ++w) {
          const int j = w * 32 + lane;
          const bool in = j < nk1;
          const unsigned alive = __ballot_sync(0xffffffffu, in && s_alive[s_clslist[in ? j : 0]]);
          const unsigned diag = in ? M[j * nw + w] : 0u;
          unsigned rem = alive & ~__shfl_sync(0xffffffffu, removed, w);
          unsigned keep = 0u;
          while (rem) {
            const int l = __ffs(rem) - 1;
            keep |= 1u << l;
            rem &= ~(__shfl_sync(0xffffffffu, diag, l) | (1u << l));
          }
          const int take = min(__popc(keep), room - rk);
          if (take < __popc(keep)) keep &= (1u << (__fns(keep, 0, take + 1) & 31)) - 1u;  // cut at top_k: drop from the (take+1)-th on
          ODK_ASSERT(rk + __popc(keep) <= kBand);
          if ((keep >> lane) & 1u) rlist[rk + __popc(keep & ((1u << lane) - 1u))] = s_clslist[j];
          rk += take;
          if (lane > w && lane < nw)
            for (unsigned kk = keep; kk; kk &= kk - 1) removed |= M[(w * 32 + __ffs(kk) - 1) * nw + lane];
        }
        if (lane == 0) s_nb = rk;
      }
      __syncthreads();
      const int rk = s_nb;
      if (tid < rk) {
        const int i = reinterpret_cast<const unsigned short*>(s_x)[tid];
        const unsigned long long key = s_band[i];
        const float4 bx = s_box[i];
        const int r = nkept2 + tid;
        ODK_ASSERT(r >= 0 && r < p.top_k);
        k2box[r] = bx;
        k2area[r] = s_area[i];
        if (p.out != nullptr) {
          float* o = p.out + ((size_t)b * p.top_k + r) * 6;
          o[0] = (float)odk_key2_class(key);
          o[1] = odk_key2_score(key);
          o[2] = bx.x; o[3] = bx.y; o[4] = bx.z; o[5] = bx.w;
          p.kept_index[(size_t)b * p.top_k + r] = (int)(key & ODK_IDX_MASK);
        }
      }
      nkept += nk1;
      nkept2 += rk;
    }
    __syncthreads();
  }
  if (STRICT) nkept = nkept2;  // rows written
  // ---- zero padding (layers.py:969) / count ---------------------------------------------------------------------
  if (p.out != nullptr) {
    for (int r = nkept + tid; r < p.top_k; r += kNmsThreads) {
      float* o = p.out + ((size_t)b * p.top_k + r) * 6;
      o[0] = o[1] = o[2] = o[3] = o[4] = o[5] = 0.0f;
      p.kept_index[(size_t)b * p.top_k + r] = -1;
    }
  }
  if (p.out_count != nullptr && tid == 0) *p.out_count = nkept;
}

}  // namespace

// ---- workspace layout ------------------------------------------------------------------------------------
struct NmsWorkspace {
  int* piece_count;               // [B, ceil(N/32)]   (only n_pieces per image are used)
  unsigned long long* cand_keys;  // [B, round_up(N, 256)]
  int* score_hist;                // [B, ODK_HIST_BINS]
  float4* spill_box;              // strict only: [B, N]
  int2* spill_meta;               // strict only: [B, N]
  size_t bytes;
};

static NmsWorkspace nms_layout(void* base, int batch, int n, bool strict = false) {
  NmsWorkspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = odk_align_up(off + bytes, 256); return o; };
  const size_t o_cnt = take(sizeof(int) * (size_t)batch * ((n + 31) / 32));
  const size_t o_ck = take(8 * (size_t)batch * odk_align_up((size_t)n, 256));
  const size_t o_hist = take(sizeof(int) * (size_t)batch * ODK_HIST_BINS);
  const size_t o_sb = strict ? take(16 * (size_t)batch * n) : 0;
  const size_t o_sm = strict ? take(8 * (size_t)batch * n) : 0;
  char* p = static_cast<char*>(base);
  w.piece_count = reinterpret_cast<int*>(p + o_cnt);
  w.cand_keys = reinterpret_cast<unsigned long long*>(p + o_ck);
  w.score_hist = reinterpret_cast<int*>(p + o_hist);
  w.spill_box = strict ? reinterpret_cast<float4*>(p + o_sb) : nullptr;
  w.spill_meta = strict ? reinterpret_cast<int2*>(p + o_sm) : nullptr;
  w.bytes = off;
  return w;
}

static int nms_check(int batch, int n, int num_classes, int top_k, const void* ws, size_t ws_bytes, bool strict = false) {
  ODK_REQUIRE(batch >= 1 && n >= 1 && num_classes >= 1 && top_k >= 0, ODK_ERR_INVALID_ARGUMENT,
              "nms: bad sizes batch=%d n=%d classes=%d top_k=%d", batch, n, num_classes, top_k);
  ODK_REQUIRE(n < (int)ODK_IDX_MASK, ODK_ERR_UNSUPPORTED, "nms: n=%d exceeds the 20-bit anchor field", n);
  ODK_REQUIRE(num_classes <= ODK_MAX_CLASSES, ODK_ERR_UNSUPPORTED, "nms: num_classes=%d > %d", num_classes, ODK_MAX_CLASSES);
  ODK_REQUIRE(top_k <= 4096, ODK_ERR_UNSUPPORTED, "nms: top_k=%d > 4096 (kept list lives in shared memory)", top_k);
  const size_t need = nms_layout(nullptr, batch, n, strict).bytes;
  ODK_REQUIRE(ws != nullptr && ws_bytes >= need, ODK_ERR_WORKSPACE, "nms: workspace of %zu bytes is smaller than the required %zu", ws_bytes,
              need);
  ODK_REQUIRE(((uintptr_t)ws & 255) == 0, ODK_ERR_INVALID_ARGUMENT, "nms: workspace must be 256-byte aligned");
  return ODK_OK;
}

// select + sort + sweep + output on already emitted candidate pieces
static int nms_run(const NmsWorkspace& w, const odk_anchors* anchors, int batch, int n, int piece, const float4* boxes, const float4* locs,
                   float iou_thr, int top_k, int limit, int iou_ge, float* out, int32_t* kept_index, int32_t* out_index, int32_t* out_count,
                   cudaStream_t stream, bool strict = false, float iou_thr2 = 0.f, bool have_hist = false) {
  NmsArgs a;
  a.score_hist = have_hist ? w.score_hist : nullptr;
  a.tl = odk_timeline_ptr();
  if (anchors != nullptr) a.av = odk_view(anchors);
  else memset(&a.av, 0, sizeof(a.av));
  a.n = n; a.piece = piece; a.n_pieces = (n + piece - 1) / piece;
  ODK_REQUIRE(a.n_pieces <= kMaxPieces, ODK_ERR_UNSUPPORTED, "nms: %d candidate pieces per image > %d", a.n_pieces, kMaxPieces);
  a.cand_keys = w.cand_keys; a.piece_count = w.piece_count;
  a.boxes = boxes; a.locs = locs;
  a.iou_thr = iou_thr;
  a.ft.enabled = (iou_thr > 1e-3f && iou_thr < 0.999f) ? 1 : 0;
  const double kk = a.ft.enabled ? (double)iou_thr / (1.0 + (double)iou_thr) : 0.0;
  a.ft.k_hi = (float)(kk * (1.0 + 3e-5));
  a.ft.k_lo = (float)(kk * (1.0 - 3e-5));
  a.iou_thr2 = iou_thr2;
  a.ft2.enabled = (strict && iou_thr2 > 1e-3f && iou_thr2 < 0.999f) ? 1 : 0;
  const double kk2 = a.ft2.enabled ? (double)iou_thr2 / (1.0 + (double)iou_thr2) : 0.0;
  a.ft2.k_hi = (float)(kk2 * (1.0 + 3e-5));
  a.ft2.k_lo = (float)(kk2 * (1.0 - 3e-5));
  a.class_cap = 2 * top_k;  // layers.py:940
  a.spill_box = w.spill_box; a.spill_meta = w.spill_meta;
  if (strict) limit = 1024;  // shared-memory part of the class-stage kept list; the rest spills to the workspace
  a.limit = limit; a.top_k = top_k; a.skip_empty = iou_ge ? 1 : 0;
  a.out = out; a.kept_index = kept_index; a.out_index = out_index; a.out_count = out_count;
  const size_t fixed = odk_align_up((size_t)(limit < 1 ? 1 : limit) * 24, 16) + odk_align_up(sizeof(int) * ((size_t)a.n_pieces + 1), 16) +
                       member_area_bytes(strict) + (strict ? odk_align_up((size_t)(top_k < 1 ? 1 : top_k) * 20, 16) : 0);
  // Key cache: what is left of a 94 KB shared-memory budget per NMS CTA (34.6 KB static + `fixed` + cache), in 512-key
  // steps.  The budget — together with the 62 registers the kernel compiles to at two CTAs per SM — is what lets an NMS
  // CTA share its SM with two score_emit CTAs (2 x 65 KB) of the next batch in flight instead of waiting for that
  // persistent kernel to end: step 41.2 -> 39.1 us (profiles/sweep_r02v.log).  VOC shape: 4 608 keys (an image with more
  // candidates reads its keys through L2; with the score histogram a round is one such pass).
  {
    const size_t budget = 94 * 1024, used = 34656 + fixed;
    const size_t room = budget > used ? (budget - used) / 8 / 512 * 512 : 0;
    a.cache_cap = (int)(room < (size_t)n ? room : (size_t)n);
  }
  static const int env_cache = getenv("ODK_NMS_CACHE") ? atoi(getenv("ODK_NMS_CACHE")) : -1;  // diagnostics: key-cache capacity
  if (env_cache >= 0 && env_cache < a.cache_cap) a.cache_cap = env_cache;
  if (fixed + (size_t)a.cache_cap * 8 > 180 * 1024) a.cache_cap = 0;
  const bool big = n >= 65536;
  if (big) a.cache_cap = 0;
  const size_t smem = fixed + (size_t)a.cache_cap * 8;
  ODK_REQUIRE(smem <= 180 * 1024, ODK_ERR_UNSUPPORTED, "nms: top_k=%d / %d pieces need %zu bytes of shared memory", top_k, a.n_pieces, smem);
  void (*kern)(const NmsArgs) = nullptr;
  if (big) kern = strict ? (iou_ge ? nms_image_kernel<true, true, true> : nms_image_kernel<false, true, true>)
                         : (iou_ge ? nms_image_kernel<true, false, true> : nms_image_kernel<false, false, true>);
  else kern = strict ? (iou_ge ? nms_image_kernel<true, true, false> : nms_image_kernel<false, true, false>)
                     : (iou_ge ? nms_image_kernel<true, false, false> : nms_image_kernel<false, false, false>);
  ODK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 180 * 1024));
  odk_prefer_max_smem(kern);
  {
    OdkProfileScope prof_(strict ? "nms_image_kernel_strict" : "nms_image_kernel", stream);
    odk_launch_high(kern, dim3(batch), dim3(kNmsThreads), smem, stream, a);
  }
  ODK_LAUNCH_CHECK("nms_image_kernel");
  return ODK_OK;
}

int odk_emit_piece_size(int num_classes);
int odk_launch_decode_score(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs,
                            int32_t* classes, float* confs, float* boxes, bool dense, float conf_thr,
                            unsigned long long* cand_keys, int* piece_count, int* score_hist, cudaStream_t stream);

extern "C" size_t odk_nms_workspace_bytes(int batch, int n, int num_classes, int top_k) {
  if (batch < 1 || n < 1 || num_classes < 1 || top_k < 0) return 0;
  return nms_layout(nullptr, batch, n).bytes;
}

static int collect(const NmsWorkspace& w, int batch, int n, int num_classes, const int32_t* classes, const float* confs, float conf_thr,
                   int take_all, cudaStream_t stream) {
  dim3 grid((n + kCollectPiece - 1) / kCollectPiece, batch);
  ODK_REQUIRE(batch <= 65535, ODK_ERR_UNSUPPORTED, "nms: batch=%d > 65535", batch);
  ODK_MAX_CARVEOUT_ONCE(collect_pieces_kernel);
  {
    OdkProfileScope prof_("collect_pieces_kernel", stream);
    collect_pieces_kernel<<<grid, kCollectPiece, 0, stream>>>(n, num_classes, classes, confs, conf_thr, take_all, w.cand_keys, w.piece_count);
  }
  ODK_LAUNCH_CHECK("collect_pieces_kernel");
  return ODK_OK;
}

static int nms_layer_impl(int batch, int n, int num_classes, const int32_t* classes, const float* confs, const float* boxes, float conf_thr,
                          float iou_thr, bool strict, float nms_all_thr, int top_k, int iou_ge, float* out, int32_t* kept_index,
                          void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = nms_check(batch, n, num_classes, top_k, workspace, workspace_bytes, strict)) return rc;
  ODK_REQUIRE(classes && confs && boxes && out && kept_index, ODK_ERR_INVALID_ARGUMENT, "nms_layer: null pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "nms_layer: boxes must be 16-byte aligned");
  NmsWorkspace w = nms_layout(workspace, batch, n, strict);
  if (int rc = collect(w, batch, n, num_classes, classes, confs, conf_thr, 0, stream)) return rc;
  return nms_run(w, nullptr, batch, n, kCollectPiece, reinterpret_cast<const float4*>(boxes), nullptr, iou_thr, top_k, top_k, iou_ge, out,
                 kept_index, nullptr, nullptr, stream, strict, nms_all_thr);
}

static int decode_nms_impl(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs, float conf_thr,
                           float iou_thr, bool strict, float nms_all_thr, int top_k, int iou_ge, float* out, int32_t* kept_index,
                           void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(a != nullptr, ODK_ERR_INVALID_ARGUMENT, "decode_nms: null anchors handle");
  if (int rc = nms_check(batch, a->n, a->num_classes, top_k, workspace, workspace_bytes, strict)) return rc;
  ODK_REQUIRE(objs && clfs && locs && out && kept_index, ODK_ERR_INVALID_ARGUMENT, "decode_nms: null pointer");
  ODK_REQUIRE((((uintptr_t)objs | (uintptr_t)clfs) & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "decode_nms: inputs must be 4-byte aligned");
  ODK_REQUIRE(((uintptr_t)locs & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "decode_nms: locs must be 16-byte aligned");
  NmsWorkspace w = nms_layout(workspace, batch, a->n, strict);
  static const bool use_hist = !(getenv("ODK_NMS_HIST") && atoi(getenv("ODK_NMS_HIST")) == 0);  // diagnostics: 0 = generic select only
  if (use_hist) ODK_CUDA(cudaMemsetAsync(w.score_hist, 0, sizeof(int) * (size_t)batch * ODK_HIST_BINS, stream));
  if (int rc = odk_launch_decode_score(a, batch, objs, clfs, locs, nullptr, nullptr, nullptr, false, conf_thr, w.cand_keys, w.piece_count,
                                       use_hist ? w.score_hist : nullptr, stream))
    return rc;
  return nms_run(w, a, batch, a->n, odk_emit_piece_size(a->num_classes), nullptr, reinterpret_cast<const float4*>(locs), iou_thr, top_k, top_k,
                 iou_ge, out, kept_index, nullptr, nullptr, stream, strict, nms_all_thr, use_hist);
}

extern "C" int odk_nms_layer(int batch, int n, int num_classes, const int32_t* classes, const float* confs, const float* boxes,
                             float conf_thr, float iou_thr, int top_k, int iou_ge, float* out, int32_t* kept_index, void* workspace,
                             size_t workspace_bytes, void* stream) {
  return nms_layer_impl(batch, n, num_classes, classes, confs, boxes, conf_thr, iou_thr, false, 0.f, top_k, iou_ge, out, kept_index, workspace,
                        workspace_bytes, stream);
}

extern "C" int odk_decode_nms(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs, float conf_thr,
                              float iou_thr, int top_k, int iou_ge, float* out, int32_t* kept_index, void* workspace,
                              size_t workspace_bytes, void* stream) {
  return decode_nms_impl(a, batch, objs, clfs, locs, conf_thr, iou_thr, false, 0.f, top_k, iou_ge, out, kept_index, workspace, workspace_bytes,
                         stream);
}

// strict_nms (od_net.py:175, layers.py:940,953-957): nms_all_threshold given
extern "C" size_t odk_nms_strict_workspace_bytes(int batch, int n, int num_classes, int top_k) {
  if (batch < 1 || n < 1 || num_classes < 1 || top_k < 0) return 0;
  return nms_layout(nullptr, batch, n, true).bytes;
}

extern "C" int odk_nms_layer_strict(int batch, int n, int num_classes, const int32_t* classes, const float* confs, const float* boxes,
                                    float conf_thr, float iou_thr, float nms_all_thr, int top_k, int iou_ge, float* out,
                                    int32_t* kept_index, void* workspace, size_t workspace_bytes, void* stream) {
  ODK_REQUIRE(nms_all_thr > 0.f, ODK_ERR_INVALID_ARGUMENT, "nms_layer_strict: nms_all_threshold must be > 0 (falsy means non-strict)");
  return nms_layer_impl(batch, n, num_classes, classes, confs, boxes, conf_thr, iou_thr, true, nms_all_thr, top_k, iou_ge, out, kept_index,
                        workspace, workspace_bytes, stream);
}

extern "C" int odk_decode_nms_strict(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs,
                                     float conf_thr, float iou_thr, float nms_all_thr, int top_k, int iou_ge, float* out,
                                     int32_t* kept_index, void* workspace, size_t workspace_bytes, void* stream) {
  ODK_REQUIRE(nms_all_thr > 0.f, ODK_ERR_INVALID_ARGUMENT, "decode_nms_strict: nms_all_threshold must be > 0 (falsy means non-strict)");
  return decode_nms_impl(a, batch, objs, clfs, locs, conf_thr, iou_thr, true, nms_all_thr, top_k, iou_ge, out, kept_index, workspace,
                         workspace_bytes, stream);
}

extern "C" int odk_nms(const float* boxes, const float* scores, int n, int top_k, float iou_thr, int iou_ge, int32_t* out_index,
                       int32_t* out_count, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(out_index && out_count, ODK_ERR_INVALID_ARGUMENT, "nms: null output pointer");
  if (n == 0) {  // ml.py:589-625 on an empty set returns an empty index array
    ODK_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t), stream));
    return ODK_OK;
  }
  if (int rc = nms_check(1, n, 1, top_k < 0 ? 0 : top_k, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(boxes && scores, ODK_ERR_INVALID_ARGUMENT, "nms: null input pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "nms: boxes must be 16-byte aligned");
  NmsWorkspace w = nms_layout(workspace, 1, n);
  if (int rc = collect(w, 1, n, 1, nullptr, scores, 0.f, 1, stream)) return rc;
  // ml.py:612-623: the length check sits only in the non-first branch -> at least two results (SURVEY A-7)
  const int limit = iou_ge ? (top_k < 2 ? 2 : top_k) : (top_k < 0 ? 0 : top_k);
  return nms_run(w, nullptr, 1, n, kCollectPiece, reinterpret_cast<const float4*>(boxes), nullptr, iou_thr, 0, limit, iou_ge, nullptr, nullptr,
                 out_index, out_count, stream);
}


// ---------------------------------------------------------------------------------------------------------------------
// odk_nms_static: odk_nms for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.  The scratch that odk_nms
// takes from the caller lives in a file-scope device array instead: ODK_NMS_STATIC_BYTES on every GPU that loads the
// library, held until the process exits.
//  * In host code `g_nms_static_scratch` is only the host-side stand-in of the symbol; the device address is looked up
//    once per device with cudaGetSymbolAddress (with that device current) and cached.
//  * Nothing in the scratch has to start at zero: collect_pieces_kernel writes the count of every piece and the keys of
//    every piece it counts before nms_image_kernel reads them, and the score histogram is not used by this path.  The
//    array keeps the previous call's contents.
//  * Every call on a device shares the one array: calls that overlap on the GPU (different streams, no ordering) race.
// ---------------------------------------------------------------------------------------------------------------------
#define ODK_NMS_STATIC_MAX_N 65536
#define ODK_NMS_STATIC_BYTES (640 * 1024)  // nms_layout(1, 65536): 8 KB counts + 512 KB keys + 4 KB histogram slot, 256-byte aligned
static __device__ __align__(256) unsigned char g_nms_static_scratch[ODK_NMS_STATIC_BYTES];

static int nms_static_scratch(void** out) {
  constexpr int kMaxDevices = 64;
  static void* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  ODK_CUDA(cudaGetDevice(&dev));
  ODK_REQUIRE(dev >= 0 && dev < kMaxDevices, ODK_ERR_UNSUPPORTED, "nms_static: device ordinal %d >= %d", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    ODK_CUDA(cudaGetSymbolAddress(&p, g_nms_static_scratch));
    ODK_REQUIRE(p != nullptr && ((uintptr_t)p & 255) == 0, ODK_ERR_CUDA, "nms_static: bad address %p for the scratch symbol", p);
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return ODK_OK;
}

extern "C" size_t odk_nms_static_max_boxes(void) { return ODK_NMS_STATIC_MAX_N; }

extern "C" int odk_nms_static(const float* boxes, const float* scores, int n, int top_k, float iou_thr, int iou_ge, int32_t* out_index,
                              int32_t* out_count, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(out_index && out_count, ODK_ERR_INVALID_ARGUMENT, "nms_static: null output pointer");
  ODK_REQUIRE(n >= 0, ODK_ERR_INVALID_ARGUMENT, "nms_static: n=%d < 0", n);
  ODK_REQUIRE(n <= ODK_NMS_STATIC_MAX_N, ODK_ERR_UNSUPPORTED,
              "nms_static: n=%d boxes exceed the built-in scratch (at most %d boxes; use odk_nms with a caller workspace)", n,
              ODK_NMS_STATIC_MAX_N);
  if (n == 0) {
    ODK_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int32_t), stream));
    return ODK_OK;
  }
  ODK_REQUIRE(nms_layout(nullptr, 1, n).bytes <= (size_t)ODK_NMS_STATIC_BYTES, ODK_ERR_UNSUPPORTED,
              "nms_static: n=%d needs %zu bytes of scratch > the built-in %d", n, nms_layout(nullptr, 1, n).bytes, ODK_NMS_STATIC_BYTES);
  void* scratch = nullptr;
  if (int rc = nms_static_scratch(&scratch)) return rc;
  if (int rc = nms_check(1, n, 1, top_k < 0 ? 0 : top_k, scratch, ODK_NMS_STATIC_BYTES)) return rc;
  ODK_REQUIRE(boxes && scores, ODK_ERR_INVALID_ARGUMENT, "nms_static: null input pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "nms_static: boxes must be 16-byte aligned");
  NmsWorkspace w = nms_layout(scratch, 1, n);
  if (int rc = collect(w, 1, n, 1, nullptr, scores, 0.f, 1, stream)) return rc;
  const int limit = iou_ge ? (top_k < 2 ? 2 : top_k) : (top_k < 0 ? 0 : top_k);  // ml.py:612-623, as in odk_nms
  return nms_run(w, nullptr, 1, n, kCollectPiece, reinterpret_cast<const float4*>(boxes), nullptr, iou_thr, 0, limit, iou_ge, nullptr, nullptr,
                 out_index, out_count, stream);
}
