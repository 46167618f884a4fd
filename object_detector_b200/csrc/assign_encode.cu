// K3 — anchor assignment: PriorBoxes._assign_boxes (dl/od_pb.py:315-363), plus the small drop-ins
// compute_iou / encode_locs.  The y_true writer that consumes the assignment (encode_truth) lives in
// encode_rows.cu.
//
// K3 — the reference's loop over ground-truth boxes is inherently sequential (each GT removes one
//   anchor), but everything that does not depend on the "taken" state is not:
//   assign_candidates_kernel — ONE WARP PER (image, GT), grid-wide.  Primary rule (od_pb.py:336-346):
//       the anchors whose pb_grid cell contains the GT centre are found through the anchor layout
//       (floor(c*m) +-1 cells x patterns x maps — at most 9*P*maps table probes, each decided by the
//       exact fp32 in_range test on the pb_grid table), cost = sum |encode_locs| in fp64 in NumPy's
//       order; the in-cell anchors are ranked by (cost, index) and the first L ranks go to the workspace.
//   assign_greedy_kernel — ONE CTA PER IMAGE: GT order = ascending fp32 area, ties by index (argsort,
//       od_pb.py:327; SURVEY A-6); the ranked lists are staged in shared memory and one warp walks the
//       GTs in that order taking the first free anchor of each list ("taken" = an N-bit bitmap in
//       shared memory).  Exhausted prefix -> exact re-evaluation over the free in-cell anchors;
//       no free in-cell anchor -> fallback (od_pb.py:347-352): CTA-wide arg-max of the exact fp32 IoU
//       over free valid anchors, first maximum wins.
//   Emits pb_indices (ascending anchor id) / gt_indices / status.
#include <float.h>

#include "odk_common.cuh"
#include "odk_sort.cuh"

namespace {


struct AssignArgs {
  AnchorView av;
  int batch, gmax;
  const float4* gt_boxes;
  const int32_t* gt_classes;  // may be null (odk_assign)
  const int32_t* gt_count;
  long long* pb_indices;
  long long* gt_indices;
  int32_t* status;
  int list_len;  // ranked candidates remembered per GT
  int use_walk;  // greedy kernel: 1 = the first ranks of every list are staged in shared memory before the transposition
  int walk_mode; // greedy kernel diagnostics (ODK_GREEDY_MODE): 0 adaptive, 1 one GT per step only, 2 batches only
  int stage_len; // greedy kernel: ranks of every GT staged in shared memory (the rest is read from the workspace on demand)
  int bitmap_words;
  int* ws_list;   // [B, Gmax, list_len] ranked in-cell anchors
  int* ws_ncand;  // [B, Gmax] in-cell valid anchors of the GT
  int* ws_nlist;  // [B, Gmax] entries stored in its ranked list
  unsigned long long* tl;  // diagnostics timeline or null
  unsigned* ws_ignore;  // [B, bitmap_words] bit a = "pb_mask[a] and some GT has IoU >= 0.5 with anchor a" (pre-zeroed) or null
};

__device__ __forceinline__ double encode_cost(const float4 box, const float4 loc, const double2 sz) {
  // sum(abs((bbox - pb_locs) / tile(pb_sizes,2) / 0.10)), fp64, left-to-right (od_pb.py:212-214, 342)
  const double d0 = fabs(__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.x, loc.x), sz.x), 0.10));
  const double d1 = fabs(__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.y, loc.y), sz.y), 0.10));
  const double d2 = fabs(__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.z, loc.z), sz.x), 0.10));
  const double d3 = fabs(__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.w, loc.w), sz.y), 0.10));
  return __dadd_rn(__dadd_rn(__dadd_rn(d0, d1), d2), d3);
}

constexpr int kGreedyThreads = 256;
constexpr int kGreedyWarps = kGreedyThreads / 32;
constexpr int kScratch = 256;  // in-cell candidates ranked per GT (VOC layout: at most 96)

// Probes the <= 9 cells around floor(c*m) of every map with the exact in_range test (od_pb.py:336) and
// calls f(anchor, ok) CONVERGENTLY (all 32 lanes, so f may use warp collectives) where ok marks the
// anchors (cell x pattern) whose cell contains the centre and whose pb_mask is set — the reference's
// `pb_mask2` before the "assignable" filter.  Returns the number of such anchors.
template <typename F>
__device__ __forceinline__ int for_each_cell_candidate(const AnchorView& av, float cx, float cy, int lane, F f) {
  int total = 0;
  const int n_probe = av.n_maps * 9;
  for (int t0 = 0; t0 < n_probe; t0 += 32) {
    const int t = t0 + lane;
    int cell_anchor = -1, m = 1;
    if (t < n_probe) {
      const int mi = t / 9, cell = t - mi * 9;
      m = av.map_size[mi];
      const int ix = (int)floorf(cx * (float)m) + (cell % 3) - 1;
      const int iy = (int)floorf(cy * (float)m) + (cell / 3) - 1;
      if (ix >= 0 && iy >= 0 && ix < m && iy < m) {
        const int a = av.map_off[mi] + iy * m + ix;  // pattern 0 of that cell; pb_grid is pattern-independent
        const float4 gr = __ldg(av.grid + a);
        if (gr.x <= cx && cx < gr.z && gr.y <= cy && cy < gr.w) cell_anchor = a;
      }
    }
    const unsigned hits = __ballot_sync(0xffffffffu, cell_anchor >= 0);
    const int n_cand = __popc(hits) * av.n_patterns;  // (hit cell, pattern) pairs, spread over the lanes
    for (int c0 = 0; c0 < n_cand; c0 += 32) {
      const int c = c0 + lane;
      const bool in_range = c < n_cand;
      const int h = in_range ? c / av.n_patterns : 0;
      const int pat = in_range ? c - h * av.n_patterns : 0;
      const int src = __fns(hits, 0, h + 1) & 31;  // lane holding the h-th hit cell (n_hit >= 1 here)
      const int base = __shfl_sync(0xffffffffu, cell_anchor, src);
      const int mm = __shfl_sync(0xffffffffu, m, src);
      const int a = base + pat * mm * mm;
      const bool ok = in_range && av.mask[a] != 0;  // od_pb.py:337
      f(a, ok);
      total += __popc(__ballot_sync(0xffffffffu, ok));
    }
  }
  return total;
}

__device__ __forceinline__ bool cand_less(double c0, int a0, double c1, int a1) {
  return c0 < c1 || (c0 == c1 && a0 < a1);  // argmin -> first minimum = lowest anchor index (od_pb.py:343)
}

// Phase A: every in-cell anchor of one GT ranked by (fp64 cost, index), ignoring "taken".
// 4 warps per CTA and <= 64 registers: all (image, GT) warps of a batch-64 x 64-GT launch are resident at once (16 CTAs
// per SM), so the kernel lasts one warp's dependent-load chain instead of two waves of it.
constexpr int kCandThreads = 128;
constexpr int kCandWarps = kCandThreads / 32;
__global__ void __launch_bounds__(kCandThreads, 8) assign_candidates_kernel(const AssignArgs p) {
  OdkTimeline tl_(ODK_TL_CANDIDATES, p.tl);
  __shared__ double s_cost[kCandWarps * kScratch];
  __shared__ int s_idx[kCandWarps * kScratch];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const long long task = (long long)blockIdx.x * kCandWarps + warp;
  if (task >= (long long)p.batch * p.gmax) return;
  const int b = (int)(task / p.gmax), g = (int)(task % p.gmax);
  const int g_count = p.gt_count[b];
  if (g_count < 0 || g_count > p.gmax || g >= g_count) return;
  const int L = p.list_len;
  double* wc = s_cost + (size_t)warp * kScratch;
  int* wi = s_idx + (size_t)warp * kScratch;
  int* list = p.ws_list + ((size_t)b * p.gmax + g) * L;

  const float4 box = p.gt_boxes[(size_t)b * p.gmax + g];
  const float cx = __fmul_rn(__fadd_rn(box.z, box.x), 0.5f);  // ml.py:485-488
  const float cy = __fmul_rn(__fadd_rn(box.w, box.y), 0.5f);
  int filled = 0;
  const int total = for_each_cell_candidate(p.av, cx, cy, lane, [&](int a, bool ok) {
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const int pos = filled + __popc(m & ((1u << lane) - 1u));
      if (pos < kScratch) {
        wc[pos] = encode_cost(box, __ldg(p.av.locs + a), p.av.sizes[a]);
        wi[pos] = a;
      }
    }
    filled += __popc(m);
  });
  __syncwarp();
  int stored = 0;
  if (total <= kScratch) {  // ranks are global only if every candidate made it into the scratch
    // rank = number of candidates that order before this one.  A lane ranks up to four entries in one pass over the
    // scratch (shared broadcast reads, unrolled), instead of one pass per entry.
    for (int e0 = 0; e0 < total; e0 += 128) {
      double ce[4];
      int ae[4], rk[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = e0 + u * 32 + lane;
        ce[u] = e < total ? wc[e] : 0.0;
        ae[u] = e < total ? wi[e] : 0;
        rk[u] = 0;
      }
#pragma unroll 4
      for (int q = 0; q < total; ++q) {
        const double cq = wc[q];
        const int aq = wi[q];
#pragma unroll
        for (int u = 0; u < 4; ++u) rk[u] += cand_less(cq, aq, ce[u], ae[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (e0 + u * 32 + lane < total && rk[u] < L) { ODK_ASSERT(rk[u] >= 0 && ae[u] >= 0 && ae[u] < p.av.n); list[rk[u]] = ae[u]; }
    }
    stored = min(total, L);
  }
  if (lane == 0) {
    p.ws_ncand[(size_t)b * p.gmax + g] = total;
    p.ws_nlist[(size_t)b * p.gmax + g] = stored;
  }
}

// Phase A': the ignore mask, scattered per GT — its own kernel (one warp per (image, GT) again) because nothing in it
// depends on the ranking above and only encode_rows consumes it: the two chains rank -> greedy walk and
// ignore -> rows run side by side (odk_assign_encode forks at entry) instead of one after the other inside one warp.
__global__ void __launch_bounds__(kCandThreads, 8) assign_ignore_kernel(const AssignArgs p) {
  OdkTimeline tl_(ODK_TL_IGNORE, p.tl);
  __shared__ int4 s_geo[kCandWarps * 32];
  __shared__ int s_end[kCandWarps * 32];
  __shared__ int s_nx[kCandWarps * 32];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const long long task = (long long)blockIdx.x * kCandWarps + warp;
  if (task >= (long long)p.batch * p.gmax) return;
  const int b = (int)(task / p.gmax), g = (int)(task % p.gmax);
  const int g_count = p.gt_count[b];
  if (g_count < 0 || g_count > p.gmax || g >= g_count) return;
  const float4 box = p.gt_boxes[(size_t)b * p.gmax + g];
  const float cx = __fmul_rn(__fadd_rn(box.z, box.x), 0.5f);  // ml.py:485-488
  const float cy = __fmul_rn(__fadd_rn(box.w, box.y), 0.5f);
  // ---- pb_valids[iou >= 0.5] = False (od_pb.py:355-357; order-independent
  //      closed form, SURVEY A-4).  Only anchors that CAN reach IoU 0.5 are visited:
  //        IoU >= 1/2  =>  I >= (A+B)/3 and min(A,B)/max(A,B) >= 1/2;  with I <= iw*min(hg,ha) and
  //        iw <= (wg+wa)/2 - |dcx|  this gives |dcx| <= (wg+wa)/6 (same in y).
  //      The cell ranges below are that bound with 2% + one cell of slack; every visited anchor is then
  //      decided by the exact fp32 expression.
  //      All (map, pattern) combinations that can reach IoU 0.5 are flattened into ONE cell index space (descriptors in
  //      the warp's scratch, offsets by a warp scan), so the anchor loads of different combinations are independent
  //      and overlap instead of costing one L2 round trip per combination.
  unsigned* ignore = p.ws_ignore + (size_t)b * p.bitmap_words;
  const float wg = __fsub_rn(box.z, box.x), hg = __fsub_rn(box.w, box.y);
  const float ga = __fmul_rn(wg, hg);
  const int n_combo = p.av.n_maps * p.av.n_patterns;
  int* d_end = s_end + warp * 32;                // [32] inclusive cell-count scan
  int4* d_geo = s_geo + warp * 32;               // [32] {base, m, ix_lo, iy_lo}
  int* d_nx = s_nx + warp * 32;                  // [32]
  for (int c0 = 0; c0 < n_combo; c0 += 32) {
    // lane i owns (map, pattern) combination c0+i: its cell-(0,0) anchor gives the size shared by all its cells
    int my_m = 1, my_base = 0;
    float4 l0 = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool have = c0 + lane < n_combo;
    if (have) {
      const int mi = (c0 + lane) / p.av.n_patterns, pat = (c0 + lane) % p.av.n_patterns;
      my_m = p.av.map_size[mi];
      my_base = p.av.map_off[mi] + pat * my_m * my_m;
      l0 = __ldg(p.av.locs + my_base);
    }
    const float wa = l0.z - l0.x, ha = l0.w - l0.y;
    const float aa = wa * ha;  // its area decides whether IoU 0.5 is reachable at all
    const float rx = (wg + wa) * 0.17f + 1e-4f, ry = (hg + ha) * 0.17f + 1e-4f;
    const float c0x = 0.5f * (l0.x + l0.z), c0y = 0.5f * (l0.y + l0.w);
    const float m = (float)my_m;
    // anchor centres sit at c0 + i/m (to fp32 rounding); 0.02 of a cell on top of the 2% absorbs that rounding
    const int ix_lo = max(0, (int)ceilf((cx - rx - c0x) * m - 0.02f));
    const int ix_hi = min(my_m - 1, (int)floorf((cx + rx - c0x) * m + 0.02f));
    const int iy_lo = max(0, (int)ceilf((cy - ry - c0y) * m - 0.02f));
    const int iy_hi = min(my_m - 1, (int)floorf((cy + ry - c0y) * m + 0.02f));
    const int nx = ix_hi - ix_lo + 1, ny = iy_hi - iy_lo + 1;
    const bool todo = have && ga >= 0.45f * aa && ga <= 2.2f * aa && nx > 0 && ny > 0;
    const int cnt = todo ? nx * ny : 0;
    int end = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, end, d);
      if (lane >= d) end += o;
    }
    const int n_cells = __shfl_sync(0xffffffffu, end, 31);
    __syncwarp();  // earlier readers of the scratch are done
    d_end[lane] = end;
    d_geo[lane] = make_int4(my_base, my_m, ix_lo, iy_lo);
    d_nx[lane] = nx > 0 ? nx : 1;
    __syncwarp();
#pragma unroll 4
    for (int t = lane; t < n_cells; t += 32) {
      int j = 0;  // first combination whose inclusive end exceeds t (ends are non-decreasing)
#pragma unroll
      for (int step = 16; step > 0; step >>= 1)
        if (d_end[j + step - 1] <= t) j += step;
      const int4 geo = d_geo[j];
      const int w = d_nx[j];
      const int cell = t - (j > 0 ? d_end[j - 1] : 0);
      const int iy = geo.w + cell / w, ix = geo.z + cell % w;
      const int a = geo.x + iy * geo.y + ix;
      const unsigned char mk = __ldg(p.av.mask + a);
      const float4 lc = __ldg(p.av.locs + a);
      ODK_ASSERT(a >= 0 && a < p.av.n);
      if (mk && odk_iou_passes<true>(box, ga, lc, odk_area(lc), 0.5f)) atomicOr(ignore + (a >> 5), 1u << (a & 31));
    }
  }
}

// Phase B: the greedy walk, one CTA per image.
//
// The reference's loop is sequential — GT r takes the best anchor no earlier GT took — but consecutive GTs rarely
// compete for the same anchor, so the walk is OPTIMISTIC: warp 0 looks at 32 consecutive GTs at once (a lane each) and
// at the best-ranked anchor of each.  Lanes whose anchor is free in the bitmap as it stood before the batch, up to the
// first lane whose anchor is taken or collides with that of a lane in front of it (one match.any), commit together:
// their choices are pairwise distinct and nothing was taken in between, so each is exactly what the sequential loop
// would have picked.  The GT the batch stopped at is then decided the way the sequential walk does it — its staged
// ranks side by side, a lane each, first free one wins — and the next batch starts behind it.  A crowded image
// degenerates to one GT per step, as before; spread-out GTs (the stress shape: 512 per image) go 32 at a time.
//
// Shared-memory layout of what the walk reads: walk_t[k][r] = rank-k anchor of the r-th GT of the walk (rank-major,
// row stride odd: lanes = consecutive r are conflict-free, and so is the transposing store), -1 beyond the list.
__global__ void __launch_bounds__(kGreedyThreads) assign_greedy_kernel(const AssignArgs p) {
  OdkTimeline tl_(ODK_TL_GREEDY, p.tl);
  extern __shared__ __align__(16) unsigned char greedy_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int b = blockIdx.x;
  const int L = p.list_len;
  unsigned char* sp = greedy_smem;
  const int SL = p.stage_len;  // ranks of every list staged in shared memory (a multiple of 32; deeper ranks come from the workspace)
  constexpr int WR = 32;       // ranks in the walk table: what a lane-per-rank step looks at
  const int Gp = p.gmax | 1;   // row stride of walk_t (odd)
  const int* glist = p.ws_list + (size_t)b * p.gmax * L;
  float4* gbox = reinterpret_cast<float4*>(sp);            sp += (size_t)p.gmax * 16;
  int* list = reinterpret_cast<int*>(sp);                  sp += p.use_walk ? (size_t)p.gmax * SL * 4 : 0;  // 16-byte aligned (SL % 8 == 0)
  float* garea = reinterpret_cast<float*>(sp);             sp += (size_t)p.gmax * 4;
  int* order = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;
  int* pick = reinterpret_cast<int*>(sp);                  sp += (size_t)p.gmax * 4;
  int* ncand = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;
  int* nlist = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;
  unsigned* taken = reinterpret_cast<unsigned*>(sp);       sp += odk_align_up((size_t)p.bitmap_words * 4, 16);
  int* walk_t = reinterpret_cast<int*>(sp);  // [WR][Gp]; before / after the walk also the scratch of the two sorts
  unsigned long long* sort_keys = reinterpret_cast<unsigned long long*>(sp);
  __shared__ int s_err;
  __shared__ float s_best_v[kGreedyWarps];
  __shared__ int s_best_a[kGreedyWarps];
  __shared__ int s_state[2];  // [0] = GT needing the CTA-wide fallback (-1: none, -2: finished), [1] = its result

  long long* out_pb = p.pb_indices + (size_t)b * p.gmax;
  long long* out_gt = p.gt_indices + (size_t)b * p.gmax;
  const int g_count = p.gt_count[b];
  const bool bad_count = g_count < 0 || g_count > p.gmax;
  const int G = bad_count ? 0 : g_count;
  if (threadIdx.x == 0) s_err = bad_count ? ODK_IMG_BAD_COUNT : ODK_IMG_OK;
  __syncthreads();

  for (int g = threadIdx.x; g < G; g += kGreedyThreads) {
    const float4 bx = p.gt_boxes[(size_t)b * p.gmax + g];
    gbox[g] = bx;
    garea[g] = odk_area(bx);
    const int nc = p.ws_ncand[(size_t)b * p.gmax + g];
    ncand[g] = nc;
    nlist[g] = p.ws_nlist[(size_t)b * p.gmax + g];
    // od_pb.py:338 (a centre inside no valid cell) / od_pb.py:331-332 (more GTs than valid anchors): whichever GT
    // trips first, the image fails as a whole, so both are decided up front
    if (nc == 0 || g >= p.av.n_valid) atomicMax(&s_err, ODK_IMG_ENCODE_ERROR);
    if (p.gt_classes != nullptr) {
      const int c = p.gt_classes[(size_t)b * p.gmax + g];
      if (c < 0 || c >= p.av.num_classes) atomicMax(&s_err, ODK_IMG_BAD_CLASS);  // od_pb.py:304
    }
  }
  for (int w = threadIdx.x; w < p.bitmap_words; w += kGreedyThreads) taken[w] = 0u;
  if (p.use_walk) {  // stage the first SL ranks of every list: 16-byte cp.async, all in flight at once (entries beyond nlist[g] are never read)
    const int per = SL / 4;
    for (int i = threadIdx.x; i < G * per; i += kGreedyThreads) {
      const int g = i / per, k4 = i - g * per;
      odk_cp_async<16>(reinterpret_cast<int4*>(list + (size_t)g * SL) + k4, reinterpret_cast<const int4*>(glist + (size_t)g * L) + k4);
    }
    odk_cp_async_commit();
  }
  __syncthreads();
  // Descending sort of 64-bit keys in the (still unused) walk_t area: used for the two rankings when G is large
  // (the all-pairs count below costs G^2 / 256 steps per thread); needs next_pow2(G) * 8 <= WR * Gp * 4 bytes: always.
  int n_sort = 2;
  while (n_sort < G) n_sort <<= 1;
  const bool by_sort = G > 128;
  if (s_err == ODK_IMG_OK) {
    // ascending-area rank, ties by index (od_pb.py:327)
    if (by_sort) {
      for (int i = threadIdx.x; i < n_sort; i += kGreedyThreads)  // descending (~area, ~index) == ascending (area, index)
        sort_keys[i] = i < G ? ((unsigned long long)(~odk_flip_score(garea[i])) << 32) | (unsigned long long)(0xffffffffu - (unsigned)i) : 0ull;
      __syncthreads();
      bitonic_sort_shared(sort_keys, n_sort, 0, 2, n_sort);
      for (int r = threadIdx.x; r < G; r += kGreedyThreads) order[r] = (int)(0xffffffffu - (unsigned)(sort_keys[r] & 0xffffffffull));
    } else {
      for (int g = threadIdx.x; g < G; g += kGreedyThreads) {
        const float ag = garea[g];
        int r = 0;
        for (int j = 0; j < G; ++j) {
          const float aj = garea[j];
          r += (aj < ag) || (aj == ag && j < g);
        }
        order[r] = g;
      }
    }
  }
  odk_cp_async_wait<0>();
  __syncthreads();
  if (s_err == ODK_IMG_OK) {
    // transpose what the walk reads: lanes k of a GT read its staged ranks (conflict-free), the store has stride Gp (odd)
    for (int i = threadIdx.x; i < G * WR; i += kGreedyThreads) {
      const int r = i / WR, k = i - r * WR;
      const int g = order[r];
      ODK_ASSERT(g >= 0 && g < G && (size_t)k * Gp + r < (size_t)WR * Gp);
      walk_t[(size_t)k * Gp + r] = (k < nlist[g]) ? (p.use_walk ? list[(size_t)g * SL + k] : __ldg(glist + (size_t)g * L + k)) : -1;
    }
  }
  __syncthreads();

  // Warp 0 walks the GTs; the other warps sleep at the barrier below and wake only to help with a
  // CTA-wide IoU arg-max when some GT has no free in-cell anchor (the crowded / stress case).
  int err = s_err;
  int r = 0;
  while (true) {
    int need = -2;
    if (warp == 0) {
      if (!err) {
        int serial = p.walk_mode == 1 ? 0x3fffffff : 0;  // GTs left to decide one at a time before the next batch attempt
        while (r < G) {
          if (serial == 0) {
            // ---- one optimistic batch: lane j looks at the BEST-ranked anchor of GT r + j ----
            const int rr = r + lane;
            const bool active = rr < G;
            const int a = active ? walk_t[rr] : -1;  // rank 0 (present whenever the GT has a ranked list)
            const bool found = a >= 0 && !((taken[a >> 5] >> (a & 31)) & 1u);
            // lanes without a free first choice get a value of their own, so they never match anybody
            const unsigned peers = __match_any_sync(0xffffffffu, found ? a : -1 - lane);
            const bool collides = found && (__ffs(peers) - 1) != lane;  // an earlier lane of the batch wants the same anchor
            const unsigned stop = __ballot_sync(0xffffffffu, active && (!found || collides));
            const unsigned act = __ballot_sync(0xffffffffu, active);
            const int n_ok = stop ? (__ffs(stop) - 1) : __popc(act);   // lanes [0, n_ok) commit
            if (lane < n_ok) {
              ODK_ASSERT(a >= 0 && a < p.av.n && rr < G && !((taken[a >> 5] >> (a & 31)) & 1u));
              atomicOr(&taken[a >> 5], 1u << (a & 31));  // several lanes may share a word
              pick[rr] = a;  // by walk position; translated to the GT index when the result is written
            }
            __syncwarp();
            r += n_ok;
            if (!stop) continue;
            // a batch that committed nothing: a crowded spot, where a batch attempt costs more than it saves
            if (n_ok == 0 && p.walk_mode != 2) serial = 8;
          }
          // ---- one GT per step: the staged ranks of GT r side by side, a lane each — the first free one wins.  Runs for
          //      the GT a batch stopped at and, in a crowded spot, for the next `serial` GTs; the ranks of the following
          //      GT are fetched a step ahead, so the chain per GT is bitmap probe -> ballot -> bitmap update ----
          {
            int x = walk_t[(size_t)lane * Gp + r];
            bool decided = true;
            while (true) {
              const int xn = (r + 1 < G) ? walk_t[(size_t)lane * Gp + r + 1] : -1;
              const bool is_free = x >= 0 && !((taken[x >> 5] >> (x & 31)) & 1u);
              const unsigned fm = __ballot_sync(0xffffffffu, is_free);
              if (!fm) { decided = false; break; }
              if (lane == __ffs(fm) - 1) {
                ODK_ASSERT(x >= 0 && x < p.av.n && r < G);
                taken[x >> 5] |= 1u << (x & 31);
                pick[r] = x;
              }
              __syncwarp();
              ++r;
              x = xn;
              if (serial == 0 || r >= G) break;
              --serial;
            }
            if (decided) continue;
          }
          // ---- no free anchor among its staged ranks: the rare serial paths ----
          const int g = order[r];
          const int nl = nlist[g];
          const int nc = ncand[g];
          int best_a = -1;
          for (int k0 = WR; k0 < nl && best_a < 0; k0 += 32) {
            const int k = k0 + lane;
            // deeper ranks (crowded spots): from the staged lists, beyond those straight from the workspace
            const int aa = (k < nl) ? ((p.use_walk && k < SL) ? list[(size_t)g * SL + k] : __ldg(glist + (size_t)g * L + k)) : -1;
            const bool is_free = aa >= 0 && !((taken[aa >> 5] >> (aa & 31)) & 1u);
            const unsigned fm2 = __ballot_sync(0xffffffffu, is_free);
            if (fm2) best_a = __shfl_sync(0xffffffffu, aa, __ffs(fm2) - 1);
          }
          if (best_a < 0 && nc > nl) {
            // the ranked prefix is exhausted but more in-cell anchors exist: exact re-evaluation over the free ones
            const float4 box = gbox[g];
            const float cx = __fmul_rn(__fadd_rn(box.z, box.x), 0.5f);
            const float cy = __fmul_rn(__fadd_rn(box.w, box.y), 0.5f);
            double mc = DBL_MAX;
            int ma = INT_MAX;
            for_each_cell_candidate(p.av, cx, cy, lane, [&](int aa, bool ok) {
              if (!ok || ((taken[aa >> 5] >> (aa & 31)) & 1u)) return;  // od_pb.py:339
              const double c = encode_cost(box, __ldg(p.av.locs + aa), p.av.sizes[aa]);
              if (cand_less(c, aa, mc, ma)) { mc = c; ma = aa; }
            });
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
              const double oc = __shfl_xor_sync(0xffffffffu, mc, d);
              const int oa = __shfl_xor_sync(0xffffffffu, ma, d);
              if (cand_less(oc, oa, mc, ma)) { mc = oc; ma = oa; }
            }
            if (ma != INT_MAX) best_a = ma;
          }
          if (best_a < 0) { need = g; break; }  // -> CTA-wide fallback
          if (lane == 0) {
            ODK_ASSERT(best_a >= 0 && best_a < p.av.n && r < G);
            taken[best_a >> 5] |= 1u << (best_a & 31);
            pick[r] = best_a;
          }
          __syncwarp();
          ++r;
        }
      }
      if (lane == 0) s_state[0] = need;
    }
    __syncthreads();
    need = s_state[0];
    if (need == -2) break;
    {
      // fallback: arg-max IoU over the free valid anchors, first maximum wins (od_pb.py:347-352)
      const float4 box = gbox[need];
      const float ga = garea[need];
      float best_v = -1.0f;
      int fa = INT_MAX;
      for (int a = threadIdx.x; a < p.av.n; a += kGreedyThreads) {
        if (!p.av.mask[a] || ((taken[a >> 5] >> (a & 31)) & 1u)) continue;
        const float4 lc = __ldg(p.av.locs + a);
        const float v = odk_iou_value(box, ga, lc, odk_area(lc));
        if (v > best_v) { best_v = v; fa = a; }  // ascending a per thread: the first maximum stays
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best_v, d);
        const int oa = __shfl_xor_sync(0xffffffffu, fa, d);
        if (ov > best_v || (ov == best_v && oa < fa)) { best_v = ov; fa = oa; }
      }
      if (lane == 0) { s_best_v[warp] = best_v; s_best_a[warp] = fa; }
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int w = 1; w < kGreedyWarps; ++w)
          if (s_best_v[w] > best_v || (s_best_v[w] == best_v && s_best_a[w] < fa)) { best_v = s_best_v[w]; fa = s_best_a[w]; }
        s_state[1] = fa;
        if (fa != INT_MAX) {
          taken[fa >> 5] |= 1u << (fa & 31);
          pick[r] = fa;  // thread 0 belongs to warp 0, whose r is the walk position of the GT in need
        }
      }
      __syncthreads();
      if (s_state[1] == INT_MAX) err = ODK_IMG_ENCODE_ERROR;
      ++r;  // warp 0 resumes with the next GT (other warps' r is unused)
    }
  }
  // every thread holds the same `err` (it changes only in the CTA-wide fallback); pick[] is complete and visible (the
  // loop's last barrier).
  if (err) {
    for (int g = threadIdx.x; g < p.gmax; g += kGreedyThreads) { out_pb[g] = -1; out_gt[g] = -1; }
    if (threadIdx.x == 0) p.status[b] = err;
    return;
  }
  // np.where(pb_assigned_gt >= 0): ascending anchor index (od_pb.py:359); picks are unique
  if (by_sort) {
    for (int i = threadIdx.x; i < n_sort; i += kGreedyThreads)  // descending ~pick == ascending pick; the GT rides in the low word
      sort_keys[i] = i < G ? ((unsigned long long)(~(unsigned)pick[i]) << 32) | (unsigned long long)(unsigned)order[i] : 0ull;
    __syncthreads();
    bitonic_sort_shared(sort_keys, n_sort, 0, 2, n_sort);
    for (int rk = threadIdx.x; rk < G; rk += kGreedyThreads) {
      const unsigned long long k = sort_keys[rk];
      out_pb[rk] = (long long)(~(unsigned)(k >> 32));
      out_gt[rk] = (long long)(unsigned)(k & 0xffffffffull);
    }
  } else {
    for (int q = threadIdx.x; q < G; q += kGreedyThreads) {  // q: walk position
      const int mine = pick[q];
      int rk = 0;
#pragma unroll 4
      for (int j = 0; j < G; ++j) rk += pick[j] < mine;
      ODK_ASSERT(rk < G && mine >= 0 && mine < p.av.n);
      out_pb[rk] = mine;
      out_gt[rk] = order[q];
    }
  }
  for (int g = G + threadIdx.x; g < p.gmax; g += kGreedyThreads) { out_pb[g] = -1; out_gt[g] = -1; }
  if (threadIdx.x == 0) p.status[b] = ODK_IMG_OK;
}

__global__ void compute_iou_kernel(const float4* __restrict__ a, int na, const float4* __restrict__ b, int nb, float* __restrict__ out) {
  const long long total = (long long)na * nb;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int ia = (int)(i / nb), ib = (int)(i % nb);
    const float4 ba = __ldg(a + ia), bb = __ldg(b + ib);
    out[i] = odk_iou_value(ba, odk_area(ba), bb, odk_area(bb));
  }
}

__global__ void encode_locs_kernel(AnchorView av, int k, const float4* __restrict__ boxes, const int32_t* __restrict__ ix,
                                   double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= k) return;
  const int a = ix[i];
  if (a < 0 || a >= av.n) {
    out[4 * i] = out[4 * i + 1] = out[4 * i + 2] = out[4 * i + 3] = __longlong_as_double(0x7ff8000000000000ll);
    return;
  }
  const float4 box = boxes[i];
  const float4 loc = av.locs[a];
  const double2 sz = av.sizes[a];
  out[4 * i + 0] = __ddiv_rn(__ddiv_rn((double)__fsub_rn(box.x, loc.x), sz.x), 0.10);
  out[4 * i + 1] = __ddiv_rn(__ddiv_rn((double)__fsub_rn(box.y, loc.y), sz.y), 0.10);
  out[4 * i + 2] = __ddiv_rn(__ddiv_rn((double)__fsub_rn(box.z, loc.z), sz.x), 0.10);
  out[4 * i + 3] = __ddiv_rn(__ddiv_rn((double)__fsub_rn(box.w, loc.w), sz.y), 0.10);
}

// Ranked in-cell anchors remembered per GT in the workspace.  The VOC layout has at most 96 anchors whose cell holds a
// centre (SURVEY A-1), so 128 ranks make the exact re-evaluation path a safeguard; only the first 32 are staged in
// shared memory by the greedy kernel.
int assign_list_len(int /*gmax*/) { return 128; }

}  // namespace

// ---- internal interface shared with encode_rows.cu ----------------------------------------------------------
size_t odk_assign_lists_bytes(int batch, int gmax) {
  return odk_align_up(sizeof(int) * (size_t)batch * gmax * ((size_t)assign_list_len(gmax) + 2), 256);
}

// Candidate ranking (+ the ignore-mask scatter when `ignore` is given: [B, ceil(N/32)] words, pre-zeroed), then
// the greedy walk; `lists_ws` holds odk_assign_lists_bytes() bytes.  Two launches so that a caller can fork other
// work between them.
static int fill_args(AssignArgs& args, const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes,
                     const int32_t* gt_count, int gmax, int64_t* pb_indices, int64_t* gt_indices, int32_t* status, void* lists_ws,
                     unsigned* ignore) {
  args.av = odk_view(a);
  args.batch = batch; args.gmax = gmax;
  args.gt_boxes = reinterpret_cast<const float4*>(gt_boxes);
  args.gt_classes = gt_classes; args.gt_count = gt_count;
  args.pb_indices = reinterpret_cast<long long*>(pb_indices);
  args.gt_indices = reinterpret_cast<long long*>(gt_indices);
  args.status = status;
  args.bitmap_words = (a->n + 31) / 32;
  args.list_len = assign_list_len(gmax);
  args.ws_list = static_cast<int*>(lists_ws);
  args.ws_ncand = args.ws_list + (size_t)batch * gmax * args.list_len;
  args.ws_nlist = args.ws_ncand + (size_t)batch * gmax;
  args.ws_ignore = ignore;
  args.use_walk = 0;
  args.walk_mode = 0;
  args.stage_len = 32;
  args.tl = odk_timeline_ptr();
  return ODK_OK;
}

int odk_launch_assign_rank(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, void* lists_ws,
                           cudaStream_t stream) {
  AssignArgs args;
  fill_args(args, a, batch, gt_boxes, nullptr, gt_count, gmax, nullptr, nullptr, nullptr, lists_ws, nullptr);
  const long long tasks = (long long)batch * gmax;
  ODK_MAX_CARVEOUT_ONCE(assign_candidates_kernel);
  {
    OdkProfileScope prof_("assign_candidates_kernel", stream);
    odk_launch_high(assign_candidates_kernel, dim3((unsigned)((tasks + kCandWarps - 1) / kCandWarps)), dim3(kCandThreads), 0, stream, args);
  }
  ODK_LAUNCH_CHECK("assign_candidates_kernel");
  return ODK_OK;
}

// `ignore`: [B, ceil(N/32)] words, zeroed here (a memset node in front of the scatter kernel).
int odk_launch_assign_ignore(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, unsigned* ignore,
                             cudaStream_t stream) {
  AssignArgs args;
  fill_args(args, a, batch, gt_boxes, nullptr, gt_count, gmax, nullptr, nullptr, nullptr, nullptr, ignore);
  ODK_CUDA(cudaMemsetAsync(ignore, 0, sizeof(unsigned) * (size_t)batch * args.bitmap_words, stream));
  const long long tasks = (long long)batch * gmax;
  ODK_MAX_CARVEOUT_ONCE(assign_ignore_kernel);
  {
    OdkProfileScope prof_("assign_ignore_kernel", stream);
    odk_launch_high(assign_ignore_kernel, dim3((unsigned)((tasks + kCandWarps - 1) / kCandWarps)), dim3(kCandThreads), 0, stream, args);
  }
  ODK_LAUNCH_CHECK("assign_ignore_kernel");
  return ODK_OK;
}

// Both on one stream (the fused loss entries): ranked lists, and the ignore bitmap when `ignore` is given.
int odk_launch_assign_candidates(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, void* lists_ws,
                                 unsigned* ignore, cudaStream_t stream) {
  if (int rc = odk_launch_assign_rank(a, batch, gt_boxes, gt_count, gmax, lists_ws, stream)) return rc;
  if (ignore == nullptr) return ODK_OK;
  return odk_launch_assign_ignore(a, batch, gt_boxes, gt_count, gmax, ignore, stream);
}

int odk_launch_assign_greedy(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                             int gmax, int64_t* pb_indices, int64_t* gt_indices, int32_t* status, void* lists_ws, cudaStream_t stream) {
  AssignArgs args;
  fill_args(args, a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, lists_ws, nullptr);
  // boxes, staged ranks, five per-GT words, taken bitmap, the rank-major walk table (also the sort scratch)
  // all 128 ranks staged while that stays under 64 KB (Gmax <= 128), fewer for larger Gmax
  args.stage_len = gmax <= 128 ? 128 : (gmax <= 256 ? 64 : 32);
  size_t smem = (size_t)gmax * (16 + 20 + 4 * (size_t)args.stage_len) + odk_align_up((size_t)args.bitmap_words * 4, 16) +
                (size_t)32 * (gmax | 1) * 4 + 16;
  static const int walk_mode = getenv("ODK_GREEDY_MODE") ? atoi(getenv("ODK_GREEDY_MODE")) : 0;
  args.walk_mode = walk_mode;
  args.use_walk = 1;  // ranked lists staged in shared memory with cp.async while the GTs are ranked by area
  if (smem > 200 * 1024) {  // very large Gmax: no staging area, the walk table is filled straight from the workspace
    args.use_walk = 0;
    smem -= (size_t)gmax * 4 * (size_t)args.stage_len;
  }
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "assign: N=%d / Gmax=%d need %zu bytes of shared memory per image", a->n, gmax, smem);
  ODK_CUDA(cudaFuncSetAttribute(assign_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  ODK_MAX_CARVEOUT_ONCE(assign_greedy_kernel);
  {
    OdkProfileScope prof_("assign_greedy_kernel", stream);
    odk_launch_high(assign_greedy_kernel, dim3(batch), dim3(kGreedyThreads), smem, stream, args);
  }
  ODK_LAUNCH_CHECK("assign_greedy_kernel");
  return ODK_OK;
}

int odk_assign_check(const odk_anchors* a, int batch, const void* gt_boxes, const void* gt_count, int gmax, const void* pb, const void* gt,
                     const void* status) {
  ODK_REQUIRE(a != nullptr, ODK_ERR_INVALID_ARGUMENT, "assign: null anchors handle");
  ODK_REQUIRE(batch >= 1 && gmax >= 1, ODK_ERR_INVALID_ARGUMENT, "assign: batch=%d gmax=%d must be >= 1", batch, gmax);
  ODK_REQUIRE(gt_boxes && gt_count && pb && gt && status, ODK_ERR_INVALID_ARGUMENT, "assign: null pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "assign: gt_boxes must be 16-byte aligned");
  return ODK_OK;
}

size_t odk_assign_workspace_bytes_impl(int n, int batch, int gmax) {
  return odk_assign_lists_bytes(batch, gmax) + odk_align_up(sizeof(unsigned) * (size_t)batch * ((n + 31) / 32), 256);
}

int odk_assign_workspace_check(const odk_anchors* a, int batch, int gmax, const void* workspace, size_t workspace_bytes) {
  const size_t need = odk_assign_workspace_bytes_impl(a->n, batch, gmax);
  ODK_REQUIRE(workspace != nullptr && workspace_bytes >= need, ODK_ERR_WORKSPACE, "assign: workspace of %zu bytes is smaller than the required %zu",
              workspace_bytes, need);
  ODK_REQUIRE(((uintptr_t)workspace & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "assign: workspace must be 16-byte aligned");
  return ODK_OK;
}

extern "C" size_t odk_assign_workspace_bytes(const odk_anchors* a, int batch, int gmax) {
  if (a == nullptr || batch < 1 || gmax < 1) return 0;
  return odk_assign_workspace_bytes_impl(a->n, batch, gmax);
}

extern "C" int odk_assign(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, int64_t* pb_indices,
                          int64_t* gt_indices, int32_t* status, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = odk_assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  if (int rc = odk_assign_workspace_check(a, batch, gmax, workspace, workspace_bytes)) return rc;
  if (int rc = odk_launch_assign_candidates(a, batch, gt_boxes, gt_count, gmax, workspace, nullptr, static_cast<cudaStream_t>(stream))) return rc;
  return odk_launch_assign_greedy(a, batch, gt_boxes, nullptr, gt_count, gmax, pb_indices, gt_indices, status, workspace,
                                  static_cast<cudaStream_t>(stream));
}

extern "C" int odk_compute_iou(const float* a, int na, const float* b, int nb, float* out, void* stream) {
  ODK_REQUIRE(na >= 1 && nb >= 1, ODK_ERR_INVALID_ARGUMENT, "compute_iou: empty input (ml.py:507-508 asserts non-empty)");
  ODK_REQUIRE(a && b && out, ODK_ERR_INVALID_ARGUMENT, "compute_iou: null pointer");
  ODK_REQUIRE((((uintptr_t)a | (uintptr_t)b) & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "compute_iou: boxes must be 16-byte aligned");
  const long long total = (long long)na * nb;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  {
    OdkProfileScope prof_("compute_iou_kernel", static_cast<cudaStream_t>(stream));
    compute_iou_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<const float4*>(a), na,
                                                                                        reinterpret_cast<const float4*>(b), nb, out);
  }
  ODK_LAUNCH_CHECK("compute_iou_kernel");
  return ODK_OK;
}

extern "C" int odk_encode_locs(const odk_anchors* a, int k, const float* boxes, const int32_t* ix, double* out, void* stream) {
  ODK_REQUIRE(a != nullptr && k >= 0, ODK_ERR_INVALID_ARGUMENT, "encode_locs: bad arguments");
  if (k == 0) return ODK_OK;
  ODK_REQUIRE(boxes && ix && out, ODK_ERR_INVALID_ARGUMENT, "encode_locs: null pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "encode_locs: boxes must be 16-byte aligned");
  {
    OdkProfileScope prof_("encode_locs_kernel", static_cast<cudaStream_t>(stream));
    encode_locs_kernel<<<(k + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(odk_view(a), k, reinterpret_cast<const float4*>(boxes), ix, out);
  }
  ODK_LAUNCH_CHECK("encode_locs_kernel");
  return ODK_OK;
}


