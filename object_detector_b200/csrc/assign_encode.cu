// K3 + K4 — training-target generation: PriorBoxes._assign_boxes + encode_truth
// (dl/od_pb.py:299-363), plus the small drop-ins compute_iou / encode_locs.
//
// K3 assign_greedy_kernel — ONE WARP PER IMAGE.  The reference's loop over ground-truth boxes is
//   inherently sequential (each GT removes one anchor), so parallelism is lanes-over-candidates and
//   warps-over-images:
//     * GT order = ascending fp32 area, ties by index (argsort, od_pb.py:327; SURVEY A-6);
//     * primary rule (od_pb.py:336-346): the anchors whose pb_grid cell contains the GT centre are
//       found through the anchor layout (floor(c*m) +-1 cells x patterns x maps — at most
//       9*P*maps table probes, each decided by the exact fp32 in_range test on the pb_grid table),
//       cost = sum |encode_locs| in fp64 in NumPy's order, warp-shuffle arg-min (lowest index wins);
//     * fallback (od_pb.py:347-352): warp-wide arg-max of the exact fp32 IoU over free valid anchors;
//     * "taken" state = an N-bit bitmap in shared memory.
//   Emits pb_indices (ascending anchor id) / gt_indices / status.
// K4 encode_rows_kernel — one CTA per (image, 128-anchor tile): computes the order-independent
//   ignore mask  pb_mask & any_g IoU(g, a) >= 0.5  (closed form of od_pb.py:355-361, SURVEY A-4)
//   with GT boxes staged in shared memory, builds the tile's y_true rows [mask, weight, obj,
//   onehot(C), locs(4)] in shared memory, patches the <= G positives, and writes the tile — a
//   contiguous 128*(3+C+4)*4-byte span of HBM — with one bulk TMA store (cp.async.bulk.global).
//   Algorithmic bytes: 4*N*(3+C+4) written per image; GT/assignment lists are a few KB.
#include <float.h>

#include "odk_common.cuh"

namespace {

constexpr int kEncodeTile = 128;

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
  int bitmap_words;
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

__global__ void __launch_bounds__(kGreedyThreads) assign_greedy_kernel(const AssignArgs p) {
  extern __shared__ __align__(16) unsigned char greedy_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int b = blockIdx.x;
  const int L = p.list_len;
  unsigned char* sp = greedy_smem;
  float4* gbox = reinterpret_cast<float4*>(sp);            sp += (size_t)p.gmax * 16;
  double* s_cost = reinterpret_cast<double*>(sp);          sp += (size_t)kGreedyWarps * kScratch * 8;
  int* s_idx = reinterpret_cast<int*>(sp);                 sp += (size_t)kGreedyWarps * kScratch * 4;
  float* garea = reinterpret_cast<float*>(sp);             sp += (size_t)p.gmax * 4;
  int* order = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;
  int* pick = reinterpret_cast<int*>(sp);                  sp += (size_t)p.gmax * 4;
  int* ncand = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;  // in-cell valid anchors of the GT
  int* nlist = reinterpret_cast<int*>(sp);                 sp += (size_t)p.gmax * 4;  // entries stored in its ranked list
  int* list = reinterpret_cast<int*>(sp);                  sp += (size_t)p.gmax * L * 4;
  unsigned* taken = reinterpret_cast<unsigned*>(sp);
  __shared__ int s_err;

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
    if (p.gt_classes != nullptr) {
      const int c = p.gt_classes[(size_t)b * p.gmax + g];
      if (c < 0 || c >= p.av.num_classes) atomicMax(&s_err, ODK_IMG_BAD_CLASS);  // od_pb.py:304
    }
  }
  for (int w = threadIdx.x; w < p.bitmap_words; w += kGreedyThreads) taken[w] = 0u;
  __syncthreads();

  if (s_err == ODK_IMG_OK) {
    // ascending-area rank, ties by index (od_pb.py:327)
    for (int g = threadIdx.x; g < G; g += kGreedyThreads) {
      const float ag = garea[g];
      int r = 0;
      for (int j = 0; j < G; ++j) {
        const float aj = garea[j];
        r += (aj < ag) || (aj == ag && j < g);
      }
      order[r] = g;
    }
    // ---- phase A (parallel over GTs, one warp each): every in-cell anchor of the GT ranked by
    //      (fp64 cost, index), ignoring "taken"; the first L ranks are kept.
    double* wc = s_cost + (size_t)warp * kScratch;
    int* wi = s_idx + (size_t)warp * kScratch;
    for (int g = warp; g < G; g += kGreedyWarps) {
      const float4 box = gbox[g];
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
        for (int e = lane; e < total; e += 32) {
          const double ce = wc[e];
          const int ae = wi[e];
          int r = 0;
          for (int q = 0; q < total; ++q) r += cand_less(wc[q], wi[q], ce, ae);
          if (r < L) list[(size_t)g * L + r] = ae;
        }
        stored = min(total, L);
      }
      if (lane == 0) { ncand[g] = total; nlist[g] = stored; }
      __syncwarp();
    }
  }
  __syncthreads();
  if (warp != 0) return;

  // ---- phase B (warp 0, sequential over GTs in area order): first free anchor of each GT's list
  int err = s_err;
  if (!err) {
    for (int r = 0; r < G; ++r) {
      const int g = order[r];
      if (r >= p.av.n_valid) { err = ODK_IMG_ENCODE_ERROR; break; }  // od_pb.py:331-332
      const int nc = ncand[g];
      if (nc == 0) { err = ODK_IMG_ENCODE_ERROR; break; }             // od_pb.py:338
      const int nl = nlist[g];
      int best_a = -1;
      for (int k0 = 0; k0 < nl && best_a < 0; k0 += 32) {
        const int k = k0 + lane;
        const int a = (k < nl) ? list[(size_t)g * L + k] : -1;
        const bool is_free = a >= 0 && !((taken[a >> 5] >> (a & 31)) & 1u);
        const unsigned fm = __ballot_sync(0xffffffffu, is_free);
        if (fm) best_a = __shfl_sync(0xffffffffu, a, __ffs(fm) - 1);
      }
      const float4 box = gbox[g];
      if (best_a < 0 && nc > nl) {
        // the ranked prefix is exhausted but more in-cell anchors exist: exact re-evaluation over the free ones
        const float cx = __fmul_rn(__fadd_rn(box.z, box.x), 0.5f);
        const float cy = __fmul_rn(__fadd_rn(box.w, box.y), 0.5f);
        double mc = DBL_MAX;
        int ma = INT_MAX;
        for_each_cell_candidate(p.av, cx, cy, lane, [&](int a, bool ok) {
          if (!ok || ((taken[a >> 5] >> (a & 31)) & 1u)) return;  // od_pb.py:339
          const double c = encode_cost(box, __ldg(p.av.locs + a), p.av.sizes[a]);
          if (cand_less(c, a, mc, ma)) { mc = c; ma = a; }
        });
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          const double oc = __shfl_xor_sync(0xffffffffu, mc, d);
          const int oa = __shfl_xor_sync(0xffffffffu, ma, d);
          if (cand_less(oc, oa, mc, ma)) { mc = oc; ma = oa; }
        }
        if (ma != INT_MAX) best_a = ma;
      }
      if (best_a < 0) {
        // fallback: arg-max IoU over the free valid anchors, first maximum wins (od_pb.py:347-352)
        float best_v = -1.0f;
        int fa = INT_MAX;
        const float ga = garea[g];
        for (int a = lane; a < p.av.n; a += 32) {
          if (!p.av.mask[a] || ((taken[a >> 5] >> (a & 31)) & 1u)) continue;
          const float4 lc = __ldg(p.av.locs + a);
          const float v = odk_iou_value(box, ga, lc, odk_area(lc));
          if (v > best_v) { best_v = v; fa = a; }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, best_v, d);
          const int oa = __shfl_xor_sync(0xffffffffu, fa, d);
          if (ov > best_v || (ov == best_v && oa < fa)) { best_v = ov; fa = oa; }
        }
        if (fa == INT_MAX) { err = ODK_IMG_ENCODE_ERROR; break; }
        best_a = fa;
      }
      if (lane == 0) {
        taken[best_a >> 5] |= 1u << (best_a & 31);
        pick[g] = best_a;
      }
      __syncwarp();
    }
  }
  if (err) {
    for (int g = lane; g < p.gmax; g += 32) { out_pb[g] = -1; out_gt[g] = -1; }
    if (lane == 0) p.status[b] = err;
    return;
  }
  // np.where(pb_assigned_gt >= 0): ascending anchor index (od_pb.py:359); picks are unique
  for (int g = lane; g < G; g += 32) {
    const int mine = pick[g];
    int r = 0;
    for (int j = 0; j < G; ++j) r += pick[j] < mine;
    out_pb[r] = mine;
    out_gt[r] = g;
  }
  for (int g = G + lane; g < p.gmax; g += 32) { out_pb[g] = -1; out_gt[g] = -1; }
  if (lane == 0) p.status[b] = ODK_IMG_OK;
}

struct EncodeArgs {
  AnchorView av;
  int batch, gmax;
  const float4* gt_boxes;
  const int32_t* gt_classes;
  const int32_t* gt_count;
  const long long* pb_indices;
  const long long* gt_indices;
  const int32_t* status;
  float* y_true;
  uint8_t* pb_valids;
};

__global__ void __launch_bounds__(kEncodeTile) encode_rows_kernel(const EncodeArgs p) {
  extern __shared__ __align__(128) unsigned char enc_smem[];
  const int C = p.av.num_classes;
  const int W = 3 + C + 4;
  const int b = blockIdx.y;
  const int a0 = blockIdx.x * kEncodeTile;
  const int rows = min(kEncodeTile, p.av.n - a0);
  const int tid = threadIdx.x;
  float* s_rows = reinterpret_cast<float*>(enc_smem);                       // [kEncodeTile * W]
  const size_t rows_bytes = odk_align_up((size_t)kEncodeTile * W * 4, 16);
  float4* s_gbox = reinterpret_cast<float4*>(enc_smem + rows_bytes);        // [gmax]
  float* s_garea = reinterpret_cast<float*>(enc_smem + rows_bytes + (size_t)p.gmax * 16);
  __shared__ unsigned char s_assigned[kEncodeTile];

  int* s_list = reinterpret_cast<int*>(enc_smem + rows_bytes + (size_t)p.gmax * 20);  // [gmax] GTs that can reach this tile
  __shared__ float s_red[6][kEncodeTile / 32];
  __shared__ int s_nlist;

  const bool ok = p.status[b] == ODK_IMG_OK;
  const int G = ok ? p.gt_count[b] : 0;
  for (int g = tid; g < G; g += kEncodeTile) {
    const float4 bx = p.gt_boxes[(size_t)b * p.gmax + g];
    s_gbox[g] = bx;
    s_garea[g] = odk_area(bx);
  }
  {  // zero the tile
    float4* z = reinterpret_cast<float4*>(s_rows);
    const int n4 = (kEncodeTile * W * 4 + 15) / 16;
    for (int i = tid; i < n4; i += kEncodeTile) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  s_assigned[tid] = 0;
  if (tid == 0) s_nlist = 0;

  // this thread's anchor, and the tile's bounding box / area range over its VALID anchors
  const int a = a0 + tid;
  const bool active = tid < rows && ok && G > 0 && p.av.mask[a];
  float4 ab = make_float4(0.f, 0.f, 0.f, 0.f);
  float aa = 0.f;
  float r0 = FLT_MAX, r1 = FLT_MAX, r2 = -FLT_MAX, r3 = -FLT_MAX, r4 = FLT_MAX, r5 = -FLT_MAX;
  if (active) {
    ab = __ldg(p.av.locs + a);
    aa = odk_area(ab);
    r0 = ab.x; r1 = ab.y; r2 = ab.z; r3 = ab.w; r4 = aa; r5 = aa;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    r0 = fminf(r0, __shfl_xor_sync(0xffffffffu, r0, d));
    r1 = fminf(r1, __shfl_xor_sync(0xffffffffu, r1, d));
    r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, d));
    r3 = fmaxf(r3, __shfl_xor_sync(0xffffffffu, r3, d));
    r4 = fminf(r4, __shfl_xor_sync(0xffffffffu, r4, d));
    r5 = fmaxf(r5, __shfl_xor_sync(0xffffffffu, r5, d));
  }
  if ((tid & 31) == 0) {
    const int w = tid >> 5;
    s_red[0][w] = r0; s_red[1][w] = r1; s_red[2][w] = r2; s_red[3][w] = r3; s_red[4][w] = r4; s_red[5][w] = r5;
  }
  __syncthreads();
#pragma unroll
  for (int w = 0; w < kEncodeTile / 32; ++w) {
    r0 = fminf(r0, s_red[0][w]); r1 = fminf(r1, s_red[1][w]); r2 = fmaxf(r2, s_red[2][w]);
    r3 = fmaxf(r3, s_red[3][w]); r4 = fminf(r4, s_red[4][w]); r5 = fmaxf(r5, s_red[5][w]);
  }
  // Conservative cull: fl(IoU) >= 0.5 needs a positive-extent overlap with some anchor of the tile
  // (=> with the tile's bounding box) and min(Ag,Aa)/max(Ag,Aa) >= IoU, i.e. Ag within [0.5,2]x the
  // anchor area; 0.49 / 2.05 leave ~2% slack against the few-ulp rounding of the fp32 areas.
  for (int g = tid; g < G; g += kEncodeTile) {
    const float4 gb = s_gbox[g];
    const float ga = s_garea[g];
    const bool reach = gb.x < r2 && gb.z > r0 && gb.y < r3 && gb.w > r1 && ga >= 0.49f * r4 && ga <= 2.05f * r5;
    if (reach) s_list[atomicAdd(&s_nlist, 1)] = g;
  }
  __syncthreads();

  const float w0 = (float)(1.0 / (double)max(1, G));  // od_pb.py:306
  bool valid = true;
  {
    const int nl = s_nlist;  // block-uniform trip count; the loop body has no divergent exits
    bool ignored = false;
    for (int k = 0; k < nl; ++k) {
      const int g = s_list[k];
      ignored |= odk_iou_passes<true>(s_gbox[g], s_garea[g], ab, aa, 0.5f);  // od_pb.py:355
    }
    valid = !(active && ignored);
    if (tid < rows && ok && valid) s_rows[tid * W] = w0;
  }
  __syncthreads();
  // positives (one per GT): od_pb.py:307-311
  if (G > 0) {
    const float w1 = (float)(1.0 / (double)G);
    for (int k = tid; k < G; k += kEncodeTile) {
      const int pa = (int)p.pb_indices[(size_t)b * p.gmax + k];
      if (pa < a0 || pa >= a0 + rows) continue;
      const int g = (int)p.gt_indices[(size_t)b * p.gmax + k];
      float* row = s_rows + (size_t)(pa - a0) * W;
      const float4 box = s_gbox[g];
      const float4 loc = __ldg(p.av.locs + pa);
      const double2 sz = p.av.sizes[pa];
      row[0] = w0;  // pb_valids[pb_indices] = True (od_pb.py:361)
      row[1] = w1;
      row[2] = 1.0f;
      row[3 + p.gt_classes[(size_t)b * p.gmax + g]] = 1.0f;
      row[3 + C + 0] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.x, loc.x), sz.x), 0.10);
      row[3 + C + 1] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.y, loc.y), sz.y), 0.10);
      row[3 + C + 2] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.z, loc.z), sz.x), 0.10);
      row[3 + C + 3] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.w, loc.w), sz.y), 0.10);
      s_assigned[pa - a0] = 1;
    }
  }
  odk_fence_proxy_async();  // generic-proxy smem writes -> visible to the bulk (async-proxy) store
  __syncthreads();
  float* dst = p.y_true + ((size_t)b * p.av.n + a0) * W;
  const uint32_t bytes = (uint32_t)rows * W * 4;
  if (((uintptr_t)dst & 15) == 0 && (bytes & 15) == 0) {
    if (tid == 0) {
      odk_bulk_s2g(dst, s_rows, bytes);
      odk_bulk_commit();
    }
  } else {
    for (int i = tid; i < rows * W; i += kEncodeTile) dst[i] = s_rows[i];
  }
  if (p.pb_valids != nullptr && tid < rows) p.pb_valids[(size_t)b * p.av.n + a] = (ok && (valid || s_assigned[tid])) ? 1 : 0;
  if (tid == 0) odk_bulk_wait_read<0>();  // smem must outlive the bulk read
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

int launch_greedy(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count, int gmax,
                  int64_t* pb_indices, int64_t* gt_indices, int32_t* status, cudaStream_t stream) {
  AssignArgs args;
  args.av = odk_view(a);
  args.batch = batch; args.gmax = gmax;
  args.gt_boxes = reinterpret_cast<const float4*>(gt_boxes);
  args.gt_classes = gt_classes; args.gt_count = gt_count;
  args.pb_indices = reinterpret_cast<long long*>(pb_indices);
  args.gt_indices = reinterpret_cast<long long*>(gt_indices);
  args.status = status;
  args.bitmap_words = (a->n + 31) / 32;
  int list_len = (int)((96 * 1024) / (4 * (size_t)gmax)) & ~7;
  list_len = list_len > 128 ? 128 : (list_len < 8 ? 8 : list_len);
  args.list_len = list_len;
  const size_t smem = odk_align_up((size_t)gmax * (16 + 20 + 4 * (size_t)list_len) + (size_t)kGreedyWarps * kScratch * 12 +
                                   (size_t)args.bitmap_words * 4, 16);
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "assign: N=%d / Gmax=%d need %zu bytes of shared memory per image", a->n, gmax, smem);
  ODK_CUDA(cudaFuncSetAttribute(assign_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {
    OdkProfileScope prof_("assign_greedy_kernel", stream);
    assign_greedy_kernel<<<batch, kGreedyThreads, smem, stream>>>(args);
  }
  ODK_LAUNCH_CHECK("assign_greedy_kernel");
  return ODK_OK;
}

int assign_check(const odk_anchors* a, int batch, const void* gt_boxes, const void* gt_count, int gmax, const void* pb, const void* gt,
                 const void* status) {
  ODK_REQUIRE(a != nullptr, ODK_ERR_INVALID_ARGUMENT, "assign: null anchors handle");
  ODK_REQUIRE(batch >= 1 && gmax >= 1, ODK_ERR_INVALID_ARGUMENT, "assign: batch=%d gmax=%d must be >= 1", batch, gmax);
  ODK_REQUIRE(gt_boxes && gt_count && pb && gt && status, ODK_ERR_INVALID_ARGUMENT, "assign: null pointer");
  ODK_REQUIRE(((uintptr_t)gt_boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "assign: gt_boxes must be 16-byte aligned");
  return ODK_OK;
}

}  // namespace

extern "C" int odk_assign(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, int64_t* pb_indices,
                          int64_t* gt_indices, int32_t* status, void* stream) {
  if (int rc = assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  return launch_greedy(a, batch, gt_boxes, nullptr, gt_count, gmax, pb_indices, gt_indices, status, static_cast<cudaStream_t>(stream));
}

extern "C" int odk_assign_encode(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                                 int gmax, float* y_true, int64_t* pb_indices, int64_t* gt_indices, uint8_t* pb_valids, int32_t* status,
                                 void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  ODK_REQUIRE(gt_classes && y_true, ODK_ERR_INVALID_ARGUMENT, "assign_encode: null pointer");
  ODK_REQUIRE(((uintptr_t)y_true & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "assign_encode: y_true must be 4-byte aligned");
  if (int rc = launch_greedy(a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, stream)) return rc;
  EncodeArgs e;
  e.av = odk_view(a);
  e.batch = batch; e.gmax = gmax;
  e.gt_boxes = reinterpret_cast<const float4*>(gt_boxes);
  e.gt_classes = gt_classes; e.gt_count = gt_count;
  e.pb_indices = reinterpret_cast<const long long*>(pb_indices);
  e.gt_indices = reinterpret_cast<const long long*>(gt_indices);
  e.status = status; e.y_true = y_true; e.pb_valids = pb_valids;
  const int W = 3 + a->num_classes + 4;
  const size_t smem = odk_align_up((size_t)kEncodeTile * W * 4, 16) + (size_t)gmax * 24;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "assign_encode: C=%d / Gmax=%d need %zu bytes of shared memory", a->num_classes, gmax, smem);
  ODK_CUDA(cudaFuncSetAttribute(encode_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  dim3 grid((a->n + kEncodeTile - 1) / kEncodeTile, batch);
  ODK_REQUIRE(batch <= 65535, ODK_ERR_UNSUPPORTED, "assign_encode: batch=%d > 65535", batch);
  {
    OdkProfileScope prof_("encode_rows_kernel", stream);
    encode_rows_kernel<<<grid, kEncodeTile, smem, stream>>>(e);
  }
  ODK_LAUNCH_CHECK("encode_rows_kernel");
  return ODK_OK;
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
