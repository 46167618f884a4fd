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
  int warp_smem_bytes;
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

__global__ void __launch_bounds__(128) assign_greedy_kernel(const AssignArgs p) {
  extern __shared__ __align__(16) unsigned char greedy_smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = (int)odk_lane();
  const int b = blockIdx.x * (blockDim.x >> 5) + warp;
  if (b >= p.batch) return;
  unsigned char* base = greedy_smem + (size_t)warp * p.warp_smem_bytes;
  float4* gbox = reinterpret_cast<float4*>(base);
  float* garea = reinterpret_cast<float*>(base + (size_t)p.gmax * 16);
  int* order = reinterpret_cast<int*>(base + (size_t)p.gmax * 20);
  int* pick = reinterpret_cast<int*>(base + (size_t)p.gmax * 24);
  unsigned* taken = reinterpret_cast<unsigned*>(base + (size_t)p.gmax * 28);

  long long* out_pb = p.pb_indices + (size_t)b * p.gmax;
  long long* out_gt = p.gt_indices + (size_t)b * p.gmax;
  const int g_count = p.gt_count[b];
  int err = ODK_IMG_OK;
  if (g_count < 0 || g_count > p.gmax) err = ODK_IMG_BAD_COUNT;
  const int G = err ? 0 : g_count;

  for (int g = lane; g < G; g += 32) {
    const float4 bx = p.gt_boxes[(size_t)b * p.gmax + g];
    gbox[g] = bx;
    garea[g] = odk_area(bx);
    if (p.gt_classes != nullptr) {
      const int c = p.gt_classes[(size_t)b * p.gmax + g];
      if (c < 0 || c >= p.av.num_classes) err = ODK_IMG_BAD_CLASS;  // od_pb.py:304
    }
  }
  for (int w = lane; w < p.bitmap_words; w += 32) taken[w] = 0u;
  err = __reduce_max_sync(0xffffffffu, err);
  __syncwarp();
  if (!err) {
    // ascending-area rank, ties by index
    for (int g = lane; g < G; g += 32) {
      const float ag = garea[g];
      int r = 0;
      for (int j = 0; j < G; ++j) {
        const float aj = garea[j];
        r += (aj < ag) || (aj == ag && j < g);
      }
      order[r] = g;
    }
    __syncwarp();
    const int per_map = 9 * p.av.n_patterns;
    const int n_probe = p.av.n_maps * per_map;
    for (int r = 0; r < G && !err; ++r) {
      const int g = order[r];
      const float4 box = gbox[g];
      const float cx = __fmul_rn(__fadd_rn(box.z, box.x), 0.5f);  // ml.py:485-488
      const float cy = __fmul_rn(__fadd_rn(box.w, box.y), 0.5f);
      if (r >= p.av.n_valid) { err = ODK_IMG_ENCODE_ERROR; break; }  // od_pb.py:331-332: no free valid anchor left
      double best_cost = DBL_MAX;
      int best_a = INT_MAX;
      bool any_cell = false;
      for (int t = lane; t < n_probe; t += 32) {
        const int mi = t / per_map;
        const int rem = t - mi * per_map;
        const int cell = rem / p.av.n_patterns;
        const int pat = rem - cell * p.av.n_patterns;
        const int m = p.av.map_size[mi];
        const int ix = (int)floorf(cx * (float)m) + (cell % 3) - 1;
        const int iy = (int)floorf(cy * (float)m) + (cell / 3) - 1;
        if (ix < 0 || iy < 0 || ix >= m || iy >= m) continue;
        const int a = p.av.map_off[mi] + pat * m * m + iy * m + ix;
        const float4 gr = __ldg(p.av.grid + a);
        // math.in_range: lo <= x < hi on both axes (od_pb.py:336), AND pb_mask (:337)
        if (!(gr.x <= cx && cx < gr.z && gr.y <= cy && cy < gr.w) || !p.av.mask[a]) continue;
        any_cell = true;
        if ((taken[a >> 5] >> (a & 31)) & 1u) continue;  // od_pb.py:339
        const double cost = encode_cost(box, __ldg(p.av.locs + a), p.av.sizes[a]);
        if (cost < best_cost || (cost == best_cost && a < best_a)) { best_cost = cost; best_a = a; }
      }
      if (!__any_sync(0xffffffffu, any_cell)) { err = ODK_IMG_ENCODE_ERROR; break; }  // od_pb.py:338
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) {
        const double oc = __shfl_xor_sync(0xffffffffu, best_cost, d);
        const int oa = __shfl_xor_sync(0xffffffffu, best_a, d);
        if (oc < best_cost || (oc == best_cost && oa < best_a)) { best_cost = oc; best_a = oa; }
      }
      if (best_a == INT_MAX) {
        // fallback: arg-max IoU over the free valid anchors, first maximum wins (od_pb.py:347-352)
        float best_v = -1.0f;
        const float ga = garea[g];
        for (int a = lane; a < p.av.n; a += 32) {
          if (!p.av.mask[a] || ((taken[a >> 5] >> (a & 31)) & 1u)) continue;
          const float4 lc = __ldg(p.av.locs + a);
          const float v = odk_iou_value(box, ga, lc, odk_area(lc));
          if (v > best_v) { best_v = v; best_a = a; }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, best_v, d);
          const int oa = __shfl_xor_sync(0xffffffffu, best_a, d);
          if (ov > best_v || (ov == best_v && oa < best_a)) { best_v = ov; best_a = oa; }
        }
        if (best_a == INT_MAX) { err = ODK_IMG_ENCODE_ERROR; break; }
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
  __syncthreads();

  const float w0 = (float)(1.0 / (double)max(1, G));  // od_pb.py:306
  bool valid = true;
  const int a = a0 + tid;
  if (tid < rows && ok) {
    if (G > 0 && p.av.mask[a]) {
      const float4 ab = __ldg(p.av.locs + a);
      const float aa = odk_area(ab);
      for (int g = 0; g < G; ++g) {
        if (odk_iou_passes<true>(s_gbox[g], s_garea[g], ab, aa, 0.5f)) { valid = false; break; }  // od_pb.py:355
      }
    }
    if (valid) s_rows[tid * W] = w0;
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
  args.warp_smem_bytes = (int)odk_align_up((size_t)gmax * 28 + (size_t)args.bitmap_words * 4, 16);
  ODK_REQUIRE(args.warp_smem_bytes <= 200 * 1024, ODK_ERR_UNSUPPORTED, "assign: N=%d / Gmax=%d need %d bytes of shared memory per image",
              a->n, gmax, args.warp_smem_bytes);
  int wpb = (200 * 1024) / args.warp_smem_bytes;
  wpb = wpb > 4 ? 4 : wpb;
  // few images: spread them over more CTAs (one warp each) so more SMs work
  if (batch <= 148 * 2) wpb = 1;
  const size_t smem = (size_t)wpb * args.warp_smem_bytes;
  ODK_CUDA(cudaFuncSetAttribute(assign_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  {
    OdkProfileScope prof_("assign_greedy_kernel", stream);
    assign_greedy_kernel<<<(batch + wpb - 1) / wpb, wpb * 32, smem, stream>>>(args);
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
  const size_t smem = odk_align_up((size_t)kEncodeTile * W * 4, 16) + (size_t)gmax * 20;
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
