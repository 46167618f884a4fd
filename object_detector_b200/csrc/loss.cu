// SURVEY §8 f3 — the consumer of y_true on the training side: PriorBoxes.loss / loss_obj / loss_clf / loss_loc and
// the rec_bg / rec_obj / acc_clf metrics (od_pb.py:365-431, losses.py:22-30), forward values per image.
//
// Two flavours of one row evaluator:
//   DENSE  (odk_loss)         y_true (B,N,3+C+4) and y_pred (B,N,3+C+4) are both streamed once: HBM-read-bound,
//                             2*4*N*(3+C+4) bytes per image.  Tiles of 128 anchor rows of both tensors are fetched
//                             with one bulk TMA copy each (cp.async.bulk, SASS UBLKCP) into shared memory; a thread
//                             then owns one row (row stride 3+C+4 floats: odd for the usual class counts, so the
//                             per-thread row walks are bank-conflict-free).
//   SPARSE (odk_assign_loss)  y_true is never materialised.  Everything the loss needs from it is known from the
//                             assignment: column 0 = 1/max(1,G) where pb_valids, and only the <= G positive rows carry
//                             a weight, a class and box targets.  A negative row therefore contributes through ONE float
//                             of y_pred (its objectness, column 2): the kernel gathers that column (one 32-byte sector
//                             per 4*(3+C+4)-byte row) and reads whole rows only for the positives.
//
// The reference evaluates these as TensorFlow float32 graph ops whose reduction order is unspecified (PARITY
// UNPINNED: tensorflow is absent from the reference tree); here the per-row arithmetic is float32 in the Keras
// formulas' order and the sums over the anchor axis are carried in float64 (tile partials reduced in a fixed order,
// so results are run-to-run deterministic).  Tolerance against the oracle: 1e-5 relative.
#include "odk_common.cuh"

namespace {

constexpr int kLossTile = 128;
constexpr int kAcc = 8;  // S_obj, S_clf, S_loc, bg_num, bg_den, obj_num, obj_den, acc_num
constexpr float kEps = 1e-7f;  // tf.keras.backend.epsilon()

struct LossArgs {
  AnchorView av;
  int batch, n, c, W, tiles_per_image;
  const float* y_true;  // DENSE
  const float* y_pred;
  // SPARSE: the assignment instead of y_true
  const unsigned* ignore;  // [B, bitmap_words]
  int bitmap_words;
  const long long* pb_indices;  // [B, gmax]
  const long long* gt_indices;  // [B, gmax]
  const float4* gt_boxes;       // [B, gmax]
  const int32_t* gt_classes;    // [B, gmax]
  const int32_t* gt_count;      // [B]
  const int32_t* status;        // [B]
  int gmax;
  double* partial;  // [B, tiles_per_image, kAcc]
  double* out;      // [B, 8]
  float* grad;            // SPARSE + gradient: [B, N, W] or null
  const float* upstream;  // [B] or null (= 1)
};

// losses.py:22-30 with alpha=0.25, gamma=2
__device__ __forceinline__ float focal(float y, float p) {
  const float om_y = __fsub_rn(1.0f, y);
  const float a_t = __fadd_rn(__fmul_rn(y, 0.25f), __fmul_rn(om_y, 0.75f));
  float p_t = __fadd_rn(__fmul_rn(y, p), __fmul_rn(om_y, __fsub_rn(1.0f, p)));
  p_t = fminf(fmaxf(p_t, kEps), __fsub_rn(1.0f, kEps));
  const float om = __fsub_rn(1.0f, p_t);
  return __fmul_rn(__fmul_rn(-a_t, __fmul_rn(om, om)), logf(p_t));
}

// d(3 * mask * focal(y, p)) / dp  (tf.clip_by_value: no gradient where p_t was clipped)
__device__ __forceinline__ float obj_grad(float y, float p, float mask) {
  const float om_y = 1.0f - y;
  const float a_t = y * 0.25f + om_y * 0.75f;
  const float p_raw = y * p + om_y * (1.0f - p);
  if (!(p_raw >= kEps && p_raw <= 1.0f - kEps)) return 0.0f;
  const float om = 1.0f - p_raw;
  return 3.0f * mask * (y - om_y) * a_t * (2.0f * om * logf(p_raw) - om * om / p_raw);
}

// Adds one anchor row's terms.  cce / mse / cls_match are that row's categorical cross-entropy (od_pb.py:419),
// mean squared box error (:429) and argmax agreement (:398).
__device__ __forceinline__ void add_row(double* acc, float gt_mask, float gt_w, float gt_obj, float pred_obj, float pbm, float cce, float mse,
                                        bool cls_match) {
  acc[0] += (double)__fmul_rn(focal(gt_obj, pred_obj), __fmul_rn(pbm, gt_mask));  // :407-410
  acc[1] += (double)__fmul_rn(cce, gt_w);                                           // :420
  acc[2] += (double)__fmul_rn(mse, gt_w);                                           // :430
  const float hit = (rintf(gt_obj) == rintf(pred_obj)) ? 1.0f : 0.0f;               // tf round: half to even
  const float gt_bg = __fmul_rn(__fsub_rn(1.0f, gt_obj), gt_mask);                  // :385
  acc[3] += (double)__fmul_rn(hit, gt_bg);
  acc[4] += (double)gt_bg;
  acc[5] += (double)__fmul_rn(hit, gt_obj);
  acc[6] += (double)gt_obj;
  acc[7] += (double)__fmul_rn(cls_match ? 1.0f : 0.0f, gt_obj);
}

// Per-row class terms against a general target vector (DENSE) — tf.keras.losses.categorical_crossentropy: the
// prediction is renormalised by its sum, clipped to [eps, 1-eps]; rows whose target is all-zero contribute 0.
__device__ __forceinline__ void class_terms(const float* gt, const float* pr, int c, bool need_argmax, float& cce, bool& match) {
  float s = 0.0f;
  for (int j = 0; j < c; ++j) s = __fadd_rn(s, pr[j]);
  float acc = 0.0f;
  int ag = 0, ap = 0;
  float bg = gt[0], bp = pr[0];
  for (int j = 0; j < c; ++j) {
    const float t = gt[j], q = pr[j];
    if (t != 0.0f) {
      const float nq = fminf(fmaxf(__fdiv_rn(q, s), kEps), __fsub_rn(1.0f, kEps));
      acc = __fadd_rn(acc, __fmul_rn(t, logf(nq)));
    }
    if (need_argmax) {
      if (t > bg) { bg = t; ag = j; }
      if (q > bp) { bp = q; ap = j; }
    }
  }
  cce = -acc;
  match = ag == ap;
}

__device__ __forceinline__ void block_reduce_store(double* acc, double* dst) {
  __shared__ double s_red[kLossTile / 32][kAcc];
  const int lane = (int)odk_lane(), warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < kAcc; ++k) {
    double v = acc[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    if (lane == 0) s_red[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < kAcc) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < kLossTile / 32; ++w) v += s_red[w][threadIdx.x];
    dst[threadIdx.x] = v;
  }
}

// ---- DENSE: one CTA per tile of kLossTile rows -----------------------------------------------------------------------
__global__ void __launch_bounds__(kLossTile) loss_dense_kernel(const LossArgs p) {
  extern __shared__ __align__(128) unsigned char loss_smem[];
  __shared__ uint64_t bar;
  const int tid = threadIdx.x;
  const int b = blockIdx.y, tile = blockIdx.x;
  const int a0 = tile * kLossTile;
  const int rows = min(kLossTile, p.n - a0);
  const int W = p.W;
  float* s_true = reinterpret_cast<float*>(loss_smem);
  float* s_pred = s_true + (size_t)kLossTile * W;
  const size_t row0 = (size_t)b * p.n + a0;
  const float* g_true = p.y_true + row0 * W;
  const float* g_pred = p.y_pred + row0 * W;
  const uint32_t bytes = (uint32_t)rows * W * 4;
  const bool tma_ok = ((bytes & 15) == 0) && ((((uintptr_t)g_true | (uintptr_t)g_pred) & 15) == 0);
  if (tma_ok) {
    if (tid == 0) {
      odk_mbar_init(&bar, 1);
      odk_mbar_fence_init();
      odk_mbar_expect_tx(&bar, 2 * bytes);
      odk_bulk_g2s(s_true, g_true, bytes, &bar);
      odk_bulk_g2s(s_pred, g_pred, bytes, &bar);
    }
    __syncthreads();  // barrier initialised before anyone polls it
    odk_mbar_wait(&bar, 0);
  } else {
    for (int i = tid; i < rows * W; i += kLossTile) {
      s_true[i] = __ldg(g_true + i);
      s_pred[i] = __ldg(g_pred + i);
    }
    __syncthreads();
  }
  double acc[kAcc];
#pragma unroll
  for (int k = 0; k < kAcc; ++k) acc[k] = 0.0;
  if (tid < rows) {
    const float* t = s_true + (size_t)tid * W;
    const float* q = s_pred + (size_t)tid * W;
    const float gt_obj = t[2];
    float cce;
    bool match;
    class_terms(t + 3, q + 3, p.c, gt_obj != 0.0f, cce, match);
    const float* tl = t + 3 + p.c;
    const float* ql = q + 3 + p.c;
    float se = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d = __fsub_rn(ql[j], tl[j]);
      se = __fadd_rn(se, __fmul_rn(d, d));
    }
    add_row(acc, t[0], t[1], gt_obj, q[2], p.av.mask[a0 + tid] ? 1.0f : 0.0f, cce, __fmul_rn(se, 0.25f), match);
  }
  block_reduce_store(acc, p.partial + ((size_t)b * p.tiles_per_image + tile) * kAcc);
}

// ---- SPARSE: the assignment stands in for y_true ---------------------------------------------------------------------
// GRAD: the gradient tile (all zero except column 2 and the positive rows) is built in shared memory and leaves with one
// bulk TMA store, so forward values and d loss / d y_pred come out of the same pass over the objectness column.
template <bool GRAD>
__global__ void __launch_bounds__(kLossTile) loss_sparse_kernel(const LossArgs p) {
  extern __shared__ __align__(128) unsigned char ls_smem[];
  __shared__ int s_pos[kLossTile];  // assignment slot + 1 of the positive anchors of this tile
  const int tid = threadIdx.x;
  const int b = blockIdx.y, tile = blockIdx.x;
  const int a0 = tile * kLossTile;
  const int rows = min(kLossTile, p.n - a0);
  const int W = p.W;
  const bool ok = p.status[b] == ODK_IMG_OK;
  const int G = ok ? p.gt_count[b] : 0;
  s_pos[tid] = 0;
  __syncthreads();
  for (int k = tid; k < G; k += kLossTile) {
    const int pa = (int)p.pb_indices[(size_t)b * p.gmax + k] - a0;
    if (pa >= 0 && pa < rows) s_pos[pa] = k + 1;
  }
  __syncthreads();
  float* s_g = reinterpret_cast<float*>(ls_smem);  // GRAD: [rows][W]
  const float up = (GRAD && p.upstream) ? p.upstream[b] : 1.0f;
  if (GRAD) {
    for (int i = tid; i < rows * W; i += kLossTile) s_g[i] = 0.0f;
    __syncthreads();
  }
  double acc[kAcc];
#pragma unroll
  for (int k = 0; k < kAcc; ++k) acc[k] = 0.0;
  if (tid < rows) {
    const int a = a0 + tid;
    const float* q = p.y_pred + ((size_t)b * p.n + a) * W;
    const int slot = s_pos[tid];
    const unsigned mbit = (__ldg(p.av.maskbits + (a >> 5)) >> (a & 31)) & 1u;
    const unsigned ibit = (p.ignore[(size_t)b * p.bitmap_words + (a >> 5)] >> (a & 31)) & 1u;
    const float w0 = __fdiv_rn(1.0f, (float)max(1, G));  // od_pb.py:306
    if (slot == 0) {
      // negative (or ignored) row: y_true = [mask, 0, 0, 0.., 0,0,0,0]; cce = 0 (all-zero target), weight 0
      const float gt_mask = (ok && !(mbit & ibit)) ? w0 : 0.0f;
      const float pr = q[2];
      add_row(acc, gt_mask, 0.0f, 0.0f, pr, mbit ? 1.0f : 0.0f, 0.0f, 0.0f, false);
      if (GRAD) s_g[tid * W + 2] = obj_grad(0.0f, pr, (mbit ? 1.0f : 0.0f) * gt_mask) * up;
    } else {
      const int g = (int)p.gt_indices[(size_t)b * p.gmax + slot - 1];
      const float4 box = p.gt_boxes[(size_t)b * p.gmax + g];
      const int cls = p.gt_classes[(size_t)b * p.gmax + g];
      const float4 ab = __ldg(p.av.locs + a);
      const double2 sz = p.av.sizes[a];
      float tl[4];  // encode_locs (od_pb.py:212-214) rounded to the float32 y_true would hold (:311)
      tl[0] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.x, ab.x), sz.x), 0.10);
      tl[1] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.y, ab.y), sz.y), 0.10);
      tl[2] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.z, ab.z), sz.x), 0.10);
      tl[3] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.w, ab.w), sz.y), 0.10);
      const float* qc = q + 3;
      float s = 0.0f, bp = qc[0];
      int ap = 0;
      for (int j = 0; j < p.c; ++j) {
        const float v = qc[j];
        s = __fadd_rn(s, v);
        if (v > bp) { bp = v; ap = j; }
      }
      const float nq = fminf(fmaxf(__fdiv_rn(qc[cls], s), kEps), __fsub_rn(1.0f, kEps));
      const float cce = -__fmul_rn(1.0f, logf(nq));
      float se = 0.0f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float d = __fsub_rn(q[3 + p.c + j], tl[j]);
        se = __fadd_rn(se, __fmul_rn(d, d));
      }
      const float wgt = __fdiv_rn(1.0f, (float)G);
      add_row(acc, w0, wgt, 1.0f, q[2], mbit ? 1.0f : 0.0f, cce, __fmul_rn(se, 0.25f), ap == cls);
      if (GRAD) {
        float* gr = s_g + tid * W;
        gr[2] = obj_grad(1.0f, q[2], (mbit ? 1.0f : 0.0f) * w0) * up;
        const float ncls = qc[cls] / s;
        const bool in = ncls >= kEps && ncls <= 1.0f - kEps;
        const float cross = in ? 1.0f : 0.0f;
        for (int j = 0; j < p.c; ++j) gr[3 + j] = -(wgt / s) * (((j == cls && in) ? 1.0f / ncls : 0.0f) - cross) * up;
#pragma unroll
        for (int j = 0; j < 4; ++j) gr[3 + p.c + j] = wgt * (q[3 + p.c + j] - tl[j]) * up;
      }
    }
  }
  if (GRAD) {
    float* g_out = p.grad + ((size_t)b * p.n + a0) * W;
    const uint32_t bytes = (uint32_t)rows * W * 4;
    if (((bytes & 15) == 0) && (((uintptr_t)g_out & 15) == 0)) {
      odk_fence_proxy_async();
      __syncthreads();
      if (tid == 0) {
        odk_bulk_s2g(g_out, s_g, bytes);
        odk_bulk_commit();
      }
    } else {
      __syncthreads();
      for (int i = tid; i < rows * W; i += kLossTile) g_out[i] = s_g[i];
    }
  }
  block_reduce_store(acc, p.partial + ((size_t)b * p.tiles_per_image + tile) * kAcc);
  if (GRAD && tid == 0) odk_bulk_wait_read<0>();  // the tile must stay in shared memory until the bulk store has read it
}

// ---- gradient of loss = loss_obj + loss_clf + loss_loc (od_pb.py:365-373) w.r.t. y_pred --------------------------------
// Same tiling as the dense forward kernel: the y_true / y_pred tiles arrive by bulk TMA, a thread differentiates its row
// in place of the y_true row in shared memory, and the finished tile leaves with one bulk TMA store.  Per row
//   d/dobj   = 3 * pb_mask * mask * (2y-1) * a_t * [2(1-p_t) log p_t - (1-p_t)^2 / p_t]      (0 where p_t was clipped)
//   d/dcls_k = -w/s * [t_k/n_k * in_k - sum_j t_j * q_j/(s*n_j) * in_j],  n = clip(q/s), in = "not clipped"
//   d/dloc_j = w * (q_j - t_j)
// (columns 0 and 1 of y_pred carry no gradient), times the upstream gradient of the image's loss when given —
// tf.clip_by_value passes the gradient only inside [eps, 1-eps], like here.
__global__ void __launch_bounds__(kLossTile) loss_grad_kernel(const LossArgs p, const float* __restrict__ upstream, float* __restrict__ grad) {
  extern __shared__ __align__(128) unsigned char lg_smem[];
  __shared__ uint64_t bar;
  const int tid = threadIdx.x;
  const int b = blockIdx.y, tile = blockIdx.x;
  const int a0 = tile * kLossTile;
  const int rows = min(kLossTile, p.n - a0);
  const int W = p.W, C = p.c;
  float* s_true = reinterpret_cast<float*>(lg_smem);
  float* s_pred = s_true + (size_t)kLossTile * W;
  const size_t row0 = (size_t)b * p.n + a0;
  const float* g_true = p.y_true + row0 * W;
  const float* g_pred = p.y_pred + row0 * W;
  float* g_out = grad + row0 * W;
  const uint32_t bytes = (uint32_t)rows * W * 4;
  const bool tma_ok = ((bytes & 15) == 0) && ((((uintptr_t)g_true | (uintptr_t)g_pred | (uintptr_t)g_out) & 15) == 0);
  if (tma_ok) {
    if (tid == 0) {
      odk_mbar_init(&bar, 1);
      odk_mbar_fence_init();
      odk_mbar_expect_tx(&bar, 2 * bytes);
      odk_bulk_g2s(s_true, g_true, bytes, &bar);
      odk_bulk_g2s(s_pred, g_pred, bytes, &bar);
    }
    __syncthreads();
    odk_mbar_wait(&bar, 0);
  } else {
    for (int i = tid; i < rows * W; i += kLossTile) {
      s_true[i] = __ldg(g_true + i);
      s_pred[i] = __ldg(g_pred + i);
    }
    __syncthreads();
  }
  const float up = upstream ? upstream[b] : 1.0f;
  if (tid < rows) {
    float* t = s_true + (size_t)tid * W;  // becomes the gradient row
    const float* q = s_pred + (size_t)tid * W;
    const float gt_mask = t[0], w = t[1], y = t[2], pr = q[2];
    // objectness
    const float om_y = 1.0f - y;
    const float a_t = y * 0.25f + om_y * 0.75f;
    const float p_raw = y * pr + om_y * (1.0f - pr);
    const bool inside = p_raw >= kEps && p_raw <= 1.0f - kEps;
    const float p_t = fminf(fmaxf(p_raw, kEps), 1.0f - kEps);
    const float om = 1.0f - p_t;
    const float dfdpt = a_t * (2.0f * om * logf(p_t) - om * om / p_t);
    const float m = (p.av.mask[a0 + tid] ? 1.0f : 0.0f) * gt_mask;
    const float g_obj = inside ? 3.0f * m * (y - om_y) * dfdpt * up : 0.0f;
    // classes
    float s = 0.0f;
    for (int j = 0; j < C; ++j) s += q[3 + j];
    float cross = 0.0f;  // sum_j t_j * in_j * n_j / nclip_j
    if (w != 0.0f) {
      for (int j = 0; j < C; ++j) {
        const float tj = t[3 + j];
        if (tj != 0.0f) {
          const float nj = q[3 + j] / s;
          if (nj >= kEps && nj <= 1.0f - kEps) cross += tj;  // n_j / clip(n_j) = 1 inside the range
        }
      }
    }
    for (int j = 0; j < C; ++j) {
      float g = 0.0f;
      if (w != 0.0f) {
        const float tj = t[3 + j];
        const float nj = q[3 + j] / s;
        const bool in = nj >= kEps && nj <= 1.0f - kEps;
        const float direct = (tj != 0.0f && in) ? tj / nj : 0.0f;
        g = -(w / s) * (direct - cross) * up;
      }
      t[3 + j] = g;
    }
    // box offsets
#pragma unroll
    for (int j = 0; j < 4; ++j) t[3 + C + j] = w * (q[3 + C + j] - t[3 + C + j]) * up;
    t[0] = 0.0f;
    t[1] = 0.0f;
    t[2] = g_obj;
  }
  if (tma_ok) {
    odk_fence_proxy_async();  // the generic-proxy writes above must be visible to the bulk store
    __syncthreads();
    if (tid == 0) {
      odk_bulk_s2g(g_out, s_true, bytes);
      odk_bulk_commit();
      odk_bulk_wait_read<0>();  // shared memory may be released once it has been read
    }
  } else {
    __syncthreads();
    for (int i = tid; i < rows * W; i += kLossTile) g_out[i] = s_true[i];
  }
}

// ---- per image: tile partials in a fixed order -> the seven reported values ---------------------------------------
__global__ void __launch_bounds__(32) loss_finish_kernel(const LossArgs p) {
  const int b = blockIdx.x, lane = threadIdx.x;
  double acc[kAcc];
#pragma unroll
  for (int k = 0; k < kAcc; ++k) acc[k] = 0.0;
  for (int t = lane; t < p.tiles_per_image; t += 32) {
    const double* src = p.partial + ((size_t)b * p.tiles_per_image + t) * kAcc;
#pragma unroll
    for (int k = 0; k < kAcc; ++k) acc[k] += src[k];
  }
#pragma unroll
  for (int k = 0; k < kAcc; ++k)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], d);
  if (lane == 0) {
    double* o = p.out + (size_t)b * 8;
    const double lo = acc[0] * 3.0, lc = acc[1], ll = acc[2] * 2.0;  // od_pb.py:411,421,431
    o[0] = lo + lc + ll;                                              // :370
    o[1] = lo;
    o[2] = lc;
    o[3] = ll;
    o[4] = acc[3] / fmax(acc[4], 1.0);  // rec_bg  :388
    o[5] = acc[5] / fmax(acc[6], 1.0);  // rec_obj :394
    o[6] = acc[7] / fmax(acc[6], 1.0);  // acc_clf :399
    o[7] = acc[6];                      // number of positive anchors
  }
}

}  // namespace

size_t odk_assign_lists_bytes(int batch, int gmax);
size_t odk_assign_workspace_bytes_impl(int n, int batch, int gmax);
int odk_assign_check(const odk_anchors* a, int batch, const void* gt_boxes, const void* gt_count, int gmax, const void* pb, const void* gt,
                     const void* status);
int odk_launch_assign_candidates(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, void* lists_ws,
                                 unsigned* ignore, cudaStream_t stream);
int odk_launch_assign_greedy(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                             int gmax, int64_t* pb_indices, int64_t* gt_indices, int32_t* status, void* workspace, cudaStream_t stream);

static size_t loss_partial_bytes(int n, int batch) {
  return odk_align_up(sizeof(double) * kAcc * (size_t)batch * ((n + kLossTile - 1) / kLossTile), 256);
}

static int loss_fill(LossArgs& p, const odk_anchors* a, int batch, const float* y_pred, double* out, void* partial) {
  ODK_REQUIRE(a != nullptr && batch >= 1, ODK_ERR_INVALID_ARGUMENT, "loss: bad arguments");
  ODK_REQUIRE(batch <= 65535, ODK_ERR_UNSUPPORTED, "loss: batch=%d > 65535", batch);
  ODK_REQUIRE(y_pred && out, ODK_ERR_INVALID_ARGUMENT, "loss: null pointer");
  ODK_REQUIRE((((uintptr_t)y_pred) & 3) == 0 && (((uintptr_t)out) & 7) == 0, ODK_ERR_INVALID_ARGUMENT, "loss: misaligned buffer");
  memset(&p, 0, sizeof(p));
  p.av = odk_view(a);
  p.batch = batch; p.n = a->n; p.c = a->num_classes; p.W = 3 + a->num_classes + 4;
  p.tiles_per_image = (a->n + kLossTile - 1) / kLossTile;
  p.y_pred = y_pred;
  p.partial = static_cast<double*>(partial);
  p.out = out;
  return ODK_OK;
}

static int loss_finish(const LossArgs& p, cudaStream_t stream) {
  {
    OdkProfileScope prof_("loss_finish_kernel", stream);
    loss_finish_kernel<<<p.batch, 32, 0, stream>>>(p);
  }
  ODK_LAUNCH_CHECK("loss_finish_kernel");
  return ODK_OK;
}

extern "C" size_t odk_loss_workspace_bytes(const odk_anchors* a, int batch) {
  if (a == nullptr || batch < 1) return 0;
  return loss_partial_bytes(a->n, batch);
}

extern "C" int odk_loss(const odk_anchors* a, int batch, const float* y_true, const float* y_pred, double* out, void* workspace,
                        size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LossArgs p;
  if (int rc = loss_fill(p, a, batch, y_pred, out, workspace)) return rc;
  ODK_REQUIRE(y_true != nullptr && (((uintptr_t)y_true) & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "loss: y_true null or misaligned");
  ODK_REQUIRE(workspace != nullptr && workspace_bytes >= loss_partial_bytes(a->n, batch) && ((uintptr_t)workspace & 7) == 0, ODK_ERR_WORKSPACE,
              "loss: workspace of %zu bytes is smaller than the required %zu (or misaligned)", workspace_bytes, loss_partial_bytes(a->n, batch));
  p.y_true = y_true;
  const size_t smem = 2 * (size_t)kLossTile * p.W * 4;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "loss: row width %d floats needs %zu bytes of shared memory per tile", p.W, smem);
  ODK_CUDA(cudaFuncSetAttribute(loss_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    OdkProfileScope prof_("loss_dense_kernel", stream);
    loss_dense_kernel<<<dim3(p.tiles_per_image, batch), kLossTile, smem, stream>>>(p);
  }
  ODK_LAUNCH_CHECK("loss_dense_kernel");
  return loss_finish(p, stream);
}

extern "C" size_t odk_assign_loss_workspace_bytes(const odk_anchors* a, int batch, int gmax) {
  if (a == nullptr || batch < 1 || gmax < 1) return 0;
  return odk_align_up(odk_assign_workspace_bytes_impl(a->n, batch, gmax), 256) + loss_partial_bytes(a->n, batch);
}

static int assign_loss_impl(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                            int gmax, const float* y_pred, const float* upstream, double* out, float* grad, int64_t* pb_indices,
                            int64_t* gt_indices, int32_t* status, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = odk_assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  ODK_REQUIRE(gt_classes != nullptr, ODK_ERR_INVALID_ARGUMENT, "assign_loss: null gt_classes");
  const size_t need = odk_assign_loss_workspace_bytes(a, batch, gmax);
  ODK_REQUIRE(workspace != nullptr && workspace_bytes >= need && ((uintptr_t)workspace & 255) == 0, ODK_ERR_WORKSPACE,
              "assign_loss: workspace of %zu bytes is smaller than the required %zu (or not 256-byte aligned)", workspace_bytes, need);
  char* ws = static_cast<char*>(workspace);
  LossArgs p;
  if (int rc = loss_fill(p, a, batch, y_pred, out, ws + odk_align_up(odk_assign_workspace_bytes_impl(a->n, batch, gmax), 256))) return rc;
  const int words = (a->n + 31) / 32;
  unsigned* ignore = reinterpret_cast<unsigned*>(ws + odk_assign_lists_bytes(batch, gmax));
  ODK_CUDA(cudaMemsetAsync(ignore, 0, sizeof(unsigned) * (size_t)batch * words, stream));
  if (int rc = odk_launch_assign_candidates(a, batch, gt_boxes, gt_count, gmax, workspace, ignore, stream)) return rc;
  if (int rc = odk_launch_assign_greedy(a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, workspace, stream))
    return rc;
  p.ignore = ignore; p.bitmap_words = words;
  p.pb_indices = reinterpret_cast<const long long*>(pb_indices);
  p.gt_indices = reinterpret_cast<const long long*>(gt_indices);
  p.gt_boxes = reinterpret_cast<const float4*>(gt_boxes);
  p.gt_classes = gt_classes; p.gt_count = gt_count; p.status = status; p.gmax = gmax;
  p.grad = grad; p.upstream = upstream;
  if (grad != nullptr) {
    const size_t smem = (size_t)kLossTile * p.W * 4;
    ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "assign_loss_grad: row width %d floats needs %zu bytes of shared memory", p.W, smem);
    ODK_CUDA(cudaFuncSetAttribute(loss_sparse_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OdkProfileScope prof_("loss_sparse_grad_kernel", stream);
    loss_sparse_kernel<true><<<dim3(p.tiles_per_image, batch), kLossTile, smem, stream>>>(p);
  } else {
    OdkProfileScope prof_("loss_sparse_kernel", stream);
    loss_sparse_kernel<false><<<dim3(p.tiles_per_image, batch), kLossTile, 0, stream>>>(p);
  }
  ODK_LAUNCH_CHECK("loss_sparse_kernel");
  return loss_finish(p, stream);
}

extern "C" int odk_assign_loss(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                               int gmax, const float* y_pred, double* out, int64_t* pb_indices, int64_t* gt_indices, int32_t* status,
                               void* workspace, size_t workspace_bytes, void* stream) {
  return assign_loss_impl(a, batch, gt_boxes, gt_classes, gt_count, gmax, y_pred, nullptr, out, nullptr, pb_indices, gt_indices, status,
                          workspace, workspace_bytes, stream);
}

extern "C" int odk_assign_loss_grad(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes,
                                    const int32_t* gt_count, int gmax, const float* y_pred, const float* upstream, double* out, float* grad,
                                    int64_t* pb_indices, int64_t* gt_indices, int32_t* status, void* workspace, size_t workspace_bytes,
                                    void* stream) {
  ODK_REQUIRE(grad != nullptr && (((uintptr_t)grad) & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "assign_loss_grad: grad null or misaligned");
  return assign_loss_impl(a, batch, gt_boxes, gt_classes, gt_count, gmax, y_pred, upstream, out, grad, pb_indices, gt_indices, status,
                          workspace, workspace_bytes, stream);
}

extern "C" int odk_loss_grad(const odk_anchors* a, int batch, const float* y_true, const float* y_pred, const float* upstream, float* grad,
                             void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(a != nullptr && batch >= 1 && batch <= 65535, ODK_ERR_INVALID_ARGUMENT, "loss_grad: bad arguments");
  ODK_REQUIRE(y_true && y_pred && grad, ODK_ERR_INVALID_ARGUMENT, "loss_grad: null pointer");
  ODK_REQUIRE((((uintptr_t)y_true | (uintptr_t)y_pred | (uintptr_t)grad) & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "loss_grad: misaligned buffer");
  LossArgs p;
  memset(&p, 0, sizeof(p));
  p.av = odk_view(a);
  p.batch = batch; p.n = a->n; p.c = a->num_classes; p.W = 3 + a->num_classes + 4;
  p.tiles_per_image = (a->n + kLossTile - 1) / kLossTile;
  p.y_true = y_true; p.y_pred = y_pred;
  const size_t smem = 2 * (size_t)kLossTile * p.W * 4;
  ODK_REQUIRE(smem <= 200 * 1024, ODK_ERR_UNSUPPORTED, "loss_grad: row width %d floats needs %zu bytes of shared memory per tile", p.W, smem);
  ODK_CUDA(cudaFuncSetAttribute(loss_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    OdkProfileScope prof_("loss_grad_kernel", stream);
    loss_grad_kernel<<<dim3(p.tiles_per_image, batch), kLossTile, smem, stream>>>(p, upstream, grad);
  }
  ODK_LAUNCH_CHECK("loss_grad_kernel");
  return ODK_OK;
}
