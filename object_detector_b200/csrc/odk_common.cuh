// Shared device/host helpers of libodk (sm_100a only).
//
// Bit-exactness rule: every arithmetic step that the reference performs in NumPy is
// written with the round-to-nearest intrinsics (__fadd_rn/__fsub_rn/__fmul_rn/__fdiv_rn), which the
// compiler never contracts into FMAs, in the reference's operation order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/odk.h"

// Device-side bounds assertions for the checks build (python -m object_detector_b200.build --checks ->
// libodk_checks.so, selected with ODK_LIB_VARIANT=checks): compute-sanitizer is closed on this GPU pool, so the index
// arithmetic of the shared-memory structures that are re-typed, bit-packed or written by several warps is guarded by
// asserts that trap (and fail the CUDA call) instead.  Compiled out of the product library.
#ifdef ODK_CHECKS
#include <assert.h>
#define ODK_ASSERT(cond) assert(cond)
#else
#define ODK_ASSERT(cond) ((void)0)
#endif

#define ODK_MAX_MAPS 8
#define ODK_IDX_BITS 20                 // anchor index field of the sort keys (N < 2^20)
#define ODK_IDX_MASK ((1u << ODK_IDX_BITS) - 1u)
#define ODK_MAX_CLASSES 4096            // class field of the sort keys (12 bits)
#define ODK_MAX_SPANS 8                 // valid-anchor runs per tile in the span table (shorter gaps are merged)

struct odk_anchors {
  int n, num_classes, n_maps, n_patterns, n_valid, device;
  int map_size[ODK_MAX_MAPS];
  int map_off[ODK_MAX_MAPS + 1];
  float4* d_locs;    // pb_locs  (x1,y1,x2,y2) unclipped
  float4* d_grid;    // pb_grid  cell bounds
  double2* d_sizes;  // pb_sizes (w,h) fp64
  float2* d_scale;   // float32(0.10 * pb_sizes)
  uint8_t* d_mask;   // pb_mask
  uint32_t* d_maskbits;  // pb_mask packed, bit (a & 31) of word a >> 5
  uint32_t* d_spans;     // [ceil(n / span_tile)][ODK_MAX_SPANS]: (first row << 16 | rows) of the runs of valid anchors in a tile
  int span_tile;         // tile height the span table was built for (= odk_piece_size(num_classes))
  int span_rows;         // anchors covered by all spans (statistics)
  int span_count;        // spans in the table (statistics)
  void* d_slab;          // the one allocation holding all of the above
};

// By-value kernel argument (pointers + layout); no device-side handle dereference.
struct AnchorView {
  int n, num_classes, n_maps, n_patterns, n_valid;
  int map_size[ODK_MAX_MAPS];
  int map_off[ODK_MAX_MAPS + 1];
  const float4* locs;
  const float4* grid;
  const double2* sizes;
  const float2* scale;
  const uint8_t* mask;
  const uint32_t* maskbits;
  const uint32_t* spans;
  int span_tile;
};

inline AnchorView odk_view(const odk_anchors* a) {
  AnchorView v;
  v.n = a->n; v.num_classes = a->num_classes; v.n_maps = a->n_maps; v.n_patterns = a->n_patterns;
  v.n_valid = a->n_valid;
  for (int i = 0; i < ODK_MAX_MAPS; ++i) v.map_size[i] = a->map_size[i];
  for (int i = 0; i <= ODK_MAX_MAPS; ++i) v.map_off[i] = a->map_off[i];
  v.locs = a->d_locs; v.grid = a->d_grid; v.sizes = a->d_sizes; v.scale = a->d_scale; v.mask = a->d_mask; v.maskbits = a->d_maskbits;
  v.spans = a->d_spans; v.span_tile = a->span_tile;
  return v;
}

// ---- host-side error plumbing ------------------------------------------------------------------
int odk_set_error(int code, const char* fmt, ...);
#define ODK_REQUIRE(cond, code, ...) \
  do { if (!(cond)) return odk_set_error((code), __VA_ARGS__); } while (0)
#define ODK_CUDA(expr) \
  do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
    return odk_set_error(ODK_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); } while (0)
#define ODK_LAUNCH_CHECK(name) \
  do { cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) \
    return odk_set_error(ODK_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e_)); } while (0)

__host__ __device__ static inline size_t odk_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
// Optional per-launch CUDA-event timing (odk_profile_enable / odk_profile_read); a no-op unless enabled.
bool odk_profile_on();  // this host thread's profiling switch (odk_profile_enable)
struct OdkProfileScope {
  OdkProfileScope(const char* name, cudaStream_t stream);
  ~OdkProfileScope();
  int slot_;
  cudaStream_t stream_;
};

static inline int odk_next_pow2(int x) { int p = 1; while (p < x) p <<= 1; return p; }

// ---- exact fp32 box arithmetic (ml.py:485-519, 605-618) ------------------------------------------
__device__ __forceinline__ float odk_area(const float4 b) {
  return __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
}

// Intersection area exactly as compute_iou: prod(rb - lt) * all(lt < rb).  `hit` tells whether the
// boxes overlap with positive extent (otherwise the product is multiplied by 0).
__device__ __forceinline__ float odk_inter(const float4 a, const float4 b, bool& hit) {
  const float lx = fmaxf(a.x, b.x), ly = fmaxf(a.y, b.y);
  const float rx = fminf(a.z, b.z), ry = fminf(a.w, b.w);
  hit = (lx < rx) && (ly < ry);
  return __fmul_rn(__fsub_rn(rx, lx), __fsub_rn(ry, ly));
}

// fl(inter / ((area_a + area_b) - inter)): the reference's IoU value.
__device__ __forceinline__ float odk_iou_value(const float4 a, float area_a, const float4 b, float area_b) {
  bool hit;
  float inter = odk_inter(a, b, hit);
  inter = hit ? inter : 0.0f;
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  return __fdiv_rn(inter, uni);
}

// Exact evaluation of  fl(inter/union) >= thr  (or > thr) with a division-free fast path: outside a
// 1e-5 relative guard band around thr*union the rounded quotient cannot cross thr (rounding is
// monotonic and thr is representable), inside it the IEEE division decides.
template <bool GE>
__device__ __forceinline__ bool odk_iou_passes(const float4 a, float area_a, const float4 b, float area_b,
                                               float thr) {
  // Written without early returns so that callers' loops stay warp-convergent (a divergent `return`
  // inside a loop body serialises the lanes for the rest of the loop on sm_70+).
  bool hit;
  const float inter = odk_inter(a, b, hit);
  const float uni = __fsub_rn(__fadd_rn(area_a, area_b), inter);
  const float p = thr * uni;
  // (hit implies both boxes have positive extent, hence uni > 0; NaNs fail both tests and take the division)
  const bool sure_yes = inter > p * 1.00001f;
  const bool sure_no = inter < p * 0.99999f;
  bool res = sure_yes;
  if (hit && !sure_yes && !sure_no) {  // rare: inside the guard band (or degenerate union)
    const float q = __fdiv_rn(inter, uni);
    res = GE ? (q >= thr) : (q > thr);
  }
  const bool miss = GE ? (0.0f >= thr) : (0.0f > thr);  // IoU is exactly 0 without overlap
  return hit ? res : miss;
}

// ---- 64-bit sort keys of the NMS pipeline -------------------------------------------------------------
// flip_score maps fp32 to uint32 so that unsigned order == float order (negative scores included).
__device__ __forceinline__ uint32_t odk_flip_score(float f) {
  const uint32_t u = __float_as_uint(f);
  return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float odk_unflip_score(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u ^ 0x80000000u) : ~u);
}
// key1 = [4095-class : 12][flip(score) : 32][anchor : 20]; descending key1 == class ascending, score
// descending, anchor index descending.
__device__ __forceinline__ unsigned long long odk_make_key1(int cls, float conf, int idx) {
  return ((unsigned long long)(ODK_MAX_CLASSES - 1 - cls) << (32 + ODK_IDX_BITS)) |
         ((unsigned long long)odk_flip_score(conf) << ODK_IDX_BITS) | (unsigned long long)idx;
}
__device__ __forceinline__ int odk_key1_class(unsigned long long k) {
  return ODK_MAX_CLASSES - 1 - (int)(k >> (32 + ODK_IDX_BITS));
}
// key2 = [flip(score) : 32][4095-class : 12][anchor : 20]
__device__ __forceinline__ unsigned long long odk_key1_to_key2(unsigned long long k) {
  const unsigned long long cls_field = k >> (32 + ODK_IDX_BITS);
  const unsigned long long score = (k >> ODK_IDX_BITS) & 0xffffffffull;
  return (score << 32) | (cls_field << ODK_IDX_BITS) | (k & ODK_IDX_MASK);
}

__device__ __forceinline__ unsigned long long odk_make_key2(int cls, float conf, int idx) {
  return ((unsigned long long)odk_flip_score(conf) << 32) |
         ((unsigned long long)(ODK_MAX_CLASSES - 1 - cls) << ODK_IDX_BITS) | (unsigned long long)idx;
}
__device__ __forceinline__ unsigned long long odk_key2_to_key1(unsigned long long k) {
  const unsigned long long score = k >> 32;
  const unsigned long long cls_field = (k >> ODK_IDX_BITS) & (ODK_MAX_CLASSES - 1);
  return (cls_field << (32 + ODK_IDX_BITS)) | (score << ODK_IDX_BITS) | (k & ODK_IDX_MASK);
}
__device__ __forceinline__ int odk_key2_class(unsigned long long k) {
  return ODK_MAX_CLASSES - 1 - (int)((k >> ODK_IDX_BITS) & (ODK_MAX_CLASSES - 1));
}
__device__ __forceinline__ float odk_key2_score(unsigned long long k) { return odk_unflip_score((uint32_t)(k >> 32)); }

// ---- per-image score histogram (filled by score_emit, read by the first select round of the NMS kernel) ------------
// 1024 bins over the fp32 bit pattern of conf: 128 bins per octave from 2^-7 up (bin = (bits >> 16) - 0x3C00, clamped),
// so the bin index is monotone in conf and a bin's lower bound is a key threshold.  conf <= 2^-7 (or negative) falls
// into bin 0, conf >= 2 into the last bin: still correct, only coarser.
#define ODK_HIST_BINS 1024
__device__ __forceinline__ int odk_score_bin(float conf) {
  const uint32_t u = __float_as_uint(conf);
  if (u >> 31) return 0;
  const int i = (int)(u >> 16) - 0x3C00;
  return i < 0 ? 0 : (i > ODK_HIST_BINS - 1 ? ODK_HIST_BINS - 1 : i);
}
// smallest key2 whose score lies in bin >= d (d >= 1); every key of a lower bin is smaller
__device__ __forceinline__ unsigned long long odk_bin_lo_key(int d) {
  return (unsigned long long)((((uint32_t)(d + 0x3C00)) << 16) | 0x80000000u) << 32;
}

// The fp32 decode of od_pb.py:216-222 as the Keras Lambda evaluates it (od_net.py:189).
__device__ __forceinline__ float4 odk_decode_box(const float4 pr, const float4 lc, const float2 sc) {
  float4 box;
  box.x = fminf(fmaxf(__fadd_rn(__fmul_rn(pr.x, sc.x), lc.x), 0.0f), 1.0f);
  box.y = fminf(fmaxf(__fadd_rn(__fmul_rn(pr.y, sc.y), lc.y), 0.0f), 1.0f);
  box.z = fminf(fmaxf(__fadd_rn(__fmul_rn(pr.z, sc.x), lc.z), 0.0f), 1.0f);
  box.w = fminf(fmaxf(__fadd_rn(__fmul_rn(pr.w, sc.y), lc.w), 0.0f), 1.0f);
  return box;
}

// ---- one shared-memory carveout for every kernel of a step -------------------------------------------------------------
// An SM's L1 / shared-memory split is reconfigured only when the SM is idle.  The step mixes kernels that need a lot of
// shared memory (score_emit ring, per-image NMS, greedy walk) with kernels that need next to none (rows, candidates,
// ignore, patch); left to the default, a CTA of the first kind cannot be placed on an SM that currently runs CTAs of the
// second kind until those have drained — the in-kernel timeline showed the greedy walk waiting 16 us for the persistent
// rows kernel to END.  Every kernel of the step therefore asks for the maximum shared-memory carveout once.
template <typename K>
inline void odk_prefer_max_smem(K kern) {
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
}
#define ODK_MAX_CARVEOUT_ONCE(kern) do { static const bool once_ = (odk_prefer_max_smem(kern), true); (void)once_; } while (0)

// ---- priority launches ------------------------------------------------------------------------------------------
// The step mixes HBM-streaming kernels with large grids (encode_rows, score_emit) and short latency-bound kernels of
// 64..1024 small CTAs (candidates, greedy walk, patch, NMS).  Launched at equal priority, the CTAs of a latency-bound
// kernel queue behind every not-yet-dispatched CTA of a streaming kernel that was issued earlier, so "concurrent"
// branches serialise.  The latency-bound kernels are therefore launched with the highest kernel priority
// (cudaLaunchAttributePriority; honoured for stream launches and captured into graph kernel nodes).
inline int odk_high_priority() {
  static int greatest = 1;
  if (greatest == 1) {
    int least = 0, g = 0;
    if (cudaDeviceGetStreamPriorityRange(&least, &g) != cudaSuccess) g = 0;
    greatest = g;
  }
  return greatest;
}
template <typename... KArgs, typename... Args>
inline cudaError_t odk_launch_high(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  static const bool no_prio = getenv("ODK_NO_PRIORITY") != nullptr;  // diagnostics
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributePriority;
  attr[0].val.priority = odk_high_priority();
  cfg.attrs = attr;
  cfg.numAttrs = no_prio ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
// (An L2 access-policy window pinning the 0.9 MB anchor-table slab for these launches was tried: no measurable change —
// with the evict-first hint on the streaming loads the tables stay resident anyway.)

// ---- optional in-kernel timeline (odk_timeline_set; diagnostics for overlapping launches inside a CUDA graph) --------
// When a buffer is installed, thread 0 of every CTA of an instrumented kernel records %globaltimer (ns) at entry
// (atomicMin into slot 2*id) and at exit (atomicMax into slot 2*id+1).  The pointer travels in the kernel arguments
// (constant bank: no load on the product path when it is null; a __device__ global cost every CTA an L2 round trip at
// entry — 20 % of the stall samples of encode_rows).  Kernel arguments are fixed at capture time: install the buffer
// before capturing a CUDA graph that should be traced.
unsigned long long* odk_timeline_ptr();  // host: the installed buffer or null (odk_api.cu)
struct OdkTimeline {
  int id_;
  unsigned long long* buf_;
  __device__ __forceinline__ static unsigned long long now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
  }
  __device__ __forceinline__ OdkTimeline(int id, unsigned long long* buf) : id_(id), buf_(buf) {
    if (buf_ != nullptr && threadIdx.x == 0) atomicMin(buf_ + 2 * id_, now());
  }
  __device__ __forceinline__ ~OdkTimeline() {
    if (buf_ != nullptr && threadIdx.x == 0) atomicMax(buf_ + 2 * id_ + 1, now());
  }
};
enum { ODK_TL_CANDIDATES = 0, ODK_TL_GREEDY = 1, ODK_TL_ROWS = 2, ODK_TL_PATCH = 3, ODK_TL_EMIT = 4, ODK_TL_NMS = 5, ODK_TL_IGNORE = 6,
       ODK_TL_SLOTS = 7 };

// ---- warp helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned odk_lane() { return threadIdx.x & 31u; }

// ---- mbarrier + 1-D bulk TMA (cp.async.bulk -> SASS UBLKCP) -------------------------------------------
__device__ __forceinline__ uint32_t odk_smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// one 16-byte shared-memory load (the compiler otherwise splits a dynamically indexed float4 read whose components
// feed a select chain into predicated scalar loads: two dependent shared-memory round trips instead of one)
__device__ __forceinline__ float4 odk_lds_f4(const float4* p) {
  float4 v;
  asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(odk_smem_u32(p)));
  return v;
}
// atomicAdd on shared memory exactly as written (the compiler wraps a predicated atomicAdd in its own warp aggregation:
// a vote, a leader election, a shuffle and a multiply-add per call)
__device__ __forceinline__ unsigned odk_atoms_add(unsigned* p, unsigned v) {
  unsigned old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(odk_smem_u32(p)), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void odk_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(odk_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void odk_mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void odk_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(odk_smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void odk_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(odk_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void odk_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(odk_smem_u32(bar)), "r"(parity)
      : "memory");
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void odk_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   odk_smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(odk_smem_u32(bar))
               : "memory");
}
// The same with an L2 evict-first policy: for data that is read exactly once (the head outputs), so that the stream does
// not push the small L2-resident anchor tables and candidate lists of the concurrently running kernels out of the cache.
__device__ __forceinline__ uint64_t odk_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void odk_bulk_g2s_stream(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                   odk_smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(odk_smem_u32(bar)), "l"(policy)
               : "memory");
}
// shared -> global bulk copy (bulk async-group completion).
__device__ __forceinline__ void odk_bulk_s2g(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(odk_smem_u32(src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void odk_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void odk_bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void odk_bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// Ampere-style per-thread async copies (SASS LDGSTS), sizes 4 / 8 / 16 bytes.
template <int BYTES>
__device__ __forceinline__ void odk_cp_async(void* dst, const void* src) {
  static_assert(BYTES == 4 || BYTES == 8 || BYTES == 16, "cp.async size");
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(odk_smem_u32(dst)), "l"(src), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void odk_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void odk_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void odk_fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
