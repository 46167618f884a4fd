// K4 — the y_true writer: PriorBoxes.encode_truth (dl/od_pb.py:299-313) on top of the assignment of
// assign_encode.cu (whose candidate kernel also scatters the ignore mask into a per-image bitmap).  Two
// kernels, arranged so that the HBM-bound stream does not wait for the latency-bound greedy walk:
//
//   encode_rows_kernel     HBM-write-bound: every row of y_true with column 0 = 1/max(1,G) where the anchor is
//                          not ignored, zeros elsewhere — produced 16 bytes at a time in registers from the
//                          packed pb_mask and ignore bitmaps and written with coalesced st.global.v4.  No shared
//                          memory, so it co-resides with the latency-bound kernels running beside it.
//                          Algorithmic bytes: 4*N*(3+C+4) written per image.
//   encode_patch_kernel    the <= G positive rows per image [w0, 1/G, 1, onehot(C), encode_locs] (od_pb.py:
//                          307-311; they are valid whatever the ignore mask says, od_pb.py:359-361), and the
//                          all-zero result of images whose assignment failed.
//
// Stream layout inside odk_assign_encode — two independent chains, forked at entry and joined with events (capturable
// into a CUDA graph):
//     caller's stream            ranked candidate lists -> greedy walk          (latency-bound, a few SMs)
//     library-owned side stream  memset -> ignore mask  -> rows kernel          (the HBM-bound write)
// then the patch kernel on the caller's stream after the join.  The rows kernel needs only the ignore bitmap and the
// walk only the ranked lists, so neither chain waits for the other's first kernel.
//
// odk_assign_encode_compact is the same assignment with a COMPACT result for callers whose consumer is on the host:
// instead of the dense (N, 3+C+4) tensor — 1.8 MB per image of which > 99.7 % is zeros and one repeated constant — it
// writes the packed pb_valids bits, the <= G positive rows and the index lists (~10 KB per image); the dense tensor is
// a trivial host-side expansion of those (object_detector_b200.od_pb.expand_truth).
#include <stdlib.h>

#include "odk_common.cuh"

// from assign_encode.cu
size_t odk_assign_lists_bytes(int batch, int gmax);
int odk_launch_assign_rank(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, void* lists_ws,
                           cudaStream_t stream);
int odk_launch_assign_ignore(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_count, int gmax, unsigned* ignore,
                             cudaStream_t stream);
int odk_launch_assign_greedy(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                             int gmax, int64_t* pb_indices, int64_t* gt_indices, int32_t* status, void* lists_ws, cudaStream_t stream);
int odk_assign_check(const odk_anchors* a, int batch, const void* gt_boxes, const void* gt_count, int gmax, const void* pb, const void* gt,
                     const void* status);
int odk_assign_workspace_check(const odk_anchors* a, int batch, int gmax, const void* workspace, size_t workspace_bytes);

namespace {

struct EncodeArgs {
  AnchorView av;
  int batch, gmax;
  const int32_t* gt_count;
  const unsigned* ignore;  // [B, bitmap_words] from assign_ignore_kernel
  int bitmap_words;
  float* y_true;
  uint8_t* pb_valids;      // optional
  unsigned long long* tl;  // diagnostics timeline or null
};

constexpr int kRowsThreads = 256;
constexpr int kRowsPerThread = 8;  // 16-byte stores per thread
#ifndef ODK_ROWS_PER_SM_DEFAULT
#define ODK_ROWS_PER_SM_DEFAULT 4  // persistent grid: profiles/sweep_r02m.log (step 41.95 -> 40.97 us with batches in flight)
#endif

// y_true of one image is a contiguous array of N*W floats, all zero except column 0.  A CTA owns a chunk of
// kRowsThreads * kRowsPerThread 16-byte pieces; every piece is stored exactly once, whole, by fully coalesced
// st.global.v4 (a warp instruction covers 512 contiguous bytes: only full sectors reach L2 — a variant that left the
// column-0 pieces to a second pass wrote half sectors and ran 50 % slower).
//   A. by ROW: the rows whose first float lies in the chunk (~kChunkPieces * 4 / W of them).  A warp takes 32
//      consecutive rows, so ONE packed pb_mask word and ONE ignore word serve the whole warp; a valid row sets the bit
//      of the piece that holds its column-0 float in a 256-byte shared-memory bitmap (one bit per piece of the chunk).
//   B. by PIECE: a piece holds a column-0 float iff its first float sits at column 0 or in the last three columns of
//      a row (W >= 8) — one add, one wrap, two compares from the running column; the bitmap word a warp needs is the
//      same for all its lanes (bit = lane), so the lookup is one broadcast shared-memory load per store.
// The first version derived (row, bitmap word, bit) per piece from global tables: ~75 instructions per 16 bytes made the
// kernel instruction-bound (ncu: math-pipe throttle 20 %, 22 us at B = 64).
template <int VEC>
__device__ __forceinline__ void encode_rows_chunk(const EncodeArgs& p, int b, unsigned chunk, unsigned step_cols, unsigned* s_bits) {
  const unsigned W = 3 + p.av.num_classes + 4;
  const unsigned n = (unsigned)p.av.n;
  const int g_count = p.gt_count[b];
  const bool count_ok = g_count >= 0 && g_count <= p.gmax;   // otherwise the image fails (patch kernel zeroes it)
  const float w0 = count_ok ? __fdiv_rn(1.0f, (float)max(1, g_count)) : 0.0f;  // float32(1/max(1,G)), od_pb.py:306
  const unsigned total = n * W;  // floats per image (< 2^31, checked by the launcher)
  float* y = p.y_true + (size_t)b * total;
  const unsigned pieces = (total + VEC - 1) / VEC;
  constexpr unsigned kChunkPieces = kRowsThreads * kRowsPerThread;
  static_assert(kChunkPieces / 32 <= kRowsThreads, "one thread per bitmap word");
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  // ---- A. piece bitmap of the chunk
  if (threadIdx.x < kChunkPieces / 32) s_bits[threadIdx.x] = 0u;
  __syncthreads();
  const unsigned f0 = chunk * kChunkPieces * VEC;
  const unsigned f1 = min(total, f0 + kChunkPieces * VEC);
  const unsigned r0 = (f0 + W - 1) / W, r1 = (f1 + W - 1) / W;  // rows with f0 <= r * W < f1
  const unsigned* ign = p.ignore + (size_t)b * p.bitmap_words;
  const unsigned* mb = p.av.maskbits;
  for (unsigned r = (r0 & ~31u) + threadIdx.x; r < r1; r += kRowsThreads) {  // aligned to 32 rows: one word of each table per warp
    if (r >= r0) {
      const unsigned bits = __ldg(mb + (r >> 5)) & ign[r >> 5];  // pb_mask and IoU >= 0.5 with some GT (od_pb.py:355-357)
      const bool valid = count_ok && !((bits >> (r & 31)) & 1u);
      if (valid) {
        const unsigned pl = (r * W - f0) / VEC;  // the piece of the chunk that holds the row's first float
        atomicOr(&s_bits[pl >> 5], 1u << (pl & 31));
      }
      if (p.pb_valids != nullptr) p.pb_valids[(size_t)b * n + r] = valid ? 1 : 0;
    }
  }
  __syncthreads();
  // ---- B. the stores.  Piece u of this thread is number u * kRowsThreads + threadIdx.x of the chunk: word u * 8 + warp, bit lane
  const unsigned i0 = chunk * kChunkPieces + threadIdx.x;
  unsigned col = (i0 * VEC) % W;  // column of the piece's first float; consecutive pieces of a thread are step_cols (mod W) apart
#pragma unroll
  for (int u = 0; u < kRowsPerThread; ++u) {
    const unsigned i = i0 + u * kRowsThreads;
    const bool has = (col == 0u) || (VEC > 1 && col > W - VEC);
    const bool on = has && ((s_bits[u * (kRowsThreads / 32) + warp] >> lane) & 1u);
    const float w = on ? w0 : 0.0f;
    const unsigned jj = (col == 0u) ? 0u : W - col;  // position of the column-0 float inside the piece
    if (i < pieces) {
      if (VEC == 4)
        reinterpret_cast<float4*>(y)[i] = make_float4(jj == 0 ? w : 0.f, jj == 1 ? w : 0.f, jj == 2 ? w : 0.f, jj == 3 ? w : 0.f);
      else
        y[i] = w;
    }
    col += step_cols;
    if (col >= W) col -= W;
  }
}

// One CTA per chunk of kRowsThreads * kRowsPerThread pieces: grid (chunks per image, batch).
template <int VEC>
__global__ void __launch_bounds__(kRowsThreads) encode_rows_kernel(const EncodeArgs p, unsigned step_cols) {
  OdkTimeline tl_(ODK_TL_ROWS, p.tl);
  __shared__ unsigned s_bits[kRowsThreads * kRowsPerThread / 32];
  encode_rows_chunk<VEC>(p, blockIdx.y, blockIdx.x, step_cols, s_bits);
}

// The same chunks walked by a bounded, persistent grid (a few CTAs per SM): the kernel then occupies a fixed share of
// every SM's thread slots and registers for its whole duration instead of queueing thousands of CTAs in front of every
// kernel launched after it, so the read-streaming kernel of the predict side and the latency-bound kernels of other
// batches in flight find room beside it.
template <int VEC>
__global__ void __launch_bounds__(kRowsThreads) encode_rows_persistent_kernel(const EncodeArgs p, unsigned step_cols, unsigned chunks_per_image,
                                                                             unsigned total_chunks) {
  OdkTimeline tl_(ODK_TL_ROWS, p.tl);
  __shared__ unsigned s_bits[2][kRowsThreads * kRowsPerThread / 32];  // alternating: a chunk's stores may still read while the next is built
  unsigned k = 0;
  for (unsigned item = blockIdx.x; item < total_chunks; item += gridDim.x, k ^= 1u) {
    const unsigned b = item / chunks_per_image;
    encode_rows_chunk<VEC>(p, (int)b, item - b * chunks_per_image, step_cols, s_bits[k]);
  }
}

struct PatchArgs {
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
  unsigned long long* tl;
};

// The y_true row of a positive: [w0, 1/G, 1, onehot(C), encode_locs] (od_pb.py:307-311); `row` is pre-zeroed.
__device__ __forceinline__ void write_positive_row(float* row, const AnchorView& av, int pa, const float4 box, int cls, int G) {
  const int C = av.num_classes;
  const float4 ab = __ldg(av.locs + pa);
  const double2 sz = av.sizes[pa];
  row[0] = __fdiv_rn(1.0f, (float)max(1, G));  // pb_valids[pb_indices] = True (od_pb.py:361)
  row[1] = __fdiv_rn(1.0f, (float)G);
  row[2] = 1.0f;
  row[3 + cls] = 1.0f;
  row[3 + C + 0] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.x, ab.x), sz.x), 0.10);
  row[3 + C + 1] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.y, ab.y), sz.y), 0.10);
  row[3 + C + 2] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.z, ab.z), sz.x), 0.10);
  row[3 + C + 3] = (float)__ddiv_rn(__ddiv_rn((double)__fsub_rn(box.w, ab.w), sz.y), 0.10);
}

// One CTA per image: the positives (od_pb.py:307-311), or an all-zero image when the assignment failed.
__global__ void __launch_bounds__(128) encode_patch_kernel(const PatchArgs p) {
  OdkTimeline tl_(ODK_TL_PATCH, p.tl);
  const int b = blockIdx.x;
  const int W = 3 + p.av.num_classes + 4;
  float* y = p.y_true + (size_t)b * p.av.n * W;
  if (p.status[b] != ODK_IMG_OK) {
    for (int a = threadIdx.x; a < p.av.n; a += blockDim.x) {
      y[(size_t)a * W] = 0.0f;
      if (p.pb_valids != nullptr) p.pb_valids[(size_t)b * p.av.n + a] = 0;
    }
    return;
  }
  const int G = p.gt_count[b];
  for (int k = threadIdx.x; k < G; k += blockDim.x) {  // assignment slot k: anchor pa owns GT g
    const int pa = (int)p.pb_indices[(size_t)b * p.gmax + k];
    const int g = (int)p.gt_indices[(size_t)b * p.gmax + k];
    write_positive_row(y + (size_t)pa * W, p.av, pa, p.gt_boxes[(size_t)b * p.gmax + g], p.gt_classes[(size_t)b * p.gmax + g], G);
    if (p.pb_valids != nullptr) p.pb_valids[(size_t)b * p.av.n + pa] = 1;
  }
}

struct CompactArgs {
  PatchArgs q;               // y_true / pb_valids unused
  const unsigned* ignore;    // [B, bitmap_words]
  int bitmap_words;
  float* pos_rows;           // [B, Gmax, W]
  unsigned* valid_bits;      // [B, bitmap_words]: bit a = pb_valids[a]
};

// One CTA per image: the compact form of encode_truth's result.
__global__ void __launch_bounds__(128) encode_compact_kernel(const CompactArgs c) {
  const PatchArgs& p = c.q;
  const int b = blockIdx.x;
  const int W = 3 + p.av.num_classes + 4;
  const int n = p.av.n;
  float* rows = c.pos_rows + (size_t)b * p.gmax * W;
  unsigned* vb = c.valid_bits + (size_t)b * c.bitmap_words;
  const bool ok = p.status[b] == ODK_IMG_OK;
  const int G = ok ? p.gt_count[b] : 0;
  for (int i = threadIdx.x; i < p.gmax * W; i += blockDim.x) rows[i] = 0.0f;
  const unsigned* ign = c.ignore + (size_t)b * c.bitmap_words;
  for (int w = threadIdx.x; w < c.bitmap_words; w += blockDim.x) {
    unsigned v = ok ? ~(__ldg(p.av.maskbits + w) & ign[w]) : 0u;  // od_pb.py:355-357
    if (w == c.bitmap_words - 1 && (n & 31)) v &= (1u << (n & 31)) - 1u;
    vb[w] = v;
  }
  __syncthreads();  // zeroed rows / plain bitmap words are in place before the positives touch them
  for (int k = threadIdx.x; k < G; k += blockDim.x) {
    const int pa = (int)p.pb_indices[(size_t)b * p.gmax + k];
    const int g = (int)p.gt_indices[(size_t)b * p.gmax + k];
    write_positive_row(rows + (size_t)k * W, p.av, pa, p.gt_boxes[(size_t)b * p.gmax + g], p.gt_classes[(size_t)b * p.gmax + g], G);
    atomicOr(vb + (pa >> 5), 1u << (pa & 31));  // od_pb.py:361
  }
}

// ---- per-host-thread side stream (fork/join with events; works under stream capture) ---------------------------
struct SideLane {
  cudaStream_t stream;
  cudaEvent_t fork, join;
  bool ready;
};

// A small pool of lanes per host thread and device, used round-robin: consecutive calls (e.g. the steps of several
// batches captured into one CUDA graph, or issued on different caller streams) must not be chained to each other
// through a shared side stream.
int side_lane(SideLane** out) {
  constexpr int kMaxDevices = 64, kLanes = 8;
  static thread_local SideLane lanes[kMaxDevices][kLanes] = {};
  static thread_local unsigned next[kMaxDevices] = {};
  int dev = 0;
  ODK_CUDA(cudaGetDevice(&dev));
  ODK_REQUIRE(dev >= 0 && dev < kMaxDevices, ODK_ERR_UNSUPPORTED, "device index %d out of range", dev);
  SideLane& l = lanes[dev][next[dev]++ % kLanes];
  if (!l.ready) {
    ODK_CUDA(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
    ODK_CUDA(cudaEventCreateWithFlags(&l.fork, cudaEventDisableTiming));
    ODK_CUDA(cudaEventCreateWithFlags(&l.join, cudaEventDisableTiming));
    l.ready = true;
  }
  *out = &l;
  return ODK_OK;
}

int rows_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

int launch_rows(const EncodeArgs& e, int n, int W, cudaStream_t stream) {
  const long long total = (long long)n * W;
  ODK_REQUIRE(total < 0x7fffffffLL, ODK_ERR_UNSUPPORTED, "assign_encode: image of %lld floats too large", total);
  const bool vec = (total % 4 == 0) && (((uintptr_t)e.y_true & 15) == 0);
  const int v = vec ? 4 : 1;
  const long long pieces = total / v;
  const long long per_cta = (long long)kRowsThreads * kRowsPerThread;
  const long long gx = (pieces + per_cta - 1) / per_cta;
  const unsigned step_cols = (unsigned)((kRowsThreads * v) % W);  // column advance between a thread's consecutive pieces
  // ODK_ROWS_PER_SM = k > 0: persistent grid of k CTAs per SM; 0: one CTA per chunk
  static const int per_sm = getenv("ODK_ROWS_PER_SM") ? atoi(getenv("ODK_ROWS_PER_SM")) : ODK_ROWS_PER_SM_DEFAULT;
  const long long chunks = gx * e.batch;
  ODK_REQUIRE(chunks < 0x7fffffffLL, ODK_ERR_UNSUPPORTED, "assign_encode: %lld row chunks", chunks);
  ODK_MAX_CARVEOUT_ONCE(encode_rows_persistent_kernel<4>); ODK_MAX_CARVEOUT_ONCE(encode_rows_persistent_kernel<1>);
  ODK_MAX_CARVEOUT_ONCE(encode_rows_kernel<4>); ODK_MAX_CARVEOUT_ONCE(encode_rows_kernel<1>);
  {
    OdkProfileScope prof_("encode_rows_kernel", stream);
    // a launch of up to ~16 chunks per persistent CTA shares the SMs with whatever runs beside it; a larger one (B = 256,
    // the stress shape) is better off with one CTA per chunk and every thread slot (profiles/bench_r02n.json: cfg3 141 vs 107 us)
    if (per_sm > 0 && chunks > (long long)rows_sm_count() * per_sm && chunks <= (long long)rows_sm_count() * per_sm * 16) {
      const unsigned grid = (unsigned)(rows_sm_count() * per_sm);
      if (vec) encode_rows_persistent_kernel<4><<<grid, kRowsThreads, 0, stream>>>(e, step_cols, (unsigned)gx, (unsigned)chunks);
      else encode_rows_persistent_kernel<1><<<grid, kRowsThreads, 0, stream>>>(e, step_cols, (unsigned)gx, (unsigned)chunks);
    } else {
      dim3 grid((unsigned)gx, (unsigned)e.batch);
      if (vec) encode_rows_kernel<4><<<grid, kRowsThreads, 0, stream>>>(e, step_cols);
      else encode_rows_kernel<1><<<grid, kRowsThreads, 0, stream>>>(e, step_cols);
    }
  }
  ODK_LAUNCH_CHECK("encode_rows_kernel");
  return ODK_OK;
}

void fill_patch(PatchArgs& q, const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                int gmax, const int64_t* pb_indices, const int64_t* gt_indices, const int32_t* status, float* y_true, uint8_t* pb_valids) {
  q.av = odk_view(a);
  q.batch = batch; q.gmax = gmax;
  q.gt_boxes = reinterpret_cast<const float4*>(gt_boxes);
  q.gt_classes = gt_classes; q.gt_count = gt_count;
  q.pb_indices = reinterpret_cast<const long long*>(pb_indices);
  q.gt_indices = reinterpret_cast<const long long*>(gt_indices);
  q.status = status; q.y_true = y_true; q.pb_valids = pb_valids; q.tl = odk_timeline_ptr();
}

}  // namespace

extern "C" int odk_assign_encode(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes, const int32_t* gt_count,
                                 int gmax, float* y_true, int64_t* pb_indices, int64_t* gt_indices, uint8_t* pb_valids, int32_t* status,
                                 void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = odk_assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  if (int rc = odk_assign_workspace_check(a, batch, gmax, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(gt_classes && y_true, ODK_ERR_INVALID_ARGUMENT, "assign_encode: null pointer");
  ODK_REQUIRE(((uintptr_t)y_true & 3) == 0, ODK_ERR_INVALID_ARGUMENT, "assign_encode: y_true must be 4-byte aligned");
  ODK_REQUIRE(batch <= 65535, ODK_ERR_UNSUPPORTED, "assign_encode: batch=%d > 65535", batch);
  const int words = (a->n + 31) / 32;
  unsigned* ignore = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + odk_assign_lists_bytes(batch, gmax));
  SideLane* lane = nullptr;
  if (int rc = side_lane(&lane)) return rc;
  static const bool single_stream = getenv("ODK_SINGLE_STREAM") != nullptr;  // debugging aid: no fork/join
  SideLane same = {stream, lane->fork, lane->join, true};
  // while per-launch event timing is on, the two chains run one after the other: a kernel's event pair then brackets that
  // kernel alone (with the fork, the rows kernel is timed while the greedy walk of the same call runs beside it)
  if (single_stream || odk_profile_on()) lane = &same;

  // fork at entry: the side stream builds the ignore mask (scattered per GT) and streams out the rows (HBM-bound) ...
  ODK_CUDA(cudaEventRecord(lane->fork, stream));
  ODK_CUDA(cudaStreamWaitEvent(lane->stream, lane->fork, 0));
  int rc_side = odk_launch_assign_ignore(a, batch, gt_boxes, gt_count, gmax, ignore, lane->stream);
  if (!rc_side) {
    EncodeArgs e;
    e.av = odk_view(a);
    e.batch = batch; e.gmax = gmax;
    e.gt_count = gt_count; e.ignore = ignore; e.bitmap_words = words;
    e.y_true = y_true; e.pb_valids = pb_valids; e.tl = odk_timeline_ptr();
    rc_side = launch_rows(e, a->n, 3 + a->num_classes + 4, lane->stream);
  }
  ODK_CUDA(cudaEventRecord(lane->join, lane->stream));
  // ... while the caller's stream ranks the in-cell anchors of every GT and walks the GTs (latency-bound, one CTA per image)
  int rc_main = rc_side;
  if (!rc_main) rc_main = odk_launch_assign_rank(a, batch, gt_boxes, gt_count, gmax, workspace, stream);
  if (!rc_main) rc_main = odk_launch_assign_greedy(a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, workspace, stream);
  // join (never leave the side stream un-joined: stream capture), then the positives
  ODK_CUDA(cudaStreamWaitEvent(stream, lane->join, 0));
  if (rc_main) return rc_main;
  PatchArgs q;
  fill_patch(q, a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, y_true, pb_valids);
  ODK_MAX_CARVEOUT_ONCE(encode_patch_kernel);
  {
    OdkProfileScope prof_("encode_patch_kernel", stream);
    odk_launch_high(encode_patch_kernel, dim3(batch), dim3(128), 0, stream, q);
  }
  ODK_LAUNCH_CHECK("encode_patch_kernel");
  return ODK_OK;
}

extern "C" int odk_assign_encode_compact(const odk_anchors* a, int batch, const float* gt_boxes, const int32_t* gt_classes,
                                         const int32_t* gt_count, int gmax, float* pos_rows, uint32_t* valid_bits, int64_t* pb_indices,
                                         int64_t* gt_indices, int32_t* status, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = odk_assign_check(a, batch, gt_boxes, gt_count, gmax, pb_indices, gt_indices, status)) return rc;
  if (int rc = odk_assign_workspace_check(a, batch, gmax, workspace, workspace_bytes)) return rc;
  ODK_REQUIRE(gt_classes && pos_rows && valid_bits, ODK_ERR_INVALID_ARGUMENT, "assign_encode_compact: null pointer");
  const int words = (a->n + 31) / 32;
  unsigned* ignore = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + odk_assign_lists_bytes(batch, gmax));
  SideLane* lane = nullptr;
  if (int rc = side_lane(&lane)) return rc;
  ODK_CUDA(cudaEventRecord(lane->fork, stream));
  ODK_CUDA(cudaStreamWaitEvent(lane->stream, lane->fork, 0));
  int rc = odk_launch_assign_ignore(a, batch, gt_boxes, gt_count, gmax, ignore, lane->stream);
  ODK_CUDA(cudaEventRecord(lane->join, lane->stream));
  if (!rc) rc = odk_launch_assign_rank(a, batch, gt_boxes, gt_count, gmax, workspace, stream);
  if (!rc) rc = odk_launch_assign_greedy(a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, workspace, stream);
  ODK_CUDA(cudaStreamWaitEvent(stream, lane->join, 0));
  if (rc) return rc;
  CompactArgs c;
  fill_patch(c.q, a, batch, gt_boxes, gt_classes, gt_count, gmax, pb_indices, gt_indices, status, nullptr, nullptr);
  c.ignore = ignore; c.bitmap_words = words; c.pos_rows = pos_rows; c.valid_bits = valid_bits;
  ODK_MAX_CARVEOUT_ONCE(encode_compact_kernel);
  {
    OdkProfileScope prof_("encode_compact_kernel", stream);
    odk_launch_high(encode_compact_kernel, dim3(batch), dim3(128), 0, stream, c);
  }
  ODK_LAUNCH_CHECK("encode_compact_kernel");
  return ODK_OK;
}

