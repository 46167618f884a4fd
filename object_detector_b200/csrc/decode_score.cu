// K1 — the streaming HBM-bound kernel of the predict side, in two flavours of one template:
//
//   DENSE  (odk_decode_score)  replaces ChannelArgMax/ChannelMax (layers.py:69-98), _objconf
//          (od_net.py:177-184) and the decode_locs Lambda (od_pb.py:216-222, od_net.py:189) — three
//          separate TF elementwise passes in the reference — with one pass over the head outputs:
//            classes = argmax_c clfs (first max wins), conf = (objs * max_c clfs) * pb_mask,
//            box = clip(locs * float32(0.10*sizes) + pb_locs, 0, 1)
//   EMIT   (first half of odk_decode_nms)  computes classes/conf only and emits one 64-bit sort key
//          per anchor with conf > conf_thr (strict, layers.py:935).  `locs` is NOT read here: the
//          NMS kernel decodes lazily only the boxes it visits, so the dense classes/confs/boxes
//          intermediates never exist.
//
// Data movement: persistent CTAs, one tile = TILE consecutive anchors of one image.  The input slabs
// of a tile (clfs TILE*C*4 B, objs TILE*4 B, pb_mask TILE B; DENSE adds locs / pb_locs / scale) are
// contiguous in HBM, so each is fetched with ONE 1-D bulk TMA copy (cp.async.bulk, SASS UBLKCP) into a
// shared-memory ring guarded by mbarriers; a thread then owns one anchor row in shared memory.  Tiles
// whose source addresses / sizes are not 16-byte multiples fall back to cooperative plain loads into
// the same ring slot.
//
// Candidate emission is atomic-free and deterministic: a tile owns the fixed slab
// cand_keys[b][tile*TILE .. +TILE) and compacts its candidates to the front of it (warp ballots + a
// per-tile scan of the warp counts); piece_count[b][tile] holds how many there are.
//
// Algorithmic HBM bytes per anchor: DENSE read 4*(1+C+4), write 4+4+16; EMIT read 4*(1+C), write 8 per
// candidate.
#include "odk_common.cuh"

namespace {

struct DecodeArgs {
  AnchorView av;
  int batch, n, c;
  const float* objs;
  const float* clfs;
  const float* locs;
  int32_t* classes;  // DENSE outputs
  float* confs;
  float4* boxes;
  float conf_thr;                 // EMIT
  unsigned long long* cand_keys;  // [B, tiles_per_image * TILE]
  int* piece_count;               // [B, tiles_per_image]
  int* score_hist;                // [B, ODK_HIST_BINS] (pre-zeroed) or null
  unsigned long long* tl;         // diagnostics timeline or null
  int tiles_per_image;
  int pieces_per_image;  // EMIT: candidate pieces per image (one per tile)
  int stages;  // depth of the shared-memory ring
  int stream_hint;  // 1: head outputs are fetched with an L2 evict-first policy
  int use_spans;    // EMIT: fetch only the runs of valid anchors of a tile (av.spans); rows outside them have pb_mask = 0
};

template <int TILE, bool DENSE>
struct StageLayout {
  // clfs | objs | pb_mask | (DENSE: locs | pb_locs | decode scale); every offset a multiple of 16
  static constexpr size_t kOffObj = 0, kOffAMask = (size_t)TILE * 4, kOffLoc = kOffAMask + (size_t)((TILE + 15) / 16 * 16),
                          kOffALoc = kOffLoc + (DENSE ? (size_t)TILE * 16 : 0), kOffAScale = kOffALoc + (DENSE ? (size_t)TILE * 16 : 0),
                          kTail = kOffAScale + (DENSE ? (size_t)TILE * 8 : 0);
  static constexpr int kMaxStages = 8;
  static size_t stage_bytes(int c) { return (size_t)TILE * c * 4 + kTail; }
};

// SPLIT threads share one anchor row (wide class vectors: each scans 1/SPLIT of the classes, the group merges by
// shuffles), so that a tile of few, long rows still fills the CTA with 256 consumer threads.
//
// Warp roles: kWarps consumer warps (TILE * SPLIT threads) + ONE producer warp.  The producer refills a ring slot the
// moment every consumer warp has released it (per-slot "empty" mbarrier, one arrival per consumer warp) — no CTA-wide
// barrier sits between a slot becoming free and its refill, and a consumer warp that is done with its 32 rows moves on
// to the next tile without waiting for the slowest warp of the CTA.
// CFIX: the class count as a compile-time constant (20 = VOC, the reference's default head) or 0 = p.c at run time.
template <int TILE, bool DENSE, int SPLIT, int CFIX>
__global__ void __launch_bounds__(TILE * SPLIT + 32, DENSE ? 1 : 1024 / (TILE * SPLIT)) decode_score_kernel(const DecodeArgs p) {
  OdkTimeline tl_(ODK_TL_EMIT, p.tl);
  using L = StageLayout<TILE, DENSE>;
  constexpr int kThreads = TILE * SPLIT;  // consumer threads
  const int kStages = p.stages;
  constexpr int kWarps = kThreads / 32;   // consumer warps; warp kWarps is the producer
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int c = CFIX ? CFIX : p.c;
  const size_t clf_bytes = (size_t)TILE * c * 4;
  const size_t stage_bytes = clf_bytes + L::kTail;
  __shared__ uint64_t full_bar[L::kMaxStages];
  __shared__ uint64_t empty_bar[L::kMaxStages];
  __shared__ unsigned s_ctr[L::kMaxStages];  // EMIT: per ring slot, (consumer warps arrived << 16) | candidates so far

  const int total_tiles = p.batch * p.tiles_per_image;  // < 2^31 (checked by the launcher): 32-bit tile arithmetic
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const unsigned lane = odk_lane();

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      odk_mbar_init(&full_bar[s], 1);
      odk_mbar_init(&empty_bar[s], kWarps);
      s_ctr[s] = 0u;
    }
    odk_mbar_fence_init();
  }
  __syncthreads();

  auto stage_ptr = [&](int s) { return smem_raw + (size_t)s * stage_bytes; };
  // (image, tile) of the CTA's tiles are tracked incrementally: t advances by gridDim.x = gq * tiles_per_image + gr
  const int tpi = p.tiles_per_image, gstep = (int)gridDim.x;
  const int gq = gstep / tpi, gr = gstep - gq * tpi;

  if (warp == kWarps) {
    // =========================================== producer warp ===========================================
    const bool span_lane = !DENSE && p.use_spans && lane < ODK_MAX_SPANS;
    auto span_word = [&](int t, int tile_of_t) -> uint32_t {
      return (span_lane && t < total_tiles) ? __ldg(p.av.spans + (size_t)tile_of_t * ODK_MAX_SPANS + lane) : 0u;
    };
    int slot = 0, fill = 0;  // `fill`: how many times the ring has wrapped (the producer waits for release number fill - 1)
    int b = (int)blockIdx.x / tpi, tile = (int)blockIdx.x - b * tpi;
    uint32_t sp = span_word((int)blockIdx.x, tile);
    for (int t = (int)blockIdx.x; t < total_tiles; t += gstep) {
      // span words of the NEXT tile are fetched before this tile's slot is waited for: no global load between a slot
      // becoming free and its refill
      int tile_next = tile + gr, b_next = b + gq;
      if (tile_next >= tpi) { tile_next -= tpi; ++b_next; }
      const uint32_t sp_next = span_word(t + gstep, tile_next);
      if (fill > 0) odk_mbar_wait(&empty_bar[slot], (uint32_t)(fill - 1) & 1u);
      const int a0 = tile * TILE;
      const int rows = min(TILE, p.n - a0);
      const size_t row0 = (size_t)b * p.n + a0;
      const float* src_clf = p.clfs + row0 * c;
      const float* src_obj = p.objs + row0;
      const float* src_loc = DENSE ? p.locs + row0 * 4 : nullptr;
      unsigned char* dst = stage_ptr(slot);
      unsigned char* tail = dst + clf_bytes;
      bool tma_ok = ((rows & 15) == 0) && ((((uintptr_t)src_clf | (uintptr_t)src_obj) & 15) == 0);
      if (DENSE) tma_ok = tma_ok && (((uintptr_t)src_loc & 15) == 0);
      const uint32_t r = (uint32_t)rows;
      if (tma_ok) {
        const uint64_t pol = odk_policy_evict_first();  // head outputs are read exactly once
        if (!DENSE && p.use_spans) {
          // lane k < ODK_MAX_SPANS fetches the k-th run of valid rows of the tile, lanes 8 / 9 the objectness and pb_mask
          // slabs of the whole tile.  Rows between the runs are never read from HBM (their pb_mask is 0, so
          // conf = x * 0 cannot pass conf_thr >= 0) and never looked at by the consumers.
          const uint32_t first = sp >> 16, len = sp & 0xffffu;
          uint32_t bytes = len * (uint32_t)c * 4u, tot = bytes;
#pragma unroll
          for (int d = 1; d < ODK_MAX_SPANS; d <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);  // lanes 0..7 hold the sum
          if (lane == 0) odk_mbar_expect_tx(&full_bar[slot], tot + r * 4 + r);
          __syncwarp();
          if (len) odk_bulk_g2s_stream(dst + (size_t)first * c * 4, src_clf + (size_t)first * c, bytes, &full_bar[slot], pol);
          if (lane == 8) odk_bulk_g2s_stream(tail + L::kOffObj, src_obj, r * 4, &full_bar[slot], pol);
          if (lane == 9) odk_bulk_g2s(tail + L::kOffAMask, p.av.mask + a0, r, &full_bar[slot]);
        } else {
          if (lane == 0) odk_mbar_expect_tx(&full_bar[slot], r * c * 4 + r * 4 + r + (DENSE ? r * 16 + r * 16 + r * 8 : 0));
          __syncwarp();
          if (lane == 0) {
            if (p.stream_hint) odk_bulk_g2s_stream(dst, src_clf, r * c * 4, &full_bar[slot], pol);
            else odk_bulk_g2s(dst, src_clf, r * c * 4, &full_bar[slot]);
          }
          if (lane == 1) {
            if (p.stream_hint) odk_bulk_g2s_stream(tail + L::kOffObj, src_obj, r * 4, &full_bar[slot], pol);
            else odk_bulk_g2s(tail + L::kOffObj, src_obj, r * 4, &full_bar[slot]);
          }
          if (lane == 2) odk_bulk_g2s(tail + L::kOffAMask, p.av.mask + a0, r, &full_bar[slot]);  // anchor tables are L2-resident
          if (DENSE) {
            if (lane == 3) odk_bulk_g2s(tail + L::kOffLoc, src_loc, r * 16, &full_bar[slot]);
            if (lane == 4) odk_bulk_g2s(tail + L::kOffALoc, p.av.locs + a0, r * 16, &full_bar[slot]);
            if (lane == 5) odk_bulk_g2s(tail + L::kOffAScale, p.av.scale + a0, r * 8, &full_bar[slot]);
          }
        }
      } else {
        // tiles whose source addresses / sizes are not 16-byte multiples: plain loads by the producer warp
        float* d0 = reinterpret_cast<float*>(dst);
        float* d2 = reinterpret_cast<float*>(tail + L::kOffObj);
        unsigned char* d5 = tail + L::kOffAMask;
        for (int i = (int)lane; i < rows * c; i += 32) d0[i] = __ldg(src_clf + i);
        for (int i = (int)lane; i < rows; i += 32) {
          d2[i] = __ldg(src_obj + i);
          d5[i] = p.av.mask[a0 + i];
        }
        if (DENSE) {
          float* d1 = reinterpret_cast<float*>(tail + L::kOffLoc);
          float4* d3 = reinterpret_cast<float4*>(tail + L::kOffALoc);
          float2* d4 = reinterpret_cast<float2*>(tail + L::kOffAScale);
          for (int i = (int)lane; i < rows * 4; i += 32) d1[i] = __ldg(src_loc + i);
          for (int i = (int)lane; i < rows; i += 32) {
            d3[i] = __ldg(p.av.locs + a0 + i);
            d4[i] = __ldg(p.av.scale + a0 + i);
          }
        }
        __syncwarp();  // the warp's plain stores are ordered before the (releasing) arrive below
        if (lane == 0) odk_mbar_arrive(&full_bar[slot]);
      }
      sp = sp_next;
      tile = tile_next; b = b_next;
      if (++slot == kStages) { slot = 0; ++fill; }
    }
    return;
  }

  // ============================================= consumer warps =============================================
  const int row_id = tid / SPLIT, part = tid % SPLIT;  // SPLIT consecutive lanes own one anchor row
  int slot = 0;
  uint32_t phase = 0;
  int b = (int)blockIdx.x / tpi, tile = (int)blockIdx.x - b * tpi;
  const unsigned char* st = smem_raw;
  const float conf_thr = p.conf_thr;
  for (int t = (int)blockIdx.x; t < total_tiles; t += gstep) {
    odk_mbar_wait(&full_bar[slot], phase);
    const int a0 = tile * TILE;
    const int rows = min(TILE, p.n - a0);
    const float* s_clf = reinterpret_cast<const float*>(st);
    const float* s_obj = reinterpret_cast<const float*>(st + clf_bytes + L::kOffObj);
    const unsigned char* s_amask = st + clf_bytes + L::kOffAMask;

    int cls = 0;
    float conf = 0.0f;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    const int a = a0 + row_id;
    const bool owner = part == 0 && row_id < rows;
    {
      const float* row = s_clf + (size_t)row_id * c;
      float best = -INFINITY;
      cls = 0x7fffffff;
      // with spans, a row with pb_mask = 0 may hold stale shared memory: it is never scanned (conf = 0, never a candidate)
      const unsigned mk = row_id < rows ? (unsigned)s_amask[row_id] : 0u;
      const float obj = row_id < rows ? s_obj[row_id] : 0.0f;
      const bool scan = row_id < rows && (DENSE || !p.use_spans || mk != 0u);
      if (scan) {
        if ((c & 3) == 0) {
          // The stream is bound by the ALU pipe (compares and selects issue at half rate), so the scan is arranged to
          // need few of them: the maximum of a float4 chunk costs two min/max instructions (FMNMX3 + FMNMX), the running
          // best one compare and two selects per CHUNK, and the position inside the winning chunk is looked up once at
          // the end.  "First maximum wins" (np.argmax): a later chunk replaces the best only if strictly greater, and
          // inside the chunk the first element equal to the maximum is taken.  NaNs never win a comparison.
          const int nchunk = c >> 2, per = (nchunk + SPLIT - 1) / SPLIT;
          const int j0 = part * per, j1 = min(nchunk, j0 + per);
          const float4* row4 = reinterpret_cast<const float4*>(row);
          int jb = -1;
          if (j1 - j0 == 5) {
            // five chunks per thread (C = 20 with a thread per row, C = 80 with four): all five loads are issued before
            // the first use, so the scan costs one shared-memory latency instead of five in a row — a consumer warp is
            // latency-bound here (2 warps per scheduler), not throughput-bound
            const float4 v0 = row4[j0], v1 = row4[j0 + 1], v2 = row4[j0 + 2], v3 = row4[j0 + 3], v4 = row4[j0 + 4];
            const float m0 = fmaxf(fmaxf(v0.x, v0.y), fmaxf(v0.z, v0.w)), m1 = fmaxf(fmaxf(v1.x, v1.y), fmaxf(v1.z, v1.w));
            const float m2 = fmaxf(fmaxf(v2.x, v2.y), fmaxf(v2.z, v2.w)), m3 = fmaxf(fmaxf(v3.x, v3.y), fmaxf(v3.z, v3.w));
            const float m4 = fmaxf(fmaxf(v4.x, v4.y), fmaxf(v4.z, v4.w));
            // fmaxf ignores NaNs, so mx is the maximum of the non-NaN values (NaN only if every value is NaN, which then
            // matches no chunk); the FIRST chunk that attains it wins, a chunk of -inf values included
            const float mx = fmaxf(fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)), m4);
            jb = (m4 == mx) ? j0 + 4 : jb;  // selects, last chunk first: the lowest matching chunk is written last
            jb = (m3 == mx) ? j0 + 3 : jb;
            jb = (m2 == mx) ? j0 + 2 : jb;
            jb = (m1 == mx) ? j0 + 1 : jb;
            jb = (m0 == mx) ? j0 : jb;
            best = (jb >= 0) ? mx : best;
          } else {
            for (int j = j0; j < j1; ++j) {
              const float4 v = row4[j];
              const float mj = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
              if (mj > best || (jb < 0 && mj == best)) { best = mj; jb = j; }  // second clause: a chunk of -inf values
            }
          }
          if (jb >= 0) {
            const float4 v = odk_lds_f4(row4 + jb);  // one 16-byte load, then selects (first element equal to the maximum)
            int pos = 3;
            pos = (v.z == best) ? 2 : pos;
            pos = (v.y == best) ? 1 : pos;
            pos = (v.x == best) ? 0 : pos;
            cls = 4 * jb + pos;
          }
        } else {
          const int per = (c + SPLIT - 1) / SPLIT;
          const int j0 = part * per, j1 = min(c, j0 + per);
          if (j0 < j1) {
            best = row[j0];
            cls = j0;
            for (int j = j0 + 1; j < j1; ++j) {
              const float v = row[j];
              if (v > best) { best = v; cls = j; }
            }
          }
        }
      }
#pragma unroll
      for (int d = 1; d < SPLIT; d <<= 1) {  // merge the parts of a row: larger value, then lower class index
        const float ob = __shfl_xor_sync(0xffffffffu, best, d);
        const int oc = __shfl_xor_sync(0xffffffffu, cls, d);
        if (ob > best || (ob == best && oc < cls)) { best = ob; cls = oc; }
      }
      if (owner && scan) {
        // a row of NaNs leaves best = -inf in every part: np.argmax returns 0 there
        if (cls == 0x7fffffff) { cls = 0; best = row[0]; }
        const float maskf = mk ? 1.0f : 0.0f;
        conf = __fmul_rn(__fmul_rn(obj, best), maskf);  // (objs*confs)*pb_mask, od_net.py:183-184
      }
      if (DENSE && owner) {
        const float4* s_loc = reinterpret_cast<const float4*>(st + clf_bytes + L::kOffLoc);
        const float4* s_aloc = reinterpret_cast<const float4*>(st + clf_bytes + L::kOffALoc);
        const float2* s_ascale = reinterpret_cast<const float2*>(st + clf_bytes + L::kOffAScale);
        box = odk_decode_box(s_loc[row_id], s_aloc[row_id], s_ascale[row_id]);
      }
    }
    if (DENSE) {
      // this warp's rows are in registers: release the ring slot (one arrival per consumer warp)
      __syncwarp();
      if (lane == 0) odk_mbar_arrive(&empty_bar[slot]);
      if (owner) {
        const size_t r = (size_t)b * p.n + a;
        p.classes[r] = cls;
        p.confs[r] = conf;
        p.boxes[r] = box;
      }
    } else {
      // Candidates of the tile are compacted to the front of the tile's slab without a CTA-wide barrier: every consumer
      // warp adds (its count | one arrival) to the slot's packed shared-memory counter with ONE atomic and gets its base
      // offset back; the warp that arrives last knows the tile's total, publishes it and clears the counter — all before
      // it releases the ring slot, so the counter is clean when the slot's next tile starts.  The order of the warps
      // inside a slab varies from run to run; the NMS kernel sorts the keys, so results do not depend on it.
      const bool is_cand = owner && (conf > conf_thr);
      const unsigned m = __ballot_sync(0xffffffffu, is_cand);
      const unsigned cnt = (unsigned)__popc(m);
      unsigned old = 0;
      if (lane == 0) old = odk_atoms_add(&s_ctr[slot], cnt | (1u << 16));  // plain atom.shared: no compiler-side warp aggregation
      old = __shfl_sync(0xffffffffu, old, 0);
      const unsigned base = old & 0xffffu;
      // the piece index of a tile is its global tile index t = b * tiles_per_image + tile
      if ((old >> 16) == (unsigned)(kWarps - 1) && lane == 0) {
        s_ctr[slot] = 0u;
        p.piece_count[t] = (int)(base + cnt);
      }
      __syncwarp();
      if (lane == 0) odk_mbar_arrive(&empty_bar[slot]);  // rows in registers, counter settled: release the ring slot
      ODK_ASSERT((old >> 16) < (unsigned)kWarps && base + cnt <= (unsigned)TILE);
      if (is_cand) {
        ODK_ASSERT(cls >= 0 && cls < c && a < p.n && t == b * tpi + tile);
        p.cand_keys[(size_t)t * TILE + base + __popc(m & ((1u << lane) - 1u))] = odk_make_key2(cls, conf, a);
        if (p.score_hist != nullptr) atomicAdd(p.score_hist + b * ODK_HIST_BINS + odk_score_bin(conf), 1);
      }
    }
    tile += gr; b += gq;
    if (tile >= tpi) { tile -= tpi; ++b; }
    st += stage_bytes;
    if (++slot == kStages) { slot = 0; phase ^= 1u; st = smem_raw; }
  }
}

// decode_locs alone (the Keras Lambda): pure elementwise float4 stream.
__global__ void __launch_bounds__(256) decode_locs_kernel(AnchorView av, long long rows, const float4* __restrict__ pred,
                                                          float4* __restrict__ out) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    const int a = (int)(r % av.n);
    const float4 pr = pred[r];
    const float4 lc = __ldg(av.locs + a);
    const float2 sc = __ldg(av.scale + a);
    out[r] = odk_decode_box(pr, lc, sc);
  }
}

// decode_locs as NumPy evaluates it (od_pb.py:216-222 with xp=np, the path check_assign.py:27 takes): pb_sizes is fp64,
// so the whole expression is fp64 — clip(pred * (0.10 * tile(pb_sizes, 2)) + pb_locs, 0, 1), one rounding per operation.
__global__ void __launch_bounds__(256) decode_locs_f64_kernel(AnchorView av, long long rows, const double* __restrict__ pred,
                                                              double* __restrict__ out) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    const int a = (int)(r % av.n);
    const float4 lc = __ldg(av.locs + a);
    const double2 sz = av.sizes[a];
    const double sx = __dmul_rn(0.10, sz.x), sy = __dmul_rn(0.10, sz.y);
    const double loc[4] = {(double)lc.x, (double)lc.y, (double)lc.z, (double)lc.w};
    const double sc[4] = {sx, sy, sx, sy};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const double v = __dadd_rn(__dmul_rn(pred[4 * r + k], sc[k]), loc[k]);
      out[4 * r + k] = (v != v) ? v : fmin(fmax(v, 0.0), 1.0);  // np.clip keeps NaN
    }
  }
}

// ChannelArgMax / ChannelMax (layers.py:69-98) on their own: one thread per row of `c` channels, first maximum wins
// (the same scan the fused kernel above performs; the hot path never materialises these tensors).
__global__ void __launch_bounds__(256) channel_reduce_kernel(long long rows, int c, const float* __restrict__ x,
                                                             long long* __restrict__ arg, float* __restrict__ mx) {
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (long long)gridDim.x * blockDim.x) {
    const float* row = x + r * c;
    float best = -INFINITY;
    int cls = -1;
    for (int j = 0; j < c; ++j) {
      const float v = __ldg(row + j);
      if (v > best || (cls < 0 && v == best)) { best = v; cls = j; }
    }
    if (cls < 0) { cls = 0; best = __ldg(row); }  // a row of NaNs: np.argmax returns 0, np.max NaN
    if (arg != nullptr) arg[r] = cls;
    if (mx != nullptr) mx[r] = best;
  }
}

int sm_count_cached() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

template <int TILE, bool DENSE, int SPLIT, int CFIX = 0>
int launch_tile(DecodeArgs& args, cudaStream_t stream) {
  using L = StageLayout<TILE, DENSE>;
  static const int env_stages = getenv("ODK_EMIT_STAGES") ? atoi(getenv("ODK_EMIT_STAGES")) : 0;
  static const int env_per_sm = getenv("ODK_EMIT_PER_SM") ? atoi(getenv("ODK_EMIT_PER_SM")) : 0;
  static const int env_hint = getenv("ODK_EMIT_HINT") ? atoi(getenv("ODK_EMIT_HINT")) : 1;
  args.stream_hint = env_hint;
  // spans need 16-byte aligned row starts (C % 4 == 0), a table built for this tile height, conf_thr >= 0 (a masked
  // anchor's conf is 0 or NaN: it can only be a candidate for a negative threshold) — else the whole tile is fetched
  static const int env_spans = getenv("ODK_EMIT_SPANS") ? atoi(getenv("ODK_EMIT_SPANS")) : 1;
  args.use_spans = (!DENSE && env_spans && (args.c & 3) == 0 && args.av.spans != nullptr && args.av.span_tile == TILE &&
                    args.conf_thr >= 0.0f) ? 1 : 0;
  // EMIT: a 3-deep ring at 2 CTAs per SM (2 x 65 KB at C = 20).  With the producer warp refilling a slot as soon as its
  // consumers release it, depth pays off (profiles/sweep_r02d.log: step 44.5 us at 2 deep x 3 CTAs, 43.6 us at
  // 3 deep x 2 CTAs, same shared memory per SM) and two CTAs per SM leave threads and registers to the kernels of the
  // other batches in flight.
  args.stages = (!DENSE && env_stages >= 1 && env_stages <= L::kMaxStages) ? env_stages : 3;
  const size_t smem = L::stage_bytes(args.c) * args.stages;
  auto kern = decode_score_kernel<TILE, DENSE, SPLIT, CFIX>;
  ODK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ODK_MAX_CARVEOUT_ONCE(kern);
  args.tiles_per_image = (args.n + TILE - 1) / TILE;
  args.pieces_per_image = args.tiles_per_image;
  const long long total = (long long)args.batch * args.tiles_per_image;
  ODK_REQUIRE(total + 16 * 2048 * 10 < 0x7fffffffLL, ODK_ERR_UNSUPPORTED, "decode_score: %lld tiles", total);
  int per_sm = (int)((220 * 1024) / (smem + 1024));
  per_sm = per_sm < 1 ? 1 : (per_sm > 2048 / (TILE * SPLIT + 32) ? 2048 / (TILE * SPLIT + 32) : per_sm);
  // EMIT runs beside the train-side kernels of the same step (other graph branch): leave them shared memory
  if (!DENSE && per_sm > (SPLIT > 1 ? 3 : 2)) per_sm = SPLIT > 1 ? 3 : 2;  // wide class vectors: more rows in flight per SM
  if (!DENSE && env_per_sm > 0) per_sm = env_per_sm;
  long long grid = (long long)sm_count_cached() * per_sm;
  if (grid > total) grid = total;
  if (grid < 1) grid = 1;
  {
    OdkProfileScope prof_(DENSE ? "decode_score_kernel" : "score_emit_kernel", stream);
    kern<<<(unsigned)grid, TILE * SPLIT + 32, smem, stream>>>(args);
  }
  ODK_LAUNCH_CHECK(DENSE ? "decode_score_kernel" : "score_emit_kernel");
  return ODK_OK;
}

}  // namespace

// Tile height used for `num_classes` channels: chosen so that the shared-memory ring stays under ~100 KB (2+ CTAs per
// SM).  The span table of the anchors handle is built for it.
int odk_piece_size(int num_classes) {
  static const int env_tile = getenv("ODK_EMIT_TILE") ? atoi(getenv("ODK_EMIT_TILE")) : 0;
  if (num_classes <= 32 && env_tile == 128) return 128;
  return num_classes <= 32 ? 256 : (num_classes <= 96 ? 64 : 32);
}

// Height of the candidate pieces score_emit writes (the NMS workspace layout depends on it): the tile.
int odk_emit_piece_size(int num_classes) { return odk_piece_size(num_classes); }

// Internal entry used by odk_decode_score (dense) and odk_decode_nms (emit).
int odk_launch_decode_score(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs,
                            int32_t* classes, float* confs, float* boxes, bool dense, float conf_thr,
                            unsigned long long* cand_keys, int* piece_count, int* score_hist, cudaStream_t stream) {
  DecodeArgs args;
  args.score_hist = score_hist;
  args.tl = odk_timeline_ptr();
  args.av = odk_view(a);
  args.batch = batch; args.n = a->n; args.c = a->num_classes;
  args.objs = objs; args.clfs = clfs; args.locs = locs;
  args.classes = classes; args.confs = confs; args.boxes = reinterpret_cast<float4*>(boxes);
  args.conf_thr = conf_thr; args.cand_keys = cand_keys; args.piece_count = piece_count;
  args.tiles_per_image = 0;
  const int c = a->num_classes;
  ODK_REQUIRE(c <= 512, ODK_ERR_UNSUPPORTED, "num_classes=%d > 512 not supported by decode_score", c);
  switch (odk_piece_size(c)) {
    case 256:
      if (c == 20) return dense ? launch_tile<256, true, 1, 20>(args, stream) : launch_tile<256, false, 1, 20>(args, stream);
      return dense ? launch_tile<256, true, 1>(args, stream) : launch_tile<256, false, 1>(args, stream);
    case 128: return dense ? launch_tile<128, true, 1>(args, stream) : launch_tile<128, false, 1>(args, stream);
    case 64: return dense ? launch_tile<64, true, 4>(args, stream) : launch_tile<64, false, 4>(args, stream);
    default: return dense ? launch_tile<32, true, 8>(args, stream) : launch_tile<32, false, 8>(args, stream);
  }
}

int odk_launch_decode_locs(const odk_anchors* a, long long rows, const float* pred, float* out, cudaStream_t stream) {
  long long blocks = (rows + 255) / 256;
  const long long cap = (long long)sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  {
    OdkProfileScope prof_("decode_locs_kernel", stream);
    decode_locs_kernel<<<(unsigned)blocks, 256, 0, stream>>>(odk_view(a), rows, reinterpret_cast<const float4*>(pred),
                                                              reinterpret_cast<float4*>(out));
  }
  ODK_LAUNCH_CHECK("decode_locs_kernel");
  return ODK_OK;
}

int odk_launch_decode_locs_f64(const odk_anchors* a, long long rows, const double* pred, double* out, cudaStream_t stream) {
  long long blocks = (rows + 255) / 256;
  const long long cap = (long long)sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  {
    OdkProfileScope prof_("decode_locs_f64_kernel", stream);
    decode_locs_f64_kernel<<<(unsigned)blocks, 256, 0, stream>>>(odk_view(a), rows, pred, out);
  }
  ODK_LAUNCH_CHECK("decode_locs_f64_kernel");
  return ODK_OK;
}

int odk_launch_channel_reduce(long long rows, int c, const float* x, long long* arg, float* mx, cudaStream_t stream) {
  long long blocks = (rows + 255) / 256;
  const long long cap = (long long)sm_count_cached() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  {
    OdkProfileScope prof_("channel_reduce_kernel", stream);
    channel_reduce_kernel<<<(unsigned)blocks, 256, 0, stream>>>(rows, c, x, arg, mx);
  }
  ODK_LAUNCH_CHECK("channel_reduce_kernel");
  return ODK_OK;
}

