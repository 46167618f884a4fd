// libodk C-ABI glue: error reporting, the anchors handle, and the thin entry points that only
// validate and launch (the kernels live in decode_score.cu / nms.cu / assign_encode.cu / voc_ap.cu).
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "odk_common.cuh"

static thread_local char g_last_error[512] = "";

int odk_set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

int odk_launch_decode_score(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs,
                            int32_t* classes, float* confs, float* boxes, bool dense, float conf_thr,
                            unsigned long long* cand_keys, int* piece_count, int* score_hist, cudaStream_t stream);
int odk_launch_decode_locs(const odk_anchors* a, long long rows, const float* pred, float* out, cudaStream_t stream);
int odk_piece_size(int num_classes);
int odk_launch_decode_locs_f64(const odk_anchors* a, long long rows, const double* pred, double* out, cudaStream_t stream);
int odk_launch_channel_reduce(long long rows, int c, const float* x, long long* arg, float* mx, cudaStream_t stream);

// ---- per-launch event timing (used by bench.py for the roofline line) -------------------------------
#include <string>
#include <vector>
namespace {
struct ProfEntry { const char* name; cudaEvent_t start, stop; };
thread_local bool g_prof_on = false;
thread_local std::vector<ProfEntry> g_prof;
}  // namespace

OdkProfileScope::OdkProfileScope(const char* name, cudaStream_t stream) : slot_(-1), stream_(stream) {
  if (!g_prof_on) return;
  ProfEntry e;
  e.name = name;
  if (cudaEventCreate(&e.start) != cudaSuccess || cudaEventCreate(&e.stop) != cudaSuccess) return;
  cudaEventRecord(e.start, stream);
  g_prof.push_back(e);
  slot_ = (int)g_prof.size() - 1;
}
OdkProfileScope::~OdkProfileScope() {
  if (slot_ >= 0) cudaEventRecord(g_prof[slot_].stop, stream_);
}

bool odk_profile_on() { return g_prof_on; }

extern "C" int odk_profile_enable(int on) {
  g_prof_on = on != 0;
  return ODK_OK;
}

// Synchronises the recorded events, aggregates by kernel name and clears the log.
extern "C" int odk_profile_read(char* names, int name_stride, float* total_ms, int32_t* launches, int cap, int32_t* n_out) {
  ODK_REQUIRE(names && total_ms && launches && n_out && name_stride >= 8 && cap >= 1, ODK_ERR_INVALID_ARGUMENT, "profile_read: bad arguments");
  std::vector<std::string> keys;
  int n = 0;
  for (auto& e : g_prof) {
    float ms = 0.f;
    cudaEventSynchronize(e.stop);
    cudaEventElapsedTime(&ms, e.start, e.stop);
    cudaEventDestroy(e.start);
    cudaEventDestroy(e.stop);
    int k = -1;
    for (int i = 0; i < n; ++i) if (keys[i] == e.name) { k = i; break; }
    if (k < 0) {
      if (n >= cap) continue;
      k = n++;
      keys.push_back(e.name);
      total_ms[k] = 0.f;
      launches[k] = 0;
      snprintf(names + (size_t)k * name_stride, name_stride, "%s", e.name);
    }
    total_ms[k] += ms;
    launches[k] += 1;
  }
  g_prof.clear();
  *n_out = n;
  return ODK_OK;
}

static unsigned long long* g_timeline = nullptr;  // process-wide, diagnostics only (include/odk.h, Conventions)
unsigned long long* odk_timeline_ptr() { return g_timeline; }
extern "C" int odk_timeline_set(unsigned long long* buf) {
  g_timeline = buf;
  return ODK_OK;
}

extern "C" const char* odk_version(void) { return "odk 0.1.0 (sm_100a)"; }
extern "C" const char* odk_last_error(void) { return g_last_error; }

extern "C" int odk_anchors_destroy(odk_anchors* a) {
  if (a == nullptr) return ODK_OK;
  cudaFree(a->d_slab);
  free(a);
  return ODK_OK;
}

extern "C" int odk_anchors_create(odk_anchors** out, int n, int num_classes, int n_maps, const int32_t* map_sizes, int n_patterns,
                                  const float* pb_locs, const float* pb_grid, const double* pb_sizes, const uint8_t* pb_mask) {
  ODK_REQUIRE(out != nullptr, ODK_ERR_INVALID_ARGUMENT, "anchors_create: null output handle");
  *out = nullptr;
  ODK_REQUIRE(n >= 1 && num_classes >= 1, ODK_ERR_INVALID_ARGUMENT, "anchors_create: n=%d num_classes=%d (od_pb.py:40-42)", n, num_classes);
  ODK_REQUIRE(n_maps >= 1 && n_maps <= ODK_MAX_MAPS && n_patterns >= 1, ODK_ERR_INVALID_ARGUMENT,
              "anchors_create: n_maps=%d (1..%d) n_patterns=%d", n_maps, ODK_MAX_MAPS, n_patterns);
  ODK_REQUIRE(map_sizes && pb_locs && pb_grid && pb_sizes && pb_mask, ODK_ERR_INVALID_ARGUMENT, "anchors_create: null table pointer");
  ODK_REQUIRE(n <= (int)ODK_IDX_MASK, ODK_ERR_UNSUPPORTED, "anchors_create: n=%d exceeds 2^20-1 anchors", n);
  ODK_REQUIRE(num_classes <= ODK_MAX_CLASSES, ODK_ERR_UNSUPPORTED, "anchors_create: num_classes=%d > %d", num_classes, ODK_MAX_CLASSES);
  odk_anchors* a = static_cast<odk_anchors*>(calloc(1, sizeof(odk_anchors)));
  ODK_REQUIRE(a != nullptr, ODK_ERR_CUDA, "anchors_create: out of host memory");
  a->n = n; a->num_classes = num_classes; a->n_maps = n_maps; a->n_patterns = n_patterns;
  long long total = 0;
  for (int i = 0; i < n_maps; ++i) {
    if (map_sizes[i] < 1) { free(a); return odk_set_error(ODK_ERR_INVALID_ARGUMENT, "anchors_create: map_sizes[%d]=%d", i, map_sizes[i]); }
    a->map_size[i] = map_sizes[i];
    a->map_off[i] = (int)total;
    total += (long long)map_sizes[i] * map_sizes[i] * n_patterns;
  }
  for (int i = n_maps; i <= ODK_MAX_MAPS; ++i) a->map_off[i] = (int)total;
  if (total != n) { free(a); return odk_set_error(ODK_ERR_INVALID_ARGUMENT, "anchors_create: n=%d but the layout holds %lld anchors", n, total); }
  // The assignment kernel locates candidate cells as floor(c*m)+{-1,0,1}; that is exhaustive iff
  // every cell [lo, hi) satisfies  ix/m - eps <= lo  and  hi <= (ix+2)/m  for its own (ix, iy).
  for (int mi = 0; mi < n_maps; ++mi) {
    const int m = a->map_size[mi];
    for (int pat = 0; pat < n_patterns; ++pat)
      for (int iy = 0; iy < m; ++iy)
        for (int ix = 0; ix < m; ++ix) {
          const float* g = pb_grid + 4 * ((size_t)a->map_off[mi] + (size_t)pat * m * m + (size_t)iy * m + ix);
          const double eps = 1e-6;
          const bool okx = g[0] >= (double)ix / m - eps && g[2] <= (double)(ix + 2) / m && g[0] < g[2];
          const bool oky = g[1] >= (double)iy / m - eps && g[3] <= (double)(iy + 2) / m && g[1] < g[3];
          if (!okx || !oky) {
            free(a);
            return odk_set_error(ODK_ERR_INVALID_ARGUMENT,
                                 "anchors_create: pb_grid row for map %d cell (%d,%d) pattern %d does not follow the PriorBoxes layout", m, ix, iy, pat);
          }
        }
  }
  int valid = 0;
  for (int i = 0; i < n; ++i) valid += pb_mask[i] ? 1 : 0;
  a->n_valid = valid;
  float* scale = static_cast<float*>(malloc(sizeof(float) * 2 * (size_t)n));
  if (scale == nullptr) { free(a); return odk_set_error(ODK_ERR_CUDA, "anchors_create: out of host memory"); }
  for (size_t i = 0; i < 2 * (size_t)n; ++i) scale[i] = (float)(0.10 * pb_sizes[i]);  // float32(_VAR_LOC * pb_sizes), od_pb.py:221
  cudaGetDevice(&a->device);
  // one device slab for all tables
  const size_t words = ((size_t)n + 31) / 32;
  uint32_t* bits = static_cast<uint32_t*>(calloc(words, sizeof(uint32_t)));
  if (bits == nullptr) { free(scale); free(a); return odk_set_error(ODK_ERR_CUDA, "anchors_create: out of host memory"); }
  for (int i = 0; i < n; ++i) if (pb_mask[i]) bits[i >> 5] |= 1u << (i & 31);
  // Valid-anchor spans per score/emit tile: anchors with pb_mask = 0 can never become candidates (conf = x * 0,
  // od_net.py:183-184), so the fused predict path fetches only the runs of valid rows of a tile.  Runs separated by at
  // most two masked rows are merged (a separate bulk copy costs more than the five sectors it saves); a tile keeps at
  // most ODK_MAX_SPANS runs (the narrowest gaps are merged first).
  static const int span_gap = getenv("ODK_SPAN_GAP") ? atoi(getenv("ODK_SPAN_GAP")) : 2;
  a->span_tile = odk_piece_size(num_classes);
  const int n_tiles = (n + a->span_tile - 1) / a->span_tile;
  std::vector<uint32_t> spans((size_t)n_tiles * ODK_MAX_SPANS, 0u);
  a->span_rows = 0; a->span_count = 0;
  for (int t = 0; t < n_tiles; ++t) {
    const int a0 = t * a->span_tile, rows = (n - a0 < a->span_tile) ? n - a0 : a->span_tile;
    std::vector<std::pair<int, int>> run;  // [first, end)
    for (int i = 0; i < rows; ++i) {
      if (!pb_mask[a0 + i]) continue;
      if (!run.empty() && i - run.back().second <= span_gap) run.back().second = i + 1;
      else run.push_back({i, i + 1});
    }
    while ((int)run.size() > ODK_MAX_SPANS) {
      size_t best = 0;
      for (size_t k = 1; k + 1 < run.size(); ++k)
        if (run[k + 1].first - run[k].second < run[best + 1].first - run[best].second) best = k;
      run[best].second = run[best + 1].second;
      run.erase(run.begin() + best + 1);
    }
    for (size_t k = 0; k < run.size(); ++k) {
      spans[(size_t)t * ODK_MAX_SPANS + k] = ((uint32_t)run[k].first << 16) | (uint32_t)(run[k].second - run[k].first);
      a->span_rows += run[k].second - run[k].first;
    }
    a->span_count += (int)run.size();
  }
  const size_t sz[7] = {16 * (size_t)n, 16 * (size_t)n, 16 * (size_t)n, 8 * (size_t)n, (size_t)n, 4 * words, 4 * spans.size()};
  const void* src[7] = {pb_locs, pb_grid, pb_sizes, scale, pb_mask, bits, spans.data()};
  size_t off[7], total_bytes = 0;
  for (int i = 0; i < 7; ++i) { off[i] = total_bytes; total_bytes = odk_align_up(total_bytes + sz[i], 256); }
  int rc = ODK_OK;
  if (cudaMalloc(&a->d_slab, total_bytes) != cudaSuccess) rc = odk_set_error(ODK_ERR_CUDA, "anchors_create: cudaMalloc of %zu bytes failed", total_bytes);
  for (int i = 0; i < 7 && !rc; ++i)
    if (cudaMemcpy(static_cast<char*>(a->d_slab) + off[i], src[i], sz[i], cudaMemcpyHostToDevice) != cudaSuccess)
      rc = odk_set_error(ODK_ERR_CUDA, "anchors_create: upload of table %d failed", i);
  if (!rc) {
    char* base = static_cast<char*>(a->d_slab);
    a->d_locs = reinterpret_cast<float4*>(base + off[0]);
    a->d_grid = reinterpret_cast<float4*>(base + off[1]);
    a->d_sizes = reinterpret_cast<double2*>(base + off[2]);
    a->d_scale = reinterpret_cast<float2*>(base + off[3]);
    a->d_mask = reinterpret_cast<uint8_t*>(base + off[4]);
    a->d_maskbits = reinterpret_cast<uint32_t*>(base + off[5]);
    a->d_spans = reinterpret_cast<uint32_t*>(base + off[6]);
  }
  free(bits);
  free(scale);
  if (rc) { odk_anchors_destroy(a); return rc; }
  *out = a;
  return ODK_OK;
}

extern "C" int odk_anchors_count(const odk_anchors* a) { return a ? a->n : 0; }
extern "C" int odk_anchors_span_stats(const odk_anchors* a, int32_t* out4) {
  ODK_REQUIRE(a && out4, ODK_ERR_INVALID_ARGUMENT, "anchors_span_stats: null pointer");
  out4[0] = a->span_tile; out4[1] = a->span_count; out4[2] = a->span_rows; out4[3] = a->n_valid;
  return ODK_OK;
}
extern "C" int odk_anchors_num_classes(const odk_anchors* a) { return a ? a->num_classes : 0; }

extern "C" int odk_decode_locs(const odk_anchors* a, int64_t rows, const float* pred, float* out, void* stream) {
  ODK_REQUIRE(a != nullptr && rows >= 0, ODK_ERR_INVALID_ARGUMENT, "decode_locs: bad arguments");
  ODK_REQUIRE(rows % a->n == 0, ODK_ERR_INVALID_ARGUMENT, "decode_locs: rows=%lld is not a multiple of N=%d", (long long)rows, a->n);
  if (rows == 0) return ODK_OK;
  ODK_REQUIRE(pred && out, ODK_ERR_INVALID_ARGUMENT, "decode_locs: null pointer");
  ODK_REQUIRE((((uintptr_t)pred | (uintptr_t)out) & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "decode_locs: buffers must be 16-byte aligned");
  return odk_launch_decode_locs(a, rows, pred, out, static_cast<cudaStream_t>(stream));
}

extern "C" int odk_decode_locs_f64(const odk_anchors* a, int64_t rows, const double* pred, double* out, void* stream) {
  ODK_REQUIRE(a != nullptr && rows >= 0, ODK_ERR_INVALID_ARGUMENT, "decode_locs_f64: bad arguments");
  ODK_REQUIRE(rows % a->n == 0, ODK_ERR_INVALID_ARGUMENT, "decode_locs_f64: rows=%lld is not a multiple of N=%d", (long long)rows, a->n);
  if (rows == 0) return ODK_OK;
  ODK_REQUIRE(pred && out, ODK_ERR_INVALID_ARGUMENT, "decode_locs_f64: null pointer");
  return odk_launch_decode_locs_f64(a, rows, pred, out, static_cast<cudaStream_t>(stream));
}

extern "C" int odk_channel_argmax_max(int64_t rows, int channels, const float* x, int64_t* arg, float* mx, void* stream) {
  ODK_REQUIRE(rows >= 0 && channels >= 1, ODK_ERR_INVALID_ARGUMENT, "channel_argmax_max: rows=%lld channels=%d", (long long)rows, channels);
  if (rows == 0) return ODK_OK;
  ODK_REQUIRE(x && (arg || mx), ODK_ERR_INVALID_ARGUMENT, "channel_argmax_max: null pointer");
  return odk_launch_channel_reduce(rows, channels, x, reinterpret_cast<long long*>(arg), mx, static_cast<cudaStream_t>(stream));
}

extern "C" int odk_decode_score(const odk_anchors* a, int batch, const float* objs, const float* clfs, const float* locs, int32_t* classes,
                                float* confs, float* boxes, void* stream) {
  ODK_REQUIRE(a != nullptr && batch >= 1, ODK_ERR_INVALID_ARGUMENT, "decode_score: bad arguments");
  ODK_REQUIRE(objs && clfs && locs && classes && confs && boxes, ODK_ERR_INVALID_ARGUMENT, "decode_score: null pointer");
  ODK_REQUIRE(((uintptr_t)boxes & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "decode_score: boxes must be 16-byte aligned");
  return odk_launch_decode_score(a, batch, objs, clfs, locs, classes, confs, boxes, true, 0.f, nullptr, nullptr, nullptr,
                                 static_cast<cudaStream_t>(stream));
}
