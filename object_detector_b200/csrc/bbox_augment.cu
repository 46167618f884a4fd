// Bounding-box side of the training augmentation that sits directly upstream of encode_truth:
// RandomZoom._padding / _crop (pytoolkit image.py:321-388), the letterbox of a failed zoom (image.py:298-314),
// RandomFlipLR / RandomFlipTB (image.py:433-465) and ObjectsAnnotation.rot90 (ml.py:102-112).
//
// The reference draws its zoom attempts one by one from a per-sample RandomState and stops at the first acceptable one
// (up to 30).  Here the host draws the attempt table for the whole batch up front (bbox_augment.py), one warp per
// sample walks its attempts in order and reports the index it accepted, and the host fast-forwards each RandomState
// by exactly that many attempts — the RNG stream, the accepted window and the surviving boxes are the reference's.
// Arithmetic is the reference's NumPy promotion: fp64 wherever an integer size array enters a product, fp32 for the
// box centre / area / 1-x of boxes that still carry their ObjectsAnnotation dtype.  No FMA contraction (-fmad=false).
#include "odk_common.cuh"

namespace {

constexpr int kAugWarps = 4;

enum { kZoomNone = 0, kZoomPad = 1, kZoomCrop = 2, kZoomLetterbox = 3 };

struct ZoomArgs {
  int batch, gmax, tries, src_f32;
  const double* boxes;      // [B,Gmax,4]
  const int32_t* count;     // [B]
  const int32_t* kind;      // [B]
  const int64_t* in_size;   // [B,2]   (w, h)
  const int64_t* try_size;  // [B,T,2] padded canvas / crop window size
  const int64_t* try_xy;    // [B,T,2] paste / crop origin
  double out_w, out_h, min_px;
  double* out;              // [B,Gmax,4]
  uint8_t* keep;            // [B,Gmax]
  int32_t* chosen;          // [B]
  uint8_t* wide;            // [B]
};

struct Box4 { double x1, y1, x2, y2; };

struct Attempt { double sw, sh, tw, th, ox, oy; };

__device__ __forceinline__ double clip01(double v) { return fmin(fmax(v, 0.0), 1.0); }

// image.py:329 / image.py:311: (xy + b * in_size) / canvas
__device__ __forceinline__ Box4 paste(const Box4& b, const Attempt& t) {
  Box4 m;
  m.x1 = (t.ox + b.x1 * t.sw) / t.tw;
  m.y1 = (t.oy + b.y1 * t.sh) / t.th;
  m.x2 = (t.ox + b.x2 * t.sw) / t.tw;
  m.y2 = (t.oy + b.y2 * t.sh) / t.th;
  return m;
}

// image.py:365-366: clip((b * in_size - xy) / window, 0, 1)
__device__ __forceinline__ Box4 cut(const Box4& b, const Attempt& t) {
  Box4 m;
  m.x1 = clip01((b.x1 * t.sw - t.ox) / t.tw);
  m.y1 = clip01((b.y1 * t.sh - t.oy) / t.th);
  m.x2 = clip01((b.x2 * t.sw - t.ox) / t.tw);
  m.y2 = clip01((b.y2 * t.sh - t.oy) / t.th);
  return m;
}

// image.py:330-331,367-368: any side shorter than min_object_px output pixels
__device__ __forceinline__ bool too_small(const Box4& m, double ow, double oh, double min_px) {
  return (m.x2 * ow - m.x1 * ow < min_px) || (m.y2 * oh - m.y1 * oh < min_px);
}

// image.py:355-363: centre inside the window and >= 30 % of the area still visible
__device__ __forceinline__ bool survives_crop(const Box4& b, const Attempt& t, bool src_f32) {
  const double wx1 = t.ox / t.sw, wy1 = t.oy / t.sh, wx2 = (t.ox + t.tw) / t.sw, wy2 = (t.oy + t.th) / t.sh;
  double cx, cy, need;
  if (src_f32) {
    const float x1 = (float)b.x1, y1 = (float)b.y1, x2 = (float)b.x2, y2 = (float)b.y2;
    cx = (double)__fdiv_rn(__fadd_rn(x2, x1), 2.0f);
    cy = (double)__fdiv_rn(__fadd_rn(y2, y1), 2.0f);
    need = (double)__fmul_rn(__fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1)), 0.3f);
  } else {
    cx = (b.x2 + b.x1) / 2.0;
    cy = (b.y2 + b.y1) / 2.0;
    need = ((b.x2 - b.x1) * (b.y2 - b.y1)) * 0.3;
  }
  const bool inside = wx1 <= cx && cx < wx2 && wy1 <= cy && cy < wy2;
  const double lx = fmax(wx1, b.x1), ly = fmax(wy1, b.y1), hx = fmin(wx2, b.x2), hy = fmin(wy2, b.y2);
  const double visible = ((hx - lx) * (hy - ly)) * ((lx < hx && ly < hy) ? 1.0 : 0.0);
  return inside && visible >= need;
}

__global__ void __launch_bounds__(kAugWarps * 32) bbox_zoom_kernel(const ZoomArgs a) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * kAugWarps + (threadIdx.x >> 5);
  if (b >= a.batch) return;
  const int g = a.count[b];
  const int kind = a.kind[b];
  const Box4* src = reinterpret_cast<const Box4*>(a.boxes) + (size_t)b * a.gmax;
  Box4* dst = reinterpret_cast<Box4*>(a.out) + (size_t)b * a.gmax;
  uint8_t* keep = a.keep + (size_t)b * a.gmax;
  Attempt t;
  t.sw = (double)a.in_size[b * 2];
  t.sh = (double)a.in_size[b * 2 + 1];
  int chosen = -1;
  const int n_try = kind == kZoomNone ? 0 : (kind == kZoomLetterbox ? 1 : a.tries);
  for (int k = 0; k < n_try && chosen < 0; ++k) {
    const size_t at = ((size_t)b * a.tries + k) * 2;
    t.tw = (double)a.try_size[at]; t.th = (double)a.try_size[at + 1];
    t.ox = (double)a.try_xy[at];   t.oy = (double)a.try_xy[at + 1];
    bool reject = false, any_keep = false;
    for (int i = lane; i < g; i += 32) {
      const Box4 box = src[i];
      if (kind == kZoomCrop) {
        any_keep |= survives_crop(box, t, a.src_f32 != 0);
        reject |= too_small(cut(box, t), a.out_w, a.out_h, a.min_px);
      } else if (kind == kZoomPad) {
        reject |= too_small(paste(box, t), a.out_w, a.out_h, a.min_px);
      }
    }
    reject = __any_sync(0xffffffffu, reject);
    any_keep = __any_sync(0xffffffffu, any_keep);
    // no survivor is tested first (image.py:357-358), so an attempt with neither is rejected either way
    if (kind == kZoomCrop ? (any_keep && !reject) : !reject) chosen = k;
  }
  for (int i = lane; i < a.gmax; i += 32) {
    Box4 m = {0.0, 0.0, 0.0, 0.0};
    uint8_t kp = 0;
    if (i < g) {
      m = src[i];
      kp = 1;
      if (chosen >= 0) {
        if (kind == kZoomCrop) {
          kp = survives_crop(m, t, a.src_f32 != 0) ? 1 : 0;
          m = cut(m, t);
        } else {
          m = paste(m, t);
        }
      }
    }
    dst[i] = m;
    keep[i] = kp;
  }
  if (lane == 0) {
    a.chosen[b] = chosen;
    a.wide[b] = (chosen >= 0 || a.src_f32 == 0) ? 1 : 0;
  }
}

struct FlipArgs {
  int batch, gmax;
  const double* boxes;     // [B,Gmax,4]
  const uint8_t* keep;     // [B,Gmax] or null
  const int32_t* classes;  // [B,Gmax]
  const int32_t* count;    // [B]
  const uint8_t* wide;     // [B] or null (all boxes still fp32)
  const int32_t* ops;      // [B] bit0 flip_lr, bit1 flip_tb, bits 2-3 quarter turns
  double* out64;           // [B,Gmax,4] or null
  float* out32;            // [B,Gmax,4]
  int32_t* classes_out;    // [B,Gmax]
  int32_t* count_out;      // [B]
};

// 1 - v in the dtype the box array has at that point of the reference pipeline
__device__ __forceinline__ double one_minus(double v, bool wide) {
  return wide ? 1.0 - v : (double)__fsub_rn(1.0f, (float)v);
}

__global__ void __launch_bounds__(kAugWarps * 32) bbox_flip_rot_kernel(const FlipArgs a) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * kAugWarps + (threadIdx.x >> 5);
  if (b >= a.batch) return;
  const int g = a.count[b];
  const bool wide = a.wide ? a.wide[b] != 0 : false;
  const int op = a.ops[b];
  const size_t base = (size_t)b * a.gmax;
  const Box4* src = reinterpret_cast<const Box4*>(a.boxes) + base;
  int written = 0;
  for (int i0 = 0; i0 < g; i0 += 32) {
    const int i = i0 + lane;
    const bool live = i < g && (a.keep == nullptr || a.keep[base + i] != 0);
    const unsigned mask = __ballot_sync(0xffffffffu, live);
    if (live) {
      Box4 m = src[i];
      if (op & 1) { const double x1 = one_minus(m.x2, wide), x2 = one_minus(m.x1, wide); m.x1 = x1; m.x2 = x2; }   // image.py:445
      if (op & 2) { const double y1 = one_minus(m.y2, wide), y2 = one_minus(m.y1, wide); m.y1 = y1; m.y2 = y2; }   // image.py:462
      const int k = (op >> 2) & 3;                                                                                 // ml.py:102-112
      Box4 r = m;
      if (k == 1)      { r.x1 = m.y1; r.y1 = one_minus(m.x2, wide); r.x2 = m.y2; r.y2 = one_minus(m.x1, wide); }
      else if (k == 2) { r.x1 = one_minus(m.x2, wide); r.y1 = one_minus(m.y2, wide); r.x2 = one_minus(m.x1, wide); r.y2 = one_minus(m.y1, wide); }
      else if (k == 3) { r.x1 = one_minus(m.y2, wide); r.y1 = m.x1; r.x2 = one_minus(m.y1, wide); r.y2 = m.x2; }
      const size_t o = base + written + __popc(mask & ((1u << lane) - 1u));
      if (a.out64) reinterpret_cast<Box4*>(a.out64)[o] = r;
      reinterpret_cast<float4*>(a.out32)[o] = make_float4((float)r.x1, (float)r.y1, (float)r.x2, (float)r.y2);
      a.classes_out[o] = a.classes[base + i];
    }
    written += __popc(mask);
  }
  for (int i = written + lane; i < a.gmax; i += 32) {
    if (a.out64) reinterpret_cast<Box4*>(a.out64)[base + i] = Box4{0.0, 0.0, 0.0, 0.0};
    reinterpret_cast<float4*>(a.out32)[base + i] = make_float4(0.f, 0.f, 0.f, 0.f);
    a.classes_out[base + i] = 0;
  }
  if (lane == 0) a.count_out[b] = written;
}

}  // namespace

extern "C" int odk_bbox_zoom(int batch, int gmax, int tries, const double* boxes, int src_f32, const int32_t* gt_count,
                             const int32_t* kind, const int64_t* in_size, const int64_t* try_size, const int64_t* try_xy,
                             int out_w, int out_h, double min_object_px, double* boxes_out, uint8_t* keep,
                             int32_t* chosen, uint8_t* wide, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(batch >= 1 && gmax >= 1 && tries >= 1, ODK_ERR_INVALID_ARGUMENT, "bbox_zoom: bad sizes batch=%d gmax=%d tries=%d", batch, gmax, tries);
  ODK_REQUIRE(out_w >= 1 && out_h >= 1, ODK_ERR_INVALID_ARGUMENT, "bbox_zoom: bad output size %dx%d", out_w, out_h);
  ODK_REQUIRE(boxes && gt_count && kind && in_size && try_size && try_xy && boxes_out && keep && chosen && wide,
              ODK_ERR_INVALID_ARGUMENT, "bbox_zoom: null pointer");
  ODK_REQUIRE((((uintptr_t)boxes | (uintptr_t)boxes_out) & 15) == 0, ODK_ERR_INVALID_ARGUMENT, "bbox_zoom: box arrays must be 16-byte aligned");
  ZoomArgs a;
  a.batch = batch; a.gmax = gmax; a.tries = tries; a.src_f32 = src_f32;
  a.boxes = boxes; a.count = gt_count; a.kind = kind; a.in_size = in_size; a.try_size = try_size; a.try_xy = try_xy;
  a.out_w = (double)out_w; a.out_h = (double)out_h; a.min_px = min_object_px;
  a.out = boxes_out; a.keep = keep; a.chosen = chosen; a.wide = wide;
  bbox_zoom_kernel<<<(batch + kAugWarps - 1) / kAugWarps, kAugWarps * 32, 0, stream>>>(a);
  ODK_CUDA(cudaGetLastError());
  return ODK_OK;
}

extern "C" int odk_bbox_flip_rot(int batch, int gmax, const double* boxes, const uint8_t* keep, const int32_t* gt_classes,
                                 const int32_t* gt_count, const uint8_t* wide, const int32_t* ops, double* boxes64_out,
                                 float* boxes_out, int32_t* classes_out, int32_t* count_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  ODK_REQUIRE(batch >= 1 && gmax >= 1, ODK_ERR_INVALID_ARGUMENT, "bbox_flip_rot: bad sizes batch=%d gmax=%d", batch, gmax);
  ODK_REQUIRE(boxes && gt_classes && gt_count && ops && boxes_out && classes_out && count_out, ODK_ERR_INVALID_ARGUMENT,
              "bbox_flip_rot: null pointer");
  ODK_REQUIRE((((uintptr_t)boxes | (uintptr_t)boxes64_out | (uintptr_t)boxes_out) & 15) == 0, ODK_ERR_INVALID_ARGUMENT,
              "bbox_flip_rot: box arrays must be 16-byte aligned");
  ODK_REQUIRE((const void*)boxes != (const void*)boxes64_out, ODK_ERR_INVALID_ARGUMENT, "bbox_flip_rot: compaction cannot run in place");
  FlipArgs a;
  a.batch = batch; a.gmax = gmax; a.boxes = boxes; a.keep = keep; a.classes = gt_classes; a.count = gt_count; a.wide = wide;
  a.ops = ops; a.out64 = boxes64_out; a.out32 = boxes_out; a.classes_out = classes_out; a.count_out = count_out;
  bbox_flip_rot_kernel<<<(batch + kAugWarps - 1) / kAugWarps, kAugWarps * 32, 0, stream>>>(a);
  ODK_CUDA(cudaGetLastError());
  return ODK_OK;
}
