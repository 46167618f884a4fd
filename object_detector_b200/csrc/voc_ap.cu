// K5 — PASCAL VOC AP (data/voc.py:165-197 -> chainercv eval_detection_voc).  Placeholder entry points
// until the kernels land: they fail loudly (no CPU fallback).
#include "odk_common.cuh"

extern "C" size_t odk_voc_ap_workspace_bytes(int, int, int, int) { return 0; }
extern "C" int odk_voc_ap(int, int, int, const float*, float, const int32_t*, const float*, const int32_t*, const uint8_t*, const int32_t*, int,
                          double*, double*, void*, size_t, void*) {
  return odk_set_error(ODK_ERR_UNSUPPORTED, "odk_voc_ap: not built yet");
}
