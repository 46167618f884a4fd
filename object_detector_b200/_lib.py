"""ctypes binding of libodk.so — the only gateway to the CUDA kernels.

No fallback of any kind: if the library is missing or a call fails, an exception is raised.
Signatures mirror include/odk.h one to one.
"""
import ctypes
import pathlib
import threading

_PKG = pathlib.Path(__file__).resolve().parent
import os
# ODK_LIB_VARIANT=checks selects the build with device-side bounds assertions (build.py --checks); diagnostics only
LIB_PATH = _PKG / ('libodk_checks.so' if os.environ.get('ODK_LIB_VARIANT') == 'checks' else 'libodk.so')

c_int, c_float, c_size_t, c_void_p, c_int64 = ctypes.c_int, ctypes.c_float, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int64
c_double = ctypes.c_double

# name -> (restype, argtypes); kept in one table so tests can check it against include/odk.h
SIGNATURES = {
    'odk_version': (ctypes.c_char_p, []),
    'odk_last_error': (ctypes.c_char_p, []),
    'odk_anchors_create': (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_void_p, c_int,
                                   c_void_p, c_void_p, c_void_p, c_void_p]),
    'odk_anchors_destroy': (c_int, [c_void_p]),
    'odk_anchors_count': (c_int, [c_void_p]),
    'odk_anchors_num_classes': (c_int, [c_void_p]),
    'odk_anchors_span_stats': (c_int, [c_void_p, c_void_p]),
    'odk_compute_iou': (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    'odk_encode_locs': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    'odk_decode_locs': (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    'odk_decode_locs_f64': (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    'odk_channel_argmax_max': (c_int, [c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    'odk_assign_workspace_bytes': (c_size_t, [c_void_p, c_int, c_int]),
    'odk_assign': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                           c_void_p]),
    'odk_assign_encode': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_assign_encode_compact': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_loss_workspace_bytes': (c_size_t, [c_void_p, c_int]),
    'odk_loss': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_loss_grad': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    'odk_assign_loss_workspace_bytes': (c_size_t, [c_void_p, c_int, c_int]),
    'odk_assign_loss': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_assign_loss_grad': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_decode_score': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
    'odk_nms_workspace_bytes': (c_size_t, [c_int, c_int, c_int, c_int]),
    'odk_nms_layer': (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_float, c_float, c_int, c_int,
                              c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_decode_nms': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_float, c_float, c_int, c_int,
                               c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_nms_strict_workspace_bytes': (c_size_t, [c_int, c_int, c_int, c_int]),
    'odk_nms_layer_strict': (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_float, c_float, c_float, c_int, c_int,
                                     c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_decode_nms_strict': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_float, c_float, c_float, c_int, c_int,
                                      c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_nms': (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p, c_size_t,
                        c_void_p]),
    'odk_nms_static_max_boxes': (c_size_t, []),
    'odk_nms_static': (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p]),
    'odk_voc_ap_workspace_bytes': (c_size_t, [c_int, c_int, c_int, c_int]),
    'odk_voc_ap': (c_int, [c_int, c_int, c_int, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p,
                           c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_coco_ap_workspace_bytes': (c_size_t, [c_int, c_int, c_int]),
    'odk_coco_ap': (c_int, [c_int, c_int, c_int, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'odk_match_counts': (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                 c_float, c_void_p, c_void_p]),
    'odk_confusion_matrix': (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p,
                                     c_void_p]),
    'odk_exact_match': (c_int, [c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p]),
    'odk_bbox_zoom': (c_int, [c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                             c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    'odk_bbox_flip_rot': (c_int, [c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p]),
    'odk_timeline_set': (c_int, [c_void_p]),
    'odk_profile_enable': (c_int, [c_int]),
    'odk_profile_read': (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
}

_lock = threading.Lock()
_lib = None


class OdkError(RuntimeError):
    """A libodk call returned a non-zero status."""

    def __init__(self, fn, code, message):
        super().__init__(f'{fn} failed with status {code}: {message}')
        self.code = code


def load():
    """Returns the loaded library; raises if it has not been built (never falls back)."""
    global _lib
    with _lock:
        if _lib is None:
            if not LIB_PATH.exists():
                raise ImportError(
                    f'{LIB_PATH} is missing: build it with `python -m object_detector_b200.build` '
                    '(there is no CPU or PyTorch fallback)')
            lib = ctypes.CDLL(str(LIB_PATH))
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def call(name, *args):
    """Calls an int-returning entry point and raises OdkError on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise OdkError(name, rc, lib.odk_last_error().decode('utf-8', 'replace'))


def version() -> str:
    return load().odk_version().decode()


def profile_enable(on: bool):
    call('odk_profile_enable', 1 if on else 0)


def profile_read(cap=32):
    """{kernel name: (total device ms, launches)} since profiling was enabled; clears the log."""
    names = ctypes.create_string_buffer(cap * 64)
    ms = (ctypes.c_float * cap)()
    cnt = (ctypes.c_int32 * cap)()
    n = ctypes.c_int32(0)
    call('odk_profile_read', ctypes.cast(names, c_void_p), 64, ctypes.cast(ms, c_void_p), ctypes.cast(cnt, c_void_p), cap,
         ctypes.cast(ctypes.byref(n), c_void_p))
    return {names.raw[i * 64:(i + 1) * 64].split(b'\0')[0].decode(): (float(ms[i]), int(cnt[i])) for i in range(n.value)}
