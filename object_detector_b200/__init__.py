"""object_detector_b200 — the B200-native (sm_100a) detection hot path of ak110/object_detector:
anchor<->GT assignment + target encoding, and decode + score + per-class NMS (+ VOC AP), behind the
reference's own Python call signatures.  All arithmetic runs in libodk.so (hand-written CUDA, reached
through ctypes); there is no CPU or PyTorch fallback."""
from . import _lib, dist, generator, layers, ml, od, od_pb, voc  # noqa: F401
from .od import ObjectDetector  # noqa: F401
from .od_pb import PriorBoxes  # noqa: F401

__version__ = '0.1.0'
