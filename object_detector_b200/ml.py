"""Mirror of the detection parts of pytoolkit/pytoolkit/ml.py: the two datatypes, ``compute_iou`` and
``non_maximum_suppression`` (ml.py:15-197, 485-519, 589-625).  The array maths runs in libodk kernels.
"""
import pathlib

import numpy as np
import torch

from . import _dev, _lib


class ObjectsAnnotation:
    """Ground truth of one image (ml.py:15-112): classes (G,) i32, bboxes (G,4) f32 in [0,1]."""

    def __init__(self, path, width, height, classes, bboxes, difficults=None, areas=None, crowdeds=None):
        assert len(classes) == len(bboxes)
        assert difficults is None or len(classes) == len(difficults)
        self.path = pathlib.Path(path)
        self.width = width
        self.height = height
        self.classes = np.asarray(classes, dtype=np.int32)
        self.bboxes = np.asarray(bboxes, dtype=np.float32).reshape((len(self.classes), 4))
        self.difficults = (np.asarray(difficults, dtype=bool) if difficults is not None
                           else np.zeros(len(self.classes), dtype=bool))
        self.areas = np.asarray(areas, dtype=np.float32) if areas is not None else None
        self.crowdeds = np.asarray(crowdeds, dtype=bool) if crowdeds is not None else None
        assert self.width >= 1 and self.height >= 1
        assert (self.bboxes >= 0).all() and (self.bboxes <= 1).all()
        assert (self.bboxes[:, :2] < self.bboxes[:, 2:]).all()
        assert self.difficults.shape == (self.num_objects,)

    @property
    def num_objects(self):
        return len(self.classes)

    @property
    def real_bboxes(self):
        """Pixel boxes: round-half-even of the f64 product (ml.py:62-65)."""
        return np.round(self.bboxes * [self.width, self.height, self.width, self.height]).astype(np.int32)

    def rot90(self, k):
        """Quarter turns of the normalised boxes (ml.py:102-112)."""
        assert 0 <= k <= 3
        b = self.bboxes
        if k == 1:
            self.bboxes = np.stack([b[:, 1], 1 - b[:, 2], b[:, 3], 1 - b[:, 0]], axis=1)
        elif k == 2:
            self.bboxes = 1 - b[:, [2, 3, 0, 1]]
        elif k == 3:
            self.bboxes = np.stack([1 - b[:, 3], b[:, 0], 1 - b[:, 1], b[:, 2]], axis=1)


class ObjectsPrediction:
    """Detections of one image (ml.py:115-197)."""

    def __init__(self, classes, confs, bboxes):
        self.classes = np.asarray(classes)
        self.confs = np.asarray(confs)
        self.bboxes = np.asarray(bboxes)
        assert self.classes.shape == (self.num_objects,)
        assert self.confs.shape == (self.num_objects,)
        assert self.bboxes.shape == (self.num_objects, 4)

    @property
    def num_objects(self):
        return len(self.classes)

    def get_real_bboxes(self, width, height):
        return np.round(self.bboxes * [width, height, width, height]).astype(np.int32)  # ml.py:184-186

    def is_match(self, classes, bboxes, conf_threshold=0, iou_threshold=0.5):
        """ml.py:159-182: True when the detections with conf >= conf_threshold match the given objects one-to-one
        (IoU >= iou_threshold, same classes).  Evaluated by ``odk_exact_match``."""
        bboxes = np.asarray(bboxes, dtype=np.float32).reshape(-1, 4)
        assert bboxes.shape == (len(classes), 4)
        return bool(_exact_matches([(np.asarray(classes), bboxes)], [self], conf_threshold, iou_threshold)[0])



def bboxes_center(bboxes):
    assert bboxes.shape[-1] == 4
    return (bboxes[..., 2:] + bboxes[..., :2]) / 2


def bboxes_size(bboxes):
    assert bboxes.shape[-1] == 4
    return bboxes[..., 2:] - bboxes[..., :2]


def bboxes_area(bboxes):
    return bboxes_size(bboxes).prod(axis=-1)


def compute_iou(bboxes_a, bboxes_b):
    """ml.py:502-519 -> (len(a), len(b)) fp32.  Inputs are converted to fp32 (the detection path's
    dtype); NumPy in -> NumPy out, CUDA tensors in -> CUDA tensor out."""
    is_np = not isinstance(bboxes_a, torch.Tensor)
    dev = _dev.device(None if is_np or not bboxes_a.is_cuda else bboxes_a.device)
    a = _dev.to_device(bboxes_a, torch.float32, dev)
    b = _dev.to_device(bboxes_b, torch.float32, dev)
    assert a.shape[0] > 0 and b.shape[0] > 0
    assert a.shape == (len(a), 4) and b.shape == (len(b), 4)
    out = torch.empty((len(a), len(b)), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_compute_iou', a.data_ptr(), len(a), b.data_ptr(), len(b), out.data_ptr(), _dev.stream_ptr(dev))
    return out.cpu().numpy() if is_np else out


def non_maximum_suppression(boxes, scores, top_k=200, iou_threshold=0.45):
    """ml.py:589-625 -> indices of the kept boxes in descending-score order (int64).
    fp32 arithmetic (inputs converted); equal scores are ordered higher-index-first
    (``argsort(kind='stable')[::-1]``, SURVEY A-6)."""
    is_np = not isinstance(boxes, torch.Tensor)
    dev = _dev.device(None if is_np or not boxes.is_cuda else boxes.device)
    b = _dev.to_device(boxes, torch.float32, dev)
    s = _dev.to_device(scores, torch.float32, dev)
    assert len(b) == len(s)
    assert b.shape == (len(s), 4)
    n, cap = len(s), max(int(top_k), 2)
    out = torch.empty((cap,), dtype=torch.int32, device=dev)
    cnt = torch.zeros((1,), dtype=torch.int32, device=dev)
    ws_bytes = _lib.load().odk_nms_workspace_bytes(1, max(n, 1), 1, int(top_k))
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_nms', b.data_ptr(), s.data_ptr(), n, int(top_k), float(iou_threshold), 1, out.data_ptr(),
                  cnt.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
    k = int(cnt.item())
    res = out[:k].to(torch.int64)
    return res.cpu().numpy() if is_np else res


def _exact_matches(gt, pred, conf_threshold, iou_threshold):
    """uint8 (n_img,) through odk_exact_match; gt = [(classes, bboxes f32)], pred = [ObjectsPrediction]."""
    from . import voc
    dev = _dev.device()
    counts = np.array([len(c) for c, _ in gt], dtype=np.int64)
    offsets = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    n_gt = int(offsets[-1])
    boxes = np.concatenate([np.asarray(b, np.float32).reshape(-1, 4) for _, b in gt]) if n_gt else np.zeros((1, 4), np.float32)
    classes = np.concatenate([np.asarray(c, np.int32) for c, _ in gt]) if n_gt else np.zeros((1,), np.int32)
    d_boxes = _dev.to_device(boxes, torch.float32, dev)
    d_classes = _dev.to_device(classes, torch.int32, dev)
    d_off = _dev.to_device(offsets, torch.int32, dev)
    dets = voc.pack_predictions(pred, dev)
    out = torch.empty((len(gt),), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_exact_match', len(gt), dets.shape[1], dets.data_ptr(), d_boxes.data_ptr(), d_classes.data_ptr(),
                  d_off.data_ptr(), float(conf_threshold), float(iou_threshold), out.data_ptr(), _dev.stream_ptr(dev))
    return out.cpu().numpy()


def od_accuracy(gt, pred, conf_threshold=0, iou_threshold=0.5):
    """ml.py:343-349: share of images whose detections match the ground truth exactly (``is_match``)."""
    assert len(gt) == len(pred)
    assert 0 < iou_threshold < 1
    assert 0 <= conf_threshold < 1
    if len(gt) == 0:
        return np.mean([])
    return np.mean(_exact_matches([(y.classes, y.bboxes) for y in gt], list(pred), conf_threshold, iou_threshold).astype(bool))


def _match_counts(gt, pred, conf_thresholds, iou_threshold, num_classes):
    """(n_thr, 3, C) int64 TP / FP / FN counts through odk_match_counts (one launch for all thresholds)."""
    from . import voc  # packing helpers shared with the VOC evaluation
    dev = _dev.device()
    g = voc.pack_ground_truth(gt, dev)
    dets = voc.pack_predictions(pred, dev)
    thr = _dev.to_device(np.asarray(conf_thresholds, dtype=np.float32), torch.float32, dev)
    counts = torch.empty((len(thr), 3, num_classes), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_match_counts', len(gt), dets.shape[1], int(num_classes), dets.data_ptr(), g['boxes'].data_ptr(),
                  g['classes'].data_ptr(), g['difficult'].data_ptr(), g['offsets'].data_ptr(), len(thr), thr.data_ptr(),
                  float(iou_threshold), counts.data_ptr(), _dev.stream_ptr(dev))
    return counts.cpu().numpy().astype(np.int64)


def _scores_from_counts(tp, fp, fn):
    supports = tp + fn
    precisions = tp.astype(float) / (tp + fp + 1e-7)
    recalls = tp.astype(float) / (supports + 1e-7)
    fscores = 2 / (1 / (precisions + 1e-7) + 1 / (recalls + 1e-7))
    return precisions, recalls, fscores, supports


def compute_scores(gt, pred, conf_threshold=0, iou_threshold=0.5, num_classes=None):
    """ml.py:352-394 -> (precisions, recalls, fscores, supports) per class."""
    assert len(gt) == len(pred)
    assert 0 < iou_threshold < 1
    assert 0 <= conf_threshold < 1
    if num_classes is None:
        num_classes = np.max(np.concatenate([y.classes for y in gt])) + 1
    c = _match_counts(gt, pred, [conf_threshold], iou_threshold, int(num_classes))[0]
    return _scores_from_counts(c[0], c[1], c[2])


def search_conf_threshold(gt, pred, iou_threshold=0.5):
    """ml.py:328-340: the conf_threshold maximising the support-weighted F1 — all 50 thresholds in one launch."""
    conf_threshold_list = np.linspace(0.01, 0.99, 50)
    num_classes = int(np.max(np.concatenate([y.classes for y in gt])) + 1)
    counts = _match_counts(gt, pred, conf_threshold_list, iou_threshold, num_classes)
    scores = []
    for c in counts:
        _, _, fscores, supports = _scores_from_counts(c[0], c[1], c[2])
        scores.append(np.average(fscores, weights=supports))
    scores = np.array(scores)
    max_scores = scores >= 1
    if max_scores.any():
        return np.mean(conf_threshold_list[max_scores])
    return conf_threshold_list[scores.argmax()]


def od_confusion_matrix(gt, pred, conf_threshold=0, iou_threshold=0.5, num_classes=None):
    """ml.py:397-451 -> (C+1, C+1) int matrix (last row: false detections, last column: missed objects)."""
    assert len(gt) == len(pred)
    assert 0 < iou_threshold < 1
    assert 0 <= conf_threshold < 1
    if num_classes is None:
        num_classes = np.max(np.concatenate([y.classes for y in gt])) + 1
    num_classes = int(num_classes)
    if len(gt) == 0:
        return np.zeros((num_classes + 1, num_classes + 1), dtype=int)
    for y_true, y_pred in zip(gt, pred):
        if y_true.num_objects > 0:
            assert len(y_pred.bboxes) > 0  # ml.compute_iou (ml.py:507-508) refuses an empty detection set
        assert all(0 <= c < num_classes for c in y_true.classes) and all(0 <= c < num_classes for c in y_pred.classes)
    from . import voc
    dev = _dev.device()
    g = voc.pack_ground_truth(gt, dev)
    dets = voc.pack_predictions(pred, dev)
    cm = torch.empty((num_classes + 1, num_classes + 1), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_confusion_matrix', len(gt), dets.shape[1], num_classes, dets.data_ptr(), g['boxes'].data_ptr(),
                  g['classes'].data_ptr(), g['offsets'].data_ptr(), float(conf_threshold), float(iou_threshold), cm.data_ptr(),
                  _dev.stream_ptr(dev))
    return cm.cpu().numpy().astype(int)
