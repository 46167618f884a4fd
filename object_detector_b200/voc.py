"""Mirror of ``pytoolkit.data.voc.evaluate`` (data/voc.py:165-197): PASCAL VOC AP / mAP, both the area
metric and the VOC2007 11-point metric, computed by libodk's ``odk_voc_ap`` kernels.

The reference delegates to ``chainercv.evaluations.eval_detection_voc`` (un-vendored, unpinned): this
path restates chainercv's published algorithm — PARITY UNPINNED against the real package.
"""
import numpy as np
import torch

from . import _dev, _lib


def pack_ground_truth(y_true, dev):
    """[ObjectsAnnotation] -> flat device arrays (img_wh, boxes, classes, difficult, offsets, max per image)."""
    counts = np.array([len(y.classes) for y in y_true], dtype=np.int32)
    offsets = np.zeros(len(y_true) + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(counts)
    total = int(offsets[-1])
    boxes = np.zeros((max(total, 1), 4), dtype=np.float32)
    classes = np.zeros((max(total, 1),), dtype=np.int32)
    difficult = np.zeros((max(total, 1),), dtype=np.uint8)
    for y, o, c in zip(y_true, offsets[:-1], counts):
        if c:
            boxes[o:o + c] = y.bboxes
            classes[o:o + c] = y.classes
            difficult[o:o + c] = np.asarray(y.difficults, dtype=np.uint8)
    wh = np.array([[y.width, y.height] for y in y_true], dtype=np.int32).reshape(len(y_true), 2)
    return {
        'img_wh': _dev.to_device(wh, torch.int32, dev), 'boxes': _dev.to_device(boxes, torch.float32, dev),
        'classes': _dev.to_device(classes, torch.int32, dev), 'difficult': _dev.to_device(difficult, torch.uint8, dev),
        'offsets': _dev.to_device(offsets, torch.int32, dev), 'gt_max': int(counts.max()) if len(counts) else 0,
        'n_img': len(y_true),
    }


def pack_predictions(y_pred, dev):
    """[ObjectsPrediction] -> (n_img, K, 6) rows [class, conf, box]; padding rows have class -1."""
    k = max(1, max((p.num_objects for p in y_pred), default=1))
    dets = np.zeros((len(y_pred), k, 6), dtype=np.float32)
    dets[:, :, 0] = -1
    for i, p in enumerate(y_pred):
        n = p.num_objects
        if n:
            dets[i, :n, 0] = p.classes
            dets[i, :n, 1] = p.confs
            dets[i, :n, 2:] = p.bboxes
    return _dev.to_device(dets, torch.float32, dev)


def evaluate_detections(dets, gt, num_classes, conf_threshold=float('-inf')):
    """Device entry: ``dets`` (n_img, K, 6) CUDA tensor (e.g. the gathered NMS-layer output), ``gt`` from
    ``pack_ground_truth``.  Rows with conf < conf_threshold (od.py:265) or class outside [0, C) are
    ignored.  Returns (ap_area, ap07) float64 CUDA tensors of length C (nan where undefined)."""
    dev = dets.device
    n_img, k = dets.shape[0], dets.shape[1]
    assert dets.dtype == torch.float32 and dets.is_contiguous() and dets.shape[2] == 6 and n_img == gt['n_img']
    ap = torch.empty((2, num_classes), dtype=torch.float64, device=dev)
    lib = _lib.load()
    ws_bytes = lib.odk_voc_ap_workspace_bytes(n_img, k, num_classes, gt['gt_max'])
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_voc_ap', n_img, k, num_classes, dets.data_ptr(), float(conf_threshold), gt['img_wh'].data_ptr(),
                  gt['boxes'].data_ptr(), gt['classes'].data_ptr(), gt['difficult'].data_ptr(), gt['offsets'].data_ptr(),
                  gt['gt_max'], ap[0].data_ptr(), ap[1].data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
    return ap[0], ap[1]


def evaluate(y_true, y_pred, num_classes=None):
    """data/voc.py:165-197: returns {'AP', 'mAP', 'AP_VOC', 'mAP_VOC'} (float64; nan = class absent).
    ``num_classes`` defaults to 1 + the largest label seen, which is chainercv's ``n_fg_class``."""
    assert len(y_true) == len(y_pred)
    dev = _dev.device()
    if num_classes is None:
        labels = [np.max(y.classes) for y in y_true if len(y.classes)] + [np.max(p.classes) for p in y_pred if p.num_objects]
        num_classes = int(max(labels)) + 1 if labels else 1
    gt = pack_ground_truth(y_true, dev)
    dets = pack_predictions(y_pred, dev)
    ap, ap07 = (t.cpu().numpy() for t in evaluate_detections(dets, gt, num_classes))
    with np.errstate(invalid='ignore'):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore', RuntimeWarning)
            return {'AP': ap, 'mAP': np.nanmean(ap), 'AP_VOC': ap07, 'mAP_VOC': np.nanmean(ap07)}
