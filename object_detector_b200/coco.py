"""Mirror of ``pytoolkit.data.coco.evaluate`` (data/coco.py:75-99): the MS COCO detection metrics, computed by
libodk's ``odk_coco_ap`` kernels.

The reference delegates to ``chainercv.evaluations.eval_detection_coco`` -> ``pycocotools.cocoeval.COCOeval``
(both un-vendored and unpinned): this path restates the published algorithm — PARITY UNPINNED against the real
packages.  The kernels produce COCOeval's ``precision[T,R,K,A,M]`` / ``recall[T,K,A,M]`` arrays; the twelve
summary numbers (chainercv ``_summarize``: nanmean over thresholds and recall points, then over classes) are
taken from those arrays on the host.
"""
import numpy as np
import torch

from . import _dev, _lib

IOU_THRS = np.linspace(.5, 0.95, int(np.round((0.95 - .5) / .05)) + 1, endpoint=True)   # pycocotools Params.setDetParams
REC_THRS = np.linspace(.0, 1.00, int(np.round((1.00 - .0) / .01)) + 1, endpoint=True)
MAX_DETS = (1, 10, 100)
AREA_RNG = np.array([[0 ** 2, 1e5 ** 2], [0 ** 2, 32 ** 2], [32 ** 2, 96 ** 2], [96 ** 2, 1e5 ** 2]], dtype=np.float64)
AREA_LBL = ('all', 'small', 'medium', 'large')
KEYS = (
    ('ap/iou=0.50:0.95/area=all/max_dets=100', True, None, 'all', 100),
    ('ap/iou=0.50/area=all/max_dets=100', True, 0, 'all', 100),
    ('ap/iou=0.75/area=all/max_dets=100', True, 5, 'all', 100),
    ('ap/iou=0.50:0.95/area=small/max_dets=100', True, None, 'small', 100),
    ('ap/iou=0.50:0.95/area=medium/max_dets=100', True, None, 'medium', 100),
    ('ap/iou=0.50:0.95/area=large/max_dets=100', True, None, 'large', 100),
    ('ar/iou=0.50:0.95/area=all/max_dets=1', False, None, 'all', 1),
    ('ar/iou=0.50:0.95/area=all/max_dets=10', False, None, 'all', 10),
    ('ar/iou=0.50:0.95/area=all/max_dets=100', False, None, 'all', 100),
    ('ar/iou=0.50:0.95/area=small/max_dets=100', False, None, 'small', 100),
    ('ar/iou=0.50:0.95/area=medium/max_dets=100', False, None, 'medium', 100),
    ('ar/iou=0.50:0.95/area=large/max_dets=100', False, None, 'large', 100),
)


def pack_ground_truth(y_true, dev):
    """[ObjectsAnnotation] -> flat device arrays.  ``areas`` None -> the pixel box area, ``crowdeds`` None -> False
    (chainercv _create_ann)."""
    counts = np.array([len(y.classes) for y in y_true], dtype=np.int32)
    offsets = np.zeros(len(y_true) + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(counts)
    total = int(offsets[-1])
    boxes = np.zeros((max(total, 1), 4), dtype=np.float32)
    classes = np.zeros((max(total, 1),), dtype=np.int32)
    areas = np.zeros((max(total, 1),), dtype=np.float64)
    crowd = np.zeros((max(total, 1),), dtype=np.uint8)
    for y, o, c in zip(y_true, offsets[:-1], counts):
        if not c:
            continue
        boxes[o:o + c] = y.bboxes
        classes[o:o + c] = y.classes
        if getattr(y, 'areas', None) is not None:
            areas[o:o + c] = y.areas
        else:
            px = np.round(np.asarray(y.bboxes, dtype=np.float32) * [y.width, y.height, y.width, y.height]).astype(np.int32)
            areas[o:o + c] = (px[:, 2] - px[:, 0]).astype(np.float64) * (px[:, 3] - px[:, 1])
        if getattr(y, 'crowdeds', None) is not None:
            crowd[o:o + c] = np.asarray(y.crowdeds, dtype=np.uint8)
    wh = np.array([[y.width, y.height] for y in y_true], dtype=np.int32).reshape(len(y_true), 2)
    return {
        'img_wh': _dev.to_device(wh, torch.int32, dev), 'boxes': _dev.to_device(boxes, torch.float32, dev),
        'classes': _dev.to_device(classes, torch.int32, dev), 'areas': _dev.to_device(areas, torch.float64, dev),
        'crowd': _dev.to_device(crowd, torch.uint8, dev), 'offsets': _dev.to_device(offsets, torch.int32, dev),
        'gt_max': int(counts.max()) if len(counts) else 0, 'n_img': len(y_true),
    }


def evaluate_detections(dets, gt, num_classes, conf_threshold=float('-inf')):
    """Device entry: ``dets`` (n_img, K, 6) CUDA tensor, ``gt`` from ``pack_ground_truth``.  Returns COCOeval's
    ``precision`` (T,R,C,A,M) and ``recall`` (T,C,A,M) as float64 CUDA tensors (-1 = undefined)."""
    dev = dets.device
    n_img, k = dets.shape[0], dets.shape[1]
    assert dets.dtype == torch.float32 and dets.is_contiguous() and dets.shape[2] == 6 and n_img == gt['n_img']
    t, r, a, m = len(IOU_THRS), len(REC_THRS), len(AREA_RNG), len(MAX_DETS)
    precision = torch.empty((t, r, num_classes, a, m), dtype=torch.float64, device=dev)
    recall = torch.empty((t, num_classes, a, m), dtype=torch.float64, device=dev)
    lib = _lib.load()
    ws_bytes = lib.odk_coco_ap_workspace_bytes(n_img, k, num_classes)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    iou, rec, rng = (np.ascontiguousarray(x, dtype=np.float64) for x in (IOU_THRS, REC_THRS, AREA_RNG))
    with torch.cuda.device(dev):
        _lib.call('odk_coco_ap', n_img, k, num_classes, dets.data_ptr(), float(conf_threshold), gt['img_wh'].data_ptr(),
                  gt['boxes'].data_ptr(), gt['classes'].data_ptr(), gt['areas'].data_ptr(), gt['crowd'].data_ptr(),
                  gt['offsets'].data_ptr(), gt['gt_max'], iou.ctypes.data, rec.ctypes.data, rng.ctypes.data,
                  precision.data_ptr(), recall.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
    return precision, recall


def summarize(precision, recall, existent_labels):
    """chainercv eval_detection_coco's result dict from the (T,R,K,A,M) / (T,K,A,M) arrays restricted to the
    ``existent_labels`` columns (pycocotools only knows categories that occur in the GT or the predictions)."""
    import warnings
    res = {}
    existent_labels = np.asarray(existent_labels, dtype=np.int64)
    n = int(existent_labels.max()) + 1 if len(existent_labels) else 0
    with warnings.catch_warnings():
        warnings.simplefilter('ignore', RuntimeWarning)
        for key, is_ap, iou_ix, area, max_det in KEYS:
            a, m = AREA_LBL.index(area), MAX_DETS.index(max_det)
            v = np.array(precision if is_ap else recall, dtype=np.float64)
            if iou_ix is not None:
                v = v[iou_ix:iou_ix + 1]
            v = v[:, :, :, a, m] if is_ap else v[:, :, a, m]
            v[v == -1] = np.nan
            v = v.reshape((-1, v.shape[-1]))
            valid = np.any(~np.isnan(v), axis=0)
            per_class = np.full(len(valid), np.nan)
            if valid.any():
                per_class[valid] = np.nanmean(v[:, valid], axis=0)
            full = np.full(n, np.nan)
            full[existent_labels] = per_class
            res[key] = full
            res['m' + key] = np.nanmean(per_class) if valid.any() else np.nan
    res['existent_labels'] = existent_labels
    return res


def evaluate(y_true, y_pred):
    """data/coco.py:75-99: chainercv's dict — per-class ``ap/...`` / ``ar/...`` arrays (nan = label absent or
    undefined), their class means ``map/...`` / ``mar/...``, and ``existent_labels``."""
    from . import voc
    assert len(y_true) == len(y_pred)
    dev = _dev.device()
    labels = sorted(set(int(c) for y in y_true for c in y.classes) | set(int(c) for p in y_pred for c in p.classes))
    num_classes = labels[-1] + 1 if labels else 1
    gt = pack_ground_truth(y_true, dev)
    dets = voc.pack_predictions(y_pred, dev)
    precision, recall = (x.cpu().numpy() for x in evaluate_detections(dets, gt, num_classes))
    return summarize(precision[:, :, labels], recall[:, labels], labels)
