"""CPU ORACLE — TEST INFRASTRUCTURE, NOT A PRODUCT PATH.

A NumPy restatement of the reference's detection hot path (ak110/object_detector,
pytoolkit submodule d542569).  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this module, and
only as the checker / timed CPU baseline.  Nothing in ``object_detector_b200/`` imports it.

Pinning (see tests/test_oracle_cpu.py and tests/golden/):
  * anchors / assign / encode / decode / IoU / NumPy-NMS: pinned against outputs of the
    unmodified reference run in the build container (``oracle/gen_golden.py`` ->
    ``tests/golden/*.npz``) and against the reference's own KATs
    (pytoolkit/tests/ml_test.py:42-53, 89-113).
  * ``nms_layer`` (in-graph ``layers.nms``): STRUCTURE restated from layers.py:925-970 with the
    reference's NumPy NMS as inner op; the real inner op (tf.image.non_max_suppression,
    tensorflow==1.10.0) is absent -> **parity unpinned** at that boundary.
  * ``eval_detection_voc``: restatement of chainercv (no version pinned by the reference,
    not vendored) -> **parity unpinned**; hand-computed micro-cases only.

Every function cites the reference lines it follows (paths relative to
/root/reference/pytoolkit/pytoolkit/).  The arithmetic (dtype of every intermediate, op order)
is the reference's as executed by NumPy 2.x (NEP 50 promotion); the code layout is not.
"""
import numpy as np

VAR_LOC = 0.10  # dl/od_pb.py:10

VOC_PATTERNS = [  # dl/od.py:20-29 (_VOC_JSON_DATA['pb_size_patterns'])
    [1.2133152484893799, 1.6041910648345947],
    [2.301241874694824, 3.812246322631836],
    [3.2387468814849854, 7.3555989265441895],
    [5.467498779296875, 4.480413436889648],
    [4.758797645568848, 11.54628849029541],
    [10.89200496673584, 5.788760185241699],
    [7.736226558685303, 8.850168228149414],
    [4.47715950012207, 16.65940284729004],
]
VOC_MAP_SIZES = [40, 20, 10]


# --------------------------------------------------------------------------------------
# math.py:8-17
def closed_interval(x, lo, hi):
    """math.between: lo <= x <= hi."""
    return (lo <= x) & (x <= hi)


def half_open_interval(x, lo, hi):
    """math.in_range: lo <= x < hi."""
    return (lo <= x) & (x < hi)


# --------------------------------------------------------------------------------------
# ml.py:485-499
def box_centers(b):
    return (b[..., 2:] + b[..., :2]) / 2


def box_areas(b):
    return (b[..., 2:] - b[..., :2]).prod(axis=-1)


def pairwise_iou(a, b):
    """ml.compute_iou (ml.py:502-519): (A,4),(Bn,4) -> (A,Bn); dtype follows the inputs."""
    assert a.ndim == 2 and a.shape[1] == 4 and a.shape[0] > 0
    assert b.ndim == 2 and b.shape[1] == 4 and b.shape[0] > 0
    lo = np.maximum(a[:, None, :2], b[None, :, :2])
    hi = np.minimum(a[:, None, 2:], b[None, :, 2:])
    overlap = np.prod(hi - lo, axis=2) * (lo < hi).all(axis=2)
    area_a = np.prod(a[:, 2:] - a[:, :2], axis=1)
    area_b = np.prod(b[:, 2:] - b[:, :2], axis=1)
    return overlap / (area_a[:, None] + area_b - overlap)


# --------------------------------------------------------------------------------------
class AnchorTables:
    """The arrays ``PriorBoxes._create_pb`` builds (dl/od_pb.py:125-184), same dtypes."""

    def __init__(self, input_size, map_sizes, pb_size_patterns, num_classes):
        self.input_size = tuple(input_size)
        self.map_sizes = np.array(map_sizes)  # from_dict keeps the given order (od_pb.py:55)
        self.pb_size_patterns = np.array(pb_size_patterns)
        self.num_classes = int(num_classes)
        px = 1 / np.array(self.input_size)  # f64 (2,)
        grids, locs, centers, sizes, masks, infos = [], [], [], [], [], []
        for m in self.map_sizes:
            cell = 1.0 / m  # np.float64 scalar
            lin = np.linspace(0.5 * cell, 1 - 0.5 * cell, m, dtype=np.float32)
            cx, cy = np.meshgrid(lin, lin)
            c4 = np.tile(np.stack([cx.ravel(), cy.ravel()], axis=1), (1, 2))  # f32 (m*m,4)
            g = c4.copy()
            g[:, :2] -= cell / 2  # f64 arithmetic, stored f32 (od_pb.py:149)
            g[:, 2:] += cell / 2 + px  # od_pb.py:150
            for k, pat in enumerate(self.pb_size_patterns):
                size = cell * pat  # f64 (2,)
                box = c4.copy()
                box[:, :2] -= size / 2  # od_pb.py:158
                box[:, 2:] += size / 2 + px  # od_pb.py:159
                cl = np.clip(box, 0, 1)
                cc = (cl[:, :2] + cl[:, 2:]) / 2
                ok = closed_interval(cc, g[:, :2], g[:, 2:]).all(axis=-1)  # od_pb.py:163
                grids.append(g)
                locs.append(box)
                centers.append(c4[:, :2])
                sizes.append(np.broadcast_to(size, (len(box), 2)))
                masks.append(ok)
                infos.append(np.full(len(box), len(infos), dtype=np.int64))
        self.pb_grid = np.concatenate(grids)
        self.pb_locs = np.concatenate(locs)
        self.pb_centers = np.concatenate(centers)
        self.pb_sizes = np.concatenate(sizes).astype(np.float64)
        self.pb_mask = np.concatenate(masks)
        self.pb_info_indices = np.concatenate(infos)
        mid = np.mean([self.pb_locs[:, 2:], self.pb_locs[:, :2]], axis=0)  # od_pb.py:197-198
        assert half_open_interval(mid, self.pb_grid[:, :2], self.pb_grid[:, 2:]).all()

    def __len__(self):
        return len(self.pb_locs)


def encode_locs(t: AnchorTables, bbox, sel):
    """dl/od_pb.py:212-214 — f32 difference, then two f64 divisions."""
    return (bbox - t.pb_locs[sel, :]) / np.tile(t.pb_sizes[sel, :], 2) / VAR_LOC


def decode_locs_f64(t: AnchorTables, pred):
    """dl/od_pb.py:216-222 with xp=np (the check_assign.py:27 use): f64 result."""
    return np.clip(pred * (VAR_LOC * np.tile(t.pb_sizes, 2)) + t.pb_locs, 0, 1)


def decode_locs_f32(t: AnchorTables, pred):
    """dl/od_pb.py:216-222 as the Keras Lambda evaluates it (od_net.py:189): the f64 constant
    ``0.10*tile(pb_sizes,2)`` is converted to the tensor dtype (f32) and mul/add/clip run in f32."""
    scale = (VAR_LOC * np.tile(t.pb_sizes, 2)).astype(np.float32)
    pred = np.asarray(pred, dtype=np.float32)
    return np.clip(pred * scale + t.pb_locs, np.float32(0), np.float32(1))


def assign_boxes(t: AnchorTables, bboxes):
    """dl/od_pb.py:315-363.  Returns (pb_indices i64 sorted, gt_indices i64, pb_valids bool)."""
    n = len(t)
    centres = box_centers(bboxes)
    owner = np.full(n, -1, dtype=np.int64)
    free = np.ones(n, dtype=bool)
    valid = np.ones(n, dtype=bool)
    order = box_areas(bboxes).argsort()  # od_pb.py:327 (default, unstable sort kind)
    for g in order:
        box, ctr = bboxes[g], centres[g]
        usable = t.pb_mask & free
        if not usable.any():
            raise AssertionError('Encode error: %s' % (ctr,))  # od_pb.py:332
        iou = pairwise_iou(box[None, :], t.pb_locs[usable, :])[0]  # od_pb.py:333
        cell = half_open_interval(ctr, t.pb_grid[:, :2], t.pb_grid[:, 2:]).all(axis=-1) & t.pb_mask
        if not cell.any():
            raise AssertionError('Encode error: %s' % (ctr,))  # od_pb.py:338
        cell &= free
        if cell.any():
            cost = np.sum(np.abs(encode_locs(t, box, cell)), axis=-1)  # od_pb.py:342
            pick = np.flatnonzero(cell)[cost.argmin()]
        else:
            pick = np.flatnonzero(usable)[iou.argmax()]  # od_pb.py:349-350
        owner[pick] = g
        free[pick] = False
        valid[np.flatnonzero(usable)[iou >= 0.5]] = False  # od_pb.py:355-357
    idx = np.flatnonzero(owner >= 0)
    if len(idx):
        valid[idx] = True  # od_pb.py:359-361
    return idx, owner[idx], valid


def encode_truth(t: AnchorTables, gt_list):
    """dl/od_pb.py:299-313.  ``gt_list`` = [(classes i32 (G,), bboxes f32 (G,4)), ...]."""
    c = t.num_classes
    out = np.zeros((len(gt_list), len(t), 3 + c + 4), dtype=np.float32)
    for i, (classes, bboxes) in enumerate(gt_list):
        classes = np.asarray(classes)
        assert half_open_interval(classes, 0, c).all()  # od_pb.py:304
        idx, gts, valid = assign_boxes(t, bboxes)
        out[i, valid, 0] = 1 / max(1, len(idx))
        for a, g in zip(idx, gts):
            out[i, a, 1] = 1 / len(idx)
            out[i, a, 2] = 1
            out[i, a, 3 + classes[g]] = 1
            out[i, a, -4:] = encode_locs(t, bboxes[g, :], a)
    assert ((out[:, :, 0] != 0) & (out[:, :, 2] != 0) == (out[:, :, 2] != 0)).all()  # od_pb.py:312
    return out


# --------------------------------------------------------------------------------------
def numpy_nms(boxes, scores, top_k=200, iou_threshold=0.45):
    """ml.non_maximum_suppression (ml.py:589-625) incl. the zero-area skip, the ``>=`` test and
    the quirk that the length check sits only in the non-first branch (SURVEY A-7)."""
    assert len(boxes) == len(scores) and boxes.shape == (len(scores), 4)
    order = scores.argsort()[::-1]
    b = boxes[order]
    area = np.prod(b[:, 2:] - b[:, :2], axis=1)
    keep = []
    for i in range(len(b)):
        if (b[i, 2:] - b[i, :2] <= 0).any():
            continue
        if keep:
            lo = np.maximum(b[i, :2], b[keep, :2])
            hi = np.minimum(b[i, 2:], b[keep, 2:])
            inter = np.prod(hi - lo, axis=1) * (lo < hi).all(axis=1)
            if (inter / (area[i] + area[keep] - inter) >= iou_threshold).any():
                continue
            keep.append(i)
            if len(keep) >= top_k:
                break
        else:
            keep.append(i)
    return order[keep]


def stable_desc_order(scores):
    """Tie policy of this build (SURVEY A-6): ``argsort(kind='stable')[::-1]`` — among equal
    scores the HIGHER original index comes first.  Equals the reference whenever scores are
    distinct (the synthetic workloads guarantee that)."""
    return np.argsort(scores, kind='stable')[::-1]


def numpy_nms_policy(boxes, scores, top_k=200, iou_threshold=0.45):
    """``numpy_nms`` with the deterministic tie policy above (what the CUDA kernel implements).
    Also returns early exactly like the reference; vectorised bookkeeping only."""
    order = stable_desc_order(scores)
    b = boxes[order]
    area = np.prod(b[:, 2:] - b[:, :2], axis=1)
    thr = scores.dtype.type(iou_threshold) if scores.dtype.kind == 'f' else iou_threshold
    keep = []
    for i in range(len(b)):
        if (b[i, 2:] - b[i, :2] <= 0).any():
            continue
        if keep:
            lo = np.maximum(b[i, :2], b[keep, :2])
            hi = np.minimum(b[i, 2:], b[keep, 2:])
            inter = np.prod(hi - lo, axis=1) * (lo < hi).all(axis=1)
            if (inter / (area[i] + area[keep] - inter) >= thr).any():
                continue
            keep.append(i)
            if len(keep) >= top_k:
                break
        else:
            keep.append(i)
    return order[keep]


def tf_nms_policy(boxes, scores, top_k=200, iou_threshold=0.45):
    """Restatement of ``tf.image.non_max_suppression`` as TensorFlow 1.10 (the version the reference pins,
    .travis.requirements.txt:13; call sites layers.py:942,954) implements it in non_max_suppression_op.cc —
    PARITY UNPINNED (TensorFlow is not in the tree; no reference test checks the op's output):
    candidates in descending score; a candidate is suppressed iff its IoU with an already selected box is
    strictly GREATER than the threshold, where the IoU of a pair is 0 when either box has area <= 0 (such boxes
    are therefore never suppressed, never suppress, and can be selected); selection stops at exactly
    ``max_output_size``.  fp32 arithmetic: inter / (area_i + area_j - inter).  Tie order among equal scores is
    unspecified in TF (a priority queue); this build's policy (``stable_desc_order``) is used."""
    order = stable_desc_order(scores)
    b = boxes[order].astype(np.float32)
    area = ((b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])).astype(np.float32)
    thr = np.float32(iou_threshold)
    keep = []
    for i in range(len(b)):
        if len(keep) >= top_k:
            break
        if keep and area[i] > 0:
            k = np.asarray(keep)
            w = np.maximum(np.minimum(b[i, 2], b[k, 2]) - np.maximum(b[i, 0], b[k, 0]), np.float32(0))
            h = np.maximum(np.minimum(b[i, 3], b[k, 3]) - np.maximum(b[i, 1], b[k, 1]), np.float32(0))
            inter = (w * h).astype(np.float32)
            with np.errstate(invalid='ignore', divide='ignore'):
                iou = np.where(area[k] > 0, inter / (area[i] + area[k] - inter), np.float32(0))
            if (iou > thr).any():
                continue
        keep.append(i)
    return order[np.asarray(keep, dtype=np.int64)]


def score_anchors(t: AnchorTables, objs, clfs):
    """layers.py:69-98 (ChannelArgMax / ChannelMax) + od_net.py:177-184 (_objconf).
    objs (B,N,1) f32, clfs (B,N,C) f32 -> classes (B,N) i64, objconfs (B,N) f32."""
    classes = np.argmax(clfs, axis=-1)
    confs = np.max(clfs, axis=-1)
    conf = objs[:, :, 0] * confs
    return classes, conf * t.pb_mask[None, :].astype(np.float32)


def greedy_nms_in_order(boxes, iou_threshold, top_k):
    """Body of ml.non_maximum_suppression (ml.py:605-624) over boxes ALREADY in visiting order:
    returns the kept positions.  Same zero-area skip, ``>=`` test and length-check placement."""
    area = np.prod(boxes[:, 2:] - boxes[:, :2], axis=1)
    thr = boxes.dtype.type(iou_threshold) if boxes.dtype.kind == 'f' else iou_threshold
    keep = []
    for i in range(len(boxes)):
        if (boxes[i, 2:] - boxes[i, :2] <= 0).any():
            continue
        if keep:
            lo = np.maximum(boxes[i, :2], boxes[keep, :2])
            hi = np.minimum(boxes[i, 2:], boxes[keep, 2:])
            inter = np.prod(hi - lo, axis=1) * (lo < hi).all(axis=1)
            if (inter / (area[i] + area[keep] - inter) >= thr).any():
                continue
            keep.append(i)
            if len(keep) >= top_k:
                break
        else:
            keep.append(i)
    return np.asarray(keep, dtype=np.int64)


def nms_layer_image(classes, confs, locs, num_classes, top_k=200, conf_threshold=0.01,
                    nms_threshold=0.45, inner=numpy_nms_policy, nms_all_threshold=None):
    """layers.py:925-970 for one image.  ``nms_all_threshold=None``: per-class NMS then global
    top_k (:958-961).  With a threshold (strict_nms, od_net.py:175): the per-class limit doubles
    (:940) and a second, class-agnostic greedy NMS runs over the class-concatenated survivors
    (:953-957).  Tie policy of this build for that second pass: equal scores are visited in
    concatenation order (lower class first) — identical to the reference when scores are distinct.
    Returns (out (top_k,6) f32, kept anchor index (top_k,) i64 with -1 padding)."""
    cat_cls, cat_conf, cat_loc, cat_idx = [], [], [], []
    class_limit = top_k * 2 if nms_all_threshold else top_k  # :940
    for c in range(num_classes):
        sel = np.flatnonzero((classes == c) & (confs > np.float32(conf_threshold)))  # :935 strict
        if len(sel) == 0:
            continue
        k = min(class_limit, len(sel))  # :941
        kept = inner(locs[sel], confs[sel], k, nms_threshold)  # :942
        cat_cls.append(np.full(len(kept), c, dtype=np.int64))
        cat_conf.append(confs[sel][kept])
        cat_loc.append(locs[sel][kept])
        cat_idx.append(sel[kept])
    out = np.zeros((top_k, 6), dtype=np.float32)
    kept_idx = np.full(top_k, -1, dtype=np.int64)
    if cat_cls:
        cls = np.concatenate(cat_cls)
        conf = np.concatenate(cat_conf)
        loc = np.concatenate(cat_loc)
        idx = np.concatenate(cat_idx)
        k = min(top_k, len(conf))  # :952
        order = np.argsort(-conf, kind='stable')  # lower position first on ties
        if nms_all_threshold:
            top = order[greedy_nms_in_order(loc[order], nms_all_threshold, k)]  # :954
            k = len(top)
        else:
            top = order[:k]  # tf.nn.top_k(sorted=True)
        out[:k, 0] = cls[top]
        out[:k, 1] = conf[top]
        out[:k, 2:] = loc[top]
        kept_idx[:k] = idx[top]
    return out, kept_idx


def predict_tail(t: AnchorTables, objs, clfs, locs, top_k=200, conf_threshold=0.01,
                 nms_threshold=0.45, inner=numpy_nms_policy, nms_all_threshold=None):
    """od_net.py:172-192: score -> decode (f32) -> NMS layer, whole batch.
    Returns (B,top_k,6) f32 and (B,top_k) i64 kept anchor ids."""
    classes, confs = score_anchors(t, objs, clfs)
    boxes = decode_locs_f32(t, locs)
    outs, idxs = [], []
    for b in range(len(objs)):
        o, k = nms_layer_image(classes[b], confs[b], boxes[b], t.num_classes, top_k,
                               conf_threshold, nms_threshold, inner, nms_all_threshold)
        outs.append(o)
        idxs.append(k)
    return np.stack(outs), np.stack(idxs)


def unpack_predictions(out, conf_threshold=0.01):
    """dl/od.py:258-266 with keep_aspect=False (offset 0, size 1): list of (classes i32, confs, boxes)."""
    res = []
    for p in out:
        m = p[:, 1] >= conf_threshold
        res.append((p[m, 0].astype(np.int32), p[m, 1], (p[m, 2:] - 0) / 1))
    return res


# --------------------------------------------------------------------------------------
def to_pixels(bboxes, width, height):
    """ml.py:62-65 / 184-186: f64 product, round-half-even, int32."""
    return np.round(bboxes * [width, height, width, height]).astype(np.int32)


def _int_iou(a, b):
    """chainercv.utils.bbox_iou on integer boxes (area as integer products, true division)."""
    lo = np.maximum(a[:, None, :2], b[None, :, :2])
    hi = np.minimum(a[:, None, 2:], b[None, :, 2:])
    inter = np.prod(hi - lo, axis=2) * (lo < hi).all(axis=2)
    aa = np.prod(a[:, 2:] - a[:, :2], axis=1)
    ab = np.prod(b[:, 2:] - b[:, :2], axis=1)
    return inter / (aa[:, None] + ab - inter)


def voc_match(pred_boxes, pred_classes, pred_scores, gt_boxes, gt_classes, gt_difficult,
              num_classes, iou_thresh=0.5):
    """chainercv calc_detection_voc_prec_rec, matching part.  Per class returns
    n_pos[c], and lists score[c], match[c] (1 TP, 0 FP, -1 ignored)."""
    n_pos = np.zeros(num_classes, dtype=np.int64)
    score = [[] for _ in range(num_classes)]
    match = [[] for _ in range(num_classes)]
    for pb, pc, ps, gb, gc, gd in zip(pred_boxes, pred_classes, pred_scores, gt_boxes, gt_classes,
                                      gt_difficult):
        for c in np.unique(np.concatenate([pc, gc]).astype(int)):
            pm = pc == c
            pbc, psc = pb[pm], ps[pm]
            order = np.argsort(psc, kind='stable')[::-1]
            pbc, psc = pbc[order], psc[order]
            gm = gc == c
            gbc, gdc = gb[gm], gd[gm]
            n_pos[c] += np.logical_not(gdc).sum()
            score[c].extend(psc)
            if len(pbc) == 0:
                continue
            if len(gbc) == 0:
                match[c].extend([0] * len(pbc))
                continue
            pbc = pbc.astype(np.int64).copy()
            pbc[:, 2:] += 1
            gbc = gbc.astype(np.int64).copy()
            gbc[:, 2:] += 1
            iou = _int_iou(pbc, gbc)
            gi = iou.argmax(axis=1)
            gi[iou.max(axis=1) < iou_thresh] = -1
            used = np.zeros(len(gbc), dtype=bool)
            for j in gi:
                if j >= 0:
                    if gdc[j]:
                        match[c].append(-1)
                    else:
                        match[c].append(0 if used[j] else 1)
                    used[j] = True
                else:
                    match[c].append(0)
    return n_pos, score, match


def voc_ap(n_pos, score, match, use_07_metric):
    """chainercv calc_detection_voc_prec_rec (curve part) + calc_detection_voc_ap."""
    num_classes = len(n_pos)
    ap = np.full(num_classes, np.nan)
    for c in range(num_classes):
        s = np.array(score[c], dtype=np.float64)
        m = np.array(match[c], dtype=np.int8)
        if n_pos[c] == 0 and len(s) == 0:
            continue  # class absent: prec/rec None -> ap nan
        order = np.argsort(s, kind='stable')[::-1]
        m = m[order]
        tp = np.cumsum(m == 1)
        fp = np.cumsum(m == 0)
        with np.errstate(divide='ignore', invalid='ignore'):
            prec = tp / (fp + tp)
            rec = tp / n_pos[c] if n_pos[c] > 0 else None
        if rec is None:
            continue
        if use_07_metric:
            a = 0.0
            for thr in np.arange(0.0, 1.1, 0.1):
                p = 0 if np.sum(rec >= thr) == 0 else np.max(np.nan_to_num(prec)[rec >= thr])
                a += p / 11
            ap[c] = a
        else:
            mpre = np.concatenate(([0], np.nan_to_num(prec), [0]))
            mrec = np.concatenate(([0], rec, [1]))
            mpre = np.maximum.accumulate(mpre[::-1])[::-1]
            i = np.where(mrec[1:] != mrec[:-1])[0]
            ap[c] = np.sum((mrec[i + 1] - mrec[i]) * mpre[i + 1])
    return ap


def voc_evaluate(gt_list, pred_list, num_classes):
    """data/voc.py:165-197.  gt_list = [(width, height, classes, bboxes f32, difficults)],
    pred_list = [(classes, confs, bboxes f32)].  Returns the reference's 4-key dict."""
    pb = [to_pixels(p[2], g[0], g[1]) for p, g in zip(pred_list, gt_list)]
    pc = [np.asarray(p[0]) for p in pred_list]
    ps = [np.asarray(p[1]) for p in pred_list]
    gb = [to_pixels(g[3], g[0], g[1]) for g in gt_list]
    gc = [np.asarray(g[2]) for g in gt_list]
    gd = [np.asarray(g[4], dtype=bool) for g in gt_list]
    n_pos, score, match = voc_match(pb, pc, ps, gb, gc, gd, num_classes)
    ap = voc_ap(n_pos, score, match, False)
    ap07 = voc_ap(n_pos, score, match, True)
    with np.errstate(invalid='ignore'):
        return {'AP': ap, 'mAP': np.nanmean(ap), 'AP_VOC': ap07, 'mAP_VOC': np.nanmean(ap07)}


# --------------------------------------------------------------------------------------
def match_counts(gt_list, pred_list, conf_threshold, iou_threshold, num_classes):
    """Counting part of ml.compute_scores (ml.py:352-389).  gt_list = [(classes, bboxes f32, difficults)],
    pred_list = [(classes, confs f32, bboxes f32)].  Returns int arrays tp, fp, fn."""
    tp = np.zeros((num_classes,), dtype=int)
    fp = np.zeros((num_classes,), dtype=int)
    fn = np.zeros((num_classes,), dtype=int)
    for (gc, gb, gd), (pc, ps, pb) in zip(gt_list, pred_list):
        enabled = ps >= conf_threshold  # f32 array vs Python float: compared in f32 (NEP 50)
        for cls, box, difficult in zip(gc, gb, gd):
            sel = np.logical_and(enabled, pc == cls)
            hit = False
            if sel.any():
                iou = pairwise_iou(box[None, :], pb[sel])[0, :]
                k = iou.argmax()
                hit = iou[k] >= iou_threshold
            if hit:
                if not difficult:
                    tp[cls] += 1
                enabled[np.flatnonzero(sel)[k]] = False
            elif not difficult:
                fn[cls] += 1
        for cls in pc[enabled]:
            fp[cls] += 1
    return tp, fp, fn


def scores_from_counts(tp, fp, fn):
    """ml.py:390-394."""
    supports = tp + fn
    precisions = tp.astype(float) / (tp + fp + 1e-7)
    recalls = tp.astype(float) / (supports + 1e-7)
    fscores = 2 / (1 / (precisions + 1e-7) + 1 / (recalls + 1e-7))
    return precisions, recalls, fscores, supports


def search_conf_threshold(gt_list, pred_list, num_classes, iou_threshold=0.5):
    """ml.py:328-340."""
    grid = np.linspace(0.01, 0.99, 50)
    scores = []
    for th in grid:
        _, _, f, s = scores_from_counts(*match_counts(gt_list, pred_list, th, iou_threshold, num_classes))
        scores.append(np.average(f, weights=s))
    scores = np.array(scores)
    full = scores >= 1
    if full.any():
        return np.mean(grid[full])
    return grid[scores.argmax()]


# --------------------------------------------------------------------------------------
# SURVEY §8 f3 — loss-side consumption of y_true (od_pb.py:365-431, losses.py:22-30).  The reference builds these
# as TensorFlow-1.10 graph ops (tf.keras.losses.*, tf.keras.backend.*; tensorflow is absent here) -> PARITY
# UNPINNED; the Keras formulas restated below are the published ones (epsilon 1e-7; categorical_crossentropy
# renormalises the prediction, clips to [eps, 1-eps]; mean_squared_error averages the last axis).  Elementwise
# arithmetic in float32 like the graph, sums in float64 (the graph's float32 reduction order is unspecified;
# tolerance 1e-5 relative, BASELINE north_star).
_K_EPS = np.float32(1e-7)


def binary_focal_loss(y_true, y_pred, alpha=0.25, gamma=2.0):
    """losses.py:22-30 (gamma=2 -> a square)."""
    one = np.float32(1)
    a_t = y_true * np.float32(alpha) + (one - y_true) * np.float32(1 - alpha)
    p_t = y_true * y_pred + (one - y_true) * (one - y_pred)
    p_t = np.clip(p_t, _K_EPS, one - _K_EPS)
    assert gamma == 2.0
    return -a_t * ((one - p_t) * (one - p_t)) * np.log(p_t)


def od_losses(t: AnchorTables, y_true, y_pred):
    """od_pb.py:365-431: the three loss terms and the three metrics, each (B,) float64.
    y_true / y_pred: (B, N, 3+C+4) float32 in the layout of encode_truth (od_pb.py:301-311)."""
    y_true = np.asarray(y_true, np.float32)
    y_pred = np.asarray(y_pred, np.float32)
    gt_mask, gt_w, gt_obj = y_true[:, :, 0], y_true[:, :, 1], y_true[:, :, 2]
    pred_obj = y_pred[:, :, 2]
    f64 = np.float64
    # loss_obj :403-411
    mask = t.pb_mask[None].astype(np.float32) * gt_mask
    loss_obj = (binary_focal_loss(gt_obj, pred_obj) * mask).astype(np.float32).sum(axis=-1, dtype=f64) * 3
    # loss_clf :413-421
    gt_cls, pred_cls = y_true[:, :, 3:-4], y_pred[:, :, 3:-4]
    norm = pred_cls / pred_cls.sum(axis=-1, keepdims=True, dtype=np.float32)
    norm = np.clip(norm, _K_EPS, np.float32(1) - _K_EPS)
    cce = -(gt_cls * np.log(norm)).sum(axis=-1, dtype=np.float32)
    loss_clf = (cce * gt_w).sum(axis=-1, dtype=f64)
    # loss_loc :423-431
    diff = y_pred[:, :, -4:] - y_true[:, :, -4:]
    mse = (diff * diff).mean(axis=-1, dtype=np.float32)
    loss_loc = (mse * gt_w).sum(axis=-1, dtype=f64) * 2
    # metrics :380-401 (tf round = half to even = np.round)
    acc = (np.round(gt_obj) == np.round(pred_obj)).astype(np.float32)
    gt_bg = (np.float32(1) - gt_obj) * gt_mask
    rec_bg = (acc * gt_bg).sum(axis=-1, dtype=f64) / np.maximum(gt_bg.sum(axis=-1, dtype=f64), 1)
    rec_obj = (acc * gt_obj).sum(axis=-1, dtype=f64) / np.maximum(gt_obj.sum(axis=-1, dtype=f64), 1)
    acc_c = (np.argmax(gt_cls, axis=-1) == np.argmax(pred_cls, axis=-1)).astype(np.float32)
    acc_clf = (acc_c * gt_obj).sum(axis=-1, dtype=f64) / np.maximum(gt_obj.sum(axis=-1, dtype=f64), 1)
    return {'loss_obj': loss_obj, 'loss_clf': loss_clf, 'loss_loc': loss_loc, 'loss': loss_obj + loss_clf + loss_loc,
            'rec_bg': rec_bg, 'rec_obj': rec_obj, 'acc_clf': acc_clf}


# --------------------------------------------------------------------------------------
def od_confusion_matrix(gt_list, pred_list, conf_threshold, iou_threshold, num_classes):
    """ml.od_confusion_matrix (ml.py:397-451).  gt_list = [(classes, bboxes f32, ...)], pred_list = [(classes,
    confs f32, bboxes f32)].  Returns the (C+1, C+1) int matrix: rows = truth (+ "nothing there"), columns =
    prediction (+ "missed").  Every detection is attached to the GT of its largest IoU (:414, first maximum) if
    that IoU reaches the threshold (:415).  NOTE the reference's indexing at :421,
    ``confs.argsort()[::-1][m]``: the boolean mask ``m`` (one flag per detection) is applied to the POSITIONS
    of the sorted index array, so the detections visited for a GT are ``order[j]`` for the positions j whose
    own detection j passes the mask — reproduced as is, it decides the reference's numbers.  Tie policy of this
    build for equal confidences: stable order reversed (the higher index first)."""
    cm = np.zeros((num_classes + 1, num_classes + 1), dtype=int)
    for g, (pc, ps, pb) in zip(gt_list, pred_list):
        gc, gb = np.asarray(g[0]), np.asarray(g[1])
        enabled = ps >= conf_threshold  # :410
        if len(gc) > 0:
            iou = pairwise_iou(gb, pb)  # asserts a non-empty detection set, like ml.compute_iou (:507-508)
            pred_gt = iou.argmax(axis=0)
            near = iou.max(axis=0) >= iou_threshold
            order = stable_desc_order(ps)
            for gt_ix, gt_class in enumerate(gc):
                m = enabled & near & (pred_gt == gt_ix)  # :419-420
                targets = order[m]                       # :421 (positional, see above)
                found = False
                for k in targets:  # :424-431
                    if pc[k] == gt_class:
                        if found:
                            cm[-1, pc[k]] += 1
                        else:
                            found = True
                            cm[gt_class, pc[k]] += 1
                        enabled[k] = False
                for k in targets:  # :433-442
                    if enabled[k] and pc[k] != gt_class:
                        if found:
                            cm[-1, pc[k]] += 1
                        else:
                            found = True
                            cm[gt_class, pc[k]] += 1
                    enabled[k] = False
                if not found:
                    cm[gt_class, -1] += 1  # :443-444
        for c in pc[enabled]:  # :446-447
            cm[-1, c] += 1
    return cm


def exact_match(gt_classes, gt_boxes, pred, conf_threshold=0, iou_threshold=0.5):
    """ObjectsPrediction.is_match (ml.py:159-182); pred = (classes, confs f32, bboxes f32)."""
    pc, ps, pb = pred
    mask = ps >= conf_threshold
    if np.sum(mask) != len(gt_classes):
        return False
    if len(gt_classes) == 0:
        return True
    iou = pairwise_iou(np.asarray(gt_boxes), pb[mask])
    if (iou.max(axis=0) < iou_threshold).any():
        return False
    pred_gt = iou.argmax(axis=0)
    if len(np.unique(pred_gt)) != len(gt_classes):
        return False
    for gt_ix, gt_class in enumerate(gt_classes):
        if gt_class != pc[mask][pred_gt == gt_ix][0]:
            return False
    return True


def od_loss_grad(t: AnchorTables, y_true, y_pred, upstream=None):
    """d(loss_obj + loss_clf + loss_loc)/d y_pred (od_pb.py:365-373), float64, analytic — what autodiff yields for
    the graph of ``od_losses``: clip_by_value passes the gradient only inside [eps, 1-eps].  Checked against
    torch.autograd in tests/test_oracle_cpu.py.  Returns (B, N, 3+C+4) float64."""
    yt = np.asarray(y_true, np.float64)
    yp = np.asarray(y_pred, np.float64)
    eps = float(_K_EPS)
    g = np.zeros_like(yp)
    mask = t.pb_mask[None].astype(np.float64) * yt[:, :, 0]
    w, y, p = yt[:, :, 1], yt[:, :, 2], yp[:, :, 2]
    a_t = y * 0.25 + (1 - y) * 0.75
    p_raw = y * p + (1 - y) * (1 - p)
    inside = (p_raw >= eps) & (p_raw <= 1 - eps)
    p_t = np.clip(p_raw, eps, 1 - eps)
    dfdpt = a_t * (2 * (1 - p_t) * np.log(p_t) - (1 - p_t) ** 2 / p_t)
    g[:, :, 2] = np.where(inside, 3 * mask * (2 * y - 1) * dfdpt, 0.0)
    q, tt = yp[:, :, 3:-4], yt[:, :, 3:-4]
    s = q.sum(axis=-1, keepdims=True)
    with np.errstate(invalid='ignore', divide='ignore'):
        n = q / s
        ok = (n >= eps) & (n <= 1 - eps)
        direct = np.where(ok & (tt != 0), tt / n, 0.0)
        cross = np.where(ok, tt, 0.0).sum(axis=-1, keepdims=True)
        gc = -(w[:, :, None] / s) * (direct - cross)
    g[:, :, 3:-4] = np.where(w[:, :, None] != 0, gc, 0.0)
    g[:, :, -4:] = w[:, :, None] * (yp[:, :, -4:] - yt[:, :, -4:])
    if upstream is not None:
        g *= np.asarray(upstream, np.float64)[:, None, None]
    return g


# ------------------------------------------------------------------------------------------------------------------
# Bounding-box side of the augmentation pipeline: RandomZoom -> RandomFlipLR -> RandomFlipTB -> RandomRotate90
# (image.py:233-388, 433-485; ml.py:102-112; per-sample RandomState, generator.py:172).  Pixel work is not restated.
AUGMENT_DEFAULTS = dict(keep_aspect=False, padding_rate=16, crop_rate=0.1, aspect_prob=0.5, max_aspect_ratio=1.5,
                        min_object_px=8, flip_lr=True, flip_tb=True, rotate90=True, tries=30)


def zoom_try_pad(boxes, in_size, padded, xy, out_size, min_px):
    """One zoom-out attempt (image.py:328-333): boxes pasted at ``xy`` inside a ``padded`` canvas; rejected when any
    box side, in output pixels, drops below ``min_px``.  Integer sizes widen the product to fp64."""
    s2, p2, x2, o2 = (np.concatenate([np.asarray(v, np.int64)] * 2) for v in (in_size, padded, xy, out_size))
    moved = (x2 + boxes * s2) / p2
    px = moved * o2
    return not bool((px[:, 2:] - px[:, :2] < min_px).any()), moved


def zoom_try_crop(boxes, in_size, wh, xy, out_size, min_px):
    """One zoom-in attempt (image.py:350-375).  Returns (accepted, boxes, keep): a box survives when its centre lies
    in the window (low <= c < high, math.py:14-17) and at least 30 % of its area does; no survivor, or any box (kept
    or not) narrower than ``min_px`` output pixels after clipping, rejects the attempt.  Centre and area are taken in
    the dtype the boxes arrive in (fp32 straight from ObjectsAnnotation), the rest in fp64."""
    in_size, wh, xy = (np.asarray(v, np.int64) for v in (in_size, wh, xy))
    window = np.concatenate([xy, xy + wh]) / np.concatenate([in_size, in_size])
    centre = (boxes[:, 2:] + boxes[:, :2]) / 2
    area = (boxes[:, 2:] - boxes[:, :2]).prod(axis=-1)
    keep = np.logical_and(window[:2] <= centre, centre < window[2:]).all(axis=-1)
    if not keep.any():
        return False, None, None
    lo = np.maximum(window[None, :2], boxes[:, :2])
    hi = np.minimum(window[None, 2:], boxes[:, 2:])
    visible = (hi - lo).prod(axis=-1) * (lo < hi).all(axis=-1)
    keep = np.logical_and(keep, visible >= area * 0.3)
    moved = np.clip((boxes * np.concatenate([in_size, in_size]) - np.concatenate([xy, xy])) / np.concatenate([wh, wh]), 0, 1)
    px = moved * np.concatenate([np.asarray(out_size, np.int64)] * 2)
    if (px[:, 2:] - px[:, :2] < min_px).any():
        return False, None, None
    return True, moved, keep


def flip_rot_boxes(boxes, flip_lr, flip_tb, k):
    """RandomFlipLR / RandomFlipTB / ObjectsAnnotation.rot90 on (G,4) boxes, in the dtype they arrive in
    (image.py:445,462; ml.py:102-112)."""
    b = boxes.copy()
    one = b.dtype.type(1)
    if flip_lr:
        b[:, [0, 2]] = one - b[:, [2, 0]]
    if flip_tb:
        b[:, [1, 3]] = one - b[:, [3, 1]]
    x1, y1, x2, y2 = (b[:, i].copy() for i in range(4))
    if k == 1:
        b = np.stack([y1, one - x2, y2, one - x1], axis=1)
    elif k == 2:
        b = np.stack([one - x2, one - y2, one - x1, one - y1], axis=1)
    elif k == 3:
        b = np.stack([one - y2, x1, one - y1, x2], axis=1)
    return b


def augment_boxes_sample(rand, width, height, boxes, classes, output_size, **cfg):
    """The whole box pipeline for one sample, drawing from ``rand`` in the reference's order (image.py:283-319, then
    :441-479).  Returns (boxes, classes, info); boxes are fp64 once a zoom or letterbox touched them, else fp32."""
    c = dict(AUGMENT_DEFAULTS, **cfg)
    out = np.array([output_size, output_size] if np.isscalar(output_size) else output_size, np.int64)
    in_size = np.array([width, height], np.int64)
    boxes = np.asarray(boxes, np.float32).reshape(-1, 4)
    classes = np.asarray(classes)
    base_ar = width / height if c['keep_aspect'] else out[0] / out[1]
    lo_ar, hi_ar = np.log(1 / c['max_aspect_ratio']), np.log(c['max_aspect_ratio'])
    rand_ar = np.exp(rand.uniform(lo_ar, hi_ar)) if rand.rand() <= c['aspect_prob'] else 1
    root = np.sqrt(base_ar * rand_ar)
    ar = np.array([root, 1 / root])
    pad = c['padding_rate'] is not None and (c['crop_rate'] is None or rand.rand() <= 0.5)
    info = {'kind': 'pad' if pad else 'crop', 'chosen': -1}
    for t in range(c['tries']):
        if pad:
            rate = rand.uniform(1, np.sqrt(c['padding_rate']))
            size = np.maximum(np.ceil(in_size * rate * ar).astype(int), in_size)
            slack = size - in_size
        else:
            rate = rand.uniform(np.sqrt(c['crop_rate']), 1)
            size = np.minimum(np.floor(in_size * rate * ar).astype(int), in_size)
            slack = in_size - size
        xy = np.array([rand.randint(0, slack[0] + 1), rand.randint(0, slack[1] + 1)])
        if pad:
            ok, moved = zoom_try_pad(boxes, in_size, size, xy, out, c['min_object_px'])
            keep = None
        else:
            ok, moved, keep = zoom_try_crop(boxes, in_size, size, xy, out, c['min_object_px'])
        if ok:
            boxes = moved if keep is None else moved[keep]
            classes = classes if keep is None else classes[keep]
            info['chosen'] = t
            break
    if info['chosen'] < 0 and c['keep_aspect'] and not (in_size == out).all():
        # every attempt failed: letterbox into the output (image.py:298-314)
        scale = (out / in_size).min()
        resized = np.floor(in_size * scale + 0.1).astype(np.int32)
        origin = (out - resized) // 2
        boxes = (np.concatenate([origin, origin]) + boxes * np.concatenate([resized, resized])) / np.concatenate([out, out])
        info['letterbox'] = (origin, resized)
    lr = c['flip_lr'] and rand.rand() <= 0.5
    tb = c['flip_tb'] and rand.rand() <= 0.5
    k = int(rand.randint(0, 4)) if c['rotate90'] else 0
    info.update(flip_lr=bool(lr), flip_tb=bool(tb), k=k)
    return flip_rot_boxes(boxes, lr, tb, k), classes, info


# --------------------------------------------------------------------------------------
# MS COCO detection metrics: data/coco.py:75-99 -> chainercv.evaluations.eval_detection_coco ->
# pycocotools.cocoeval.COCOeval(iouType='bbox').  Neither package is vendored or pinned by the reference
# (README.md:20 lists chainercv without a version; pycocotools is its optional dependency), so this follows
# the published algorithm (COCOeval.evaluate / evaluateImg / accumulate, chainercv _create_ann / _summarize)
# — PARITY UNPINNED against the real packages.
COCO_IOU_THRS = np.linspace(.5, 0.95, int(np.round((0.95 - .5) / .05)) + 1, endpoint=True)
COCO_REC_THRS = np.linspace(.0, 1.00, int(np.round((1.00 - .0) / .01)) + 1, endpoint=True)
COCO_MAX_DETS = (1, 10, 100)
COCO_AREA_RNG = ((0 ** 2, 1e5 ** 2), (0 ** 2, 32 ** 2), (32 ** 2, 96 ** 2), (96 ** 2, 1e5 ** 2))
COCO_AREA_LBL = ('all', 'small', 'medium', 'large')
COCO_KEYS = (
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


def _coco_box_iou(d, g, crowd):
    """maskApi bbIou on [x, y, w, h] doubles: crowd GT -> union = area of the detection."""
    out = np.zeros((len(d), len(g)))
    for i, a in enumerate(d):
        for j, b in enumerate(g):
            w = min(a[0] + a[2], b[0] + b[2]) - max(a[0], b[0])
            h = min(a[1] + a[3], b[1] + b[3]) - max(a[1], b[1])
            if w <= 0 or h <= 0:
                continue
            inter = w * h
            u = a[2] * a[3] if crowd[j] else a[2] * a[3] + b[2] * b[3] - inter
            out[i, j] = inter / u
    return out


def coco_evaluate_image(dt_xywh, dt_score, gt_xywh, gt_area, gt_crowd, area_rng, max_det=100):
    """COCOeval.computeIoU + evaluateImg for one (image, category, area range).  Returns (scores sorted desc,
    dtMatched (T,D) bool, dtIgnore (T,D) bool, gtIgnore (G,) bool) for the first ``max_det`` detections."""
    order = np.argsort(-np.asarray(dt_score, dtype=np.float64), kind='mergesort')[:max_det]
    dt = np.asarray(dt_xywh, dtype=np.float64).reshape(-1, 4)[order]
    sc = np.asarray(dt_score, dtype=np.float64)[order]
    gt_xywh = np.asarray(gt_xywh, dtype=np.float64).reshape(-1, 4)
    g_ign = np.array([bool(c) or a < area_rng[0] or a > area_rng[1] for a, c in zip(gt_area, gt_crowd)], dtype=bool)
    gtind = np.argsort(g_ign.astype(np.int64), kind='mergesort')
    gt = gt_xywh[gtind]
    crowd = np.asarray(gt_crowd, dtype=bool)[gtind] if len(gtind) else np.zeros(0, dtype=bool)
    g_ign = g_ign[gtind]
    ious = _coco_box_iou(dt, gt, crowd)
    T, D, G = len(COCO_IOU_THRS), len(dt), len(gt)
    gtm = np.zeros((T, G), dtype=bool)
    dtm = np.zeros((T, D), dtype=bool)
    dt_ig = np.zeros((T, D), dtype=bool)
    for ti, t in enumerate(COCO_IOU_THRS):
        for di in range(D):
            iou = min(t, 1 - 1e-10)
            m = -1
            for gi in range(G):
                if gtm[ti, gi] and not crowd[gi]:
                    continue
                if m > -1 and not g_ign[m] and g_ign[gi]:
                    break
                if ious[di, gi] < iou:
                    continue
                iou = ious[di, gi]
                m = gi
            if m == -1:
                continue
            dt_ig[ti, di] = g_ign[m]
            dtm[ti, di] = True
            gtm[ti, m] = True
    d_area = dt[:, 2] * dt[:, 3]
    out_of_range = np.logical_or(d_area < area_rng[0], d_area > area_rng[1])
    dt_ig = np.logical_or(dt_ig, np.logical_and(~dtm, out_of_range[None, :]))
    return sc, dtm, dt_ig, g_ign


def coco_accumulate(per_image, max_det):
    """COCOeval.accumulate for one (category, area range, max_det): ``per_image`` = list of coco_evaluate_image
    results.  Returns (precision (T,R), recall (T,)), all -1 when there is no non-ignored GT."""
    T, R = len(COCO_IOU_THRS), len(COCO_REC_THRS)
    precision, recall = -np.ones((T, R)), -np.ones(T)
    if not per_image:
        return precision, recall
    scores = np.concatenate([e[0][:max_det] for e in per_image])
    inds = np.argsort(-scores, kind='mergesort')
    dtm = np.concatenate([e[1][:, :max_det] for e in per_image], axis=1)[:, inds]
    dt_ig = np.concatenate([e[2][:, :max_det] for e in per_image], axis=1)[:, inds]
    npig = int(np.count_nonzero(~np.concatenate([e[3] for e in per_image])))
    if npig == 0:
        return precision, recall
    tp_sum = np.cumsum(np.logical_and(dtm, ~dt_ig), axis=1).astype(np.float64)
    fp_sum = np.cumsum(np.logical_and(~dtm, ~dt_ig), axis=1).astype(np.float64)
    for t in range(T):
        tp, fp = tp_sum[t], fp_sum[t]
        nd = len(tp)
        rc = tp / npig
        pr = tp / (fp + tp + np.spacing(1))
        q = np.zeros(R)
        recall[t] = rc[-1] if nd else 0
        pr = pr.tolist()
        for i in range(nd - 1, 0, -1):
            if pr[i] > pr[i - 1]:
                pr[i - 1] = pr[i]
        for ri, pi in enumerate(np.searchsorted(rc, COCO_REC_THRS, side='left')):
            if pi >= nd:
                break
            q[ri] = pr[pi]
        precision[t] = q
    return precision, recall


def coco_summarize(precision, recall, existent_labels):
    """chainercv eval_detection_coco's result dict from precision (T,R,K,A,M) / recall (T,K,A,M) over the
    ``existent_labels`` (K of them, sorted)."""
    res = {}
    n = int(np.max(existent_labels)) + 1 if len(existent_labels) else 0
    for key, is_ap, iou_ix, area, max_det in COCO_KEYS:
        a, m = COCO_AREA_LBL.index(area), COCO_MAX_DETS.index(max_det)
        v = (precision if is_ap else recall).copy()
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
    res['existent_labels'] = np.asarray(existent_labels)
    return res


def coco_evaluate(gt_list, pred_list):
    """data/coco.py:75-99.  gt_list = [(width, height, classes, bboxes f32, areas, crowdeds)], pred_list =
    [(classes, confs, bboxes f32)].  Boxes go through real_bboxes / get_real_bboxes (ml.py:62-65,184-186) to int32
    pixels; chainercv reads them as (y_min, x_min, y_max, x_max), which only swaps the axes consistently."""
    labels = sorted(set(int(c) for g in gt_list for c in g[2]) | set(int(c) for p in pred_list for c in p[0]))
    T, R, K, A, M = len(COCO_IOU_THRS), len(COCO_REC_THRS), len(labels), len(COCO_AREA_RNG), len(COCO_MAX_DETS)
    precision, recall = -np.ones((T, R, K, A, M)), -np.ones((T, K, A, M))

    def xywh(px):
        px = np.asarray(px, dtype=np.float64).reshape(-1, 4)
        return np.stack([px[:, 0], px[:, 1], px[:, 2] - px[:, 0], px[:, 3] - px[:, 1]], axis=1)

    gts = [(np.asarray(g[2]).astype(int), xywh(to_pixels(np.asarray(g[3], dtype=np.float32).reshape(-1, 4), g[0], g[1])),
            np.asarray(g[4], dtype=np.float64), np.asarray(g[5], dtype=bool)) for g in gt_list]
    dts = [(np.asarray(p[0]).astype(int), np.asarray(p[1]), xywh(to_pixels(np.asarray(p[2], dtype=np.float32).reshape(-1, 4), g[0], g[1])))
           for p, g in zip(pred_list, gt_list)]
    for k, c in enumerate(labels):
        for a, rng in enumerate(COCO_AREA_RNG):
            per_image = []
            for (gc, gb, ga, gw), (dc, ds, db) in zip(gts, dts):
                gm, dm = gc == c, dc == c
                if not gm.any() and not dm.any():
                    continue  # COCOeval.evaluateImg returns None
                per_image.append(coco_evaluate_image(db[dm], ds[dm], gb[gm], ga[gm], gw[gm], rng))
            for m, max_det in enumerate(COCO_MAX_DETS):
                precision[:, :, k, a, m], recall[:, k, a, m] = coco_accumulate(per_image, max_det)
    res = coco_summarize(precision, recall, labels)
    res['precision'], res['recall'] = precision, recall
    return res
