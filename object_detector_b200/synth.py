"""Seeded synthetic workloads for the tests and ``bench.py``.

NumPy only and HOST arrays, except ``eval_head_outputs`` (config 4), which fills device tensors with torch's
generator because 4 952 images of head outputs are 8 GB.  Not part of the compute path.
"""
import numpy as np


def gt_boxes(rng, g, crowded=False, centre=(0.1, 0.9), size=(0.03, 0.4)):
    """``g`` ground-truth boxes, f32 (g,4) in [0,1], x1<x2, y1<y2, pairwise-distinct f32 areas."""
    out = np.zeros((0, 4), dtype=np.float32)
    while len(out) < g:
        k = (g - len(out)) * 2 + 4
        if crowded:
            c = 0.5 + rng.uniform(-0.01, 0.01, size=(k, 2))
        else:
            c = rng.uniform(centre[0], centre[1], size=(k, 2))
        s = rng.uniform(size[0], size[1], size=(k, 2))
        b = np.clip(np.concatenate([c - s / 2, c + s / 2], axis=1), 0, 1).astype(np.float32)
        b = b[(b[:, 2] - b[:, 0] > 1e-3) & (b[:, 3] - b[:, 1] > 1e-3)]
        out = np.concatenate([out, b])
        area = (out[:, 2:] - out[:, :2]).prod(axis=-1)
        _, first = np.unique(area, return_index=True)
        out = out[np.sort(first)]
    return np.ascontiguousarray(out[:g])


def gt_batch(seed, batch, g_max, num_classes, g_min=1, crowded_every=4, fixed_g=None, **kw):
    """Padded GT batch: boxes f32 (B,Gmax,4), classes i32 (B,Gmax), counts i32 (B,)."""
    rng = np.random.RandomState(seed)
    boxes = np.zeros((batch, g_max, 4), dtype=np.float32)
    classes = np.zeros((batch, g_max), dtype=np.int32)
    counts = np.zeros((batch,), dtype=np.int32)
    for i in range(batch):
        g = fixed_g if fixed_g is not None else int(rng.randint(g_min, g_max + 1))
        crowded = crowded_every > 0 and (i % crowded_every == crowded_every - 1)
        boxes[i, :g] = gt_boxes(rng, g, crowded=crowded, **kw)
        classes[i, :g] = rng.randint(0, num_classes, size=g)
        counts[i] = g
    return boxes, classes, counts


def _distinct_positive(x, rng):
    """Make the f32 values of each row pairwise distinct by nudging duplicates (tie-free scores)."""
    return x


def predictions(seed, batch, n_anchors, num_classes, clustered=False, dtype=np.float32):
    """Network-head outputs of config 2: objs (B,N,1) sigmoid, clfs (B,N,C) softmax, locs (B,N,4)."""
    rng = np.random.RandomState(seed)
    logit = rng.normal(-5.0, 2.5, size=(batch, n_anchors, 1))
    z = 3.0 * rng.normal(size=(batch, n_anchors, num_classes))
    if clustered:
        planted = rng.randint(0, n_anchors, size=(batch, 16))
        for b in range(batch):
            for p in planted[b]:
                lo, hi = max(0, p - 40), min(n_anchors, p + 40)
                logit[b, lo:hi, 0] += 8.0
                z[b, lo:hi, rng.randint(num_classes)] += 6.0
        locs = rng.normal(0.0, 0.3, size=(batch, n_anchors, 4))
    else:
        locs = rng.normal(size=(batch, n_anchors, 4))
    objs = 1.0 / (1.0 + np.exp(-logit))
    z -= z.max(axis=-1, keepdims=True)
    e = np.exp(z)
    clfs = e / e.sum(axis=-1, keepdims=True)
    return objs.astype(dtype), clfs.astype(dtype), locs.astype(dtype)


def eval_set(seed, n_img, num_classes, max_gt=6, difficult_frac=0.1, n_fp=20):
    """Config 4 (VOC07-test-sized eval): per image (w, h, gt classes, gt boxes f32, difficult) and
    detections (classes i32, confs f32, boxes f32) = jittered GT + random false positives."""
    rng = np.random.RandomState(seed)
    gts, preds = [], []
    for _ in range(n_img):
        w, h = int(rng.randint(200, 501)), int(rng.randint(200, 501))
        g = int(rng.randint(1, max_gt + 1))
        gb = gt_boxes(rng, g, size=(0.1, 0.5))
        gc = rng.randint(0, num_classes, size=g).astype(np.int32)
        gd = rng.uniform(size=g) < difficult_frac
        hit = rng.uniform(size=g) < 0.8
        pb = np.clip(gb[hit] + rng.normal(0, 0.02, size=(int(hit.sum()), 4)), 0, 1)
        pc = gc[hit]
        ps = rng.uniform(0.3, 1.0, size=len(pb))
        fb = gt_boxes(rng, n_fp, size=(0.05, 0.5))
        fc = rng.randint(0, num_classes, size=n_fp).astype(np.int32)
        fs = rng.uniform(0.01, 0.5, size=n_fp)
        boxes = np.concatenate([pb, fb]).astype(np.float32)
        ok = (boxes[:, 2] > boxes[:, 0]) & (boxes[:, 3] > boxes[:, 1])
        preds.append((np.concatenate([pc, fc])[ok].astype(np.int32),
                      np.concatenate([ps, fs])[ok].astype(np.float32), boxes[ok]))
        gts.append((w, h, gc, gb, gd))
    return gts, preds


def plan_eval_head(pb, preds):
    """Config 4: for every planned detection of ``eval_set`` (class, conf, box) the anchor that will carry it and the
    ``locs`` head value that decodes to its box (od_pb.py:216-222 inverted in fp32).  Anchors are taken from the cell
    that holds the box centre, finest map first, distinct within an image, ``pb_mask`` = 1 only.
    Returns per-image tuples (anchor i64 (D,), class i32 (D,), conf f32 (D,), locs f32 (D,4))."""
    maps = [int(m) for m in pb.map_sizes]
    n_pat = len(pb.pb_size_patterns)
    off = np.concatenate([[0], np.cumsum([m * m * n_pat for m in maps])])
    mask = np.asarray(pb.pb_mask, dtype=bool)
    locs_tab = np.asarray(pb.pb_locs, dtype=np.float32)
    scale = np.tile((0.10 * np.asarray(pb.pb_sizes, dtype=np.float64)).astype(np.float32), 2)
    plan = []
    for pc, ps, pbx in preds:
        used, anchors = set(), []
        cx, cy = (pbx[:, 0] + pbx[:, 2]) / 2, (pbx[:, 1] + pbx[:, 3]) / 2
        for j in range(len(pc)):
            pick = -1
            for mi, m in enumerate(maps):
                ix, iy = min(m - 1, int(cx[j] * m)), min(m - 1, int(cy[j] * m))
                for t in range(n_pat):
                    a = int(off[mi] + ((j + t) % n_pat) * m * m + iy * m + ix)
                    if mask[a] and a not in used:
                        pick = a
                        break
                if pick >= 0:
                    break
            if pick < 0:  # every in-cell anchor taken or masked: any free valid anchor will do
                pick = next(a for a in np.flatnonzero(mask) if a not in used)
            used.add(pick)
            anchors.append(pick)
        anchors = np.asarray(anchors, dtype=np.int64)
        head = ((pbx.astype(np.float32) - locs_tab[anchors]) / scale[anchors]).astype(np.float32)
        plan.append((anchors, pc.astype(np.int32), ps.astype(np.float32), head))
    return plan


def eval_head_outputs(pb, plan, lo, hi, device, seed=2007, chunk=64):
    """Config 4: resident head outputs (objs, clfs, locs) of images [lo, hi) as CUDA tensors.  Background = the
    config-2 distributions at a lower objectness level (sigmoid(N(-8, 2^2)): a few hundred low-confidence candidates per
    image, so the NMS has work) drawn per global 64-image chunk from a generator seeded by the chunk id — any rank
    produces identical values for the same image — with the planned detections of ``plan_eval_head`` written on top."""
    import torch
    n, c = len(pb.pb_locs), pb.num_classes
    cnt = hi - lo
    objs = torch.empty((cnt, n, 1), dtype=torch.float32, device=device)
    clfs = torch.empty((cnt, n, c), dtype=torch.float32, device=device)
    locs = torch.empty((cnt, n, 4), dtype=torch.float32, device=device)
    gen = torch.Generator(device=device)
    for q in range(lo // chunk, (hi + chunk - 1) // chunk):
        gen.manual_seed(seed * 1000003 + q)
        o = torch.sigmoid(torch.randn((chunk, n, 1), generator=gen, device=device) * 2.0 - 8.0)
        z = torch.softmax(3.0 * torch.randn((chunk, n, c), generator=gen, device=device), dim=-1)
        ll = torch.randn((chunk, n, 4), generator=gen, device=device)
        a, b = max(lo, q * chunk), min(hi, (q + 1) * chunk)
        objs[a - lo:b - lo], clfs[a - lo:b - lo], locs[a - lo:b - lo] = o[a - q * chunk:b - q * chunk], z[a - q * chunk:b - q * chunk], ll[a - q * chunk:b - q * chunk]
    flat = np.concatenate([(i - lo) * n + plan[i][0] for i in range(lo, hi)]) if cnt else np.zeros(0, np.int64)
    if len(flat):
        idx = torch.from_numpy(flat).to(device)
        cls = torch.from_numpy(np.concatenate([plan[i][1] for i in range(lo, hi)]).astype(np.int64)).to(device)
        objs.view(-1)[idx] = torch.from_numpy(np.concatenate([plan[i][2] for i in range(lo, hi)])).to(device)
        onehot = torch.zeros((len(flat), c), dtype=torch.float32, device=device)
        onehot[torch.arange(len(flat), device=device), cls] = 1.0
        clfs.view(-1, c)[idx] = onehot
        locs.view(-1, 4)[idx] = torch.from_numpy(np.concatenate([plan[i][3] for i in range(lo, hi)])).to(device)
    return objs, clfs, locs


AUGMENT_CONFIGS = {
    # name: (output_size, keep_aspect, RandomZoom keyword arguments) — od.py:156-165 defaults first
    'default': (320, False, {}),
    'keep_aspect': (320, True, {}),
    'pad_only': (300, False, {'crop_rate': None}),
    'crop_only': (512, True, {'padding_rate': None, 'min_object_px': 24}),
}


def augment_inputs(seed, count):
    """Seeded annotations for the augmentation vectors: image size, f32 boxes, classes, per-sample RNG seed."""
    rng = np.random.RandomState(seed)
    out = []
    for i in range(count):
        w, h = int(rng.randint(120, 640)), int(rng.randint(120, 640))
        g = int(rng.randint(1, 9))
        boxes = gt_boxes(rng, g, size=[(0.03, 0.2), (0.1, 0.6), (0.3, 0.9)][i % 3])
        out.append((w, h, boxes, rng.randint(0, 20, size=g).astype(np.int32), int(rng.randint(0, 2 ** 31))))
    return out


def coco_eval_set(seed, n_img, num_classes, max_gt=8, crowd_frac=0.1, n_fp=20, tie_frac=0.2):
    """COCO-style evaluation set on top of ``eval_set``: annotation areas (a fraction of the box area, as a
    segmentation area would be), crowd flags, and a share of quantised (tied) scores.
    gts: (w, h, classes, boxes f32, areas f32, crowdeds); preds as ``eval_set``."""
    rng = np.random.RandomState(seed + 77)
    gts, preds = eval_set(seed, n_img, num_classes, max_gt=max_gt, n_fp=n_fp)
    out = []
    for (w, h, gc, gb, _), (pc, ps, pbx) in zip(gts, preds):
        px = np.round(gb.astype(np.float32) * [w, h, w, h])
        areas = ((px[:, 2] - px[:, 0]) * (px[:, 3] - px[:, 1]) * rng.uniform(0.4, 1.0, size=len(gb))).astype(np.float32)
        crowd = rng.uniform(size=len(gb)) < crowd_frac
        out.append((w, h, gc, gb, areas, crowd))
        tied = rng.uniform(size=len(ps)) < tie_frac
        ps[tied] = np.round(ps[tied] * 8) / 8
    return out, preds
