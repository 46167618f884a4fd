"""Generates tests/golden/*.npz by running the UNMODIFIED reference (via oracle/ref_loader.py).

Run in the build container only (needs /root/reference):  python oracle/gen_golden.py
The fixtures are committed; the GPU box never runs this script.
"""
import hashlib
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from object_detector_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402
from oracle.restated import VOC_PATTERNS  # noqa: E402

OUT = ROOT / 'tests' / 'golden'
tk = ref_loader.load()
PriorBoxes = tk.dl.od_pb.PriorBoxes
Ann = tk.ml.ObjectsAnnotation


def sha(a):
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.tobytes()).hexdigest()


def ref_pb(input_size, map_sizes, num_classes, patterns=VOC_PATTERNS):
    pb = PriorBoxes(input_size, map_sizes, num_classes)
    pb.from_dict({'map_sizes': map_sizes, 'pb_size_patterns': patterns})
    return pb


ANCHOR_SETS = {
    'voc320': ((320, 320), [40, 20, 10], 20),
    'voc640': ((640, 640), [40, 20, 10], 20),
    'tiny128': ((128, 128), [8], 2),
    'stress1120': ((1120, 1120), [140, 70, 35], 80),
}
TABLES = ['pb_grid', 'pb_locs', 'pb_centers', 'pb_sizes', 'pb_mask', 'pb_info_indices']


def gen_anchors():
    d = {}
    for name, (inp, maps, c) in ANCHOR_SETS.items():
        pb = ref_pb(inp, maps, c)
        for k in TABLES:
            a = getattr(pb, k)
            d[f'{name}/{k}/sha256'] = np.array(sha(a))
            d[f'{name}/{k}/dtype'] = np.array(str(a.dtype))
            d[f'{name}/{k}/shape'] = np.array(a.shape)
            d[f'{name}/{k}/head'] = a[:5]
            d[f'{name}/{k}/tail'] = a[-5:]
        d[f'{name}/valid'] = np.array(int(pb.pb_mask.sum()))
    pb = ref_pb(*ANCHOR_SETS['tiny128'])
    for k in TABLES:
        d[f'tiny128/{k}/full'] = getattr(pb, k)
    np.savez_compressed(OUT / 'anchors.npz', **d)


def assign_case(pb, boxes, classes):
    ann = Ann('.', 300, 300, classes, boxes)
    idx, gts, valid = pb._assign_boxes(ann.bboxes)
    y = pb.encode_truth([ann])[0]
    pos = np.flatnonzero(y[:, 2] != 0)
    assert np.array_equal(pos, idx)
    return {
        'boxes': ann.bboxes, 'classes': ann.classes,
        'pb_indices': idx, 'gt_indices': gts, 'pb_valids': np.packbits(valid),
        'y_sha256': np.array(sha(y)), 'y_pos_rows': y[pos],
        'y_ch0_value': np.array(y[valid, 0].max() if valid.any() else 0, dtype=np.float32),
    }


def gen_assign():
    d = {}

    def put(tag, anchors, case):
        d[f'{tag}/anchors'] = np.array(anchors)
        for k, v in case.items():
            d[f'{tag}/{k}'] = v

    pb640 = ref_pb(*ANCHOR_SETS['voc640'])
    pb320 = ref_pb(*ANCHOR_SETS['voc320'])
    tiny = ref_pb(*ANCHOR_SETS['tiny128'])
    rng = np.random.RandomState(11)
    for g in (0, 1, 5, 64):
        b = synth.gt_boxes(rng, g) if g else np.zeros((0, 4), np.float32)
        put(f'voc640_g{g}', 'voc640', assign_case(pb640, b, rng.randint(0, 20, size=g)))
    b = synth.gt_boxes(rng, 64, crowded=True)
    put('voc640_crowded64', 'voc640', assign_case(pb640, b, rng.randint(0, 20, size=64)))
    # more GTs than one cell neighbourhood has valid anchors -> forces the IoU-argmax fallback
    b = synth.gt_boxes(rng, 120, crowded=True, size=(0.02, 0.6))
    put('voc640_crowded120', 'voc640', assign_case(pb640, b, rng.randint(0, 20, size=120)))
    # centres exactly on cell borders of the 40/20/10 grids
    ctr = np.array([[0.5, 0.5], [0.25, 0.75], [0.1, 0.1], [0.9, 0.5], [0.525, 0.475], [0.3, 0.6]])
    half = np.array([[0.05, 0.1], [0.11, 0.07], [0.06, 0.04], [0.08, 0.2], [0.2, 0.21], [0.02, 0.03]])
    b = np.concatenate([ctr - half, ctr + half], axis=1).astype(np.float32)
    put('voc640_border', 'voc640', assign_case(pb640, b, np.arange(6) % 20))
    for s in range(3):
        g = int(rng.randint(1, 8))
        put(f'voc320_cfg1_{s}', 'voc320', assign_case(pb320, synth.gt_boxes(rng, g), rng.randint(0, 20, size=g)))
    # the reference's own XML fixtures (pytoolkit/tests/data/od/Annotations, data_voc_test.py:21)
    xml = np.array([[203, 255, 601, 355]], dtype=np.float64) / [768, 614, 768, 614]
    put('tiny128_xml', 'tiny128', assign_case(tiny, xml.astype(np.float32), np.array([1])))
    b = synth.gt_boxes(rng, 12, centre=(0.2, 0.8))
    put('tiny128_g12', 'tiny128', assign_case(tiny, b, rng.randint(0, 2, size=12)))
    y = tiny.encode_truth([Ann('.', 300, 300, d['tiny128_g12/classes'], d['tiny128_g12/boxes'])])
    d['tiny128_g12/y_full'] = y[0]
    np.savez_compressed(OUT / 'assign.npz', **d)

    # stress shape: one image, G=512 (≈25 s in the reference)
    stress = ref_pb(*ANCHOR_SETS['stress1120'])
    rng = np.random.RandomState(5)
    b = synth.gt_boxes(rng, 512, centre=(0.05, 0.95), size=(0.01, 0.3))
    case = assign_case(stress, b, rng.randint(0, 80, size=512))
    np.savez_compressed(OUT / 'assign_stress.npz', **{f'stress_g512/{k}': v for k, v in case.items()})


def gen_iou_nms_decode():
    d = {}
    # ml_test.py:42-53 verbatim inputs, outputs from the reference
    a = np.array([[0, 0, 200, 200], [1000, 1000, 1001, 1001]])
    b = np.array([[100, 100, 300, 300]])
    d['iou_kat/a'], d['iou_kat/b'], d['iou_kat/out'] = a, b, tk.ml.compute_iou(a, b)
    rng = np.random.RandomState(3)
    fa, fb = synth.gt_boxes(rng, 7), synth.gt_boxes(rng, 300)
    d['iou_f32/a'], d['iou_f32/b'], d['iou_f32/out'] = fa, fb, tk.ml.compute_iou(fa, fb)
    # ml_test.py:89-113 verbatim
    kb = np.array([[200, 0, 200, 200], [0, 201, 200, 200], [0, 0, 200, 200], [0, 0, 201, 201],
                   [0, 0, 202, 202], [1000, 1000, 1001, 1001]])
    ks = np.array([0.9, 0.9, 0.7, 0.8, 0.6, 0.85])
    d['nms_kat/boxes'], d['nms_kat/scores'] = kb, ks
    d['nms_kat/out'] = tk.ml.non_maximum_suppression(kb, ks, 200, 0.45)
    for i, (n, spread, topk) in enumerate([(50, 0.3, 200), (400, 0.15, 200), (400, 0.15, 7),
                                           (1500, 0.1, 200), (300, 0.02, 1), (64, 0.05, 2)]):
        c = 0.5 + rng.uniform(-spread, spread, size=(n, 2))
        s = rng.uniform(0.02, 0.3, size=(n, 2))
        bx = np.clip(np.concatenate([c - s / 2, c + s / 2], 1), 0, 1).astype(np.float32)
        bx[::17, 2] = bx[::17, 0]  # zero-area boxes (skipped by the reference)
        sc = rng.permutation(n).astype(np.float32) / n * 0.9 + 0.05  # pairwise distinct
        d[f'nms_f32_{i}/boxes'], d[f'nms_f32_{i}/scores'] = bx, sc
        d[f'nms_f32_{i}/top_k'] = np.array(topk)
        d[f'nms_f32_{i}/out'] = tk.ml.non_maximum_suppression(bx, sc, topk, 0.45)
    # decode_locs(xp=np): f64 path of check_assign.py:27
    tiny = ref_pb(*ANCHOR_SETS['tiny128'])
    pred = rng.normal(size=(2, len(tiny.pb_locs), 4)).astype(np.float32)
    d['decode_tiny/pred'], d['decode_tiny/out_f64'] = pred, tiny.decode_locs(pred, xp=np)
    pb640 = ref_pb(*ANCHOR_SETS['voc640'])
    pred = np.random.RandomState(99).normal(size=(1, len(pb640.pb_locs), 4)).astype(np.float32)
    out = pb640.decode_locs(pred, xp=np)
    d['decode_voc640/seed'] = np.array(99)
    d['decode_voc640/out_f64_sha256'] = np.array(sha(out))
    d['decode_voc640/out_f64_rows'] = out[0, ::997]
    # encode_locs for a few anchors
    box = np.array([0.2, 0.3, 0.55, 0.8], dtype=np.float32)
    ix = np.array([0, 5, 1700, 9000, 16799])
    d['encode_locs/box'], d['encode_locs/ix'], d['encode_locs/out'] = box, ix, pb640.encode_locs(box, ix)
    np.savez_compressed(OUT / 'iou_nms_decode.npz', **d)


def gen_scores():
    """ml.compute_scores / search_conf_threshold (ml.py:328-394) of the unmodified reference on a seeded set."""
    d = {}
    gts, preds = synth.eval_set(77, 60, 5, max_gt=5, n_fp=6)
    y_true = [Ann('.', w, h, c, b, dif) for (w, h, c, b, dif) in gts]
    y_pred = [tk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in preds]
    d['seed'] = np.array([77, 60, 5, 5, 6])
    for i, th in enumerate([0.0, 0.25, 0.6]):
        p, r, f, s = tk.ml.compute_scores(y_true, y_pred, th, 0.5, num_classes=5)
        d[f'scores_{i}/conf_threshold'] = np.array(th)
        d[f'scores_{i}/precisions'], d[f'scores_{i}/recalls'], d[f'scores_{i}/fscores'], d[f'scores_{i}/supports'] = p, r, f, s
    d['search_conf_threshold'] = np.array(tk.ml.search_conf_threshold(y_true, y_pred, 0.5))
    # ml.od_confusion_matrix (ml.py:397-451); the eval set has several detections per GT and class confusions
    for i, th in enumerate([0.0, 0.25, 0.6]):
        d[f'cm_{i}/conf_threshold'] = np.array(th)
        d[f'cm_{i}/out'] = tk.ml.od_confusion_matrix(y_true, y_pred, th, 0.5, num_classes=5)
    gts2, preds2 = synth.eval_set(78, 40, 3, max_gt=6, n_fp=12)
    y_true2 = [Ann('.', w, h, c, b, dif) for (w, h, c, b, dif) in gts2]
    y_pred2 = [tk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in preds2]
    d['cm_dense/seed'] = np.array([78, 40, 3, 6, 12])
    d['cm_dense/out'] = tk.ml.od_confusion_matrix(y_true2, y_pred2, 0.1, 0.3, num_classes=3)
    np.savez_compressed(OUT / 'scores.npz', **d)


def gen_augment():
    """Bounding-box side of RandomZoom -> RandomFlipLR -> RandomFlipTB -> RandomRotate90 (image.py:233-485,
    ml.py:102-112) of the unmodified reference, per-sample RandomState(seed) as generator.py:172 makes it."""
    import types
    tk = ref_loader.load_augment()
    ctx = types.SimpleNamespace(do_augmentation=lambda rand, probability=1: probability >= 1 or rand.rand() <= probability)
    d = {}
    for name, (size, keep_aspect, kw) in synth.AUGMENT_CONFIGS.items():
        ops = [tk.image.RandomZoom(probability=1, output_size=(size, size), keep_aspect=keep_aspect, **kw),
               tk.image.RandomFlipLR(probability=0.5), tk.image.RandomFlipTB(probability=0.5),
               tk.image.RandomRotate90(probability=1)]
        boxes, classes, counts, nxt = [], [], [], []
        for w, h, b, c, seed in synth.augment_inputs(911, 160):
            ann = tk.ml.ObjectsAnnotation('.', w, h, c, b)
            rand = np.random.RandomState(seed)
            x, y = np.zeros((h, w, 3), np.float32), ann
            for op in ops:
                x, y, _ = op.execute(x, y, None, rand, ctx)
            assert x.shape[:2] == (size, size)
            boxes.append(np.asarray(y.bboxes, dtype=np.float64).reshape(-1, 4))
            classes.append(np.asarray(y.classes, dtype=np.int32))
            counts.append(len(y.classes))
            nxt.append(rand.rand())
        d[name + '_boxes'] = np.concatenate(boxes)
        d[name + '_classes'] = np.concatenate(classes)
        d[name + '_counts'] = np.asarray(counts, np.int32)
        d[name + '_next_draw'] = np.asarray(nxt, np.float64)
    np.savez_compressed(OUT / 'augment.npz', **d)
    print('augment.npz', {k: v.shape for k, v in d.items() if k.endswith('_counts')},
          'dropped', {n: int(sum(len(i[3]) for i in synth.augment_inputs(911, 160)) - d[n + '_counts'].sum()) for n in synth.AUGMENT_CONFIGS})


if __name__ == '__main__':
    OUT.mkdir(parents=True, exist_ok=True)
    gen_anchors()
    gen_iou_nms_decode()
    gen_assign()
    gen_scores()
    gen_augment()
    for p in sorted(OUT.glob('*.npz')):
        print(p.name, p.stat().st_size)
