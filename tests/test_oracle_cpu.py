"""The CPU oracle (oracle/restated.py) against the golden vectors generated from the unmodified
reference (oracle/gen_golden.py) and the reference's own known-answer tests."""
import hashlib

import numpy as np
import pytest

from object_detector_b200 import synth
from oracle import ref_loader, restated as R

TABLES = ['pb_grid', 'pb_locs', 'pb_centers', 'pb_sizes', 'pb_mask', 'pb_info_indices']


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize('name', ['voc320', 'voc640', 'tiny128', 'stress1120'])
def test_anchor_tables_match_reference(golden, oracle_tables, name):
    g, t = golden['anchors'], oracle_tables(name)
    for k in TABLES:
        a = getattr(t, k)
        assert str(a.dtype) == str(g[f'{name}/{k}/dtype'])
        assert tuple(a.shape) == tuple(g[f'{name}/{k}/shape'])
        assert sha(a) == str(g[f'{name}/{k}/sha256']), (name, k)
    assert int(t.pb_mask.sum()) == int(g[f'{name}/valid'])


def test_anchor_counts_from_survey(oracle_tables):
    # N = 16 800, valid 12 546 @320 / 12 512 @640; stress N = 205 800, valid 189 682
    assert (len(oracle_tables('voc320')), int(oracle_tables('voc320').pb_mask.sum())) == (16800, 12546)
    assert (len(oracle_tables('voc640')), int(oracle_tables('voc640').pb_mask.sum())) == (16800, 12512)
    assert (len(oracle_tables('stress1120')), int(oracle_tables('stress1120').pb_mask.sum())) == (205800, 189682)
    assert 0 < oracle_tables("voc320").pb_grid[0, 0] < 2e-10  # 1.8626e-10, not 0


def _assign_cases(g):
    return sorted({k.split('/')[0] for k in g.files})


def test_assign_and_encode_match_reference(golden, oracle_tables):
    g = golden['assign']
    for case in _assign_cases(g):
        t = oracle_tables(str(g[f'{case}/anchors']))
        boxes, classes = g[f'{case}/boxes'], g[f'{case}/classes']
        idx, gts, valid = R.assign_boxes(t, boxes)
        assert np.array_equal(idx, g[f'{case}/pb_indices']), case
        assert np.array_equal(gts, g[f'{case}/gt_indices']), case
        assert np.array_equal(np.packbits(valid), g[f'{case}/pb_valids']), case
        y = R.encode_truth(t, [(classes, boxes)])[0]
        assert sha(y) == str(g[f'{case}/y_sha256']), case
        assert np.array_equal(y[idx], g[f'{case}/y_pos_rows']), case
    assert np.array_equal(R.encode_truth(oracle_tables('tiny128'), [(g['tiny128_g12/classes'], g['tiny128_g12/boxes'])])[0],
                          g['tiny128_g12/y_full'])


def test_fallback_branch_is_exercised(golden, oracle_tables):
    """crowded120 has more GTs than its cell neighbourhood has anchors -> the IoU arg-max fallback
    (od_pb.py:347-352) must have produced some assignments outside the centre cells."""
    g, t = golden['assign'], oracle_tables('voc640')
    boxes = g['voc640_crowded120/boxes']
    idx, gts = g['voc640_crowded120/pb_indices'], g['voc640_crowded120/gt_indices']
    ctr = R.box_centers(boxes)
    outside = 0
    for a, gt in zip(idx, gts):
        inside = R.half_open_interval(ctr[gt], t.pb_grid[a, :2], t.pb_grid[a, 2:]).all()
        outside += not inside
    assert outside > 0


def test_iou_kat_and_golden(golden):
    g = golden['iou_nms_decode']
    out = R.pairwise_iou(g['iou_kat/a'], g['iou_kat/b'])
    assert out.shape == (2, 1)
    assert out[0, 0] == pytest.approx(100 * 100 / (200 * 200 * 2 - 100 * 100))  # ml_test.py:52
    assert out[1, 0] == 0
    assert np.array_equal(out, g['iou_kat/out'])
    out = R.pairwise_iou(g['iou_f32/a'], g['iou_f32/b'])
    assert out.dtype == np.float32 and np.array_equal(out, g['iou_f32/out'])


def test_nms_kat_and_golden(golden):
    g = golden['iou_nms_decode']
    # ml_test.py:89-113
    assert np.array_equal(R.numpy_nms(g['nms_kat/boxes'], g['nms_kat/scores'], 200, 0.45), [5, 3])
    assert np.array_equal(R.numpy_nms(np.array([[0, 0, 200, 200]]), np.array([0.2]), 200, 0.45), [0])
    assert R.numpy_nms(np.array([[200, 0, 200, 200]]), np.array([0.1]), 200, 0.45).shape == (0,)
    for i in range(6):
        b, s, k = g[f'nms_f32_{i}/boxes'], g[f'nms_f32_{i}/scores'], int(g[f'nms_f32_{i}/top_k'])
        ref = g[f'nms_f32_{i}/out']
        assert np.array_equal(R.numpy_nms(b, s, k, 0.45), ref)
        assert np.array_equal(R.numpy_nms_policy(b, s, k, 0.45), ref)  # distinct scores: policy == reference
    # the top_k quirk (SURVEY A-7): top_k=1 can return 2
    assert len(g['nms_f32_4/out']) == 2


def test_decode_and_encode_locs_golden(golden, oracle_tables):
    g = golden['iou_nms_decode']
    tiny, voc = oracle_tables('tiny128'), oracle_tables('voc640')
    out = R.decode_locs_f64(tiny, g['decode_tiny/pred'])
    assert out.dtype == np.float64 and np.array_equal(out, g['decode_tiny/out_f64'])
    np.testing.assert_allclose(R.decode_locs_f32(tiny, g['decode_tiny/pred']), out, rtol=0, atol=2e-7)
    pred = np.random.RandomState(int(g['decode_voc640/seed'])).normal(size=(1, len(voc), 4)).astype(np.float32)
    assert sha(R.decode_locs_f64(voc, pred)) == str(g['decode_voc640/out_f64_sha256'])
    assert np.array_equal(R.encode_locs(voc, g['encode_locs/box'], g['encode_locs/ix']), g['encode_locs/out'])


def test_nms_layer_structure():
    """layers.py:925-970 restated: per-class NMS, class-order concat, global top-k, zero pad."""
    boxes = np.array([[0, 0, .5, .5], [0, 0, .5, .52], [.5, .5, 1, 1], [0, 0, .5, .5], [.1, .1, .2, .2]], np.float32)
    classes = np.array([1, 1, 0, 0, 1])
    confs = np.array([.9, .8, .7, .6, .005], np.float32)
    out, idx = R.nms_layer_image(classes, confs, boxes, 2, top_k=4)
    assert idx.tolist() == [0, 2, 3, -1]  # box 1 suppressed by 0 (same class); 3 kept (other class); 4 below 0.01
    assert out[:3, 0].tolist() == [1, 0, 0] and np.allclose(out[:3, 1], [.9, .7, .6]) and (out[3] == 0).all()


def test_voc_ap_micro_cases():
    """Hand-computed cases for the chainercv restatement (PARITY UNPINNED: chainercv absent)."""
    gt = [(100, 100, np.array([0]), np.array([[.1, .1, .5, .5]], np.float32), np.array([False]))]
    perfect = [(np.array([0]), np.array([.9], np.float32), np.array([[.1, .1, .5, .5]], np.float32))]
    r = R.voc_evaluate(gt, perfect, 2)
    assert r['AP'][0] == 1.0 and r['AP_VOC'][0] == pytest.approx(1.0) and np.isnan(r['AP'][1]) and r['mAP'] == 1.0
    miss = [(np.array([0]), np.array([.9], np.float32), np.array([[.6, .6, .9, .9]], np.float32))]
    r = R.voc_evaluate(gt, miss, 2)
    assert r['AP'][0] == 0.0 and r['AP_VOC'][0] == 0.0
    dup = [(np.array([0, 0]), np.array([.9, .8], np.float32), np.array([[.1, .1, .5, .5], [.1, .1, .5, .5]], np.float32))]
    r = R.voc_evaluate(gt, dup, 2)  # TP then FP: prec [1, .5], rec [1, 1] -> AP 1 (both metrics)
    assert r['AP'][0] == 1.0 and r['AP_VOC'][0] == pytest.approx(1.0)
    fp_first = [(np.array([0, 0]), np.array([.9, .8], np.float32), np.array([[.6, .6, .9, .9], [.1, .1, .5, .5]], np.float32))]
    r = R.voc_evaluate(gt, fp_first, 2)  # FP then TP: prec [0, .5], rec [0, 1] -> AP .5
    assert r['AP'][0] == pytest.approx(0.5) and r['AP_VOC'][0] == pytest.approx(0.5)
    gt_d = [(100, 100, np.array([0, 0]), np.array([[.1, .1, .5, .5], [.6, .6, .9, .9]], np.float32), np.array([True, False]))]
    r = R.voc_evaluate(gt_d, perfect, 2)  # only match is the difficult GT -> ignored; one non-difficult GT missed
    assert r['AP'][0] == 0.0


def test_pixel_rounding_half_even():
    assert R.to_pixels(np.array([[0.5, 0.25, 0.75, 1.0]], np.float32), 5, 10).tolist() == [[2, 2, 4, 10]]


@pytest.mark.skipif(not ref_loader.available(), reason='reference tree not mounted (GPU box)')
def test_restatement_against_live_reference(oracle_tables):
    """Only in the build container: random GT sets through the real reference vs the restatement."""
    from object_detector_b200 import synth
    tk = ref_loader.load()
    pb = tk.dl.od_pb.PriorBoxes((320, 320), [40, 20, 10], 20)
    pb.from_dict({'map_sizes': [40, 20, 10], 'pb_size_patterns': R.VOC_PATTERNS})
    t = oracle_tables('voc320')
    boxes, classes, counts = synth.gt_batch(21, 6, 24, 20)
    for i in range(6):
        g = counts[i]
        for x, y in zip(pb._assign_boxes(boxes[i, :g]), R.assign_boxes(t, boxes[i, :g])):
            assert np.array_equal(x, y)
    rng = np.random.RandomState(2)
    b = synth.gt_boxes(rng, 300, size=(0.05, 0.3))
    s = rng.permutation(300).astype(np.float32) / 300
    assert np.array_equal(tk.ml.non_maximum_suppression(b, s, 50, 0.45), R.numpy_nms_policy(b, s, 50, 0.45))


def _score_sets():
    from object_detector_b200 import synth
    gts, preds = synth.eval_set(77, 60, 5, max_gt=5, n_fp=6)
    return [(c, b, d) for (_, _, c, b, d) in gts], preds


def test_compute_scores_golden(golden):
    """ml.compute_scores / search_conf_threshold (ml.py:328-394) vs outputs of the unmodified reference."""
    g = golden['scores']
    gt, pred = _score_sets()
    for i in range(3):
        th = float(g[f'scores_{i}/conf_threshold'])
        p, r, f, s = R.scores_from_counts(*R.match_counts(gt, pred, th, 0.5, 5))
        assert np.array_equal(s, g[f'scores_{i}/supports'])
        assert np.array_equal(p, g[f'scores_{i}/precisions']) and np.array_equal(r, g[f'scores_{i}/recalls'])
        assert np.array_equal(f, g[f'scores_{i}/fscores'])
    assert R.search_conf_threshold(gt, pred, 5) == float(g['search_conf_threshold'])


# ----------------------------------------------------------------------------------------- strict_nms (layers.py:940,953-957)
def _strict_two_pass(classes, confs, boxes, num_classes, top_k, thr, thr_all):
    """Independent statement of the strict branch: per-class NumPy NMS with limit 2*top_k, concat, then one
    more NumPy NMS over everything (scores distinct, so the reference's own argsort order is unambiguous)."""
    cat = []
    for c in range(num_classes):
        sel = np.flatnonzero((classes == c) & (confs > np.float32(0.01)))
        if len(sel):
            cat.append(sel[R.numpy_nms(boxes[sel], confs[sel], min(2 * top_k, len(sel)), thr)])
    if not cat:
        return np.zeros(0, np.int64)
    cat = np.concatenate(cat)
    return cat[R.numpy_nms(boxes[cat], confs[cat], min(top_k, len(cat)), thr_all)]


def test_strict_layer_equals_two_independent_passes():
    rng = np.random.RandomState(5)
    for trial in range(6):
        n, c = 400, 5
        ctr = rng.uniform(0.2, 0.8, size=(12, 2))[rng.randint(0, 12, size=n)] + rng.normal(0, 0.03, size=(n, 2))
        wh = rng.uniform(0.1, 0.3, size=(n, 2))
        boxes = np.clip(np.concatenate([ctr - wh / 2, ctr + wh / 2], axis=1), 0, 1).astype(np.float32)
        confs = rng.permutation(n).astype(np.float32) / np.float32(n) * np.float32(0.9) + np.float32(0.02)
        classes = rng.randint(0, c, size=n)
        top_k = [200, 10, 3][trial % 3]
        out, kept = R.nms_layer_image(classes, confs, boxes, c, top_k=top_k, nms_all_threshold=0.5)
        ref = _strict_two_pass(classes, confs, boxes, c, top_k, 0.45, 0.5)
        k = int((kept >= 0).sum())
        assert np.array_equal(kept[:k], ref)
        assert np.array_equal(out[:k, 1], confs[ref]) and np.array_equal(out[:k, 0], classes[ref].astype(np.float32))
        assert (out[k:] == 0).all()
        # the second pass can only remove boxes relative to the doubled-limit class stage
        loose, lk = R.nms_layer_image(classes, confs, boxes, c, top_k=2 * top_k * c)
        assert set(kept[:k]) <= set(lk[lk >= 0])


def test_strict_class_limit_cuts_before_all_class_pass(strict_cap_case):
    """top_k=2 -> per-class limit 4 (layers.py:940): four mutually compatible class-0 boxes around a stronger
    class-1 box use the limit up and are then all removed by the class-agnostic pass; the fifth class-0 box
    was already cut by the limit, so only the class-1 box survives."""
    boxes, classes, confs = strict_cap_case
    out, kept = R.nms_layer_image(classes, confs, boxes, 2, top_k=2, nms_all_threshold=0.5)
    assert kept.tolist() == [0, -1]
    assert np.array_equal(kept[:1], _strict_two_pass(classes, confs, boxes, 2, 2, 0.45, 0.5))


# ----------------------------------------------------------------------------------------- losses (od_pb.py:365-431; parity unpinned: TF absent)
def test_od_losses_closed_form_cases(oracle_tables):
    """Hand-checkable cases of the restated loss: a perfect prediction gives ~0 loss and metric 1; the focal term of
    a single negative row equals -0.75 * p^2 * log(1-p) * mask; the weights sum to 1 over the positives."""
    t = oracle_tables('tiny128')
    rng = np.random.RandomState(0)
    boxes = np.array([[0.1, 0.1, 0.5, 0.6], [0.55, 0.3, 0.9, 0.8]], np.float32)
    classes = np.array([0, 1], np.int32)
    y_true = R.encode_truth(t, [(classes, boxes)])
    assert np.isclose(y_true[0, :, 1].sum(), 1.0)
    perfect = y_true.copy()
    perfect[0, y_true[0, :, 2] == 0, 3:-4] = 0.5  # a softmax output is never all-zero (0/0 would be NaN, as in TF)
    res = R.od_losses(t, y_true, perfect)
    assert res['loss'][0] < 1e-4 and res['rec_bg'][0] == 1 and res['rec_obj'][0] == 1 and res['acc_clf'][0] == 1
    # one negative valid row predicts objectness p, everything else perfect
    neg = np.flatnonzero((y_true[0, :, 2] == 0) & (y_true[0, :, 0] > 0) & t.pb_mask)[3]
    pred = perfect.copy()
    p = np.float32(0.7)
    pred[0, neg, 2] = p
    res2 = R.od_losses(t, y_true, pred)
    want = 3 * (-0.75 * float(p) ** 2 * np.log(1 - float(p))) * float(y_true[0, neg, 0])
    assert np.isclose(res2['loss_obj'][0] - res['loss_obj'][0], want, rtol=1e-5)
    assert res2['rec_bg'][0] < 1
    # random predictions: finite, positive
    pred = rng.uniform(0.01, 0.99, size=y_true.shape).astype(np.float32)
    r3 = R.od_losses(t, y_true, pred)
    assert all(np.isfinite(v).all() for v in r3.values()) and r3['loss'][0] > 0


def test_confusion_matrix_golden(golden):
    """ml.od_confusion_matrix (ml.py:397-451) vs outputs of the unmodified reference, incl. its positional use of the
    per-detection mask on the sorted index array (ml.py:421)."""
    from object_detector_b200 import synth
    g = golden['scores']
    gts, preds = synth.eval_set(77, 60, 5, max_gt=5, n_fp=6)
    gt = [(c, b, d) for (_, _, c, b, d) in gts]
    for i in range(3):
        th = float(g[f'cm_{i}/conf_threshold'])
        cm = R.od_confusion_matrix(gt, preds, th, 0.5, 5)
        assert np.array_equal(cm, g[f'cm_{i}/out'])
        assert cm[:-1].sum() == sum(len(x[0]) for x in gt)  # every GT lands in exactly one cell of its row
    gts, preds = synth.eval_set(78, 40, 3, max_gt=6, n_fp=12)
    gt = [(c, b, d) for (_, _, c, b, d) in gts]
    assert np.array_equal(R.od_confusion_matrix(gt, preds, 0.1, 0.3, 3), g['cm_dense/out'])


def test_tf_nms_restatement_micro_cases():
    """The TF-1.10 NonMaxSuppression restatement (PARITY UNPINNED) on hand-computed cases: '>' instead of '>=', zero-area
    boxes selectable and inert, exact max_output_size; equals the NumPy rule when no pair sits at the threshold and no
    box is degenerate."""
    boxes = np.array([[0, 0, 1, 1], [0, 0, 1, 0.5], [0.3, 0.3, 0.3, 0.8], [0.3, 0.3, 0.3, 0.8], [0, 0, 0.9, 1]], np.float32)
    confs = np.array([0.9, 0.8, 0.7, 0.6, 0.5], np.float32)
    assert R.tf_nms_policy(boxes, confs, 5, 0.5).tolist() == [0, 1, 2, 3]
    assert R.tf_nms_policy(boxes, confs, 3, 0.5).tolist() == [0, 1, 2] and R.tf_nms_policy(boxes, confs, 1, 0.5).tolist() == [0]
    assert R.numpy_nms_policy(boxes, confs, 5, 0.5).tolist() == [0]
    rng = np.random.RandomState(3)
    c = rng.uniform(0.2, 0.8, size=(300, 2))
    s = rng.uniform(0.05, 0.3, size=(300, 2))
    b = np.concatenate([c - s / 2, c + s / 2], axis=1).astype(np.float32)
    sc = rng.permutation(300).astype(np.float32) / 300
    assert R.tf_nms_policy(b, sc, 200, 0.45).tolist() == R.numpy_nms_policy(b, sc, 200, 0.45).tolist()


def test_annotation_helpers_match_reference():
    """rot90 / real_bboxes of the mirror datatypes against the unmodified reference classes
    (ml.py:62-65, 102-112), and the oracle's is_match against the reference's own (ml.py:159-182) on random cases."""
    if not ref_loader.available():
        pytest.skip('reference tree not mounted')
    import object_detector_b200.ml as M  # host-side helpers only (no GPU call in this test)
    from object_detector_b200 import synth
    tk = ref_loader.load()
    gts, preds = synth.eval_set(3, 30, 4, max_gt=5, n_fp=4)
    for (w, h, c, b, d), (pc, ps, pb) in zip(gts, preds):
        for k in range(4):
            mine, ref = M.ObjectsAnnotation('.', w, h, c, b, d), tk.ml.ObjectsAnnotation('.', w, h, c, b, d)
            mine.rot90(k)
            ref.rot90(k)
            assert np.array_equal(mine.bboxes, ref.bboxes) and mine.bboxes.dtype == ref.bboxes.dtype
        mine, ref = M.ObjectsAnnotation('.', w, h, c, b, d), tk.ml.ObjectsAnnotation('.', w, h, c, b, d)
        assert np.array_equal(mine.real_bboxes, ref.real_bboxes)
        rp = tk.ml.ObjectsPrediction(pc, ps, pb)
        for thr in (0.0, 0.3, 0.6):
            assert R.exact_match(c, b, (pc, ps, pb), thr, 0.5) == rp.is_match(c, b, thr, 0.5)
        keep = ps >= 0.3  # a prediction set that does match: the jittered GT copies only
        sub = (pc[keep][:len(c)], ps[keep][:len(c)], pb[keep][:len(c)])
        assert R.exact_match(c, b, sub, 0.0, 0.5) == tk.ml.ObjectsPrediction(*sub).is_match(c, b, 0.0, 0.5)


def test_loss_gradient_oracle_vs_torch_autograd(oracle_tables):
    """The analytic gradient of the oracle equals torch.autograd of the Keras formulas written in float64 torch
    (clamp passes gradients only inside its range, like tf.clip_by_value)."""
    import torch
    t = oracle_tables('tiny128')
    rng = np.random.RandomState(4)
    boxes = np.array([[0.1, 0.1, 0.5, 0.6], [0.55, 0.3, 0.9, 0.8], [0.2, 0.5, 0.45, 0.95]], np.float32)
    y_true = R.encode_truth(t, [(np.array([0, 1, 1], np.int32), boxes), (np.array([1], np.int32), boxes[:1])])
    y_pred = rng.uniform(0.02, 0.98, size=y_true.shape).astype(np.float32)
    y_pred[0, 5, 2] = 1.0          # p_t = 0 for a negative row -> clipped -> no gradient
    y_pred[0, 7, 2] = 0.0          # p_t = 1 -> clipped
    up = np.array([1.0, 0.37])
    yt, yp = torch.tensor(y_true, dtype=torch.float64), torch.tensor(y_pred, dtype=torch.float64, requires_grad=True)
    eps = 1e-7
    y, p = yt[:, :, 2], yp[:, :, 2]
    a_t = y * 0.25 + (1 - y) * 0.75
    p_t = torch.clamp(y * p + (1 - y) * (1 - p), float(np.float32(eps)), 1 - float(np.float32(eps)))
    mask = torch.tensor(t.pb_mask.astype(np.float64))[None] * yt[:, :, 0]
    loss_obj = 3 * ((-a_t * (1 - p_t) ** 2 * torch.log(p_t)) * mask).sum(-1)
    q = yp[:, :, 3:-4]
    n = torch.clamp(q / q.sum(-1, keepdim=True), float(np.float32(eps)), 1 - float(np.float32(eps)))
    loss_clf = (-(yt[:, :, 3:-4] * torch.log(n)).sum(-1) * yt[:, :, 1]).sum(-1)
    loss_loc = 2 * (((yp[:, :, -4:] - yt[:, :, -4:]) ** 2).mean(-1) * yt[:, :, 1]).sum(-1)
    loss = loss_obj + loss_clf + loss_loc
    (loss * torch.tensor(up)).sum().backward()
    got = R.od_loss_grad(t, y_true, y_pred, up)
    np.testing.assert_allclose(got, yp.grad.numpy(), rtol=1e-9, atol=1e-12)
    assert got[0, 5, 2] == 0 and got[0, 7, 2] == 0 and np.abs(got[:, :, 2]).max() > 0
    ref = R.od_losses(t, y_true, y_pred)
    np.testing.assert_allclose(ref['loss'], loss.detach().numpy(), rtol=1e-5)


# ----------------------------------------------------------------------------------------- box side of the augmentation (f2)
def _golden_augment_rows(golden, name):
    g = golden['augment']
    ends = np.cumsum(g[name + '_counts'])
    for i, e in enumerate(ends):
        s = e - g[name + '_counts'][i]
        yield g[name + '_boxes'][s:e], g[name + '_classes'][s:e], g[name + '_next_draw'][i]


@pytest.mark.parametrize('name', sorted(synth.AUGMENT_CONFIGS))
def test_box_augmentation_matches_reference(golden, name):
    """RandomZoom -> FlipLR -> FlipTB -> Rotate90 on boxes (image.py:233-485): the restatement reproduces the
    unmodified reference bit for bit on 160 seeded samples per configuration — surviving boxes, their classes, and
    the position the per-sample RandomState is left at (checked through its next draw)."""
    size, keep_aspect, kw = synth.AUGMENT_CONFIGS[name]
    failed = 0
    for (w, h, b, c, seed), (gb, gc, nxt) in zip(synth.augment_inputs(911, 160), _golden_augment_rows(golden, name)):
        rand = np.random.RandomState(seed)
        ob, oc, info = R.augment_boxes_sample(rand, w, h, b, c, size, keep_aspect=keep_aspect, **kw)
        assert np.array_equal(ob.astype(np.float64), gb) and np.array_equal(oc, gc)
        assert rand.rand() == nxt
        failed += info['chosen'] < 0
    if name != 'default':
        assert failed >= 3      # the all-attempts-rejected branch (and its letterbox) is in the vectors


def test_flip_rot_boxes_keeps_fp32_rounding():
    """1-(1-x) is not x in fp32: boxes that no zoom widened must be flipped in their own dtype (image.py:445,
    ml.py:102-112 operate in place / by fancy indexing on the annotation's fp32 array)."""
    rng = np.random.RandomState(5)
    b = synth.gt_boxes(rng, 64)
    twice = R.flip_rot_boxes(R.flip_rot_boxes(b, True, False, 0), False, False, 2)
    assert twice.dtype == np.float32
    wide = R.flip_rot_boxes(R.flip_rot_boxes(b.astype(np.float64), True, False, 0), False, False, 2)
    assert (twice.astype(np.float64)[:, [0, 2]] != wide[:, [0, 2]]).any()
    for k in range(4):   # four quarter turns compose to the identity up to rounding; areas are preserved
        r = R.flip_rot_boxes(b.astype(np.float64), False, False, k)
        assert np.allclose((r[:, 2] - r[:, 0]) * (r[:, 3] - r[:, 1]), (b[:, 2] - b[:, 0]).astype(np.float64) * (b[:, 3] - b[:, 1]), rtol=1e-6)


def test_coco_ap_micro_cases():
    """Hand-computed cases for the pycocotools / chainercv restatement (PARITY UNPINNED: both packages absent)."""
    f = np.float32
    key = 'ap/iou=0.50:0.95/area=all/max_dets=100'

    def gt(boxes, crowd=None, areas=None, classes=None):
        boxes = np.array(boxes, f)
        n = len(boxes)
        px = boxes * 100
        a = np.array(areas, np.float64) if areas is not None else (px[:, 2] - px[:, 0]) * (px[:, 3] - px[:, 1])
        return [(100, 100, np.zeros(n, int) if classes is None else np.array(classes), boxes, a,
                 np.zeros(n, bool) if crowd is None else np.array(crowd))]

    def pred(boxes, scores, classes=None):
        n = len(boxes)
        return [(np.zeros(n, int) if classes is None else np.array(classes), np.array(scores, f), np.array(boxes, f).reshape(-1, 4))]

    one = gt([[.1, .1, .5, .5]])  # 40x40 px = 1600 px^2: "medium"
    r = R.coco_evaluate(one, pred([[.1, .1, .5, .5]], [.9]))
    assert r['m' + key] == pytest.approx(1.0, abs=1e-12) and r['mar/iou=0.50:0.95/area=all/max_dets=1'] == pytest.approx(1.0, abs=1e-12)
    assert r['map/iou=0.50:0.95/area=medium/max_dets=100'] == pytest.approx(1.0, abs=1e-12)
    assert np.isnan(r['map/iou=0.50:0.95/area=small/max_dets=100']) and np.isnan(r['map/iou=0.50:0.95/area=large/max_dets=100'])
    r = R.coco_evaluate(one, pred([[.6, .6, .9, .9]], [.9]))
    assert r['m' + key] == 0.0 and r['mar/iou=0.50:0.95/area=all/max_dets=100'] == 0.0
    # FP (higher score) then TP: precision 0.5 at recall 1, and the envelope carries it to every recall threshold
    r = R.coco_evaluate(one, pred([[.6, .6, .9, .9], [.1, .1, .5, .5]], [.9, .8]))
    assert r['m' + key] == pytest.approx(0.5) and r['mar/iou=0.50:0.95/area=all/max_dets=1'] == 0.0
    assert r['mar/iou=0.50:0.95/area=all/max_dets=10'] == pytest.approx(1.0, abs=1e-12)
    # IoU 0.625 with the GT: a true positive at thresholds 0.50 .. 0.60 only -> 3 of 10
    r = R.coco_evaluate(one, pred([[.1, .1, .5, .35]], [.9]))
    assert r['m' + key] == pytest.approx(0.3) and r['map/iou=0.50/area=all/max_dets=100'] == pytest.approx(1.0, abs=1e-12)
    assert r['map/iou=0.75/area=all/max_dets=100'] == 0.0
    # the only GT is a crowd: it is ignored, the detection matching it is ignored too -> nothing to score
    r = R.coco_evaluate(gt([[.1, .1, .5, .5]], crowd=[True]), pred([[.1, .1, .5, .5]], [.9]))
    assert np.isnan(r['m' + key])
    # a crowd GT absorbs any number of detections; the plain GT is matched once, the duplicate is a false positive
    r = R.coco_evaluate(gt([[.1, .1, .5, .5], [.6, .6, .9, .9]], crowd=[True, False]),
                        pred([[.1, .1, .5, .5], [.1, .1, .5, .5], [.6, .6, .9, .9], [.6, .6, .9, .9]], [.9, .8, .7, .6]))
    assert r['m' + key] == pytest.approx(1.0, abs=1e-12)  # TP at rank 3 (ranks 1-2 ignored), FP after it does not lower the envelope at recall 1
    # labels absent between 0 and the maximum get nan, existent_labels lists the rest
    r = R.coco_evaluate(gt([[.1, .1, .5, .5]], classes=[2]), pred([[.1, .1, .5, .5]], [.9], classes=[2]))
    assert r['existent_labels'].tolist() == [2] and np.isnan(r[key][:2]).all() and r[key][2] == pytest.approx(1.0, abs=1e-12)
