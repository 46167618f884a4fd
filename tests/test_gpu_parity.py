"""Parity of the CUDA path (through the C-ABI / Python mirror) against the CPU oracle and the golden
vectors generated from the reference.  Integer/index/mask outputs: bit-exact.  fp32 coordinates:
bit-exact against the fp32 restatement (tolerance stated where the reference computes in fp64)."""
import hashlib

import numpy as np
import pytest
import torch

import object_detector_b200 as odk
from object_detector_b200 import _lib, synth
from oracle import restated as R

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def assign_cases(g):
    return sorted({k.split('/')[0] for k in g.files})


# ----------------------------------------------------------------------------------------- anchors / small ops
def test_library_loaded_and_gpu_is_blackwell():
    assert _lib.version().startswith('odk ')
    assert torch.cuda.get_device_capability(0)[0] >= 10


def test_compute_iou_kat_and_golden(golden):
    g = golden['iou_nms_decode']
    out = odk.ml.compute_iou(g['iou_kat/a'], g['iou_kat/b'])
    assert out.shape == (2, 1) and out[1, 0] == 0
    assert out[0, 0] == pytest.approx(100 * 100 / (200 * 200 * 2 - 100 * 100))  # ml_test.py:52
    out = odk.ml.compute_iou(g['iou_f32/a'], g['iou_f32/b'])
    assert np.array_equal(out, g['iou_f32/out'])  # bit-exact fp32


def test_encode_locs_bit_exact(golden, prior_boxes):
    g, pb = golden['iou_nms_decode'], prior_boxes('voc640')
    out = pb.encode_locs(g['encode_locs/box'], g['encode_locs/ix'])
    assert out.dtype == np.float64 and np.array_equal(out, g['encode_locs/out'])
    assert np.array_equal(pb.encode_locs(g['encode_locs/box'], 5), g['encode_locs/out'][1])
    mask = np.zeros(len(pb), bool)
    mask[g['encode_locs/ix']] = True
    assert np.array_equal(pb.encode_locs(g['encode_locs/box'], mask), g['encode_locs/out'])


def test_decode_locs(golden, prior_boxes, oracle_tables):
    g = golden['iou_nms_decode']
    pb, t = prior_boxes('tiny128'), oracle_tables('tiny128')
    out = pb.decode_locs(g['decode_tiny/pred'])
    assert out.dtype == np.float32
    assert np.array_equal(out, R.decode_locs_f32(t, g['decode_tiny/pred']))  # fp32 graph semantics: bit-exact
    # the reference's xp=np path is fp64: agree within 1e-5 relative (north_star) -> abs 2e-7 on [0,1] values
    np.testing.assert_allclose(out, g['decode_tiny/out_f64'], rtol=1e-5, atol=2e-7)
    # xp=np (check_assign.py:27): float64, bit-for-bit the unmodified reference's output
    out64 = pb.decode_locs(g['decode_tiny/pred'], xp=np)
    assert out64.dtype == np.float64 and np.array_equal(out64, g['decode_tiny/out_f64'])
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    pred = np.random.RandomState(int(g['decode_voc640/seed'])).normal(size=(1, len(pb), 4)).astype(np.float32)
    out = pb.decode_locs(pred)
    assert np.array_equal(out, R.decode_locs_f32(t, pred))
    np.testing.assert_allclose(out[0, ::997], g['decode_voc640/out_f64_rows'], rtol=1e-5, atol=2e-7)
    out64 = pb.decode_locs(pred, xp=np)
    assert out64.dtype == np.float64 and np.array_equal(out64[0, ::997], g['decode_voc640/out_f64_rows'])
    assert hashlib.sha256(np.ascontiguousarray(out64).tobytes()).hexdigest() == str(g['decode_voc640/out_f64_sha256'])
    # the exact call of check_assign.py:27: zeros in -> the clipped prior boxes themselves, float64
    zero = pb.decode_locs(np.zeros((len(pb), 4)), xp=np)
    assert zero.dtype == np.float64 and np.array_equal(zero, np.clip(pb.pb_locs.astype(np.float64), 0, 1))


def test_channel_argmax_max_layers(prior_boxes):
    """layers.channel_argmax / channel_max (layers.py:69-98) run through odk_channel_argmax_max: NumPy argmax / max
    semantics (first maximum wins), int64 indices, ties and an all-NaN row included."""
    rng = np.random.RandomState(8)
    x = rng.normal(size=(3, 1000, 20)).astype(np.float32)
    x[0, 5, :] = 0.25            # all equal: index 0
    x[1, 7, 3] = x[1, 7, 11] = 9.0  # tie: the lower index
    x[2, 9, :] = np.nan
    arg = odk.layers.channel_argmax()()(x)
    mx = odk.layers.channel_max()()(x)
    assert arg.dtype == np.int64 and arg.shape == (3, 1000) and mx.dtype == np.float32
    ok = ~np.isnan(x).any(axis=-1)
    assert np.array_equal(arg[ok], x.argmax(axis=-1)[ok]) and np.array_equal(mx[ok], x.max(axis=-1)[ok])
    assert arg[2, 9] == 0 and np.isnan(mx[2, 9]) and arg[0, 5] == 0 and arg[1, 7] == 3
    t = torch.from_numpy(x).cuda()
    assert torch.equal(odk.layers.channel_argmax()()(t).cpu(), torch.from_numpy(arg))


# ----------------------------------------------------------------------------------------- assign / encode
def test_assign_boxes_golden_bit_exact(golden, prior_boxes):
    g = golden['assign']
    for case in assign_cases(g):
        pb = prior_boxes(str(g[f'{case}/anchors']))
        idx, gts, valid = pb._assign_boxes(g[f'{case}/boxes'])
        assert idx.dtype == np.int64 and gts.dtype == np.int64 and valid.dtype == bool
        assert np.array_equal(idx, g[f'{case}/pb_indices']), case
        assert np.array_equal(gts, g[f'{case}/gt_indices']), case
        assert np.array_equal(np.packbits(valid), g[f'{case}/pb_valids']), case


def test_encode_truth_golden_bit_exact(golden, prior_boxes):
    g = golden['assign']
    for case in assign_cases(g):
        pb = prior_boxes(str(g[f'{case}/anchors']))
        ann = odk.ml.ObjectsAnnotation('.', 300, 300, g[f'{case}/classes'], g[f'{case}/boxes'])
        y = pb.encode_truth([ann])
        assert y.dtype == np.float32 and y.shape == (1, len(pb), 3 + pb.num_classes + 4)
        assert sha(y[0]) == str(g[f'{case}/y_sha256']), case  # whole tensor, bit for bit
        assert np.array_equal(y[0][g[f'{case}/pb_indices']], g[f'{case}/y_pos_rows']), case
    assert np.array_equal(pb_tiny_full(prior_boxes, g), g['tiny128_g12/y_full'])


def pb_tiny_full(prior_boxes, g):
    pb = prior_boxes('tiny128')
    return pb.encode_truth([odk.ml.ObjectsAnnotation('.', 300, 300, g['tiny128_g12/classes'], g['tiny128_g12/boxes'])])[0]


def test_stress_assign_g512_bit_exact(golden, prior_boxes):
    g, pb = golden['assign_stress'], prior_boxes('stress1120')
    idx, gts, valid = pb._assign_boxes(g['stress_g512/boxes'])
    assert np.array_equal(idx, g['stress_g512/pb_indices'])
    assert np.array_equal(gts, g['stress_g512/gt_indices'])
    assert np.array_equal(np.packbits(valid), g['stress_g512/pb_valids'])
    ann = odk.ml.ObjectsAnnotation('.', 300, 300, g['stress_g512/classes'], g['stress_g512/boxes'])
    assert sha(pb.encode_truth([ann])[0]) == str(g['stress_g512/y_sha256'])


def test_encode_truth_batch_vs_oracle(prior_boxes, oracle_tables):
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    boxes, classes, counts = synth.gt_batch(7, 6, 64, 20)
    counts[2] = 0  # an image without objects: all-valid mask, weight 1 (od_pb.py:306 max(1, 0))
    res = pb.encode_truth_batch(boxes, classes, counts, want_valids=True)
    y = res['y_true'].cpu().numpy()
    assert (res['status'].cpu().numpy() == 0).all()
    ref = R.encode_truth(t, [(classes[i, :counts[i]], boxes[i, :counts[i]]) for i in range(6)])
    assert np.array_equal(y, ref)
    for i in range(6):
        idx, gts, valid = R.assign_boxes(t, boxes[i, :counts[i]])
        assert np.array_equal(res['pb_indices'][i, :counts[i]].cpu().numpy(), idx)
        assert np.array_equal(res['gt_indices'][i, :counts[i]].cpu().numpy(), gts)
        assert (res['pb_indices'][i, counts[i]:].cpu().numpy() == -1).all()
        assert np.array_equal(res['pb_valids'][i].cpu().numpy(), valid)


def test_encode_truth_errors_like_reference(prior_boxes):
    pb = prior_boxes('voc640')
    ok = odk.ml.ObjectsAnnotation('.', 10, 10, [3], [[0.1, 0.1, 0.5, 0.5]])
    bad_class = odk.ml.ObjectsAnnotation('.', 10, 10, [20], [[0.1, 0.1, 0.5, 0.5]])
    with pytest.raises(AssertionError):
        pb.encode_truth([ok, bad_class])  # od_pb.py:304
    res = pb.encode_truth_batch(np.zeros((1, 2, 4), np.float32), np.zeros((1, 2), np.int32), np.array([5], np.int32))
    assert int(res['status'][0]) == 3 and float(res['y_true'].abs().sum()) == 0.0
    assert pb.encode_truth([]).shape == (0, len(pb), 27)


def test_encode_full_size_properties(prior_boxes):
    """BASELINE config 3 at full size (B=256, <=64 GT): size-independent invariants (SURVEY A-10)."""
    pb = prior_boxes('voc640')
    boxes, classes, counts = synth.gt_batch(7, 256, 64, 20)
    res = pb.encode_truth_batch(boxes, classes, counts)
    y, status = res['y_true'], res['status']
    assert int(status.abs().sum()) == 0
    cnt = torch.from_numpy(counts).cuda()
    assert torch.equal((y[:, :, 2] == 1).sum(dim=1).int(), cnt)  # exactly G positives per image
    assert torch.equal(y[:, :, 3:23].sum(dim=(1, 2)).int(), cnt)  # one-hot rows only on positives
    pos = y[:, :, 2] == 1
    assert bool((y[:, :, 0][pos] > 0).all())  # obj => valid (od_pb.py:312)
    w = (1.0 / np.maximum(1, counts)).astype(np.float32)
    ch0 = y[:, :, 0]
    assert bool(((ch0 == 0) | (ch0 == torch.from_numpy(w).cuda()[:, None])).all())
    assert bool((y[:, :, 1][pos] == torch.from_numpy(w).cuda()[:, None].expand_as(pos)[pos]).all())
    assert bool((y[:, :, 1:][~pos] == 0).all())
    # idempotence / determinism: a second run gives the same bytes
    res2 = pb.encode_truth_batch(boxes, classes, counts)
    assert torch.equal(res2['y_true'], y) and torch.equal(res2['pb_indices'], res['pb_indices'])
    idx = res['pb_indices']
    srt = torch.where(idx >= 0, idx, torch.full_like(idx, 1 << 40))
    assert bool((srt[:, 1:] >= srt[:, :-1]).all())  # ascending anchor ids, -1 padding last


# ----------------------------------------------------------------------------------------- decode / score / NMS
@pytest.mark.parametrize('name,batch', [('tiny128', 3), ('voc640', 2)])
def test_decode_score_bit_exact(prior_boxes, oracle_tables, name, batch):
    pb, t = prior_boxes(name), oracle_tables(name)
    objs, clfs, locs = synth.predictions(5, batch, len(pb), pb.num_classes)
    classes, confs, boxes = pb.decode_score(objs, clfs, locs)
    rc, rs = R.score_anchors(t, objs, clfs)
    assert np.array_equal(classes.cpu().numpy(), rc)
    assert np.array_equal(confs.cpu().numpy(), rs)
    assert np.array_equal(boxes.cpu().numpy(), R.decode_locs_f32(t, locs))


def test_decode_score_unaligned_views(prior_boxes, oracle_tables):
    """Inputs that are not 16-byte aligned take the plain-load path of the ring; same results."""
    pb, t = prior_boxes('tiny128'), oracle_tables('tiny128')
    objs, clfs, locs = synth.predictions(6, 2, len(pb), pb.num_classes)

    def shifted(x):
        buf = torch.empty(x.size + 1, dtype=torch.float32, device='cuda')
        v = buf[1:].view(x.shape)
        v.copy_(torch.from_numpy(x))
        return v
    classes, confs, boxes = pb.decode_score(shifted(objs), shifted(clfs), shifted(locs))
    rc, rs = R.score_anchors(t, objs, clfs)
    assert np.array_equal(classes.cpu().numpy(), rc) and np.array_equal(confs.cpu().numpy(), rs)
    assert np.array_equal(boxes.cpu().numpy(), R.decode_locs_f32(t, locs))


def test_nms_kat_from_reference_tests(golden):
    g = golden['iou_nms_decode']  # ml_test.py:89-113
    idx = odk.ml.non_maximum_suppression(g['nms_kat/boxes'], g['nms_kat/scores'], 200, 0.45)
    assert idx.dtype == np.int64 and (idx == np.array([5, 3])).all()
    idx = odk.ml.non_maximum_suppression(np.array([[0, 0, 200, 200]]), np.array([0.2]), 200, 0.45)
    assert (idx == np.array([0])).all()
    idx = odk.ml.non_maximum_suppression(np.array([[200, 0, 200, 200]]), np.array([0.1]), 200, 0.45)
    assert idx.shape == (0,)
    assert odk.ml.non_maximum_suppression(np.zeros((0, 4)), np.zeros((0,))).shape == (0,)


def test_nms_golden_sets_bit_exact(golden):
    g = golden['iou_nms_decode']
    for i in range(6):
        out = odk.ml.non_maximum_suppression(g[f'nms_f32_{i}/boxes'], g[f'nms_f32_{i}/scores'], int(g[f'nms_f32_{i}/top_k']), 0.45)
        assert np.array_equal(out, g[f'nms_f32_{i}/out']), i


def test_nms_large_set_uses_global_sort_path():
    """> 16384 candidates in one segment -> hybrid shared/global bitonic sort; kept set vs oracle."""
    rng = np.random.RandomState(8)
    n = 40000
    c = rng.uniform(0.05, 0.95, size=(n, 2))
    s = rng.uniform(0.01, 0.08, size=(n, 2))
    boxes = np.clip(np.concatenate([c - s / 2, c + s / 2], 1), 0, 1).astype(np.float32)
    scores = (rng.permutation(n).astype(np.float32) + 1) / (n + 1)
    out = odk.ml.non_maximum_suppression(boxes, scores, 300, 0.3)
    assert np.array_equal(out, R.numpy_nms_policy(boxes, scores, 300, 0.3))


def test_nms_tie_policy():
    """Equal scores: higher original index first (argsort(stable)[::-1], SURVEY A-6)."""
    boxes = np.array([[0, 0, 1, 1], [0, 0, 1, 1], [2, 2, 3, 3]], np.float32)
    scores = np.array([0.5, 0.5, 0.5], np.float32)
    assert odk.ml.non_maximum_suppression(boxes, scores).tolist() == R.numpy_nms_policy(boxes, scores).tolist() == [2, 1]


@pytest.mark.parametrize('clustered', [False, True])
def test_nms_layer_vs_oracle(prior_boxes, oracle_tables, clustered):
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    objs, clfs, locs = synth.predictions(1234, 2, len(pb), 20, clustered=clustered)
    classes, confs = R.score_anchors(t, objs, clfs)
    boxes = R.decode_locs_f32(t, locs)
    layer = odk.layers.nms()(20, len(pb))
    out, kept = layer([classes, confs, boxes], return_index=True)
    for b in range(2):
        ro, rk = R.nms_layer_image(classes[b], confs[b], boxes[b], 20)
        assert np.array_equal(kept[b], rk), b   # kept-index sets AND order: bit-exact
        assert np.array_equal(out[b], ro), b    # rows [class, conf, box]: bit-exact fp32
    assert layer.compute_output_shape([(2, len(pb)), (2, len(pb)), (2, len(pb), 4)]) == (2, 200, 6)


@pytest.mark.parametrize('name,batch,top_k', [('voc640', 3, 200), ('tiny128', 4, 200), ('tiny128', 2, 5), ('voc320', 1, 200)])
def test_fused_decode_nms_vs_oracle(prior_boxes, oracle_tables, name, batch, top_k):
    pb, t = prior_boxes(name), oracle_tables(name)
    objs, clfs, locs = synth.predictions(77, batch, len(pb), pb.num_classes, clustered=(name == 'voc320'))
    out, kept = pb.decode_nms(objs, clfs, locs, top_k=top_k, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs, top_k=top_k)
    assert np.array_equal(kept.cpu().numpy(), rk)
    assert np.array_equal(out.cpu().numpy(), ro)


def test_fused_equals_unfused_full_batch(prior_boxes):
    """BASELINE config 2 at full size (B=64): fused kernel chain == dense decode + NMS layer; output
    invariants: descending scores, zero padding, boxes inside [0,1], kept anchors unique per image."""
    pb = prior_boxes('voc640')
    objs, clfs, locs = synth.predictions(1234, 64, len(pb), 20)
    d = [torch.from_numpy(x).cuda() for x in (objs, clfs, locs)]
    out, kept = pb.decode_nms(*d, return_index=True)
    classes, confs, boxes = pb.decode_score(*d)
    out2, kept2 = odk.layers.nms()(20, len(pb))([classes, confs, boxes], return_index=True)
    assert torch.equal(out, out2) and torch.equal(kept, kept2)
    conf = out[:, :, 1]
    assert bool((conf[:, 1:] <= conf[:, :-1]).all())
    assert bool(((out[:, :, 2:] >= 0) & (out[:, :, 2:] <= 1)).all())
    pad = kept < 0
    assert bool((out[pad] == 0).all()) and bool((conf[~pad] > 0.01).all())
    for b in range(64):
        k = kept[b][kept[b] >= 0]
        assert k.unique().numel() == k.numel()
        assert torch.equal(confs[b][k.long()], conf[b][: k.numel()])
    # idempotence: NMS of the kept set keeps everything (per class)
    again = pb.decode_nms(*d)
    assert torch.equal(again, out)


def test_full_batch_config2_sampled_vs_oracle(prior_boxes, oracle_tables):
    """BASELINE config 2 at full size (B=64, the bench's own first input set): 16 seeded images of the batch — most of
    them with batch index >= 32 — against the oracle: kept anchor ids and output rows bit-exact.  Plus four images of
    the clustered variant of the same batch."""
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    for seed, clustered, picks in ((1234, False, [0, 7, 21, 31, 32, 33, 38, 40, 45, 47, 48, 52, 55, 59, 62, 63]),
                                   (4321, True, [5, 33, 50, 63])):
        objs, clfs, locs = synth.predictions(seed, 64, len(pb), 20, clustered=clustered)
        out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
        out, kept = out.cpu().numpy(), kept.cpu().numpy()
        ro, rk = R.predict_tail(t, objs[picks], clfs[picks], locs[picks])
        for j, b in enumerate(picks):
            assert np.array_equal(kept[b], rk[j]), (clustered, b)
            assert np.array_equal(out[b], ro[j]), (clustered, b)


def test_full_batch_config3_sampled_vs_oracle(prior_boxes, oracle_tables):
    """BASELINE config 3 at full size (B=256, <=64 GT, every 4th image crowded): 16 seeded images — batch indices up to
    255, crowded ones included — against the oracle: the whole y_true slab of each image, the index lists and the
    valid mask, bit for bit."""
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    boxes, classes, counts = synth.gt_batch(7, 256, 64, 20)
    res = pb.encode_truth_batch(boxes, classes, counts, want_valids=True)
    picks = [0, 3, 31, 32, 35, 63, 64, 100, 127, 128, 131, 190, 199, 223, 254, 255]
    assert sum(1 for b in picks if b % 4 == 3) >= 6 and int(res['status'].abs().sum()) == 0
    idx = torch.tensor(picks, device='cuda')
    y = res['y_true'][idx].cpu().numpy()
    pi, gi, pv = (res[k][idx].cpu().numpy() for k in ('pb_indices', 'gt_indices', 'pb_valids'))
    for j, b in enumerate(picks):
        g = int(counts[b])
        ref = R.encode_truth(t, [(classes[b, :g], boxes[b, :g])])[0]
        assert np.array_equal(y[j], ref), b
        ri, rg, rv = R.assign_boxes(t, boxes[b, :g])
        assert np.array_equal(pi[j, :g], ri) and np.array_equal(gi[j, :g], rg) and np.array_equal(pv[j], rv), b
        assert (pi[j, g:] == -1).all()


def test_tf_style_layer_vs_restated_tf_nms(prior_boxes, oracle_tables):
    """iou_ge=0 against the restatement of TF-1.10's NonMaxSuppression (PARITY UNPINNED: TensorFlow absent): the whole
    layer on random and clustered head outputs, then the cases where the two rules differ — an IoU exactly at the
    threshold (kept by '>', dropped by '>='), zero-area boxes (selectable, never suppressing), and the exact
    max_output_size cut."""
    pb, t = prior_boxes('voc320'), oracle_tables('voc320')
    for clustered in (False, True):
        objs, clfs, locs = synth.predictions(91, 2, len(pb), 20, clustered=clustered)
        classes, confs = R.score_anchors(t, objs, clfs)
        boxes = R.decode_locs_f32(t, locs)
        out, kept = odk.layers.nms()(20, len(pb), iou_ge=False)([classes, confs, boxes], return_index=True)
        for b in range(2):
            ro, rk = R.nms_layer_image(classes[b], confs[b], boxes[b], 20, inner=R.tf_nms_policy)
            assert np.array_equal(kept[b], rk) and np.array_equal(out[b], ro), (clustered, b)
    boxes = np.array([[0, 0, 1, 1], [0, 0, 1, 0.5], [0.3, 0.3, 0.3, 0.8], [0.3, 0.3, 0.3, 0.8], [0, 0, 0.9, 1]], np.float32)
    confs = np.array([[0.9, 0.8, 0.7, 0.6, 0.5]], np.float32)
    classes = np.zeros((1, 5), np.int32)
    for top_k in (5, 3, 1):
        tf_kept = odk.layers.nms()(1, 5, top_k=top_k, nms_threshold=0.5, iou_ge=False)([classes, confs, boxes[None]], return_index=True)[1][0]
        want = R.tf_nms_policy(boxes, confs[0], top_k, 0.5)
        assert tf_kept[tf_kept >= 0].tolist() == want.tolist()
    assert R.tf_nms_policy(boxes, confs[0], 5, 0.5).tolist() == [0, 1, 2, 3]       # IoU(0,1) == 0.5 is not > 0.5; both zero-area boxes selected
    assert R.numpy_nms_policy(boxes, confs[0], 5, 0.5).tolist() == [0]              # the NumPy rule drops all four others


def test_tf_style_mode_runs_and_differs_only_by_rule(prior_boxes):
    """iou_ge=0 (TF-1.10 style, PARITY UNPINNED): strict '>' and no zero-area skip."""
    boxes = np.array([[0, 0, 1, 1], [0, 0, 1, 0.45], [0.2, 0.2, 0.2, 0.9]], np.float32)  # IoU(0,1)=0.45 exactly? no: fp
    classes = np.zeros((1, 3), np.int32)
    confs = np.array([[0.9, 0.8, 0.7]], np.float32)
    np_mode = odk.layers.nms()(1, 3, top_k=3)([classes, confs, boxes[None]], return_index=True)[1]
    tf_mode = odk.layers.nms()(1, 3, top_k=3, iou_ge=False)([classes, confs, boxes[None]], return_index=True)[1]
    assert np_mode[0].tolist()[0] == 0 and 2 not in np_mode[0].tolist()  # zero-area box dropped (ml.py:610)
    assert 2 in tf_mode[0].tolist()  # kept in TF style


def test_predict_from_head_matches_oracle_unpack(prior_boxes, oracle_tables):
    od = odk.ObjectDetector.load_voc((640, 640))
    t = oracle_tables('voc640')
    objs, clfs, locs = synth.predictions(3, 2, len(od.pb), 20)
    preds = od.predict_from_head(objs, clfs, locs, conf_threshold=0.25, strict_nms=False)
    ro, _ = R.predict_tail(t, objs, clfs, locs)
    ref = R.unpack_predictions(ro, 0.25)
    for p, (rc, rs, rb) in zip(preds, ref):
        assert p.classes.dtype == np.int32 and np.array_equal(p.classes, rc)
        assert np.array_equal(p.confs, rs) and np.array_equal(p.bboxes, rb)


def test_nms_guard_band_pairs_exact():
    """Pairs whose IoU sits within a few ulps of the threshold: the division-free fast test must hand
    them to the exact fp32 expression (ml.py:615-619), in both the kept-list and intra-chunk paths."""
    rng = np.random.RandomState(42)
    n_img, thr = 512, np.float32(0.45)
    boxes = np.zeros((n_img, 3, 4), np.float32)
    for i in range(n_img):
        w = np.float32(rng.uniform(0.2, 0.9))
        h = np.float32(rng.uniform(0.2, 0.9))
        # B = A cut at height t*h: IoU = t exactly in real arithmetic; nudge t by a few ulps around thr
        t = np.nextafter(thr, np.float32(1 if rng.rand() < 0.5 else 0), dtype=np.float32) if i % 3 else thr
        for _ in range(int(rng.randint(0, 4))):
            t = np.nextafter(t, np.float32(rng.choice([0, 1])), dtype=np.float32)
        boxes[i, 0] = [0, 0, w, h]
        boxes[i, 1] = [0, 0, w, np.float32(h * t)]
        boxes[i, 2] = [0, 0, np.float32(w * t), h]
    confs = np.tile(np.array([[0.9, 0.8, 0.7]], np.float32), (n_img, 1))
    classes = np.zeros((n_img, 3), np.int32)
    out, kept = odk.layers.nms()(1, 3, top_k=3)([classes, confs, boxes], return_index=True)
    decided_both_ways = set()
    for i in range(n_img):
        ro, rk = R.nms_layer_image(classes[i], confs[i], boxes[i], 1, top_k=3)
        assert np.array_equal(kept[i], rk), (i, boxes[i])
        decided_both_ways.add(int((rk >= 0).sum()))
    assert len(decided_both_ways) >= 2  # the set really straddles the threshold


# ----------------------------------------------------------------------------------------- VOC AP (parity unpinned: chainercv absent)
def _to_objects(gts, preds):
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, c, b, d) for (w, h, c, b, d) in gts]
    y_pred = [odk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in preds]
    return y_true, y_pred


def _assert_ap_equal(res, ref):
    for k in ('AP', 'AP_VOC'):
        assert np.array_equal(np.isnan(res[k]), np.isnan(ref[k])), k
        np.testing.assert_allclose(res[k], ref[k], rtol=1e-5, atol=1e-12, equal_nan=True)  # north_star: AP within 1e-5 relative
    for k in ('mAP', 'mAP_VOC'):
        np.testing.assert_allclose(res[k], ref[k], rtol=1e-5, equal_nan=True)


def test_voc_ap_micro_cases_on_gpu():
    f = np.float32
    gt = [(100, 100, np.array([0]), np.array([[.1, .1, .5, .5]], f), np.array([False]))]
    cases = [
        [(np.array([0]), np.array([.9], f), np.array([[.1, .1, .5, .5]], f))],                                  # perfect
        [(np.array([0]), np.array([.9], f), np.array([[.6, .6, .9, .9]], f))],                                  # miss
        [(np.array([0, 0]), np.array([.9, .8], f), np.array([[.1, .1, .5, .5], [.1, .1, .5, .5]], f))],         # duplicate
        [(np.array([0, 0]), np.array([.9, .8], f), np.array([[.6, .6, .9, .9], [.1, .1, .5, .5]], f))],         # FP first
        [(np.zeros((0,), np.int32), np.zeros((0,), f), np.zeros((0, 4), f))],                                   # no detections
    ]
    for preds in cases:
        y_true, y_pred = _to_objects(gt, preds)
        _assert_ap_equal(odk.voc.evaluate(y_true, y_pred, num_classes=2), R.voc_evaluate(gt, preds, 2))
    gt_d = [(100, 100, np.array([0, 0]), np.array([[.1, .1, .5, .5], [.6, .6, .9, .9]], f), np.array([True, False]))]
    y_true, y_pred = _to_objects(gt_d, cases[0])
    res = odk.voc.evaluate(y_true, y_pred, num_classes=2)
    _assert_ap_equal(res, R.voc_evaluate(gt_d, cases[0], 2))
    assert res['AP'][0] == 0.0 and np.isnan(res['AP'][1])


@pytest.mark.parametrize('n_img,classes,seed', [(300, 20, 2007), (4952, 20, 2007), (4952, 2, 11)])
def test_voc_ap_vs_oracle(n_img, classes, seed):
    """Config 4 shape (4952 images); C=2 pushes > 16384 detections into one class -> hybrid sort path."""
    gts, preds = synth.eval_set(seed, n_img, classes)
    y_true, y_pred = _to_objects(gts, preds)
    res = odk.voc.evaluate(y_true, y_pred, num_classes=classes)
    ref = R.voc_evaluate(gts, preds, classes)
    _assert_ap_equal(res, ref)
    assert 0.05 < res['mAP'] < 1.0 and 0.05 < res['mAP_VOC'] < 1.0


def test_voc_ap_from_nms_layer_output(prior_boxes, oracle_tables):
    """The evaluation path end to end on one GPU: head outputs -> decode+NMS -> AP, vs the oracle chain."""
    pb, t = prior_boxes('voc320'), oracle_tables('voc320')
    objs, clfs, locs = synth.predictions(9, 4, len(pb), 20, clustered=True)
    gts, _ = synth.eval_set(5, 4, 20)
    out = pb.decode_nms(objs, clfs, locs)
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, c, b, d) for (w, h, c, b, d) in gts]
    gt = odk.voc.pack_ground_truth(y_true, out.device)
    ap, ap07 = (x.cpu().numpy() for x in odk.voc.evaluate_detections(out, gt, 20, conf_threshold=0.01))
    ro, _ = R.predict_tail(t, objs, clfs, locs)
    ref = R.voc_evaluate(gts, R.unpack_predictions(ro, 0.01), 20)
    np.testing.assert_allclose(ap, ref['AP'], rtol=1e-5, atol=1e-12, equal_nan=True)
    np.testing.assert_allclose(ap07, ref['AP_VOC'], rtol=1e-5, atol=1e-12, equal_nan=True)
