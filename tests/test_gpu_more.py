"""More GPU parity cases: class counts that select the other kernel templates, ties, extreme ground-truth
sizes for the scattered ignore mask, multi-round / uncached NMS, CUDA-graph replay and the host pipeline."""
import numpy as np
import pytest
import torch

import object_detector_b200 as odk
from object_detector_b200 import _lib, pipeline, synth
from oracle import restated as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('name', ['mid80', 'mid100', 'odd7'])
def test_other_class_counts_full_path(prior_boxes, oracle_tables, name):
    """C=80 (64-row pieces), C=100 (32-row pieces and encode tiles), C=7 (scalar arg-max, odd row width)."""
    pb, t = prior_boxes(name), oracle_tables(name)
    c, n = pb.num_classes, len(pb)
    objs, clfs, locs = synth.predictions(31, 3, n, c)
    classes, confs, boxes = pb.decode_score(objs, clfs, locs)
    rc, rs = R.score_anchors(t, objs, clfs)
    assert np.array_equal(classes.cpu().numpy(), rc) and np.array_equal(confs.cpu().numpy(), rs)
    assert np.array_equal(boxes.cpu().numpy(), R.decode_locs_f32(t, locs))
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs)
    assert np.array_equal(kept.cpu().numpy(), rk) and np.array_equal(out.cpu().numpy(), ro)
    gtb, gtc, gtn = synth.gt_batch(5, 4, 9, c, size=(0.1, 0.5))
    res = pb.encode_truth_batch(gtb, gtc, gtn)
    ref = R.encode_truth(t, [(gtc[i, :gtn[i]], gtb[i, :gtn[i]]) for i in range(4)])
    assert (res['status'].cpu().numpy() == 0).all()
    assert np.array_equal(res['y_true'].cpu().numpy(), ref)


def test_argmax_ties_first_maximum_wins(prior_boxes, oracle_tables):
    """Quantised class scores produce many exact ties: np.argmax semantics (layers.py:80) in both kernels."""
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    rng = np.random.RandomState(3)
    objs, clfs, locs = synth.predictions(9, 2, len(pb), 20)
    clfs = (rng.randint(0, 4, size=clfs.shape) / 4.0).astype(np.float32)
    objs = rng.uniform(0.05, 1.0, size=objs.shape).astype(np.float32)
    classes, confs, _ = pb.decode_score(objs, clfs, locs)
    rc, rs = R.score_anchors(t, objs, clfs)
    assert np.array_equal(classes.cpu().numpy(), rc) and np.array_equal(confs.cpu().numpy(), rs)
    # fused path on the same tied data: compare with the unfused layer fed by the dense kernel
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    out2, kept2 = odk.layers.nms()(20, len(pb))(list(pb.decode_score(objs, clfs, locs)), return_index=True)
    assert torch.equal(out, out2) and torch.equal(kept, kept2)


def _closed_form_valids(t, boxes, assigned):
    """SURVEY A-4: pb_valids = not(pb_mask and any_g IoU(g, a) >= 0.5) or assigned[a]."""
    iou = R.pairwise_iou(boxes, t.pb_locs)
    v = ~(t.pb_mask & (iou >= np.float32(0.5)).any(axis=0))
    v[assigned] = True
    return v


@pytest.mark.parametrize('name,seed', [('voc640', 1), ('voc320', 2), ('mid80', 3), ('stress1120', 4)])
def test_ignore_mask_scatter_extreme_sizes(prior_boxes, oracle_tables, name, seed):
    """The ignore mask is scattered per GT over a cell range derived from a necessary condition for
    IoU >= 0.5; it must equal the dense closed form for boxes from a few pixels up to the whole image,
    including boxes that match an anchor exactly."""
    pb, t = prior_boxes(name), oracle_tables(name)
    rng = np.random.RandomState(seed)
    batch, g = 6, 48
    gtb = np.zeros((batch, g, 4), np.float32)
    for i in range(batch):
        lo, hi = [(0.004, 0.05), (0.02, 0.3), (0.2, 1.0)][i % 3]
        # anchors themselves (clipped), jittered by up to ~1%: IoU close to 1 and around the cell borders
        pick = rng.choice(np.flatnonzero(t.pb_mask), 8, replace=False)
        a = np.clip(t.pb_locs[pick] + rng.uniform(-0.003, 0.003, size=(8, 4)).astype(np.float32), 0, 1)
        a = a[(a[:, 2] - a[:, 0] > 1e-3) & (a[:, 3] - a[:, 1] > 1e-3)]
        gtb[i] = np.concatenate([synth.gt_boxes(rng, g - len(a), size=(lo, hi)), a]).astype(np.float32)
    gtn = np.full((batch,), g, np.int32)
    gtc = rng.randint(0, pb.num_classes, size=(batch, g)).astype(np.int32)
    res = pb.encode_truth_batch(gtb, gtc, gtn, want_valids=True)
    status = res['status'].cpu().numpy()
    valids = res['pb_valids'].cpu().numpy()
    pbi = res['pb_indices'].cpu().numpy()
    checked = 0
    for i in range(batch):
        if status[i] != 0:  # a centre inside no valid cell: the reference raises as well
            with pytest.raises(AssertionError):
                R.assign_boxes(t, gtb[i])
            continue
        assert np.array_equal(valids[i], _closed_form_valids(t, gtb[i], pbi[i])), i
        checked += 1
    assert checked >= 3


def test_nms_many_rounds_and_large_top_k():
    """Heavy suppression with top_k above the band size: several select/sort/sweep rounds per image."""
    rng = np.random.RandomState(21)
    n = 6000
    c = 0.5 + rng.uniform(-0.08, 0.08, size=(n, 2))
    s = rng.uniform(0.05, 0.25, size=(n, 2))
    boxes = np.clip(np.concatenate([c - s / 2, c + s / 2], 1), 0, 1).astype(np.float32)
    scores = (rng.permutation(n).astype(np.float32) + 1) / (n + 1)
    for top_k, thr in [(200, 0.45), (700, 0.6), (1, 0.45), (2, 0.3)]:
        out = odk.ml.non_maximum_suppression(boxes, scores, top_k, thr)
        assert np.array_equal(out, R.numpy_nms_policy(boxes, scores, top_k, thr)), (top_k, thr)


def test_nms_equal_scores_need_low_key_bits():
    """All scores equal: the radix select has to descend into the class / anchor bits of the keys."""
    rng = np.random.RandomState(5)
    n = 3000
    c = rng.uniform(0.05, 0.95, size=(n, 2))
    boxes = np.clip(np.concatenate([c - 0.01, c + 0.01], 1), 0, 1).astype(np.float32)
    scores = np.full((n,), 0.5, np.float32)
    out = odk.ml.non_maximum_suppression(boxes, scores, 300, 0.45)
    assert np.array_equal(out, R.numpy_nms_policy(boxes, scores, 300, 0.45))


def test_nms_layer_single_class_flood(prior_boxes, oracle_tables):
    """Every candidate in ONE class (the worst case for the per-class warps) and > 8192 candidates per image
    (keys do not fit the shared-memory cache)."""
    pb, t = prior_boxes('voc640'), oracle_tables('voc640')
    objs, clfs, locs = synth.predictions(13, 1, len(pb), 20)
    clfs[:] = 0.01
    clfs[:, :, 7] = 0.81
    objs[:] = np.random.RandomState(1).uniform(0.2, 1.0, size=objs.shape).astype(np.float32)
    locs *= 0.3
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs)
    assert (rk >= 0).sum() > 50
    assert np.array_equal(kept.cpu().numpy(), rk) and np.array_equal(out.cpu().numpy(), ro)


def test_graph_replay_equals_eager(prior_boxes):
    pb = prior_boxes('voc640')
    gt = [torch.from_numpy(x).cuda() for x in synth.gt_batch(3, 8, 16, 20)]
    head = [torch.from_numpy(x).cuda() for x in synth.predictions(4, 8, len(pb), 20)]
    y_ref = pb.encode_truth_batch(*gt)['y_true']
    d_ref = pb.decode_nms(*head)
    y, d = torch.zeros_like(y_ref), torch.zeros_like(d_ref)
    for two in (True, False):
        g = pipeline.capture_step(pb, gt, head, y, d, two_streams=two)
        for _ in range(3):
            y.fill_(-1)
            d.fill_(-1)
            g.replay()
        torch.cuda.synchronize()
        assert torch.equal(y, y_ref) and torch.equal(d, d_ref)
    # new inputs in the same buffers: the graph reads them
    gt2 = synth.gt_batch(33, 8, 16, 20)
    for dst, src in zip(gt, gt2):
        dst.copy_(torch.from_numpy(src))
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(y, pb.encode_truth_batch(*gt)['y_true'])


@pytest.mark.parametrize('use_graph', [True, False])
def test_host_pipeline_results_in_order(prior_boxes, use_graph):
    pb = prior_boxes('voc320')
    batch, gmax = 4, 8
    pipe = pipeline.HostPipeline(pb, batch, gmax, use_graph=use_graph)
    sets = [synth.gt_batch(100 + i, batch, gmax, 20) + synth.predictions(200 + i, batch, len(pb), 20) for i in range(5)]
    got = []
    prev = None
    for i, s in enumerate(sets):
        src = [torch.from_numpy(x).pin_memory() for x in s] if i % 2 == 0 else list(s)  # pinned and pageable inputs
        slot = pipe.submit(*src)
        if prev is not None:
            y, d = pipe.result(prev)
            got.append((y.clone(), d.clone()))
        prev = slot
    y, d = pipe.result(prev)
    got.append((y.clone(), d.clone()))
    for s, (y, d) in zip(sets, got):
        assert torch.equal(y, pb.encode_truth_batch(*s[:3])['y_true'].cpu())
        assert torch.equal(d, pb.decode_nms(*s[3:]).cpu())
    assert pipe.h2d_bytes == sum(x.nbytes for x in sets[0])


def test_encode_truth_compact_expands_to_dense(prior_boxes):
    """odk_assign_encode_compact: valid bits + positive rows + indices, expanded on the host by od_pb.expand_truth, give
    the dense y_true of odk_assign_encode bit for bit — incl. crowded images, G = 0, a failed image and the bits beyond N."""
    from object_detector_b200.od_pb import expand_truth
    for name, batch, gmax in (('voc640', 12, 24), ('odd7', 5, 6), ('tiny128', 3, 3)):
        pb = prior_boxes(name)
        boxes, classes, counts = synth.gt_batch(61, batch, gmax, pb.num_classes)
        counts[1] = 0
        classes[2, 0] = pb.num_classes + 3  # bad class: the image fails as a whole (od_pb.py:304)
        dense = pb.encode_truth_batch(boxes, classes, counts, want_valids=True)
        comp = pb.encode_truth_compact(boxes, classes, counts)
        assert torch.equal(comp['status'], dense['status']) and int(dense['status'][2]) == 2
        assert torch.equal(comp['pb_indices'], dense['pb_indices']) and torch.equal(comp['gt_indices'], dense['gt_indices'])
        h = {k: v.cpu().numpy() for k, v in comp.items()}
        y = expand_truth(h['pos_rows'], h['valid_bits'], h['pb_indices'], counts, len(pb))
        assert np.array_equal(y, dense['y_true'].cpu().numpy())
        n = len(pb)
        bits = np.unpackbits(h['valid_bits'].view(np.uint8).reshape(batch, -1), axis=1, bitorder='little')
        assert np.array_equal(bits[:, :n].astype(bool), dense['pb_valids'].cpu().numpy()) and not bits[:, n:].any()
        assert not h['pos_rows'][2].any() and not bits[2].any()


@pytest.mark.parametrize('mode', ['compact', 'fused'])
def test_host_pipeline_compact_and_fused_modes(prior_boxes, mode):
    """HostPipeline's other two result contracts against the direct calls: 'compact' (valid bits + positive rows ->
    expand_truth == dense y_true) and 'fused' (GT in, loss terms + detections out, the network's outputs resident)."""
    from object_detector_b200.od_pb import expand_truth
    pb = prior_boxes('voc320')
    batch, gmax, n = 4, 8, len(pb)
    sets = [synth.gt_batch(300 + i, batch, gmax, 20) + synth.predictions(400 + i, batch, n, 20) for i in range(3)]
    resident = None
    if mode == 'fused':
        rng = np.random.RandomState(5)
        y_pred = torch.from_numpy(_head_like(rng, (batch, n, 27), 20)).cuda()
        resident = {'y_pred': y_pred, 'head': tuple(torch.from_numpy(x).cuda() for x in sets[0][3:])}
    pipe = pipeline.HostPipeline(pb, batch, gmax, mode=mode, resident=resident)
    assert pipe.d2h_bytes < 0.02 * batch * n * 27 * 4
    for s in sets:
        slot = pipe.submit(*(s if mode == 'compact' else s[:3]))
        a, d = pipe.result(slot)
        if mode == 'compact':
            y = expand_truth(a['pos_rows'].numpy(), a['valid_bits'].numpy(), a['pb_indices'].numpy(), s[2], n)
            assert np.array_equal(y, pb.encode_truth_batch(*s[:3])['y_true'].cpu().numpy())
            assert torch.equal(d, pb.decode_nms(*s[3:]).cpu())
        else:
            ref = pb.assign_loss_batch(*s[:3], resident['y_pred'])
            assert torch.equal(a, ref['terms'].cpu()) and torch.equal(d, pb.decode_nms(*resident['head']).cpu())
    ceiling = pipeline.HostPipeline(pb, batch, gmax, mode=mode, resident=resident, kernels=False)
    ceiling.result(ceiling.submit(*(sets[0] if mode == 'compact' else sets[0][:3])))  # copies only: runs, returns buffers


class _Guarded:
    """Device buffer with 4 KB canary zones (0xA5) on both sides of the payload: an out-of-bounds WRITE by a kernel into
    its neighbourhood shows up as a changed guard byte (compute-sanitizer is closed on this pool)."""
    PAD = 4096

    def __init__(self, shape, dtype, fill=None):
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        self.raw = torch.full((n + 2 * self.PAD,), 0xA5, dtype=torch.uint8, device='cuda')
        self.t = self.raw[self.PAD:self.PAD + n].view(dtype).view(shape)
        if fill is not None:
            self.t.copy_(torch.as_tensor(fill).to(dtype).reshape(shape))

    def intact(self):
        return bool((self.raw[:self.PAD] == 0xA5).all()) and bool((self.raw[-self.PAD:] == 0xA5).all())


@pytest.mark.parametrize('name,batch,gmax', [('voc640', 5, 33), ('odd7', 3, 5), ('mid80', 2, 9), ('tiny128', 7, 3), ('mid100', 3, 4)])
def test_c_abi_writes_stay_inside_caller_buffers(prior_boxes, name, batch, gmax):
    """Both fused entry points called straight through the C-ABI with every caller-owned buffer — inputs, outputs and the
    WORKSPACES, sized exactly as odk_*_workspace_bytes says — between canary zones: no guard byte may change, and the
    results equal the mirror's.  Odd batch sizes, ragged GT counts (incl. 0 and a crowded image), partial last tiles."""
    pb = prior_boxes(name)
    dev = torch.device('cuda', 0)
    h, lib = pb._handle(dev), _lib.load()
    n, c = len(pb), pb.num_classes
    w = 3 + c + 4
    boxes, classes, counts = synth.gt_batch(123, batch, gmax, c, crowded_every=2)
    counts[0] = 0
    objs, clfs, locs = synth.predictions(321, batch, n, c, clustered=True)
    g = {k: _Guarded(sh, dt, fill) for k, (sh, dt, fill) in {
        'boxes': ((batch, gmax, 4), torch.float32, boxes), 'classes': ((batch, gmax), torch.int32, classes), 'counts': ((batch,), torch.int32, counts),
        'y': ((batch, n, w), torch.float32, None), 'pbi': ((batch, gmax), torch.int64, None), 'gti': ((batch, gmax), torch.int64, None),
        'valids': ((batch, n), torch.uint8, None), 'status': ((batch,), torch.int32, None),
        'ws_a': ((lib.odk_assign_workspace_bytes(h, batch, gmax),), torch.uint8, None),
        'rows': ((batch, gmax, w), torch.float32, None), 'vbits': ((batch, (n + 31) // 32), torch.int32, None),
        'objs': ((batch, n, 1), torch.float32, objs), 'clfs': ((batch, n, c), torch.float32, clfs), 'locs': ((batch, n, 4), torch.float32, locs),
        'out': ((batch, 50, 6), torch.float32, None), 'kept': ((batch, 50), torch.int32, None),
        'ws_n': ((lib.odk_nms_strict_workspace_bytes(batch, n, c, 50),), torch.uint8, None)}.items()}
    st = torch.cuda.current_stream().cuda_stream
    P = lambda k: g[k].t.data_ptr()  # noqa: E731
    _lib.call('odk_assign_encode', h, batch, P('boxes'), P('classes'), P('counts'), gmax, P('y'), P('pbi'), P('gti'), P('valids'), P('status'),
              P('ws_a'), g['ws_a'].t.numel(), st)
    ref = pb.encode_truth_batch(boxes, classes, counts, want_valids=True)
    assert torch.equal(g['y'].t, ref['y_true']) and torch.equal(g['pbi'].t, ref['pb_indices']) and torch.equal(g['valids'].t.bool(), ref['pb_valids'])
    _lib.call('odk_assign_encode_compact', h, batch, P('boxes'), P('classes'), P('counts'), gmax, P('rows'), P('vbits'), P('pbi'), P('gti'),
              P('status'), P('ws_a'), g['ws_a'].t.numel(), st)
    for strict in (False, True):
        args = (h, batch, P('objs'), P('clfs'), P('locs'), 0.01, 0.45) + ((0.5,) if strict else ()) + (50, 1, P('out'), P('kept'), P('ws_n'),
                                                                                                 g['ws_n'].t.numel(), st)
        _lib.call('odk_decode_nms_strict' if strict else 'odk_decode_nms', *args)
        want = pb.decode_nms(objs, clfs, locs, top_k=50, nms_all_threshold=0.5 if strict else None)
        assert torch.equal(g['out'].t, want)
    torch.cuda.synchronize()
    bad = [k for k, v in g.items() if not v.intact()]
    assert not bad, f'guard zones overwritten around: {bad}'


def test_repeated_launches_are_bit_identical(prior_boxes):
    """Race detector of last resort (racecheck is closed on this pool): the kernels whose warps share shared-memory state
    — packed slot counters and empty barriers in score_emit, the taken bitmap of the batched greedy walk, the re-typed
    buffers of the strict NMS — must give bit-identical results on 25 back-to-back launches over the same inputs, eager
    and with several launches in flight on different streams."""
    pb = prior_boxes('voc640')
    gt = [torch.from_numpy(x).cuda() for x in synth.gt_batch(9, 24, 48, 20, crowded_every=3)]
    head = [torch.from_numpy(x).cuda() for x in synth.predictions(9, 24, len(pb), 20, clustered=True)]
    y0 = pb.encode_truth_batch(*gt)
    d0, k0 = pb.decode_nms(*head, return_index=True)
    s0 = pb.decode_nms(*head, nms_all_threshold=0.5)
    streams = [torch.cuda.Stream() for _ in range(3)]
    for it in range(25):
        with torch.cuda.stream(streams[it % 3]):
            y = pb.encode_truth_batch(*gt)
            d, k = pb.decode_nms(*head, return_index=True)
            s = pb.decode_nms(*head, nms_all_threshold=0.5)
        streams[it % 3].synchronize()
        assert torch.equal(y['y_true'], y0['y_true']) and torch.equal(y['pb_indices'], y0['pb_indices']), it
        assert torch.equal(d, d0) and torch.equal(k, k0) and torch.equal(s, s0), it


def _score_objects(seed, n_img, classes, **kw):
    gts, preds = synth.eval_set(seed, n_img, classes, **kw)
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, c, b, d) for (w, h, c, b, d) in gts]
    y_pred = [odk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in preds]
    return y_true, y_pred, [(c, b, d) for (_, _, c, b, d) in gts], preds


def test_compute_scores_golden_from_reference(golden):
    g = golden['scores']
    y_true, y_pred, _, _ = _score_objects(77, 60, 5, max_gt=5, n_fp=6)
    for i in range(3):
        p, r, f, s = odk.ml.compute_scores(y_true, y_pred, float(g[f'scores_{i}/conf_threshold']), 0.5, num_classes=5)
        assert np.array_equal(s, g[f'scores_{i}/supports'])
        assert np.array_equal(p, g[f'scores_{i}/precisions']) and np.array_equal(r, g[f'scores_{i}/recalls'])
        assert np.array_equal(f, g[f'scores_{i}/fscores'])
    assert odk.ml.search_conf_threshold(y_true, y_pred) == float(g['search_conf_threshold'])


def test_compute_scores_vs_oracle_voc_sized():
    """4952 images x 50 thresholds in one launch; counts are integers -> bit-exact scores."""
    y_true, y_pred, gt, pred = _score_objects(2007, 4952, 20)
    for th in (0.0, 0.31, 0.9):
        res = odk.ml.compute_scores(y_true, y_pred, th, 0.5)
        ref = R.scores_from_counts(*R.match_counts(gt, pred, th, 0.5, 20))
        for a, b in zip(res, ref):
            assert np.array_equal(a, b), th
    assert odk.ml.search_conf_threshold(y_true[:400], y_pred[:400]) == R.search_conf_threshold(gt[:400], pred[:400], 20)


def test_encode_truth_hook_coalesces_threads(prior_boxes):
    """The per-sample hook of the data pipeline (dl/od.py:169) called from worker threads like the reference's
    generator does (generator.py:160-166): same rows as the batched call, few launches."""
    import threading
    from object_detector_b200 import generator
    od = odk.ObjectDetector.load_voc((320, 320))
    boxes, classes, counts = synth.gt_batch(5, 24, 6, 20, crowded_every=0)
    anns = [odk.ml.ObjectsAnnotation('.', 100, 100, classes[i, :counts[i]], boxes[i, :counts[i]]) for i in range(24)]
    ref = od.pb.encode_truth(anns)
    hook = generator.encode_truth_hook(od, max_batch=16, max_wait_ms=100.0)
    out = [None] * 24

    def work(i):
        out[i] = hook(anns[i])
    threads = [threading.Thread(target=work, args=(i,)) for i in range(24)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert all(np.array_equal(out[i], ref[i]) for i in range(24))
    assert hook.coalescer.batches < 24 and hook.coalescer.items == 24
    assert np.array_equal(generator.collate_truth(od.pb, anns), ref)
    hook.coalescer.close()



def test_confusion_matrix_golden_and_oracle(golden):
    """ml.od_confusion_matrix (ml.py:397-451): golden matrices of the unmodified reference, then a VOC-sized set and
    equal-confidence ties against the oracle; an image with ground truth but no detection raises like ml.compute_iou."""
    g = golden['scores']
    y_true, y_pred, _, _ = _score_objects(77, 60, 5, max_gt=5, n_fp=6)
    for i in range(3):
        cm = odk.ml.od_confusion_matrix(y_true, y_pred, float(g[f'cm_{i}/conf_threshold']), 0.5, num_classes=5)
        assert cm.dtype == int and np.array_equal(cm, g[f'cm_{i}/out'])
    y_true, y_pred, _, _ = _score_objects(78, 40, 3, max_gt=6, n_fp=12)
    assert np.array_equal(odk.ml.od_confusion_matrix(y_true, y_pred, 0.1, 0.3, num_classes=3), g['cm_dense/out'])
    y_true, y_pred, gt, pred = _score_objects(2007, 1500, 20)
    for th, iou in ((0.0, 0.5), (0.31, 0.5), (0.2, 0.25)):
        assert np.array_equal(odk.ml.od_confusion_matrix(y_true, y_pred, th, iou), R.od_confusion_matrix(gt, pred, th, iou, 20))
    # quantised confidences -> many ties
    y_true, y_pred, gt, pred = _score_objects(5, 200, 4, max_gt=6, n_fp=10)
    pred = [(c, np.round(s * 8) / np.float32(8), b) for (c, s, b) in pred]
    y_pred = [odk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in pred]
    assert np.array_equal(odk.ml.od_confusion_matrix(y_true, y_pred, 0.2, 0.4, 4), R.od_confusion_matrix(gt, pred, 0.2, 0.4, 4))
    empty = odk.ml.ObjectsPrediction(np.zeros(0, np.int32), np.zeros(0, np.float32), np.zeros((0, 4), np.float32))
    with pytest.raises(AssertionError):
        odk.ml.od_confusion_matrix(y_true[:1], [empty], 0.0, 0.5, 4)


# ----------------------------------------------------------------------------------------- strict_nms
@pytest.mark.parametrize('name,batch,top_k,clustered', [('voc640', 3, 200, False), ('voc640', 2, 200, True),
                                                        ('voc320', 2, 20, True), ('tiny128', 4, 5, False),
                                                        ('mid80', 2, 200, False)])
def test_strict_nms_fused_vs_oracle(prior_boxes, oracle_tables, name, batch, top_k, clustered):
    """layers.py:940,953-957 (nms_all_threshold=0.5 as od_net.py:175 sets it): kept anchors and rows bit-exact."""
    pb, t = prior_boxes(name), oracle_tables(name)
    objs, clfs, locs = synth.predictions(99, batch, len(pb), pb.num_classes, clustered=clustered)
    out, kept = pb.decode_nms(objs, clfs, locs, top_k=top_k, nms_all_threshold=0.5, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs, top_k=top_k, nms_all_threshold=0.5)
    assert np.array_equal(kept.cpu().numpy(), rk)
    assert np.array_equal(out.cpu().numpy(), ro)
    # and it differs from the non-strict result where suppression across classes happens at all
    loose = pb.decode_nms(objs, clfs, locs, top_k=top_k)
    if clustered:
        assert not torch.equal(loose, out)


def test_strict_nms_layer_class_limit_and_spill(strict_cap_case):
    """(1) the per-class limit 2*top_k cuts before the all-class pass; (2) > 1024 class-stage survivors (the
    shared-memory part of that list) with heavy cross-class suppression: 1500 classes x 4 places."""
    boxes, classes, confs = strict_cap_case
    layer = odk.layers.nms()(2, len(boxes), top_k=2, nms_all_threshold=0.5)
    out, kept = layer([classes[None].astype(np.int32), confs[None], boxes[None]], return_index=True)
    ro, rk = R.nms_layer_image(classes, confs, boxes, 2, top_k=2, nms_all_threshold=0.5)
    assert kept[0].tolist() == rk.tolist() == [0, -1] and np.array_equal(out[0], ro)

    rng = np.random.RandomState(8)
    c, places = 1500, 4
    n = c * places
    base = np.array([[0.05, 0.05, 0.3, 0.3], [0.6, 0.1, 0.9, 0.35], [0.1, 0.6, 0.4, 0.9], [0.55, 0.55, 0.95, 0.95]], np.float32)
    boxes = (np.repeat(base, c, axis=0) + rng.normal(0, 0.004, size=(n, 4))).astype(np.float32)
    classes = np.tile(np.arange(c), places).astype(np.int32)
    confs = (rng.permutation(n).astype(np.float32) + 100) / np.float32(n + 200)
    layer = odk.layers.nms()(c, n, top_k=200, nms_all_threshold=0.5)
    out, kept = layer([classes[None], confs[None], boxes[None]], return_index=True)
    ro, rk = R.nms_layer_image(classes, confs, boxes, c, top_k=200, nms_all_threshold=0.5)
    assert np.array_equal(kept[0], rk) and np.array_equal(out[0], ro)
    assert 1 <= int((rk >= 0).sum()) <= 40  # a handful per place survive the class-agnostic pass


def test_strict_nms_full_band_of_class_survivors_multi_round():
    """Regression for an overrun the checks build found (csrc/nms.cu, suppression bit matrix of the all-class stage): 600
    places x 4 classes with identical boxes — the class stage keeps every box (no same-class overlap), so a band holds
    up to 512 class-stage survivors (a 32 KB bit matrix), and the all-class stage keeps one box per place, so reaching
    top_k = 200 takes several rounds, whose keys come from the shared-memory cache the matrix used to overwrite."""
    rng = np.random.RandomState(17)
    places = 600
    gx, gy = np.meshgrid(np.arange(30), np.arange(20))
    c = np.stack([gx.ravel(), gy.ravel()], axis=1)[:places] / np.array([30.0, 20.0]) + 0.01
    one = np.concatenate([c, c + np.array([0.025, 0.04])], axis=1).astype(np.float32)
    boxes = np.repeat(one, 4, axis=0)[None]                      # (1, 2400, 4): four identical boxes per place
    classes = np.tile(np.arange(4), places)[None].astype(np.int32)
    confs = rng.permutation(places * 4).astype(np.float32)[None] / (places * 4) * 0.9 + 0.05
    out, kept = odk.layers.nms()(4, places * 4, nms_all_threshold=0.5)([classes, confs, boxes], return_index=True)
    ro, rk = R.nms_layer_image(classes[0], confs[0], boxes[0], 4, nms_all_threshold=0.5)
    assert int((rk >= 0).sum()) == 200
    assert np.array_equal(kept[0], rk) and np.array_equal(out[0], ro)


def test_predict_from_head_strict_flag(prior_boxes, oracle_tables):
    od = odk.ObjectDetector.load_voc((320, 320))
    assert od.strict_nms is True  # od.py:81 default
    t = oracle_tables('voc320')
    objs, clfs, locs = synth.predictions(4, 2, len(od.pb), 20, clustered=True)
    preds = od.predict_from_head(objs, clfs, locs, conf_threshold=0.25)  # default: what load_voc() was given = strict
    ro, _ = R.predict_tail(t, objs, clfs, locs, nms_all_threshold=0.5)
    for p, (rc, rs, rb) in zip(preds, R.unpack_predictions(ro, 0.25)):
        assert np.array_equal(p.classes, rc) and np.array_equal(p.confs, rs) and np.array_equal(p.bboxes, rb)


# ----------------------------------------------------------------------------------------- losses
def _head_like(rng, shape, c):
    """A y_pred in the network's output layout: sigmoid objectness, softmax classes, free box offsets."""
    y = rng.normal(size=shape).astype(np.float32)
    y[..., :3] = 1 / (1 + np.exp(-y[..., :3] * 2))
    e = np.exp(y[..., 3:3 + c] * 2)
    y[..., 3:3 + c] = e / e.sum(axis=-1, keepdims=True)
    return y.astype(np.float32)


@pytest.mark.parametrize('name,batch,gmax', [('voc640', 4, 64), ('voc320', 3, 7), ('mid80', 2, 9), ('odd7', 3, 5), ('tiny128', 5, 3)])
def test_loss_terms_dense_and_fused_vs_oracle(prior_boxes, oracle_tables, name, batch, gmax):
    """od_pb.py:365-431.  Floating point: 1e-5 relative (BASELINE north_star) against the float32/float64 oracle;
    the fused assign+loss path (no y_true) must agree with the dense one to the same bar, its indices exactly."""
    pb, t = prior_boxes(name), oracle_tables(name)
    c = pb.num_classes
    gtb, gtc, gtn = synth.gt_batch(11, batch, gmax, c, size=(0.1, 0.5))
    gtn[0] = 0  # an image without ground truth: weights 1/max(1,0)
    enc = pb.encode_truth_batch(gtb, gtc, gtn)
    assert (enc['status'].cpu().numpy() == 0).all()
    y_true = enc['y_true']
    rng = np.random.RandomState(2)
    y_pred = _head_like(rng, tuple(y_true.shape), c)
    ref = R.od_losses(t, y_true.cpu().numpy(), y_pred)
    terms = pb.loss_terms(y_true, y_pred).cpu().numpy()
    fused = pb.assign_loss_batch(gtb, gtc, gtn, y_pred)
    assert torch.equal(fused['pb_indices'], enc['pb_indices']) and torch.equal(fused['gt_indices'], enc['gt_indices'])
    ft = fused['terms'].cpu().numpy()
    for k, col in enumerate(pb.LOSS_COLUMNS[:7]):
        np.testing.assert_allclose(terms[:, k], ref[col], rtol=1e-5, atol=1e-7, err_msg=col)
        np.testing.assert_allclose(ft[:, k], ref[col], rtol=1e-5, atol=1e-7, err_msg='fused ' + col)
    assert np.array_equal(terms[:, 7], gtn.astype(np.float64)) and np.array_equal(ft[:, 7], terms[:, 7])
    # the reference's call signatures
    np.testing.assert_allclose(pb.loss(y_true, y_pred).cpu().numpy(), ref['loss'], rtol=2e-5)
    got = [m(y_true, y_pred).cpu().numpy() for m in pb.metrics]
    for g, col in zip(got, ('loss_obj', 'loss_clf', 'loss_loc', 'rec_bg', 'rec_obj', 'acc_clf')):
        np.testing.assert_allclose(g, ref[col], rtol=2e-5, atol=1e-7, err_msg=col)


def test_loss_full_batch_properties(prior_boxes):
    """BASELINE config 3 at full size (B=256, <= 64 GT): perfect prediction -> ~0 loss and unit metrics; linear in
    the weights (scaling a term's inputs); deterministic across runs."""
    pb = prior_boxes('voc640')
    gtb, gtc, gtn = synth.gt_batch(7, 256, 64, 20)
    enc = pb.encode_truth_batch(gtb, gtc, gtn)
    ok = (enc['status'] == 0)
    y_true = enc['y_true']
    y_perfect = y_true.clone()
    y_perfect[:, :, 3:-4] += (y_true[:, :, 2:3] == 0) * 0.05  # negatives: a uniform softmax row instead of all-zero
    perfect = pb.loss_terms(y_true, y_perfect)
    assert bool((perfect[ok][:, 0] < 1e-3).all()) and bool((perfect[ok][:, 4:7] == 1).all())
    rng = np.random.RandomState(5)
    y_pred = torch.from_numpy(_head_like(rng, (256, len(pb), 27), 20)).cuda()
    a = pb.loss_terms(y_true, y_pred)
    assert torch.equal(a, pb.loss_terms(y_true, y_pred))  # fixed reduction order
    f = pb.assign_loss_batch(gtb, gtc, gtn, y_pred)['terms']
    assert torch.allclose(f[ok], a[ok], rtol=1e-9, atol=1e-12)
    assert bool(torch.isfinite(a).all()) and bool((a[:, 0] > 0).all())


# ----------------------------------------------------------------------------------------- BASELINE config 5 (dense-anchor stress)
def test_stress_config5_predict_side(prior_boxes, oracle_tables):
    """200k anchors x 80 classes, NMS IoU 0.45, top-k 200: two images against the oracle (kept anchors and rows
    bit-exact, candidate lists beyond the shared-memory key cache), then B=8 through size-independent properties."""
    pb, t = prior_boxes('stress1120'), oracle_tables('stress1120')
    n, c = len(pb), 80
    assert n == 205800
    objs, clfs, locs = synth.predictions(55, 2, n, c)
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs)
    assert np.array_equal(kept.cpu().numpy(), rk) and np.array_equal(out.cpu().numpy(), ro)
    so, sk = pb.decode_nms(objs, clfs, locs, nms_all_threshold=0.5, return_index=True)
    ro2, rk2 = R.predict_tail(t, objs, clfs, locs, nms_all_threshold=0.5)
    assert np.array_equal(sk.cpu().numpy(), rk2) and np.array_equal(so.cpu().numpy(), ro2)
    objs, clfs, locs = (torch.from_numpy(x).cuda() for x in synth.predictions(56, 8, n, c))
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    classes, confs, boxes = pb.decode_score(objs, clfs, locs)
    out2, kept2 = odk.layers.nms()(c, n)([classes, confs, boxes], return_index=True)
    assert torch.equal(out, out2) and torch.equal(kept, kept2)            # fused == dense decode + layer
    conf = out[:, :, 1]
    assert bool((conf[:, 1:] <= conf[:, :-1]).all()) and bool((kept >= 0).all())  # sorted, 200 kept everywhere
    for b in range(8):
        k = kept[b].long()
        assert k.unique().numel() == 200
        assert torch.equal(confs[b][k], conf[b]) and torch.equal(classes[b][k].float(), out[b, :, 0])
        assert torch.equal(boxes[b][k], out[b, :, 2:])
    assert torch.equal(pb.decode_nms(objs, clfs, locs), out)               # deterministic


# ----------------------------------------------------------------------------------------- empty and ragged inputs
def test_empty_and_ragged_inputs(prior_boxes, oracle_tables):
    """No ground truth, no candidate above the confidence threshold, an empty box set, one-element sets, a batch mixing
    G = 0 / 1 / Gmax: the reference's results for the degenerate shapes (od_pb.py:306 weight 1/max(1,0); layers.py:969
    all-zero rows; ml.py:589-625 on empty / single inputs)."""
    pb, t = prior_boxes('voc320'), oracle_tables('voc320')
    n = len(pb)
    # train side: ragged counts incl. zero
    gtb, gtc, gtn = synth.gt_batch(3, 5, 16, 20, crowded_every=0, fixed_g=16)
    gtn[:] = [0, 1, 16, 0, 7]
    res = pb.encode_truth_batch(gtb, gtc, gtn, want_valids=True)
    ref = R.encode_truth(t, [(gtc[i, :gtn[i]], gtb[i, :gtn[i]]) for i in range(5)])
    assert (res['status'].cpu().numpy() == 0).all() and np.array_equal(res['y_true'].cpu().numpy(), ref)
    y0 = res['y_true'][0].cpu().numpy()
    assert (y0[:, 0] == 1).all() and (y0[:, 1:] == 0).all() and bool(res['pb_valids'][0].all())  # G = 0: everything valid, weight 1
    assert (res['pb_indices'][0] == -1).all() and (res['pb_indices'][1, 1:] == -1).all()
    idx, gts, valid = pb._assign_boxes(np.zeros((0, 4), np.float32))
    assert idx.shape == (0,) and gts.shape == (0,) and valid.all() and idx.dtype == np.int64
    assert pb.encode_truth([]).shape == (0, n, 27)
    # predict side: nothing above the threshold
    objs, clfs, locs = synth.predictions(8, 2, n, 20)
    objs[:] = 0.001
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    assert bool((out == 0).all()) and bool((kept == -1).all())
    so = pb.decode_nms(objs, clfs, locs, nms_all_threshold=0.5)
    assert bool((so == 0).all())
    # exactly one candidate in one image
    objs[1, 777, 0] = 0.9
    out, kept = pb.decode_nms(objs, clfs, locs, return_index=True)
    ro, rk = R.predict_tail(t, objs, clfs, locs)
    assert np.array_equal(out.cpu().numpy(), ro) and np.array_equal(kept.cpu().numpy(), rk) and int((rk >= 0).sum()) in (0, 1)
    # NumPy NMS mirror on empty / single / all-degenerate sets
    assert odk.ml.non_maximum_suppression(np.zeros((0, 4), np.float32), np.zeros(0, np.float32)).shape == (0,)
    one = np.array([[0.1, 0.1, 0.4, 0.5]], np.float32)
    assert odk.ml.non_maximum_suppression(one, np.array([0.3], np.float32)).tolist() == [0]
    flat = np.array([[0.1, 0.1, 0.1, 0.5], [0.2, 0.3, 0.6, 0.3]], np.float32)  # zero width / zero height: skipped (ml.py:610)
    assert odk.ml.non_maximum_suppression(flat, np.array([0.3, 0.2], np.float32)).shape == (0,)
    assert R.numpy_nms(flat, np.array([0.3, 0.2], np.float32)).shape == (0,)


# ----------------------------------------------------------------------------------------- is_match / od_accuracy (ml.py:159-182, 343-349)
def test_od_accuracy_reference_kat_and_oracle():
    """The reference's own test (pytoolkit/tests/ml_test.py:116-155) verbatim, then random sets against the oracle."""
    gt = [odk.ml.ObjectsAnnotation('.', 100, 100, [0, 1], [[0.00, 0.00, 0.05, 0.05], [0.25, 0.25, 0.75, 0.75]])] * 6
    P = odk.ml.ObjectsPrediction
    pred = [
        P([1, 0], [1, 1], [[0.25, 0.25, 0.75, 0.75], [0.00, 0.00, 0.05, 0.05]]),
        P([1, 0, 0], [1, 0, 1], [[0.25, 0.25, 0.75, 0.75], [0.00, 0.00, 0.05, 0.05], [0.00, 0.00, 0.05, 0.05]]),
        P([1, 1], [1, 1], [[0.25, 0.25, 0.75, 0.75], [0.00, 0.00, 0.05, 0.05]]),
        P([1, 0, 0], [1, 1, 1], [[0.25, 0.25, 0.75, 0.75], [0.00, 0.00, 0.05, 0.05], [0.00, 0.00, 0.05, 0.05]]),
        P([1], [1], [[0.25, 0.25, 0.75, 0.75]]),
        P([1, 0], [1, 1], [[0.25, 0.25, 0.75, 0.75], [0.90, 0.90, 0.95, 0.95]]),
    ]
    for yt, yp, m in zip(gt, pred, [True, True, False, False, False, False]):
        assert yp.is_match(yt.classes, yt.bboxes, conf_threshold=0.5) == m
    assert odk.ml.od_accuracy(gt, pred, conf_threshold=0.5) == pytest.approx(2 / 6)
    # the reference's confusion-matrix KAT (ml_test.py:158-193)
    assert (odk.ml.od_confusion_matrix(np.array([]), np.array([]), num_classes=3) == np.zeros((4, 4), dtype=int)).all()
    cm = odk.ml.od_confusion_matrix(
        [odk.ml.ObjectsAnnotation('.', 100, 100, [1], [[0.25, 0.25, 0.75, 0.75]])],
        [P([0, 2, 1, 1, 2], [0, 1, 1, 1, 1], [[0.25, 0.25, 0.75, 0.75]] * 4 + [[0.95, 0.95, 0.99, 0.99]])],
        conf_threshold=0.5, num_classes=3)
    assert (cm == np.array([[0, 0, 0, 0], [0, 1, 0, 0], [0, 0, 0, 0], [0, 1, 2, 0]])).all()
    # random sets: mostly mismatches plus exact copies of the ground truth
    y_true, y_pred, gt_l, pred_l = _score_objects(9, 300, 6, max_gt=4, n_fp=2)
    for i in range(0, 300, 3):
        c, b, _ = gt_l[i]
        pred_l[i] = (c.copy(), np.full(len(c), 0.9, np.float32), b.copy())
        y_pred[i] = P(*pred_l[i])
    for thr in (0.0, 0.4):
        want = [R.exact_match(c, b, p, thr, 0.5) for (c, b, _), p in zip(gt_l, pred_l)]
        assert 0.2 < np.mean(want) < 0.9
        assert odk.ml.od_accuracy(y_true, y_pred, thr, 0.5) == np.mean(want)


@pytest.mark.parametrize('name,batch,gmax', [('voc640', 3, 24), ('mid80', 2, 9), ('odd7', 3, 5), ('tiny128', 4, 3)])
def test_loss_gradient_vs_oracle_and_autograd(prior_boxes, oracle_tables, name, batch, gmax):
    """odk_loss_grad against the float64 analytic oracle (itself checked against torch.autograd on the CPU): fp32
    kernel arithmetic -> 2e-5 relative to the largest gradient magnitude of each column group; clipped entries give an
    exact zero; the autograd wrapper hands the same tensor to backward()."""
    pb, t = prior_boxes(name), oracle_tables(name)
    c = pb.num_classes
    gtb, gtc, gtn = synth.gt_batch(13, batch, gmax, c, size=(0.1, 0.5))
    enc = pb.encode_truth_batch(gtb, gtc, gtn)
    y_true = enc['y_true']
    rng = np.random.RandomState(6)
    y_pred = _head_like(rng, tuple(y_true.shape), c)
    y_pred[0, 3, 2] = 1.0  # negative row with p_t clipped to eps: no gradient
    up = rng.uniform(0.5, 2.0, size=batch).astype(np.float32)
    ref = R.od_loss_grad(t, y_true.cpu().numpy(), y_pred, up)
    got = pb.loss_grad(y_true, y_pred, up).cpu().numpy()
    assert got.dtype == np.float32 and got.shape == ref.shape
    assert (got[:, :, :2] == 0).all() and got[0, 3, 2] == 0
    for sl in (slice(2, 3), slice(3, 3 + c), slice(3 + c, None)):
        scale = np.abs(ref[:, :, sl]).max()
        assert scale > 0
        np.testing.assert_allclose(got[:, :, sl], ref[:, :, sl], rtol=2e-5, atol=2e-5 * scale)
    # fused: assignment + forward + gradient without y_true
    fused = pb.assign_loss_batch(gtb, gtc, gtn, y_pred, with_grad=True, upstream=up)
    fg = fused['grad'].cpu().numpy()
    for sl in (slice(2, 3), slice(3, 3 + c), slice(3 + c, None)):
        scale = np.abs(ref[:, :, sl]).max()
        np.testing.assert_allclose(fg[:, :, sl], ref[:, :, sl], rtol=2e-5, atol=2e-5 * scale)
    assert (fg[:, :, :2] == 0).all()
    assert torch.equal(fused['terms'], pb.assign_loss_batch(gtb, gtc, gtn, y_pred)['terms'])
    yp = torch.from_numpy(y_pred).cuda().requires_grad_(True)
    loss = pb.loss_with_grad(y_true, yp)
    (loss * torch.from_numpy(up).cuda()).sum().backward()
    assert torch.equal(yp.grad, torch.from_numpy(got).cuda())
    np.testing.assert_allclose(loss.detach().cpu().numpy(), R.od_losses(t, y_true.cpu().numpy(), y_pred)['loss'], rtol=2e-5)


# ----------------------------------------------------------------------------------------- concurrency and randomised sweeps
def test_concurrent_host_threads_on_own_streams(prior_boxes, oracle_tables):
    """The reference calls encode_truth from a pool of host threads (generator.py:160-166).  Eight threads, each on its
    own CUDA stream, hammer both sides of the path with different inputs; every result must equal the serial one
    (no global mutable state in the library, per-thread side stream and error slot)."""
    import threading
    pb = prior_boxes('voc320')
    n = len(pb)
    jobs = []
    for k in range(8):
        gt = synth.gt_batch(100 + k, 6, 12, 20, crowded_every=3)
        head = synth.predictions(200 + k, 3, n, 20, clustered=bool(k % 2))
        jobs.append((gt, head))
    want = [(pb.encode_truth_batch(*gt)['y_true'].clone(), pb.decode_nms(*head).clone()) for gt, head in jobs]
    torch.cuda.synchronize()
    got, errors = [None] * len(jobs), []

    def work(k):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(5):
                    y = pb.encode_truth_batch(*jobs[k][0])['y_true']
                    d = pb.decode_nms(*jobs[k][1], nms_all_threshold=0.5 if k == 3 else None)
                    d = pb.decode_nms(*jobs[k][1])
                st.synchronize()
            got[k] = (y, d)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=work, args=(k,)) for k in range(len(jobs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for (wy, wd), (gy, gd) in zip(want, got):
        assert torch.equal(wy, gy) and torch.equal(wd, gd)


def test_randomised_small_cases_vs_oracle(prior_boxes, oracle_tables):
    """A few hundred small random problems per side: odd ground-truth counts, tiny and border-touching boxes, random
    thresholds and top_k, equal boxes — every assignment and every kept set equals the oracle's."""
    rng = np.random.RandomState(20261104)
    pb, t = prior_boxes('tiny128'), oracle_tables('tiny128')
    checked = 0
    for trial in range(120):
        g = int(rng.randint(1, 9))
        size = [(0.02, 0.1), (0.05, 0.5), (0.3, 0.95)][trial % 3]
        boxes = synth.gt_boxes(rng, g, size=size)
        if trial % 5 == 0:
            boxes[0, :2] = 0.0      # touches the top-left border
        if trial % 7 == 0:
            boxes[-1, 2:] = 1.0     # touches the bottom-right border
        boxes = boxes.astype(np.float32)
        try:
            ri, rg, rv = R.assign_boxes(t, boxes)
        except AssertionError:
            with pytest.raises(AssertionError):
                pb._assign_boxes(boxes)
            continue
        gi, gg, gv = pb._assign_boxes(boxes)
        assert np.array_equal(gi, ri) and np.array_equal(gg, rg) and np.array_equal(gv, rv), trial
        checked += 1
    assert checked >= 60
    for trial in range(120):
        n = int(rng.randint(1, 400))
        hubs = rng.uniform(0.2, 0.8, size=(int(rng.randint(1, 6)), 2))
        ctr = hubs[rng.randint(0, len(hubs), size=n)]
        ctr = ctr + rng.normal(0, [0.01, 0.05, 0.2][trial % 3], size=(n, 2))
        wh = rng.uniform(0.02, 0.4, size=(n, 2))
        b = np.clip(np.concatenate([ctr - wh / 2, ctr + wh / 2], 1), 0, 1).astype(np.float32)
        if n > 3:
            b[1] = b[0]             # identical boxes
            b[2, 2] = b[2, 0]       # zero width
        s = (rng.permutation(n).astype(np.float32) + 1) / np.float32(n + 1)
        top_k = int(rng.choice([1, 2, 5, 50, 200]))
        thr = float(rng.choice([0.3, 0.45, 0.5, 0.7]))
        assert np.array_equal(odk.ml.non_maximum_suppression(b, s, top_k, thr), R.numpy_nms_policy(b, s, top_k, thr)), trial


# ----------------------------------------------------------------------------------------- box side of the augmentation (f2)
def _augment_batch(samples):
    gmax = max(len(s[3]) for s in samples)
    boxes = np.zeros((len(samples), gmax, 4), np.float32)
    classes = np.zeros((len(samples), gmax), np.int32)
    count = np.zeros(len(samples), np.int32)
    for i, (_, _, b, c, _) in enumerate(samples):
        boxes[i, :len(c)], classes[i, :len(c)], count[i] = b, c, len(c)
    sizes = np.array([[s[0], s[1]] for s in samples], np.int64)
    return boxes, classes, count, sizes


@pytest.mark.parametrize('name', sorted(synth.AUGMENT_CONFIGS))
def test_box_augmenter_matches_reference_vectors(golden, name):
    """odk_bbox_zoom + odk_bbox_flip_rot over a whole batch against the vectors of the unmodified reference
    (tests/golden/augment.npz): surviving boxes bit-exact in fp64, classes, counts, and every sample's RandomState left
    where the reference leaves it."""
    from object_detector_b200.bbox_augment import BoxAugmenter
    size, keep_aspect, kw = synth.AUGMENT_CONFIGS[name]
    samples = synth.augment_inputs(911, 160)
    boxes, classes, count, sizes = _augment_batch(samples)
    rands = [np.random.RandomState(s[4]) for s in samples]
    out = BoxAugmenter(size, keep_aspect=keep_aspect, **kw)(boxes, classes, count, sizes, rands)
    g = golden['augment']
    assert np.array_equal(out['gt_count'].cpu().numpy(), g[name + '_counts'])
    b64, b32, cls = out['boxes64'].cpu().numpy(), out['gt_boxes'].cpu().numpy(), out['gt_classes'].cpu().numpy()
    at = 0
    for i, n in enumerate(g[name + '_counts']):
        assert np.array_equal(b64[i, :n], g[name + '_boxes'][at:at + n]), (name, i)
        assert np.array_equal(cls[i, :n], g[name + '_classes'][at:at + n])
        assert np.array_equal(b32[i, :n], g[name + '_boxes'][at:at + n].astype(np.float32))
        assert not b64[i, n:].any() and not cls[i, n:].any()
        assert rands[i].rand() == g[name + '_next_draw'][i]
        at += n


def test_box_augmenter_vs_oracle_and_feeds_encode_truth(prior_boxes, oracle_tables):
    """Fresh seeds (not in the vectors), larger ground-truth counts (more than one warp pass), then straight into
    encode_truth_batch on the device: every augmented image still encodes without an error status."""
    from object_detector_b200.bbox_augment import BoxAugmenter
    rng = np.random.RandomState(31)
    samples = []
    for i in range(96):
        g = int(rng.randint(1, 70))
        samples.append((int(rng.randint(200, 700)), int(rng.randint(200, 700)), synth.gt_boxes(rng, g, size=(0.05, 0.3)),
                        rng.randint(0, 20, size=g).astype(np.int32), int(rng.randint(0, 2 ** 31))))
    boxes, classes, count, sizes = _augment_batch(samples)
    rands = [np.random.RandomState(s[4]) for s in samples]
    aug = BoxAugmenter(320, keep_aspect=True, min_object_px=4)
    out = aug(boxes, classes, count, sizes, rands)
    cnt, b64, cls = out['gt_count'].cpu().numpy(), out['boxes64'].cpu().numpy(), out['gt_classes'].cpu().numpy()
    for i, (w, h, b, c, seed) in enumerate(samples):
        rand = np.random.RandomState(seed)
        ob, oc, info = R.augment_boxes_sample(rand, w, h, b, c, 320, keep_aspect=True, min_object_px=4)
        assert cnt[i] == len(oc) and np.array_equal(b64[i, :cnt[i]], ob.astype(np.float64)) and np.array_equal(cls[i, :cnt[i]], oc)
        assert info['chosen'] == out['plan']['chosen'][i] and info['k'] == out['plan']['k'][i]
        assert rand.rand() == rands[i].rand()
    pb = prior_boxes('voc320')
    res = pb.encode_truth_batch(out['gt_boxes'], out['gt_classes'], out['gt_count'])
    status = res['status'].cpu().numpy()
    live = cnt > 0
    assert (status[live] == 0).mean() > 0.9     # a zoom can push a centre out of every cell; the reference raises there too
    ok = np.nonzero(live & (status == 0))[0][:8]
    t = oracle_tables('voc320')
    for i in ok:
        want = R.encode_truth(t, [(cls[i, :cnt[i]], out['gt_boxes'][i, :cnt[i]].cpu().numpy())])[0]
        assert np.array_equal(res['y_true'][i].cpu().numpy(), want)


@pytest.mark.gpu
@pytest.mark.parametrize('use_graph,lanes,group', [(True, 4, 1), (True, 2, 2), (True, 1, 4), (False, 2, 2)])
def test_step_ring_batches_in_flight(use_graph, lanes, group):
    """pipeline.StepRing: several batches in flight (groups of slots captured into one graph with a branch per slot,
    groups alternating on streams) give, slot by slot, exactly the results of the two separate calls (every slot
    owns its outputs and workspaces; reuse of a slot is ordered by its lane).  13 steps: whole groups + a remainder."""
    import torch
    import object_detector_b200 as odk
    from object_detector_b200 import pipeline, synth
    B, GMAX, TOPK, SLOTS = 8, 32, 200, 4
    pb = odk.ObjectDetector.load_voc((640, 640)).pb
    n, c = len(pb.pb_locs), pb.num_classes
    dev = torch.device('cuda', 0)
    slots, want = [], []
    for k in range(SLOTS):
        gt = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(500 + k, B, GMAX, c))
        head = tuple(torch.from_numpy(x).to(dev) for x in synth.predictions(600 + k, B, n, c))
        want.append((pb.encode_truth_batch(*gt)['y_true'].clone(), pb.decode_nms(*head, top_k=TOPK).clone()))
        slots.append((gt, head, torch.full((B, n, 3 + c + 4), 7.0, device=dev), torch.full((B, TOPK, 6), 7.0, device=dev)))
    ring = pipeline.StepRing(pb, slots, lanes=lanes, group=group, top_k=TOPK, use_graph=use_graph)
    ring.fork()
    ring.run(3 * SLOTS + 1)
    ring.join()
    torch.cuda.synchronize()
    for (gt, head, y, d), (wy, wd) in zip(slots, want):
        assert torch.equal(y, wy) and torch.equal(d, wd)
    if use_graph:
        # the same run as ONE graph (capture_run): a single host call, same results in every slot
        for _, _, y, d in slots:
            y.fill_(7.0)
            d.fill_(7.0)
        ring.capture_run(3 * SLOTS + 1)
        ring.fork()
        ring.run(3 * SLOTS + 1)
        ring.join()
        torch.cuda.synchronize()
        for (gt, head, y, d), (wy, wd) in zip(slots, want):
            assert torch.equal(y, wy) and torch.equal(d, wd)
    if use_graph and group == 2:
        # two groups chained inside one graph (one host launch per 4 batches, 2 in flight): 11 steps = 2 chained graphs + 1 group + 1 step
        for chain_barrier in (True, False):  # False: no barrier between the groups of a chain (a chain of steps per branch)
            for _, _, y, d in slots:
                y.fill_(7.0)
                d.fill_(7.0)
            chained = pipeline.StepRing(pb, slots, lanes=1, group=2, chain=2, top_k=TOPK, chain_barrier=chain_barrier)
            chained.fork()
            chained.run(2 * SLOTS + 3)
            chained.join()
            torch.cuda.synchronize()
            for (gt, head, y, d), (wy, wd) in zip(slots, want):
                assert torch.equal(y, wy) and torch.equal(d, wd)


# ----------------------------------------------------------------------------------------- COCO metrics (parity unpinned)
def _coco_objects(gts, preds):
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, c, b, areas=a, crowdeds=cr) for (w, h, c, b, a, cr) in gts]
    y_pred = [odk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in preds]
    return y_true, y_pred


def _assert_coco_equal(res, ref):
    from object_detector_b200 import coco
    assert res['existent_labels'].tolist() == list(ref['existent_labels'])
    for key, *_ in coco.KEYS:
        assert np.array_equal(np.isnan(res[key]), np.isnan(ref[key])), key
        np.testing.assert_allclose(res[key], ref[key], rtol=1e-5, atol=1e-12, equal_nan=True)  # north_star: AP within 1e-5 relative
        np.testing.assert_allclose(res['m' + key], ref['m' + key], rtol=1e-5, equal_nan=True)


@pytest.mark.gpu
@pytest.mark.parametrize('n_img,classes,seed,n_fp', [(60, 5, 1, 20), (400, 20, 2017, 20), (40, 2, 3, 150)])
def test_coco_ap_vs_oracle(n_img, classes, seed, n_fp):
    """odk_coco_ap (matching per image x class x area range x IoU threshold, per-class sort, precision / recall
    arrays) vs the pycocotools restatement, incl. crowd GT, annotation areas, tied scores and — n_fp=150 with two
    classes — more than maxDets = 100 detections of one class in an image."""
    import torch
    from object_detector_b200 import coco, voc
    R = pytest.importorskip('oracle.restated')
    gts, preds = synth.coco_eval_set(seed, n_img, classes, n_fp=n_fp)
    y_true, y_pred = _coco_objects(gts, preds)
    ref = R.coco_evaluate(gts, preds)
    res = coco.evaluate(y_true, y_pred)
    _assert_coco_equal(res, ref)
    assert 0.02 < res['map/iou=0.50:0.95/area=all/max_dets=100'] < 1.0
    # the raw COCOeval arrays, bit for bit (same fp64 operations in the same order)
    dev = torch.device('cuda', 0)
    labels = [int(x) for x in ref['existent_labels']]
    prec, rec = coco.evaluate_detections(voc.pack_predictions(y_pred, dev), coco.pack_ground_truth(y_true, dev), labels[-1] + 1)
    assert np.array_equal(prec.cpu().numpy()[:, :, labels], ref['precision'])
    assert np.array_equal(rec.cpu().numpy()[:, labels], ref['recall'])


@pytest.mark.gpu
def test_coco_ap_micro_cases_on_gpu():
    from object_detector_b200 import coco
    R = pytest.importorskip('oracle.restated')
    f = np.float32
    gts = [(100, 100, np.array([0, 0]), np.array([[.1, .1, .5, .5], [.6, .6, .9, .9]], f), np.array([1600, 900], f), np.array([True, False]))]
    cases = [
        [(np.array([0, 0, 0, 0]), np.array([.9, .8, .7, .6], f), np.array([[.1, .1, .5, .5], [.1, .1, .5, .5], [.6, .6, .9, .9], [.6, .6, .9, .9]], f))],
        [(np.array([0]), np.array([.9], f), np.array([[.6, .6, .9, .8]], f))],
        [(np.zeros((0,), np.int32), np.zeros((0,), f), np.zeros((0, 4), f))],
        [(np.array([3]), np.array([.9], f), np.array([[.6, .6, .9, .9]], f))],
    ]
    for preds in cases:
        y_true, y_pred = _coco_objects(gts, preds)
        _assert_coco_equal(coco.evaluate(y_true, y_pred), R.coco_evaluate(gts, preds))
    # areas / crowdeds absent (VOC-style annotations): box area, no crowd
    y_true = [odk.ml.ObjectsAnnotation('.', 100, 100, gts[0][2], gts[0][3])]
    y_pred = [odk.ml.ObjectsPrediction(*cases[0][0])]
    ref = R.coco_evaluate([(100, 100, gts[0][2], gts[0][3], np.array([1600., 900.]), np.array([False, False]))], cases[0])
    _assert_coco_equal(coco.evaluate(y_true, y_pred), ref)
