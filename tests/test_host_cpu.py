"""Host-side logic that runs without a GPU: table construction, the C-ABI surface, loud failure."""
import ctypes
import hashlib
import pathlib
import re

import numpy as np
import pytest

import object_detector_b200 as odk
from object_detector_b200 import _lib, build

ROOT = pathlib.Path(__file__).resolve().parent.parent
TABLES = ['pb_grid', 'pb_locs', 'pb_centers', 'pb_sizes', 'pb_mask', 'pb_info_indices']


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope='module', autouse=True)
def built():
    build.build()


def header_symbols():
    text = (ROOT / 'include' / 'odk.h').read_text()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(odk_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = header_symbols()
    assert len(names) >= 18
    for name in names:
        assert hasattr(lib, name), f'{name} declared in include/odk.h but not exported by libodk.so'


def test_ctypes_table_covers_header():
    assert sorted(_lib.SIGNATURES) == header_symbols()
    assert _lib.version().startswith('odk ')


def test_workspace_size_query_needs_no_gpu():
    lib = _lib.load()
    small = lib.odk_nms_workspace_bytes(1, 16800, 20, 200)
    big = lib.odk_nms_workspace_bytes(64, 16800, 20, 200)
    assert 0 < small < big and big % 256 == 0
    assert lib.odk_nms_workspace_bytes(0, 16800, 20, 200) == 0


@pytest.mark.parametrize('name', ['voc320', 'voc640', 'tiny128', 'stress1120'])
def test_product_tables_bit_identical_to_reference(golden, prior_boxes, name):
    g, pb = golden['anchors'], prior_boxes(name)
    for k in TABLES:
        a = getattr(pb, k)
        assert str(a.dtype) == str(g[f'{name}/{k}/dtype']), (name, k)
        assert sha(a) == str(g[f'{name}/{k}/sha256']), (name, k)


def test_from_dict_errors_like_reference():
    pb = odk.PriorBoxes((320, 320), [40, 20, 10], 20)
    with pytest.raises(ValueError, match='PriorBoxes load error'):
        pb.from_dict({'map_sizes': [40]})  # od_pb.py:57-58
    assert pb.map_sizes.tolist() == [40]
    with pytest.raises(AssertionError):
        odk.PriorBoxes((320,), [40], 20)


def test_load_voc_and_version_migration():
    od = odk.ObjectDetector.load_voc((640, 640))
    assert len(od.pb) == 16800 and int(od.pb.pb_mask.sum()) == 12512
    old = {'version': '0.0.1', 'input_size': [320, 320], 'map_sizes': [40, 20, 10], 'num_classes': 20,
           'pb': {'pb_size_patterns': odk.od._VOC_JSON_DATA['pb_size_patterns']}}
    od1 = odk.ObjectDetector.load_from_dict(old)  # od.py:70-71
    assert len(od1.pb) == 16800
    assert set(od1.to_dict()) >= {'version', 'input_size', 'map_sizes', 'num_classes', 'pb_size_patterns'}


def test_annotation_types_validate_like_reference():
    a = odk.ml.ObjectsAnnotation('.', 768, 614, [1], np.array([[203, 255, 601, 355]]) / [768, 614, 768, 614])
    assert a.real_bboxes.tolist() == [[203, 255, 601, 355]]  # data_voc_test.py:21
    with pytest.raises(AssertionError):
        odk.ml.ObjectsAnnotation('.', 10, 10, [0], [[0.5, 0.5, 0.4, 0.6]])
    p = odk.ml.ObjectsPrediction([0], [0.9], [[0.5, 0.25, 0.75, 1.0]])
    assert p.get_real_bboxes(5, 10).tolist() == [[2, 2, 4, 10]]
    e = odk.ml.ObjectsAnnotation('.', 10, 10, [], [])
    assert e.bboxes.shape == (0, 4)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    pb = odk.ObjectDetector.load_voc().pb
    with pytest.raises(RuntimeError, match='no CPU path'):
        pb.encode_truth([odk.ml.ObjectsAnnotation('.', 10, 10, [0], [[0.1, 0.1, 0.5, 0.5]])])
    with pytest.raises(RuntimeError, match='no CPU path'):
        odk.ml.non_maximum_suppression(np.zeros((1, 4), np.float32), np.zeros((1,), np.float32))


def test_product_never_imports_oracle():
    for p in (ROOT / 'object_detector_b200').rglob('*.py'):
        assert 'oracle' not in re.sub(r'#.*', '', p.read_text()), f'{p} mentions the oracle'


def test_batch_coalescer_groups_concurrent_calls():
    """generator.BatchCoalescer: per-item calls from many threads are served by few batched calls, each caller
    gets its own result, and an exception of the batched function reaches every caller of that batch."""
    import threading
    from object_detector_b200.generator import BatchCoalescer
    seen = []

    def fn(items):
        seen.append(len(items))
        if any(i < 0 for i in items):
            raise ValueError('bad item')
        return [i * i for i in items]
    co = BatchCoalescer(fn, max_batch=16, max_wait_ms=50.0)
    out = {}

    def work(i):
        out[i] = co(i)
    threads = [threading.Thread(target=work, args=(i,)) for i in range(40)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert out == {i: i * i for i in range(40)}
    assert sum(seen) == 40 and max(seen) <= 16 and len(seen) < 40
    with pytest.raises(ValueError):
        co(-1)
    assert co(3) == 9  # still alive after an error
    co.close()


def test_batch_coalescer_per_item_failure_reaches_only_its_caller():
    """A ``_Failed`` entry in the batched result (what encode_truth_hook produces for a bad sample after retrying the
    micro-batch singly) raises in the caller of that item only — like the reference, where only the bad sample's
    encode_truth call asserts (od_pb.py:304,332,338)."""
    import threading
    from object_detector_b200.generator import BatchCoalescer, _Failed

    def fn(items):
        return [_Failed(AssertionError(f'bad {i}')) if i % 5 == 0 else i + 1 for i in items]
    co = BatchCoalescer(fn, max_batch=16, max_wait_ms=50.0)
    out, err = {}, {}

    def work(i):
        try:
            out[i] = co(i)
        except AssertionError as e:
            err[i] = str(e)
    threads = [threading.Thread(target=work, args=(i,)) for i in range(20)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    co.close()
    assert err == {i: f'bad {i}' for i in range(0, 20, 5)} and out == {i: i + 1 for i in range(20) if i % 5}


def test_coco_summarize_host_logic_matches_oracle():
    """coco.summarize (the host part of the COCO mirror: chainercv _summarize over COCOeval's arrays) against the
    oracle's summary of the same arrays, incl. absent labels and undefined (-1) entries; and the pycocotools
    parameter tables the mirror hands to odk_coco_ap equal the oracle's."""
    import numpy as np
    from object_detector_b200 import coco, synth
    from oracle import restated as R
    assert np.array_equal(coco.IOU_THRS, R.COCO_IOU_THRS) and np.array_equal(coco.REC_THRS, R.COCO_REC_THRS)
    assert np.array_equal(coco.AREA_RNG, np.array(R.COCO_AREA_RNG, dtype=np.float64)) and coco.MAX_DETS == R.COCO_MAX_DETS
    assert [k[0] for k in coco.KEYS] == [k[0] for k in R.COCO_KEYS]
    gts, preds = synth.coco_eval_set(5, 40, 6)
    gts = [(w, h, c + 2 * (c >= 3), b, a, cr) for (w, h, c, b, a, cr) in gts]      # labels 3, 4 never occur
    preds = [(c + 2 * (c >= 3), s, b) for (c, s, b) in preds]
    ref = R.coco_evaluate(gts, preds)
    res = coco.summarize(ref['precision'], ref['recall'], ref['existent_labels'])
    assert res['existent_labels'].tolist() == list(ref['existent_labels']) and 3 not in res['existent_labels']
    for key, *_ in coco.KEYS:
        np.testing.assert_array_equal(res[key], ref[key])
        np.testing.assert_array_equal(res['m' + key], ref['m' + key])
    assert np.isnan(res['ap/iou=0.50:0.95/area=all/max_dets=100'][3:5]).all()
