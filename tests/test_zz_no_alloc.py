"""odk_nms_static — the allocation-free variant of odk_nms (scratch in a static device array inside libodk.so).

No counterpart in the reference: the entry point was added for callers that cannot
call cudaMalloc.  Collected last (file name), so that a failure here cannot hide another test under ``pytest -x``.
Every check goes through the C-ABI and compares with the existing entry point ``odk_nms`` on the same inputs; the work
is index work, so the bar is bit-exact (as in the other NMS tests), on two calls in a row.
"""
import numpy as np
import pytest
import torch

from object_detector_b200 import _dev, _lib
from oracle import restated as R

pytestmark = pytest.mark.gpu


def _boxes(seed, n, clustered):
    rng = np.random.RandomState(seed)
    if clustered:  # groups of near-duplicates: many suppressions, several select rounds
        k = max(n // 40, 1)
        cxy = rng.uniform(0.15, 0.85, size=(k, 2))[rng.randint(0, k, size=n)] + rng.normal(0, 0.01, size=(n, 2))
        wh = rng.uniform(0.08, 0.3, size=(k, 2))[rng.randint(0, k, size=n)] * rng.uniform(0.9, 1.1, size=(n, 2))
    else:
        cxy = rng.uniform(0.05, 0.95, size=(n, 2))
        wh = rng.uniform(0.02, 0.25, size=(n, 2))
    boxes = np.concatenate([cxy - wh / 2, cxy + wh / 2], axis=1).astype(np.float32)
    scores = rng.permutation(n).astype(np.float32) / np.float32(n) + np.float32(0.001)  # distinct
    return boxes, scores.astype(np.float32)


def _run(entry, boxes, scores, top_k, iou_thr, dev):
    """One call of `entry` ('odk_nms' with a caller workspace, or 'odk_nms_static') -> kept indices (int32 array)."""
    b = torch.from_numpy(boxes).to(dev)
    s = torch.from_numpy(scores).to(dev)
    n, cap = len(scores), max(int(top_k), 2)
    out = torch.full((cap,), -7, dtype=torch.int32, device=dev)
    cnt = torch.full((1,), -7, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        if entry == 'odk_nms':
            ws_bytes = _lib.load().odk_nms_workspace_bytes(1, max(n, 1), 1, int(top_k))
            ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
            _lib.call('odk_nms', b.data_ptr(), s.data_ptr(), n, int(top_k), float(iou_thr), 1, out.data_ptr(), cnt.data_ptr(),
                      ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
        else:
            _lib.call('odk_nms_static', b.data_ptr(), s.data_ptr(), n, int(top_k), float(iou_thr), 1, out.data_ptr(),
                      cnt.data_ptr(), _dev.stream_ptr(dev))
        torch.cuda.synchronize(dev)
    k = int(cnt.item())
    assert 0 <= k <= cap
    return out[:k].cpu().numpy()


def test_nms_static_matches_odk_nms_two_calls_in_a_row():
    dev = _dev.device()
    assert _lib.load().odk_nms_static_max_boxes() == 65536
    # second call: fewer boxes than the first, so the array still holds keys and counts of the first call behind them;
    # third: the first input again (the result must not depend on what the array held)
    cases = [(11, 20000, True, 200, 0.45), (12, 3000, False, 50, 0.3), (11, 20000, True, 200, 0.45), (13, 1, False, 200, 0.45)]
    first = None
    for seed, n, clustered, top_k, thr in cases:
        boxes, scores = _boxes(seed, n, clustered)
        ref = _run('odk_nms', boxes, scores, top_k, thr, dev)
        got = _run('odk_nms_static', boxes, scores, top_k, thr, dev)
        assert got.dtype == ref.dtype and np.array_equal(got, ref), (seed, n)
        if n <= 3000:  # and the oracle (the reference's NumPy NMS policy) on the sizes it finishes quickly
            assert got.tolist() == R.numpy_nms_policy(boxes, scores, top_k, thr).tolist()
        if first is None:
            first = got
        elif seed == 11:
            assert np.array_equal(got, first)
    assert len(first) > 2


def test_nms_static_back_to_back_on_one_stream_and_no_device_allocation():
    """Two calls queued back to back on one stream without a synchronisation in between (ordered, hence no race), and the
    free device memory does not move across the calls."""
    dev = _dev.device()
    b1, s1 = _boxes(21, 16800, True)
    b2, s2 = _boxes(22, 9000, False)
    ref1, ref2 = _run('odk_nms', b1, s1, 200, 0.45, dev), _run('odk_nms', b2, s2, 200, 0.45, dev)
    _run('odk_nms_static', b1, s1, 200, 0.45, dev)  # warm-up: lazy module loading, symbol address look-up
    tb1, ts1, tb2, ts2 = (torch.from_numpy(x).to(dev) for x in (b1, s1, b2, s2))
    out = torch.full((2, 200), -7, dtype=torch.int32, device=dev)
    cnt = torch.full((2,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)
    free_before, _ = torch.cuda.mem_get_info(dev)
    with torch.cuda.device(dev):
        st = _dev.stream_ptr(dev)
        _lib.call('odk_nms_static', tb1.data_ptr(), ts1.data_ptr(), len(s1), 200, 0.45, 1, out[0].data_ptr(), cnt[0:].data_ptr(), st)
        _lib.call('odk_nms_static', tb2.data_ptr(), ts2.data_ptr(), len(s2), 200, 0.45, 1, out[1].data_ptr(), cnt[1:].data_ptr(), st)
        torch.cuda.synchronize(dev)
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after >= free_before
    k = cnt.cpu().numpy()
    assert np.array_equal(out[0, :k[0]].cpu().numpy(), ref1) and np.array_equal(out[1, :k[1]].cpu().numpy(), ref2)


def test_nms_static_limits_and_empty_input():
    dev = _dev.device()
    cap = _lib.load().odk_nms_static_max_boxes()
    n = cap + 1
    b = torch.zeros((n, 4), dtype=torch.float32, device=dev)
    s = torch.zeros((n,), dtype=torch.float32, device=dev)
    out = torch.zeros((200,), dtype=torch.int32, device=dev)
    cnt = torch.full((1,), -7, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        with pytest.raises(_lib.OdkError) as e:
            _lib.call('odk_nms_static', b.data_ptr(), s.data_ptr(), n, 200, 0.45, 1, out.data_ptr(), cnt.data_ptr(), _dev.stream_ptr(dev))
        assert e.value.code != 0 and 'at most 65536 boxes' in str(e.value)
        assert int(cnt.item()) == -7  # nothing was launched
        # n = 0: ml.py:589-625 on an empty set -> empty result, as odk_nms
        _lib.call('odk_nms_static', b.data_ptr(), s.data_ptr(), 0, 200, 0.45, 1, out.data_ptr(), cnt.data_ptr(), _dev.stream_ptr(dev))
        assert int(cnt.item()) == 0
    # the limit itself is accepted
    boxes, scores = _boxes(31, cap, False)
    assert np.array_equal(_run('odk_nms_static', boxes, scores, 100, 0.45, dev), _run('odk_nms', boxes, scores, 100, 0.45, dev))
