"""``PriorBoxes`` — drop-in mirror of pytoolkit/pytoolkit/dl/od_pb.py for the detection hot path.

Same constructor, ``from_dict``/``to_dict``, ``encode_truth``, ``_assign_boxes``, ``encode_locs`` and
``decode_locs`` signatures, argument meaning, output layouts and error behaviour as the reference
class (od_pb.py:13-363); the arithmetic runs in libodk's sm_100a kernels.  Anchor tables are built
on the host exactly as ``_create_pb`` does (od_pb.py:125-184) because the kernels compare against
them bit for bit, then uploaded once per device.

Out of scope here: ``fit`` (offline KMeans anchor fitting, od_pb.py:63-123) and
``check_prior_boxes`` (a logging diagnostic, od_pb.py:224-297).  The ``loss*`` / ``metrics`` functions
(od_pb.py:365-431) are mirrored as device calls (``loss_terms`` and friends).

Besides the reference's host-array API there is a batched device API (``encode_truth_batch``,
``decode_score``, ``decode_nms``) that takes / returns torch CUDA tensors and never synchronises.
"""
import ctypes
import threading

import numpy as np
import torch

from . import _dev, _lib

_VAR_LOC = 0.10  # od_pb.py:10


class PriorBoxes:
    """Anchor ("prior box") set + target encoding / prediction decoding (od_pb.py:13)."""

    def __init__(self, input_size, map_sizes, num_classes):
        self.input_size = tuple(input_size)
        self.map_sizes = np.array(sorted(map_sizes)[::-1])  # od_pb.py:20
        self.num_classes = num_classes
        self.pb_size_patterns = np.empty((0,))
        self.pb_grid = np.empty((0,))
        self.pb_locs = np.empty((0,))
        self.pb_info_indices = np.empty((0,))
        self.pb_centers = np.empty((0,))
        self.pb_sizes = np.empty((0,))
        self.pb_mask = np.empty((0,))
        assert np.array(self.input_size).shape == (2,)
        assert len(self.map_sizes.shape) == 1 and len(self.map_sizes) >= 1
        assert self.num_classes >= 1
        self._handles = {}  # device index -> odk_anchors*
        self._lock = threading.Lock()

    # ------------------------------------------------------------------ persistence (od_pb.py:44-60)
    def to_dict(self):
        return {'map_sizes': self.map_sizes, 'pb_size_patterns': self.pb_size_patterns}

    def from_dict(self, data: dict):
        try:
            self.map_sizes = np.array(data['map_sizes'])
            self.pb_size_patterns = np.array(data['pb_size_patterns'])
        except KeyError as e:
            raise ValueError(f'PriorBoxes load error: data={data}') from e
        self._create_pb()

    # ------------------------------------------------------------------ tables (od_pb.py:125-198)
    def _create_pb(self):
        """Host-side table construction.  The kernels compare GT centres against these tables bit for bit, so every
        rounding step of od_pb.py:133-177 is kept: f32 ``linspace`` ticks, f64 offsets added in place into f32 arrays
        (``+=`` keeps the array dtype), f64 ``pb_sizes``, closed-interval validity test on the clipped box centre.
        The per-(map, pattern) ``pb_info`` dicts of the reference only feed ``check_prior_boxes`` (out of scope);
        ``pb_info_indices`` — which (map, pattern) plane an anchor belongs to — is kept as an output table."""
        self._release()
        inv_px = 1 / np.array(self.input_size)
        planes = []  # one entry per (map, pattern): grid, locs, centres, sizes, mask
        for map_size in self.map_sizes:
            cell = 1.0 / map_size
            ticks = np.linspace(0.5 * cell, 1 - 0.5 * cell, map_size, dtype=np.float32)
            xs, ys = np.meshgrid(ticks, ticks)
            centre4 = np.tile(np.stack([xs.ravel(), ys.ravel()], axis=1), (1, 2))
            grid = centre4.copy()
            grid[:, :2] -= cell / 2
            grid[:, 2:] += cell / 2 + inv_px
            for pattern in self.pb_size_patterns:
                size = cell * pattern
                box = centre4.copy()
                box[:, :2] -= size / 2
                box[:, 2:] += size / 2 + inv_px
                clipped = np.clip(box, 0, 1)
                cc = (clipped[:, :2] + clipped[:, 2:]) / 2
                inside = ((grid[:, :2] <= cc) & (cc <= grid[:, 2:])).all(axis=-1)  # math.between (closed)
                planes.append((grid, box, centre4[:, :2], np.tile(size, (len(box), 1)), inside))
        self.pb_grid, self.pb_locs, self.pb_centers, self.pb_sizes, self.pb_mask = (
            np.concatenate([pl[k] for pl in planes]) for k in range(5))
        self.pb_info_indices = np.repeat(np.arange(len(planes), dtype=np.int64), [len(pl[1]) for pl in planes])
        self._check_state()

    def _check_state(self):
        """Consistency of the tables (the reference's od_pb.py:186-198 checks): one row per anchor in every table, and
        every unclipped anchor centre lies in its own half-open grid cell — the property ``for_each_cell_candidate``
        in csrc/assign_encode.cu relies on."""
        n = len(self.pb_locs)
        expect = {'pb_locs': (n, 4), 'pb_grid': (n, 4), 'pb_info_indices': (n,), 'pb_centers': (n, 2),
                  'pb_sizes': (n, 2), 'pb_mask': (n,)}
        for name, shape in expect.items():
            got = getattr(self, name).shape
            assert got == shape, f'{name}: expected {shape}, got {got}'
        mid = np.mean([self.pb_locs[:, 2:], self.pb_locs[:, :2]], axis=0)
        assert ((self.pb_grid[:, :2] <= mid) & (mid < self.pb_grid[:, 2:])).all(), 'anchor centre outside its grid cell'

    def __len__(self):
        return len(self.pb_locs)

    # ------------------------------------------------------------------ device handle
    def _handle(self, dev: torch.device):
        """odk_anchors* for ``dev`` (uploaded on first use; od_pb tables are immutable after _create_pb)."""
        key = dev.index if dev.index is not None else torch.cuda.current_device()
        with self._lock:
            h = self._handles.get(key)
            if h is None:
                if len(self.pb_locs) == 0:
                    raise RuntimeError('PriorBoxes has no tables yet: call from_dict() first')
                maps = np.ascontiguousarray(self.map_sizes, dtype=np.int32)
                locs = np.ascontiguousarray(self.pb_locs, dtype=np.float32)
                grid = np.ascontiguousarray(self.pb_grid, dtype=np.float32)
                sizes = np.ascontiguousarray(self.pb_sizes, dtype=np.float64)
                mask = np.ascontiguousarray(self.pb_mask, dtype=np.uint8)
                out = ctypes.c_void_p()
                with torch.cuda.device(key):
                    _lib.call('odk_anchors_create', ctypes.byref(out), len(locs), int(self.num_classes), len(maps),
                              maps.ctypes.data, len(self.pb_size_patterns), locs.ctypes.data, grid.ctypes.data,
                              sizes.ctypes.data, mask.ctypes.data)
                h = out.value
                self._handles[key] = h
        return h

    def _release(self):
        for key, h in list(getattr(self, '_handles', {}).items()):
            try:
                with torch.cuda.device(key):
                    _lib.load().odk_anchors_destroy(h)
            except Exception:  # noqa: BLE001 - interpreter shutdown
                pass
        self._handles = {}

    def __del__(self):
        self._release()

    # ------------------------------------------------------------------ encode / decode (od_pb.py:212-222)
    def encode_locs(self, bbox, pb_ix):
        """od_pb.py:212-214: (bbox - pb_locs[pb_ix]) / tile(pb_sizes[pb_ix], 2) / 0.10 -> float64.
        ``pb_ix`` may be an int, an index array or a boolean mask, as in the reference."""
        dev = _dev.device()
        ix = np.arange(len(self.pb_locs))[pb_ix]
        scalar = np.ndim(ix) == 0
        ix = np.atleast_1d(ix).astype(np.int32)
        boxes = np.ascontiguousarray(np.broadcast_to(np.asarray(bbox, dtype=np.float32), (len(ix), 4)))
        out = torch.empty((len(ix), 4), dtype=torch.float64, device=dev)
        d_boxes = _dev.to_device(boxes, torch.float32, dev)
        d_ix = _dev.to_device(ix, torch.int32, dev)
        with torch.cuda.device(dev):
            _lib.call('odk_encode_locs', self._handle(dev), len(ix), d_boxes.data_ptr(), d_ix.data_ptr(), out.data_ptr(),
                      _dev.stream_ptr(dev))
        res = out.cpu().numpy()
        return res[0] if scalar else res

    def decode_locs(self, pred, xp=None):
        """od_pb.py:216-222.  ``xp=None`` is the in-graph use (od_net.py:189): the fp32 Keras Lambda
        clip(pred * float32(0.10*tile(pb_sizes,2)) + pb_locs, 0, 1); NumPy in -> NumPy fp32 out, torch CUDA tensor
        in -> torch tensor out (no sync).  ``xp=np`` is the NumPy use (check_assign.py:27): pb_sizes is fp64 there,
        so the expression is evaluated in fp64 and a float64 array comes back (``odk_decode_locs_f64``)."""
        is_np = not isinstance(pred, torch.Tensor)
        dev = _dev.device(None if is_np or not pred.is_cuda else pred.device)
        f64 = xp is np
        t = _dev.to_device(pred, torch.float64 if f64 else torch.float32, dev)
        if t.shape[-1] != 4 or t.shape[-2] != len(self.pb_locs):
            raise ValueError(f'decode_locs: expected (..., {len(self.pb_locs)}, 4), got {tuple(t.shape)}')
        out = torch.empty_like(t)
        with torch.cuda.device(dev):
            _lib.call('odk_decode_locs_f64' if f64 else 'odk_decode_locs', self._handle(dev), t.numel() // 4, t.data_ptr(),
                      out.data_ptr(), _dev.stream_ptr(dev))
        return out.cpu().numpy() if is_np else out

    # ------------------------------------------------------------------ training targets (od_pb.py:299-363)
    @staticmethod
    def _pad_gt(y_gt):
        counts = np.array([len(y.classes) for y in y_gt], dtype=np.int32)
        gmax = max(1, int(counts.max()) if len(counts) else 1)
        boxes = np.zeros((len(y_gt), gmax, 4), dtype=np.float32)
        classes = np.zeros((len(y_gt), gmax), dtype=np.int32)
        for i, y in enumerate(y_gt):
            g = counts[i]
            if g:
                boxes[i, :g] = np.asarray(y.bboxes, dtype=np.float32).reshape(g, 4)
                classes[i, :g] = np.asarray(y.classes, dtype=np.int32)
        return boxes, classes, counts

    def encode_truth_batch(self, gt_boxes, gt_classes, gt_count, y_true=None, want_valids=False):
        """Batched device entry (no host sync): padded GT (B,Gmax,4) f32 / (B,Gmax) i32 / (B,) i32 on the
        GPU (host arrays are uploaded) -> dict of CUDA tensors: ``y_true`` (B,N,3+C+4) f32,
        ``pb_indices``/``gt_indices`` (B,Gmax) i64 (-1 padded), ``status`` (B,) i32 (odk.h ODK_IMG_*),
        optionally ``pb_valids`` (B,N) bool."""
        dev = _dev.device(gt_boxes.device if isinstance(gt_boxes, torch.Tensor) and gt_boxes.is_cuda else None)
        boxes = _dev.to_device(gt_boxes, torch.float32, dev)
        classes = _dev.to_device(gt_classes, torch.int32, dev)
        count = _dev.to_device(gt_count, torch.int32, dev)
        b, gmax = boxes.shape[0], boxes.shape[1]
        assert boxes.shape == (b, gmax, 4) and classes.shape == (b, gmax) and count.shape == (b,)
        n, w = len(self.pb_locs), 3 + self.num_classes + 4
        if y_true is None:
            y_true = torch.empty((b, n, w), dtype=torch.float32, device=dev)
        assert y_true.shape == (b, n, w) and y_true.dtype == torch.float32 and y_true.is_contiguous()
        pb_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        gt_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        status = torch.empty((b,), dtype=torch.int32, device=dev)
        valids = torch.empty((b, n), dtype=torch.uint8, device=dev) if want_valids else None
        ws_bytes = _lib.load().odk_assign_workspace_bytes(self._handle(dev), b, gmax)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.call('odk_assign_encode', self._handle(dev), b, boxes.data_ptr(), classes.data_ptr(), count.data_ptr(), gmax,
                      y_true.data_ptr(), pb_idx.data_ptr(), gt_idx.data_ptr(), _dev.ptr(valids), status.data_ptr(),
                      ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
        res = {'y_true': y_true, 'pb_indices': pb_idx, 'gt_indices': gt_idx, 'status': status}
        if want_valids:
            res['pb_valids'] = valids.view(torch.bool)
        return res

    def encode_truth_compact(self, gt_boxes, gt_classes, gt_count, out=None):
        """``encode_truth_batch`` with the compact result of ``odk_assign_encode_compact`` (no host sync): dict of CUDA
        tensors ``pos_rows`` (B,Gmax,3+C+4) f32, ``valid_bits`` (B,ceil(N/32)) i32 (bit a of word a>>5 = pb_valids[a]),
        ``pb_indices`` / ``gt_indices`` (B,Gmax) i64, ``status`` (B,) i32 — ~10 KB per image instead of the 1.8 MB dense
        tensor.  ``expand_truth`` turns the host copies of these into the dense array.  ``out``: a dict of preallocated
        tensors with those keys (fixed addresses for CUDA-graph capture)."""
        dev = _dev.device(gt_boxes.device if isinstance(gt_boxes, torch.Tensor) and gt_boxes.is_cuda else None)
        boxes = _dev.to_device(gt_boxes, torch.float32, dev)
        classes = _dev.to_device(gt_classes, torch.int32, dev)
        count = _dev.to_device(gt_count, torch.int32, dev)
        b, gmax = boxes.shape[0], boxes.shape[1]
        assert boxes.shape == (b, gmax, 4) and classes.shape == (b, gmax) and count.shape == (b,)
        n, w = len(self.pb_locs), 3 + self.num_classes + 4
        res = out if out is not None else {
            'pos_rows': torch.empty((b, gmax, w), dtype=torch.float32, device=dev),
            'valid_bits': torch.empty((b, (n + 31) // 32), dtype=torch.int32, device=dev),
            'pb_indices': torch.empty((b, gmax), dtype=torch.int64, device=dev),
            'gt_indices': torch.empty((b, gmax), dtype=torch.int64, device=dev),
            'status': torch.empty((b,), dtype=torch.int32, device=dev)}
        assert res['pos_rows'].shape == (b, gmax, w) and res['valid_bits'].shape == (b, (n + 31) // 32)
        ws_bytes = _lib.load().odk_assign_workspace_bytes(self._handle(dev), b, gmax)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.call('odk_assign_encode_compact', self._handle(dev), b, boxes.data_ptr(), classes.data_ptr(), count.data_ptr(), gmax,
                      res['pos_rows'].data_ptr(), res['valid_bits'].data_ptr(), res['pb_indices'].data_ptr(),
                      res['gt_indices'].data_ptr(), res['status'].data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
        return res

    def span_stats(self, dev=None):
        """{tile, spans, rows_covered, valid} of the valid-anchor span table the fused predict path fetches through."""
        dev = _dev.device(dev)
        out = (ctypes.c_int32 * 4)()
        _lib.call('odk_anchors_span_stats', self._handle(dev), ctypes.cast(out, ctypes.c_void_p))
        return {'tile': int(out[0]), 'spans': int(out[1]), 'rows_covered': int(out[2]), 'valid': int(out[3]), 'anchors': len(self.pb_locs)}

    # ---- loss side (od_pb.py:365-431) --------------------------------------------------------------------
    LOSS_COLUMNS = ('loss', 'loss_obj', 'loss_clf', 'loss_loc', 'rec_bg', 'rec_obj', 'acc_clf', 'n_positive')

    def loss_terms(self, y_true, y_pred):
        """All of od_pb.py:365-431 in one pass over (B,N,3+C+4) ``y_true`` / ``y_pred``: a (B,8) float64 CUDA
        tensor with the columns ``LOSS_COLUMNS`` (forward values; ``odk_loss``)."""
        dev = _dev.device(y_pred.device if isinstance(y_pred, torch.Tensor) and y_pred.is_cuda else None)
        yt = _dev.to_device(y_true, torch.float32, dev)
        yp = _dev.to_device(y_pred, torch.float32, dev)
        n, w = len(self.pb_locs), 3 + self.num_classes + 4
        assert yt.dim() == 3 and yt.shape[1:] == (n, w) and yp.shape == yt.shape
        b = yt.shape[0]
        out = torch.empty((b, 8), dtype=torch.float64, device=dev)
        ws_bytes = _lib.load().odk_loss_workspace_bytes(self._handle(dev), b)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.call('odk_loss', self._handle(dev), b, yt.data_ptr(), yp.data_ptr(), out.data_ptr(), ws.data_ptr(), ws_bytes,
                      _dev.stream_ptr(dev))
        return out

    def assign_loss_batch(self, gt_boxes, gt_classes, gt_count, y_pred, with_grad=False, upstream=None):
        """``encode_truth`` followed by ``loss`` without ever writing y_true (``odk_assign_loss``): padded GT as in
        ``encode_truth_batch`` + the network output (B,N,3+C+4) -> dict with ``terms`` (B,8) float64 (columns
        ``LOSS_COLUMNS``), ``pb_indices``, ``gt_indices``, ``status``; ``with_grad`` adds ``grad`` (B,N,3+C+4) float32 =
        ``upstream[b]`` * d loss / d y_pred from the same pass (``odk_assign_loss_grad``)."""
        dev = _dev.device(y_pred.device if isinstance(y_pred, torch.Tensor) and y_pred.is_cuda else None)
        boxes = _dev.to_device(gt_boxes, torch.float32, dev)
        classes = _dev.to_device(gt_classes, torch.int32, dev)
        count = _dev.to_device(gt_count, torch.int32, dev)
        yp = _dev.to_device(y_pred, torch.float32, dev)
        b, gmax = boxes.shape[0], boxes.shape[1]
        n, w = len(self.pb_locs), 3 + self.num_classes + 4
        assert boxes.shape == (b, gmax, 4) and classes.shape == (b, gmax) and count.shape == (b,) and yp.shape == (b, n, w)
        out = torch.empty((b, 8), dtype=torch.float64, device=dev)
        pb_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        gt_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        status = torch.empty((b,), dtype=torch.int32, device=dev)
        ws_bytes = _lib.load().odk_assign_loss_workspace_bytes(self._handle(dev), b, gmax)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        res = {'terms': out, 'pb_indices': pb_idx, 'gt_indices': gt_idx, 'status': status}
        with torch.cuda.device(dev):
            if with_grad:
                up = _dev.to_device(upstream, torch.float32, dev) if upstream is not None else None
                assert up is None or up.shape == (b,)
                res['grad'] = torch.empty_like(yp)
                _lib.call('odk_assign_loss_grad', self._handle(dev), b, boxes.data_ptr(), classes.data_ptr(), count.data_ptr(), gmax,
                          yp.data_ptr(), _dev.ptr(up), out.data_ptr(), res['grad'].data_ptr(), pb_idx.data_ptr(), gt_idx.data_ptr(),
                          status.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
            else:
                _lib.call('odk_assign_loss', self._handle(dev), b, boxes.data_ptr(), classes.data_ptr(), count.data_ptr(), gmax,
                          yp.data_ptr(), out.data_ptr(), pb_idx.data_ptr(), gt_idx.data_ptr(), status.data_ptr(), ws.data_ptr(),
                          ws_bytes, _dev.stream_ptr(dev))
        return res

    def loss_grad(self, y_true, y_pred, upstream=None):
        """d ``loss`` / d ``y_pred`` (B,N,3+C+4) float32 — the gradient TensorFlow derives from od_pb.py:365-431
        (``odk_loss_grad``).  ``upstream`` (B,) scales each image's gradient (the incoming dL/dloss)."""
        dev = _dev.device(y_pred.device if isinstance(y_pred, torch.Tensor) and y_pred.is_cuda else None)
        yt = _dev.to_device(y_true, torch.float32, dev)
        yp = _dev.to_device(y_pred, torch.float32, dev)
        n, w = len(self.pb_locs), 3 + self.num_classes + 4
        assert yt.dim() == 3 and yt.shape[1:] == (n, w) and yp.shape == yt.shape
        up = _dev.to_device(upstream, torch.float32, dev) if upstream is not None else None
        assert up is None or up.shape == (yt.shape[0],)
        grad = torch.empty_like(yp)
        with torch.cuda.device(dev):
            _lib.call('odk_loss_grad', self._handle(dev), yt.shape[0], yt.data_ptr(), yp.data_ptr(), _dev.ptr(up), grad.data_ptr(),
                      _dev.stream_ptr(dev))
        return grad

    def loss_with_grad(self, y_true, y_pred):
        """``loss`` as a differentiable torch op: forward ``odk_loss``, backward ``odk_loss_grad`` — for training
        loops whose network lives in torch (the reference relies on TensorFlow's autodiff here)."""
        pb = self

        class _Loss(torch.autograd.Function):
            @staticmethod
            def forward(ctx, yt, yp):
                ctx.save_for_backward(yt, yp)
                return pb.loss_terms(yt, yp)[:, 0].to(torch.float32)

            @staticmethod
            def backward(ctx, grad_out):
                yt, yp = ctx.saved_tensors
                return None, pb.loss_grad(yt, yp, grad_out.contiguous())

        return _Loss.apply(y_true, y_pred)

    def _loss_column(self, y_true, y_pred, k):
        return self.loss_terms(y_true, y_pred)[:, k].to(torch.float32)

    def loss(self, y_true, y_pred):
        """od_pb.py:365-373: (B,) float32 = loss_obj + loss_clf + loss_loc."""
        return self._loss_column(y_true, y_pred, 0)

    def loss_obj(self, y_true, y_pred):
        """od_pb.py:403-411 (binary focal loss over pb_mask * gt_mask, times 3)."""
        return self._loss_column(y_true, y_pred, 1)

    def loss_clf(self, y_true, y_pred):
        """od_pb.py:413-421 (categorical cross-entropy weighted by channel 1)."""
        return self._loss_column(y_true, y_pred, 2)

    def loss_loc(self, y_true, y_pred):
        """od_pb.py:423-431 (mean squared box error weighted by channel 1, times 2)."""
        return self._loss_column(y_true, y_pred, 3)

    @property
    def metrics(self):
        """od_pb.py:375-401: [loss_obj, loss_clf, loss_loc, rec_bg, rec_obj, acc_clf] as callables (y_true, y_pred)."""
        def column(k, name):
            def fn(y_true, y_pred):
                return self._loss_column(y_true, y_pred, k)
            fn.__name__ = name
            return fn
        return [self.loss_obj, self.loss_clf, self.loss_loc, column(4, 'rec_bg'), column(5, 'rec_obj'), column(6, 'acc_clf')]

    @staticmethod
    def _raise_for_status(status, y_gt):
        """Maps device status words to the reference's AssertionErrors (od_pb.py:304,332,338)."""
        for i, s in enumerate(status):
            if s == 1:
                raise AssertionError(f'Encode error: image {i}: {_centres(y_gt[i].bboxes)}')
            if s == 2:
                raise AssertionError(f'class id out of range [0, num_classes) in image {i}')
            if s != 0:
                raise AssertionError(f'bad ground-truth count in image {i}')

    def encode_truth(self, y_gt):
        """od_pb.py:299-313: list of annotations (``.classes``, ``.bboxes``) -> np.float32
        (len(y_gt), N, 3+C+4) = [mask, weight, obj, onehot, locs]."""
        if len(y_gt) == 0:
            return np.zeros((0, len(self.pb_locs), 3 + self.num_classes + 4), dtype=np.float32)
        boxes, classes, counts = self._pad_gt(y_gt)
        res = self.encode_truth_batch(boxes, classes, counts)
        self._raise_for_status(res['status'].cpu().numpy(), y_gt)
        return res['y_true'].cpu().numpy()

    def _assign_boxes(self, bboxes):
        """od_pb.py:315-363 -> (pb_indices i64 ascending, assigned gt index i64, pb_valids bool (N,))."""
        bboxes = np.asarray(bboxes, dtype=np.float32).reshape(-1, 4)
        g = len(bboxes)
        boxes = np.zeros((1, max(1, g), 4), dtype=np.float32)
        boxes[0, :g] = bboxes
        res = self.encode_truth_batch(boxes, np.zeros((1, max(1, g)), np.int32), np.array([g], np.int32), want_valids=True)
        status = res['status'].cpu().numpy()
        if status[0] != 0:
            raise AssertionError(f'Encode error: {_centres(bboxes)}')
        return (res['pb_indices'][0, :g].cpu().numpy(), res['gt_indices'][0, :g].cpu().numpy(),
                res['pb_valids'][0].cpu().numpy())

    def assign_batch(self, gt_boxes, gt_count):
        """``_assign_boxes`` for a padded batch without building y_true (``odk_assign``): -> ``pb_indices``,
        ``gt_indices`` (B,Gmax) i64 (-1 padded) and ``status`` (B,) i32 CUDA tensors."""
        dev = _dev.device(gt_boxes.device if isinstance(gt_boxes, torch.Tensor) and gt_boxes.is_cuda else None)
        boxes = _dev.to_device(gt_boxes, torch.float32, dev)
        count = _dev.to_device(gt_count, torch.int32, dev)
        b, gmax = boxes.shape[0], boxes.shape[1]
        assert boxes.shape == (b, gmax, 4) and count.shape == (b,)
        pb_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        gt_idx = torch.empty((b, gmax), dtype=torch.int64, device=dev)
        status = torch.empty((b,), dtype=torch.int32, device=dev)
        ws_bytes = _lib.load().odk_assign_workspace_bytes(self._handle(dev), b, gmax)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.call('odk_assign', self._handle(dev), b, boxes.data_ptr(), count.data_ptr(), gmax, pb_idx.data_ptr(),
                      gt_idx.data_ptr(), status.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
        return {'pb_indices': pb_idx, 'gt_indices': gt_idx, 'status': status}

    # ------------------------------------------------------------------ prediction side (od_net.py:172-192)
    def _head_inputs(self, objs, clfs, locs):
        dev = _dev.device(objs.device if isinstance(objs, torch.Tensor) and objs.is_cuda else None)
        objs = _dev.to_device(objs, torch.float32, dev)
        clfs = _dev.to_device(clfs, torch.float32, dev)
        locs = _dev.to_device(locs, torch.float32, dev)
        b, n, c = clfs.shape[0], len(self.pb_locs), self.num_classes
        if clfs.shape != (b, n, c) or locs.shape != (b, n, 4) or objs.numel() != b * n:
            raise ValueError(f'head outputs must be objs (B,{n},1), clfs (B,{n},{c}), locs (B,{n},4); got '
                             f'{tuple(objs.shape)}, {tuple(clfs.shape)}, {tuple(locs.shape)}')
        return dev, b, objs, clfs, locs

    def decode_score(self, objs, clfs, locs):
        """ChannelArgMax/ChannelMax/_objconf/decode_locs in one pass (od_net.py:186-189):
        -> classes (B,N) i32, objconfs (B,N) f32, boxes (B,N,4) f32 CUDA tensors."""
        dev, b, objs, clfs, locs = self._head_inputs(objs, clfs, locs)
        n = len(self.pb_locs)
        classes = torch.empty((b, n), dtype=torch.int32, device=dev)
        confs = torch.empty((b, n), dtype=torch.float32, device=dev)
        boxes = torch.empty((b, n, 4), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.call('odk_decode_score', self._handle(dev), b, objs.data_ptr(), clfs.data_ptr(), locs.data_ptr(),
                      classes.data_ptr(), confs.data_ptr(), boxes.data_ptr(), _dev.stream_ptr(dev))
        return classes, confs, boxes

    def decode_nms(self, objs, clfs, locs, top_k=200, conf_threshold=0.01, nms_threshold=0.45, iou_ge=True,
                   out=None, return_index=False, nms_all_threshold=None):
        """The whole prediction tail (od_net.py:186-190) fused: head outputs -> (B, top_k, 6)
        [class, conf, x1, y1, x2, y2], zero padded (layers.py:963-970).  ``out`` may be a slice of a
        larger (e.g. all-gather) buffer.  ``nms_all_threshold`` = the strict_nms second pass
        (od_net.py:175: 0.5 when ``strict_nms``)."""
        dev, b, objs, clfs, locs = self._head_inputs(objs, clfs, locs)
        n, c = len(self.pb_locs), self.num_classes
        if out is None:
            out = torch.empty((b, top_k, 6), dtype=torch.float32, device=dev)
        assert out.shape == (b, top_k, 6) and out.dtype == torch.float32 and out.is_contiguous()
        kept = torch.empty((b, top_k), dtype=torch.int32, device=dev)
        strict = bool(nms_all_threshold)
        ws_bytes = getattr(_lib.load(), 'odk_nms_strict_workspace_bytes' if strict else 'odk_nms_workspace_bytes')(b, n, c, top_k)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
        thr = (float(nms_threshold), float(nms_all_threshold)) if strict else (float(nms_threshold),)
        with torch.cuda.device(dev):
            _lib.call('odk_decode_nms_strict' if strict else 'odk_decode_nms', self._handle(dev), b, objs.data_ptr(),
                      clfs.data_ptr(), locs.data_ptr(), float(conf_threshold), *thr, int(top_k), 1 if iou_ge else 0,
                      out.data_ptr(), kept.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
        return (out, kept) if return_index else out


def expand_truth(pos_rows, valid_bits, pb_indices, gt_count, n_anchors, out=None):
    """Host-side expansion of ``encode_truth_compact``'s result (NumPy arrays) into the dense ``encode_truth`` array
    (B, N, 3+C+4) float32 of od_pb.py:299-313: zeros, column 0 = float32(1/max(1,G)) where ``valid_bits`` is set, and the
    G positive rows at ``pb_indices``.  A memset and two scatters — no arithmetic beyond the reference's own 1/G."""
    pos_rows, valid_bits = np.asarray(pos_rows), np.asarray(valid_bits)
    pb_indices, gt_count = np.asarray(pb_indices), np.asarray(gt_count)
    b, gmax, w = pos_rows.shape
    y = np.zeros((b, n_anchors, w), dtype=np.float32) if out is None else out
    if out is not None:
        y[...] = 0
    bits = np.unpackbits(valid_bits.view(np.uint8).reshape(b, -1), axis=1, bitorder='little')[:, :n_anchors].astype(bool)
    for i in range(b):
        g = int(gt_count[i])
        y[i, bits[i], 0] = np.float32(1 / max(1, g))  # od_pb.py:306
        if bits[i].any() and g:
            y[i, pb_indices[i, :g]] = pos_rows[i, :g]
    return y


def _centres(bboxes):
    b = np.asarray(bboxes, dtype=np.float32).reshape(-1, 4)
    return (b[:, 2:] + b[:, :2]) / 2
