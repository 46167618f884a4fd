"""Bounding-box side of the training augmentation, batched on the GPU.

Mirrors what happens to ``ObjectsAnnotation.bboxes`` / ``classes`` between the loader and ``encode_truth`` in the
reference's generator (od.py:155-169): ``image.RandomZoom`` (image.py:233-388), ``RandomFlipLR`` / ``RandomFlipTB``
(image.py:433-465) and ``RandomRotate90`` -> ``ObjectsAnnotation.rot90`` (image.py:468-485, ml.py:102-112), with the
per-sample ``RandomState(seed)`` of generator.py:172.  Pixels are not touched here: the returned ``plan`` carries, per
sample, the window / flips / turns the image side has to apply.

The reference retries a zoom up to 30 times per sample and stops at the first acceptable attempt, so how far a
sample's RNG advances depends on its boxes.  ``BoxAugmenter`` keeps that stream intact: it draws the whole attempt
table from a *copy* of each RandomState, lets ``odk_bbox_zoom`` pick the first acceptable attempt for the whole batch
in one launch, then advances the real RandomState by exactly the attempts the reference would have consumed before
drawing the flip / rotation decisions.  Host work is integer bookkeeping; all box arithmetic is in the kernels.
"""
import numpy as np
import torch

from . import _dev, _lib

ZOOM_NONE, ZOOM_PAD, ZOOM_CROP, ZOOM_LETTERBOX = 0, 1, 2, 3


class BoxAugmenter:
    """Arguments follow ``ObjectDetector.fit`` (od.py:90-130) / ``image.RandomZoom.__init__`` (image.py:267-283)."""

    def __init__(self, output_size, keep_aspect=False, padding_rate=16, crop_rate=0.1, aspect_prob=0.5,
                 max_aspect_ratio=1.5, min_object_px=8, flip_h=True, flip_v=True, rotate90=True, tries=30):
        assert max_aspect_ratio >= 1
        assert padding_rate is None or padding_rate > 1
        assert crop_rate is None or 0 < crop_rate < 1
        self.output_size = np.array([output_size, output_size] if np.isscalar(output_size) else output_size, np.int64)
        self.keep_aspect = keep_aspect
        self.padding_rate, self.crop_rate = padding_rate, crop_rate
        self.aspect_prob, self.max_aspect_ratio = aspect_prob, max_aspect_ratio
        self.min_object_px = float(min_object_px)
        self.flip_h, self.flip_v, self.rotate90 = flip_h, flip_v, rotate90
        self.tries = int(tries)

    # ---- host side: the reference's RNG draws (image.py:283-293, 324-327, 343-348) ---------------------------
    def _draw_attempt(self, rand, pad, in_size, ar):
        if pad:
            rate = rand.uniform(1, np.sqrt(self.padding_rate))
            size = np.maximum(np.ceil(in_size * rate * ar).astype(int), in_size)
            slack = size - in_size
        else:
            rate = rand.uniform(np.sqrt(self.crop_rate), 1)
            size = np.minimum(np.floor(in_size * rate * ar).astype(int), in_size)
            slack = in_size - size
        return size, (rand.randint(0, slack[0] + 1), rand.randint(0, slack[1] + 1))

    def _plan_zoom(self, rand, in_size, size_row, xy_row):
        """Consumes the aspect / branch draws from ``rand`` and fills the attempt table from a copy of it."""
        out = self.output_size
        base_ar = in_size[0] / in_size[1] if self.keep_aspect else out[0] / out[1]
        lo, hi = np.log(1 / self.max_aspect_ratio), np.log(self.max_aspect_ratio)
        rand_ar = np.exp(rand.uniform(lo, hi)) if rand.rand() <= self.aspect_prob else 1
        root = np.sqrt(base_ar * rand_ar)
        ar = np.array([root, 1 / root])
        if self.padding_rate is not None and (self.crop_rate is None or rand.rand() <= 0.5):
            kind = ZOOM_PAD
        elif self.crop_rate is not None:
            kind = ZOOM_CROP
        else:
            return ZOOM_NONE, ar
        ahead = np.random.RandomState()
        ahead.set_state(rand.get_state())
        for t in range(self.tries):
            size_row[t], xy_row[t] = self._draw_attempt(ahead, kind == ZOOM_PAD, in_size, ar)
        return kind, ar

    def _letterbox(self, in_size):
        """image.py:298-309: resize keeping the aspect ratio, centre in the output."""
        scale = (self.output_size / in_size).min()
        resized = np.floor(in_size * scale + 0.1).astype(np.int32).astype(np.int64)
        assert (resized <= self.output_size).all() and (resized == self.output_size).any()
        return (self.output_size - resized) // 2, resized

    # ---- the batched pipeline --------------------------------------------------------------------------------
    def __call__(self, gt_boxes, gt_classes, gt_count, image_sizes, rands, device=None):
        """``gt_boxes`` (B,Gmax,4) f32, ``gt_classes`` (B,Gmax) i32, ``gt_count`` (B,) — host arrays or CUDA tensors;
        ``image_sizes`` (B,2) host ints (width, height); ``rands`` one ``np.random.RandomState`` per sample (advanced
        exactly as the reference advances it).  Returns a dict of CUDA tensors ready for ``encode_truth_batch`` —
        ``gt_boxes`` f32, ``gt_classes``, ``gt_count`` (compacted, zero beyond the count) and ``boxes64`` (the
        reference's unrounded values) — plus ``plan``: host arrays ``kind``, ``chosen``, ``size``, ``xy`` (accepted
        window, -1 if none), ``flip_lr``, ``flip_tb``, ``k``."""
        dev = _dev.device(device if device is not None else
                          (gt_boxes.device if isinstance(gt_boxes, torch.Tensor) and gt_boxes.is_cuda else None))
        boxes = _dev.to_device(gt_boxes, torch.float32, dev).to(torch.float64)
        classes = _dev.to_device(gt_classes, torch.int32, dev)
        count = _dev.to_device(gt_count, torch.int32, dev)
        b, gmax = boxes.shape[0], boxes.shape[1]
        sizes = np.asarray(image_sizes, np.int64).reshape(b, 2)
        assert boxes.shape == (b, gmax, 4) and classes.shape == (b, gmax) and count.shape == (b,) and len(rands) == b
        t = self.tries
        kind = np.zeros(b, np.int32)
        try_size = np.ones((b, t, 2), np.int64)
        try_xy = np.zeros((b, t, 2), np.int64)
        ars = [None] * b
        for i in range(b):
            kind[i], ars[i] = self._plan_zoom(rands[i], sizes[i], try_size[i], try_xy[i])
        stream = _dev.stream_ptr(dev)
        zoomed = torch.empty_like(boxes)
        keep = torch.empty((b, gmax), dtype=torch.uint8, device=dev)
        chosen_d = torch.empty((b,), dtype=torch.int32, device=dev)
        wide = torch.empty((b,), dtype=torch.uint8, device=dev)

        def zoom(src, src_f32, kinds, in_sizes, tsize, txy, n_try, dst, keep_d, chosen_out, wide_d):
            # the uploaded tables must stay referenced until the launch is queued (a dropped tensor's block is
            # handed to the next allocation at once)
            tabs = [_dev.to_device(kinds, torch.int32, dev)] + [_dev.to_device(v, torch.int64, dev) for v in (in_sizes, tsize, txy)]
            with torch.cuda.device(dev):
                _lib.call('odk_bbox_zoom', b, gmax, n_try, src.data_ptr(), src_f32, count.data_ptr(),
                          tabs[0].data_ptr(), tabs[1].data_ptr(), tabs[2].data_ptr(), tabs[3].data_ptr(),
                          int(self.output_size[0]), int(self.output_size[1]), self.min_object_px,
                          dst.data_ptr(), keep_d.data_ptr(), chosen_out.data_ptr(), wide_d.data_ptr(), stream)

        zoom(boxes, 1, kind, sizes, try_size, try_xy, t, zoomed, keep, chosen_d, wide)
        chosen = chosen_d.cpu().numpy()        # the one host sync: how many attempts each sample's RNG has consumed
        plan = {'kind': kind.copy(), 'chosen': chosen, 'size': np.full((b, 2), -1, np.int64), 'xy': np.full((b, 2), -1, np.int64)}
        lb_kind = np.zeros(b, np.int32)
        lb_in, lb_size, lb_xy = sizes.copy(), np.ones((b, 1, 2), np.int64), np.zeros((b, 1, 2), np.int64)
        for i in range(b):
            if kind[i] == ZOOM_NONE:
                continue
            used = t if chosen[i] < 0 else chosen[i] + 1
            for _ in range(used):           # the draws the reference's loop made before it stopped
                self._draw_attempt(rands[i], kind[i] == ZOOM_PAD, sizes[i], ars[i])
            if chosen[i] >= 0:
                plan['size'][i], plan['xy'][i] = try_size[i, chosen[i]], try_xy[i, chosen[i]]
        failed = (chosen < 0) | (kind == ZOOM_NONE)
        if self.keep_aspect and failed.any():
            for i in np.nonzero(failed)[0]:
                if (sizes[i] == self.output_size).all():
                    continue
                lb_kind[i] = ZOOM_LETTERBOX
                lb_xy[i, 0], lb_in[i] = self._letterbox(sizes[i])
                lb_size[i, 0] = self.output_size
            if lb_kind.any():
                # letterboxed samples failed every attempt, so their boxes are still the fp32 input; one more launch
                # over the batch rewrites just those rows (kind 0 rows pass through)
                again, keep2 = torch.empty_like(boxes), torch.empty_like(keep)
                chosen2, wide2 = torch.empty_like(chosen_d), torch.empty_like(wide)
                zoom(zoomed, 1, lb_kind, lb_in, lb_size, lb_xy, 1, again, keep2, chosen2, wide2)
                wide = torch.maximum(wide, wide2)
                zoomed = again
                plan['letterbox'] = {int(i): (lb_xy[i, 0].copy(), lb_in[i].copy()) for i in np.nonzero(lb_kind)[0]}
        ops = np.zeros(b, np.int32)
        plan['flip_lr'], plan['flip_tb'], plan['k'] = np.zeros(b, bool), np.zeros(b, bool), np.zeros(b, np.int32)
        for i in range(b):
            lr = self.flip_h and rands[i].rand() <= 0.5
            tb = self.flip_v and rands[i].rand() <= 0.5
            k = int(rands[i].randint(0, 4)) if self.rotate90 else 0
            plan['flip_lr'][i], plan['flip_tb'][i], plan['k'][i] = lr, tb, k
            ops[i] = int(lr) | (int(tb) << 1) | (k << 2)
        out64 = torch.empty_like(boxes)
        out32 = torch.empty((b, gmax, 4), dtype=torch.float32, device=dev)
        classes_out, count_out = torch.empty_like(classes), torch.empty_like(count)
        ops_d = _dev.to_device(ops, torch.int32, dev)
        with torch.cuda.device(dev):
            _lib.call('odk_bbox_flip_rot', b, gmax, zoomed.data_ptr(), keep.data_ptr(), classes.data_ptr(), count.data_ptr(),
                      wide.data_ptr(), ops_d.data_ptr(), out64.data_ptr(), out32.data_ptr(),
                      classes_out.data_ptr(), count_out.data_ptr(), stream)
        return {'gt_boxes': out32, 'gt_classes': classes_out, 'gt_count': count_out, 'boxes64': out64, 'plan': plan}
