"""Mirror of the prediction-tail layers of pytoolkit/pytoolkit/dl/layers.py: ``nms()``,
``channel_argmax()``, ``channel_max()`` (layers.py:69-98, 876-972) as plain callables over CUDA
tensors (there is no Keras graph here)."""
import torch

from . import _dev, _lib


def nms():
    """Factory, like the reference (layers.py:876): returns the ``NMS`` class."""

    class NMS:
        """layers.py:890-970.  ``NMS(num_classes, prior_boxes, ...)([classes, confs, locs]) -> (B, top_k, 6)``
        rows [class, conf, x1, y1, x2, y2], zero padded.  ``nms_all_threshold`` (strict all-class NMS,
        od_net.py:175, layers.py:940,953-957) runs both stages in the same per-image sweep
        (``odk_nms_layer_strict``)."""

        def __init__(self, num_classes, prior_boxes, top_k=200, conf_threshold=0.01, nms_threshold=0.45,
                     nms_all_threshold=None, iou_ge=True, **kwargs):
            del kwargs
            self.num_classes = num_classes
            self.prior_boxes = prior_boxes
            self.top_k = top_k
            self.conf_threshold = conf_threshold
            self.nms_threshold = nms_threshold
            self.nms_all_threshold = nms_all_threshold
            self.iou_ge = iou_ge  # True: reference NumPy-NMS semantics (pinned); False: TF-1.10 style (unpinned)

        def compute_output_shape(self, input_shape):
            assert input_shape and len(input_shape) >= 2
            return (input_shape[0][0], self.top_k, 1 + 1 + 4)

        def get_config(self):
            return {
                'num_classes': self.num_classes, 'prior_boxes': self.prior_boxes, 'top_k': self.top_k,
                'conf_threshold': self.conf_threshold, 'nms_threshold': self.nms_threshold,
                'nms_all_threshold': self.nms_all_threshold,
            }

        def call(self, inputs, return_index=False, **kwargs):
            del kwargs
            classes, confs, locs = inputs
            is_np = not isinstance(confs, torch.Tensor)
            dev = _dev.device(None if is_np or not confs.is_cuda else confs.device)
            n = self.prior_boxes
            classes = _dev.to_device(classes, torch.int32, dev).reshape(-1, n)
            confs = _dev.to_device(confs, torch.float32, dev).reshape(-1, n)
            locs = _dev.to_device(locs, torch.float32, dev).reshape(-1, n, 4)
            b = confs.shape[0]
            out = torch.empty((b, self.top_k, 6), dtype=torch.float32, device=dev)
            kept = torch.empty((b, self.top_k), dtype=torch.int32, device=dev)
            strict = bool(self.nms_all_threshold)  # layers.py:940,953: truthiness decides
            ws_fn = 'odk_nms_strict_workspace_bytes' if strict else 'odk_nms_workspace_bytes'
            ws_bytes = getattr(_lib.load(), ws_fn)(b, n, self.num_classes, self.top_k)
            ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
            thr = (float(self.nms_threshold), float(self.nms_all_threshold)) if strict else (float(self.nms_threshold),)
            with torch.cuda.device(dev):
                _lib.call('odk_nms_layer_strict' if strict else 'odk_nms_layer', b, n, self.num_classes, classes.data_ptr(),
                          confs.data_ptr(), locs.data_ptr(), float(self.conf_threshold), *thr, int(self.top_k),
                          1 if self.iou_ge else 0, out.data_ptr(), kept.data_ptr(), ws.data_ptr(), ws_bytes, _dev.stream_ptr(dev))
            if is_np:
                out, kept = out.cpu().numpy(), kept.cpu().numpy()
            return (out, kept) if return_index else out

        __call__ = call

    return NMS


def _channel_reduce(inputs, want_arg):
    is_np = not isinstance(inputs, torch.Tensor)
    dev = _dev.device(None if is_np or not inputs.is_cuda else inputs.device)
    x = _dev.to_device(inputs, torch.float32, dev)
    rows, c = x.numel() // x.shape[-1], x.shape[-1]
    out = torch.empty(x.shape[:-1], dtype=torch.int64 if want_arg else torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.call('odk_channel_argmax_max', rows, c, x.data_ptr(), out.data_ptr() if want_arg else 0,
                  0 if want_arg else out.data_ptr(), _dev.stream_ptr(dev))
    return out.cpu().numpy() if is_np else out


def channel_argmax():
    """layers.py:69-82: argmax over the last axis, int64, first maximum wins (``odk_channel_argmax_max``)."""

    class ChannelArgMax:
        def compute_output_shape(self, input_shape):
            return input_shape[:-1]

        def __call__(self, inputs):
            return _channel_reduce(inputs, True)

    return ChannelArgMax


def channel_max():
    """layers.py:85-98: max over the last axis (``odk_channel_argmax_max``)."""

    class ChannelMax:
        def compute_output_shape(self, input_shape):
            return input_shape[:-1]

        def __call__(self, inputs):
            return _channel_reduce(inputs, False)

    return ChannelMax
