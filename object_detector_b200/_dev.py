"""Device-buffer plumbing: torch tensors are used ONLY as HBM allocations + streams."""
import numpy as np
import torch


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError('object_detector_b200 needs a CUDA device (B200, sm_100a); there is no CPU path')


def device(dev=None) -> torch.device:
    require_cuda()
    if dev is None:
        return torch.device('cuda', torch.cuda.current_device())
    return torch.device(dev)


def stream_ptr(dev) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def to_device(x, dtype, dev, non_blocking=False) -> torch.Tensor:
    """Host array / tensor -> contiguous device tensor of ``dtype`` (no copy if already there)."""
    if isinstance(x, torch.Tensor):
        t = x
    else:
        x = np.ascontiguousarray(x)
        t = torch.from_numpy(x if x.flags.writeable else x.copy())
    if t.dtype != dtype:
        t = t.to(dtype)
    if t.device != dev:
        t = t.to(dev, non_blocking=non_blocking)
    return t.contiguous()


def ptr(t):
    return 0 if t is None else t.data_ptr()
