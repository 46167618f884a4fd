"""Multi-GPU plumbing for the evaluation path: images are sharded contiguously over the
ranks (one process per GPU), every rank runs decode+NMS on its shard straight into its slice of one
(R*per_rank, top_k, 6) buffer, a single IN-PLACE ``all_gather_into_tensor`` (NCCL over NVLink: the send
buffer of rank r is ``recv + r*per`` of the same allocation, which is NCCL's in-place all-gather)
completes the buffer on every rank, and AP is computed from it.  Padding slabs are all-zero with class -1
so they can never count as detections.  The reference has no eval-time collective (rank-0 only,
pytoolkit/od_voc.py:35-39); assign/encode/decode/NMS need none.
"""
import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of ``n_items`` for ``rank``; every rank owns ceil(n/world) slots."""
    per = (n_items + world - 1) // world
    lo = min(n_items, rank * per)
    return lo, min(n_items, lo + per), per


def new_gather_buffer(n_items: int, world: int, top_k: int, device):
    """(world*per, top_k, 6) fp32 buffer whose padding rows are marked invalid (class -1, conf 0)."""
    per = (n_items + world - 1) // world
    buf = torch.zeros((world * per, top_k, 6), dtype=torch.float32, device=device)
    buf[:, :, 0] = -1
    return buf


def local_slab(buf: torch.Tensor, rank: int, world: int):
    per = buf.shape[0] // world
    return buf[rank * per:(rank + 1) * per]


def all_gather_detections(buf: torch.Tensor, rank: int, world: int, group=None):
    """In-place all-gather: rank r contributes buf[r*per:(r+1)*per] — the slab its NMS kernel wrote, no staging
    copy; afterwards buf is identical on all ranks."""
    if world == 1:
        return buf
    dist.all_gather_into_tensor(buf, local_slab(buf, rank, world), group=group)
    return buf


def decode_nms_shard(pb, objs, clfs, locs, out, batch=256, top_k=200, **kw):
    """decode + per-class NMS of this rank's resident head outputs (n_local, N, ...) in batches of ``batch`` images,
    written straight into ``out`` (n_local, top_k, 6) — normally a prefix of ``local_slab(buf, rank, world)``."""
    n_local = objs.shape[0]
    for s in range(0, n_local, batch):
        e = min(n_local, s + batch)
        pb.decode_nms(objs[s:e], clfs[s:e], locs[s:e], top_k=top_k, out=out[s:e], **kw)
    return out


def sharded_decode_nms_eval(pb, head_fn, y_true, num_classes, rank, world, batch=64, top_k=200, conf_threshold=0.01, group=None):
    """Config 4 (VOC07-test-sized eval): ``head_fn(lo, hi)`` yields this rank's head outputs
    (objs, clfs, locs) for images [lo, hi); returns the reference's 4-key dict on every rank."""
    from . import voc
    import numpy as np
    dev = torch.device('cuda', torch.cuda.current_device())
    n = len(y_true)
    lo, hi, per = shard_range(n, rank, world)
    buf = new_gather_buffer(n, world, top_k, dev)
    mine = local_slab(buf, rank, world)
    for s in range(lo, hi, batch):
        e = min(hi, s + batch)
        objs, clfs, locs = head_fn(s, e)
        pb.decode_nms(objs, clfs, locs, top_k=top_k, out=mine[s - lo:e - lo])  # the kernel writes into the send slab
    all_gather_detections(buf, rank, world, group)
    # slabs are rank-major with `per` slots each: image i lives at slot (i // per) * per + i % per == i
    gt = voc.pack_ground_truth(y_true, dev)
    ap, ap07 = (t.cpu().numpy() for t in voc.evaluate_detections(buf[:n], gt, num_classes, conf_threshold))
    with np.errstate(invalid='ignore'):
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore', RuntimeWarning)
            return {'AP': ap, 'mAP': np.nanmean(ap), 'AP_VOC': ap07, 'mAP_VOC': np.nanmean(ap07)}
