"""BASELINE config 4: VOC07-test-sized evaluation (4952 synthetic images) sharded over N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/eval_multi_gpu.py [--images 4952]

Each rank: decode + per-class NMS on its contiguous image shard, written straight into its slab of the
all-gather buffer; one NCCL all-gather; AP on every rank.  Rank 0 prints one JSON line and, with
--check, compares the result against a single-rank run of the whole set (bit-identical detections).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk  # noqa: E402
from object_detector_b200 import dist as odist, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--images', type=int, default=4952)
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--check', action='store_true')
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', '0'), ('WORLD_SIZE', '1'), ('LOCAL_RANK', '0')))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    od = odk.ObjectDetector.load_voc((640, 640))
    n, c = len(od.pb), 20
    gts, _ = synth.eval_set(2007, args.images, c)
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, cl, b, d) for (w, h, cl, b, d) in gts]

    def head_fn(lo, hi):  # synthetic head outputs, a function of the image index only (rank-independent)
        parts = [synth.predictions(50000 + i, 1, n, c, clustered=(i % 2 == 0)) for i in range(lo, hi)]
        return tuple(torch.from_numpy(np.concatenate([p[k] for p in parts])).to(dev) for k in range(3))

    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = odist.sharded_decode_nms_eval(od.pb, head_fn, y_true, c, rank, world, batch=args.batch)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    line = {'workload': f'cfg4: {args.images} images, {world} GPU(s), decode+NMS+allgather+AP (includes synthetic input generation on host)',
            'ms': float(ms.item()), 'mAP': float(res['mAP']), 'mAP_VOC': float(res['mAP_VOC']), 'n_gpus': world}
    if args.check:
        ref = odist.sharded_decode_nms_eval(od.pb, head_fn, y_true, c, 0, 1, batch=args.batch)
        same = all(np.array_equal(res[k], ref[k], equal_nan=True) for k in ('AP', 'AP_VOC'))
        line['matches_single_rank'] = bool(same)
        assert same, 'sharded evaluation differs from the single-rank evaluation'
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
