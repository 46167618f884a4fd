"""Loss-side kernels, device time per launch from CUDA-graph replays: dense loss over y_true+y_pred,
encode_truth + dense loss, and the fused assign+loss that never writes y_true (diagnostic; prints JSON lines)."""
import json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth
N, C, GMAX, ROT = 16800, 20, 64, 3
W = 3 + C + 4
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb


def timed(graphs, steps=100):
    for i in range(6): graphs[i % len(graphs)].replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(steps): graphs[i % len(graphs)].replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e3


def cap(fn):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): fn()
    return g


for B in (64, 256):
    gts = [tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7 + r, B, GMAX, C)) for r in range(ROT)]
    preds = [torch.rand((B, N, W), device=dev) * 0.98 + 0.01 for _ in range(ROT)]
    ys = [torch.empty((B, N, W), device=dev) for _ in range(ROT)]
    for g, y in zip(gts, ys): pb.encode_truth_batch(*g, y_true=y)
    res = {'B': B}
    res['loss_dense_us'] = timed([cap(lambda y=y, p=p: pb.loss_terms(y, p)) for y, p in zip(ys, preds)])
    res['loss_dense_gbs'] = round(2 * B * N * W * 4 / (res['loss_dense_us'] * 1e-6) / 1e9)
    res['encode_us'] = timed([cap(lambda g=g, y=y: pb.encode_truth_batch(*g, y_true=y)) for g, y in zip(gts, ys)])
    res['encode_then_loss_us'] = timed([cap(lambda g=g, y=y, p=p: (pb.encode_truth_batch(*g, y_true=y), pb.loss_terms(y, p)))
                                        for g, y, p in zip(gts, ys, preds)])
    res['assign_loss_fused_us'] = timed([cap(lambda g=g, p=p: pb.assign_loss_batch(*g, p)) for g, p in zip(gts, preds)])
    res['loss_grad_dense_us'] = timed([cap(lambda y=y, p=p: pb.loss_grad(y, p)) for y, p in zip(ys, preds)])
    res['loss_grad_dense_gbs'] = round(3 * B * N * W * 4 / (res['loss_grad_dense_us'] * 1e-6) / 1e9)
    res['train_step_dense_us'] = timed([cap(lambda g=g, y=y, p=p: (pb.encode_truth_batch(*g, y_true=y), pb.loss_terms(y, p), pb.loss_grad(y, p)))
                                        for g, y, p in zip(gts, ys, preds)])
    res['train_step_fused_us'] = timed([cap(lambda g=g, p=p: pb.assign_loss_batch(*g, p, with_grad=True)) for g, p in zip(gts, preds)])
    a = pb.loss_terms(ys[0], preds[0]); f = pb.assign_loss_batch(*gts[0], preds[0])['terms']
    res['fused_vs_dense_max_rel'] = float(((a - f).abs() / a.abs().clamp_min(1e-12)).max())
    print(json.dumps({k: (round(v, 2) if isinstance(v, float) and v > 1e-3 else v) for k, v in res.items()}), flush=True)
