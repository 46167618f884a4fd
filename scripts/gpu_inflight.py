"""Throughput decomposition with several batches in flight: predict side only, train side only, both (graphs of
consecutive batches replayed round-robin on LANES streams).  Usage: gpurun -- 'python scripts/gpu_inflight.py'"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth
B, N, C, GMAX, TOPK = 64, 16800, 20, 64, 200
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
ROT = 4
sets = [tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7 + r, B, GMAX, C) + synth.predictions(1234 + r, B, N, C)) for r in range(ROT)]
ys = [torch.empty((B, N, 27), device=dev) for _ in range(ROT)]
ds = [torch.empty((B, TOPK, 6), device=dev) for _ in range(ROT)]
pb.encode_truth_batch(*sets[0][:3], y_true=ys[0]); pb.decode_nms(*sets[0][3:], out=ds[0]); torch.cuda.synchronize()


def capture(fn):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g


variants = {
    'predict': lambda k: pb.decode_nms(*sets[k][3:], out=ds[k], top_k=TOPK),
    'train': lambda k: pb.encode_truth_batch(*sets[k][:3], y_true=ys[k]),
}
for name, fn in variants.items():
    gs = [capture(lambda k=k: fn(k)) for k in range(ROT)]
    for S in (1, 2, 4):
        streams = [torch.cuda.Stream(device=dev) for _ in range(S)]
        def run(steps):
            for i in range(steps):
                with torch.cuda.stream(streams[i % S]):
                    gs[i % ROT].replay()
        run(20); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 400
        e0.record()
        for s in streams: s.wait_event(e0)
        run(steps)
        for s in streams: torch.cuda.current_stream().wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(json.dumps({'side': name, 'in_flight': S, 'us_per_step': round(ms / steps * 1e3, 2), 'img_s': round(B * steps / ms * 1e3)}), flush=True)
