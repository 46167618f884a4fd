"""Step time of the bench workload under different launch arrangements (diagnostic)."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import pipeline, synth
B, N, C, GMAX, TOPK, ROT = 64, 16800, 20, 64, 200, 3
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
sets = []
for r in range(ROT):
    s = synth.gt_batch(7 + r, B, GMAX, C) + synth.predictions(1234 + r, B, N, C)
    sets.append(tuple(torch.from_numpy(x).to(dev) for x in s))
y = torch.empty((B, N, 27), device=dev); d = torch.empty((B, TOPK, 6), device=dev)
def timed(fn, steps=200):
    for i in range(10): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e3
def eager(i):
    s = sets[i % ROT]; pb.encode_truth_batch(*s[:3], y_true=y); pb.decode_nms(*s[3:], top_k=TOPK, out=d)
def train_only(i):
    s = sets[i % ROT]; pb.encode_truth_batch(*s[:3], y_true=y)
def pred_only(i):
    s = sets[i % ROT]; pb.decode_nms(*s[3:], top_k=TOPK, out=d)
for i in range(3): eager(i)
res = {'eager': timed(eager), 'eager_train': timed(train_only), 'eager_pred': timed(pred_only)}
for two in (False, True):
    gs = [pipeline.capture_step(pb, s[:3], s[3:], y, d, top_k=TOPK, two_streams=two) for s in sets]
    res['graph_two' if two else 'graph_one'] = timed(lambda i: gs[i % ROT].replay())
# train-only and predict-only graphs
def cap(fn):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): fn()
    return g
gt = [cap(lambda s=s: pb.encode_truth_batch(*s[:3], y_true=y)) for s in sets]
gp = [cap(lambda s=s: pb.decode_nms(*s[3:], top_k=TOPK, out=d)) for s in sets]
res['graph_train'] = timed(lambda i: gt[i % ROT].replay())
res['graph_pred'] = timed(lambda i: gp[i % ROT].replay())
print(json.dumps({k: round(v, 1) for k, v in res.items()}))
