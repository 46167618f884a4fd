"""Train-side latency probe (diagnostic): encode_truth graph time and per-kernel event times for GT mixes with and
without the crowded images, and for fixed GT counts."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth, _lib
B, N, C, GMAX = 64, 16800, 20, 64
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
y = torch.empty((B, N, 27), device=dev)
for name, kw in [('mix_crowded4', {}), ('mix_nocrowd', {'crowded_every': 0}), ('all_crowded', {'crowded_every': 1}),
                 ('g64_nocrowd', {'crowded_every': 0, 'fixed_g': 64}), ('g8_nocrowd', {'crowded_every': 0, 'fixed_g': 8})]:
    g = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7, B, GMAX, C, **kw))
    for _ in range(3): pb.encode_truth_batch(*g, y_true=y)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    for _ in range(50): pb.encode_truth_batch(*g, y_true=y)
    torch.cuda.synchronize()
    _lib.profile_enable(False)
    prof = {k: round(ms / c * 1e3, 1) for k, (ms, c) in _lib.profile_read().items()}
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr): pb.encode_truth_batch(*g, y_true=y)
    for _ in range(5): gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(100): gr.replay()
    e1.record(); torch.cuda.synchronize()
    print(json.dumps({'case': name, 'graph_us': round(e0.elapsed_time(e1) * 10, 1), 'kernels_us': prof}), flush=True)
