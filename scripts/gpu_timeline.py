"""Timeline of one graph replay of the bench step (both branches): per kernel first-CTA entry / last-CTA exit in
microseconds relative to the earliest entry, from the in-kernel %globaltimer stamps (odk_timeline_set)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import pipeline, synth, _lib
B, N, C, GMAX, TOPK = 64, 16800, 20, 64, 200
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
sets = [tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7 + r, B, GMAX, C) + synth.predictions(1234 + r, B, N, C)) for r in range(3)]
y = torch.empty((B, N, 27), device=dev); d = torch.empty((B, TOPK, 6), device=dev)
names = ['candidates', 'greedy', 'rows', 'patch', 'emit', 'nms', 'ignore']
buf = torch.zeros(14, dtype=torch.int64, device=dev)
_lib.call('odk_timeline_set', buf.data_ptr())  # launch argument: must be installed before the graphs are captured
for two in (True, False):
    gs = [pipeline.capture_step(pb, s[:3], s[3:], y, d, top_k=TOPK, two_streams=two) for s in sets]
    for i in range(6): gs[i % 3].replay()
    torch.cuda.synchronize()
    init = torch.tensor([2**63 - 1, 0] * 7, dtype=torch.int64, device=dev)
    acc = {}
    for rep in range(5):
        buf.copy_(init)
        gs[rep % 3].replay()
        torch.cuda.synchronize()
        t = buf.cpu().numpy().reshape(7, 2).astype('float64')
        t0 = t[:, 0].min()
        for k, n in enumerate(names):
            acc.setdefault(n, []).append(((t[k, 0] - t0) / 1e3, (t[k, 1] - t0) / 1e3))
    out = {n: [round(sorted(x[0] for x in v)[2], 1), round(sorted(x[1] for x in v)[2], 1)] for n, v in acc.items()}
    print(json.dumps({'two_streams': two, 'start_end_us': out}), flush=True)
