"""Device time of the assignment kernels alone (odk_assign: candidates + greedy walk) for uniform, crowded and stress
ground truth: python scripts/gpu_greedy_probe.py   (ODK_GREEDY_MODE=0/1/2 selects the walk mode)."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import _lib, synth

def run(name, pb, gt, reps=20):
    boxes, counts = (torch.from_numpy(gt[0]).cuda(), torch.from_numpy(gt[2]).cuda())
    for _ in range(3):
        pb.assign_batch(boxes, counts)
    torch.cuda.synchronize()
    _lib.profile_enable(True)
    for _ in range(reps):
        pb.assign_batch(boxes, counts)
    torch.cuda.synchronize()
    _lib.profile_enable(False)
    prof = _lib.profile_read()
    print(json.dumps({'case': name, 'mode': os.environ.get('ODK_GREEDY_MODE', '0'), **{k: round(ms / c * 1e3, 1) for k, (ms, c) in prof.items()}}), flush=True)

pb = odk.ObjectDetector.load_voc((640, 640)).pb
run('voc uniform B=64 G<=64', pb, synth.gt_batch(7, 64, 64, 20, crowded_every=0))
run('voc mixed B=64 (every 4th crowded)', pb, synth.gt_batch(7, 64, 64, 20))
run('voc all crowded B=64', pb, synth.gt_batch(7, 64, 64, 20, crowded_every=1))
run('voc uniform B=64 G=64', pb, synth.gt_batch(7, 64, 64, 20, crowded_every=0, fixed_g=64))
pb5 = odk.PriorBoxes((1120, 1120), [140, 70, 35], 80)
pb5.from_dict({'map_sizes': [140, 70, 35], 'pb_size_patterns': odk.od._VOC_JSON_DATA['pb_size_patterns']})
run('stress B=8 G=512', pb5, synth.gt_batch(5, 8, 512, 80, fixed_g=512, crowded_every=0, centre=(0.05, 0.95), size=(0.01, 0.3)), reps=10)
