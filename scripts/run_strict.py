"""Driver for profiling the strict NMS path: a few fused strict decode+NMS calls on the config-2 workload."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth
pb = odk.ObjectDetector.load_voc((640, 640)).pb
h = tuple(torch.from_numpy(x).cuda() for x in synth.predictions(1234, 64, len(pb), 20))
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    pb.decode_nms(*h, nms_all_threshold=0.5)
torch.cuda.synchronize()
print('ok')
