"""assign_candidates_kernel phase timing via ODK_DEBUG_SKIP (1: no ranking, 2: no ignore scatter, 4: exit after the box load)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json, torch
sys.path.insert(0, %r)
import object_detector_b200 as odk
from object_detector_b200 import synth, _lib
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
g = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7, 64, 64, 20))
y = torch.empty((64, 16800, 27), device=dev)
for _ in range(3): pb.encode_truth_batch(*g, y_true=y)
torch.cuda.synchronize()
_lib.profile_enable(True)
for _ in range(50): pb.encode_truth_batch(*g, y_true=y)
torch.cuda.synchronize()
_lib.profile_enable(False)
print(json.dumps({k: round(ms / c * 1e3, 1) for k, (ms, c) in _lib.profile_read().items()}))
''' % ROOT
for skip in ('0', '1', '2', '3', '4'):
    e = dict(os.environ); e['ODK_DEBUG_SKIP'] = skip
    r = subprocess.run([sys.executable, '-c', CHILD], env=e, capture_output=True, text=True, timeout=300)
    print('skip', skip, r.stdout.strip() or r.stderr.strip()[-300:], flush=True)
