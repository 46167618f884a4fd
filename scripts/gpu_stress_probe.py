"""Config-5 (dense-anchor stress) per-kernel device times from the library's event hooks (diagnostic)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth, _lib
from oracle.restated import VOC_PATTERNS
dev = torch.device('cuda', 0)
sp = odk.PriorBoxes((1120, 1120), [140, 70, 35], 80)
sp.from_dict({'map_sizes': [140, 70, 35], 'pb_size_patterns': VOC_PATTERNS})
NS, CS = len(sp), 80
head = tuple(torch.from_numpy(x).to(dev) for x in synth.predictions(55, 8, NS, CS))
gt = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(9, 8, 512, CS, fixed_g=512, crowded_every=0, size=(0.01, 0.3)))
d = torch.empty((8, 200, 6), device=dev); y = torch.empty((8, NS, 87), device=dev)
for _ in range(3):
    sp.decode_nms(*head, out=d); sp.encode_truth_batch(*gt, y_true=y)
torch.cuda.synchronize()
_lib.profile_enable(True)
for _ in range(20):
    sp.decode_nms(*head, out=d); sp.encode_truth_batch(*gt, y_true=y)
torch.cuda.synchronize()
_lib.profile_enable(False)
print(json.dumps({k: round(ms / c * 1e3, 1) for k, (ms, c) in _lib.profile_read().items()}))
