"""score_emit_kernel tuning sweep (diagnostic): ring depth / CTAs per SM / tile height via the ODK_EMIT_* environment
knobs, one subprocess per setting.  Prints per-launch time from the library's event hooks and the predict-side graph time."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json, torch
sys.path.insert(0, %r)
import object_detector_b200 as odk
from object_detector_b200 import synth, _lib
B = int(sys.argv[1]); N, C, ROT = 16800, 20, 3
dev = torch.device('cuda', 0)
pb = odk.ObjectDetector.load_voc((640, 640)).pb
sets = [tuple(torch.from_numpy(x).to(dev) for x in synth.predictions(1234 + r, min(B, 64), N, C)) for r in range(ROT)]
if B > 64: sets = [tuple(x.repeat(B // 64, 1, 1) for x in s) for s in sets]
d = torch.empty((B, 200, 6), device=dev)
def run(i): pb.decode_nms(*sets[i %% ROT], out=d)
for i in range(5): run(i)
torch.cuda.synchronize()
_lib.profile_enable(True)
for i in range(100): run(i)
torch.cuda.synchronize()
_lib.profile_enable(False)
prof = _lib.profile_read()
gs = []
for s in sets:
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): pb.decode_nms(*s, out=d)
    gs.append(g)
for i in range(10): gs[i %% ROT].replay()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for i in range(200): gs[i %% ROT].replay()
e1.record(); torch.cuda.synchronize()
us = {k: round(ms / c * 1e3, 2) for k, (ms, c) in prof.items()}
emit = us.get('score_emit_kernel', 0)
print(json.dumps({'B': B, 'us': us, 'emit_gbs': round(B * N * 4 * (1 + C) / (emit * 1e-6) / 1e9, 0) if emit else None,
                  'graph_pred_us': round(e0.elapsed_time(e1) / 200 * 1e3, 2)}))
''' % ROOT
settings = [{}, {'ODK_EMIT_STAGES': '4'}, {'ODK_EMIT_PER_SM': '3'}, {'ODK_EMIT_STAGES': '2', 'ODK_EMIT_PER_SM': '4'},
            {'ODK_EMIT_TILE': '128', 'ODK_EMIT_STAGES': '4', 'ODK_EMIT_PER_SM': '4'},
            {'ODK_EMIT_TILE': '128', 'ODK_EMIT_STAGES': '6', 'ODK_EMIT_PER_SM': '4'},
            {'ODK_EMIT_TILE': '128', 'ODK_EMIT_STAGES': '3', 'ODK_EMIT_PER_SM': '6'},
            {'ODK_EMIT_TILE': '128', 'ODK_EMIT_STAGES': '3', 'ODK_EMIT_PER_SM': '8'}]
for batch in (64, 256):
    for env in settings:
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, '-c', CHILD, str(batch)], env=e, capture_output=True, text=True, timeout=300)
        print(json.dumps(env), r.stdout.strip() or r.stderr.strip()[-400:], flush=True)
