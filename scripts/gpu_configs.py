"""Device time of each BASELINE.json config on one GPU (CUDA-graph replays, inputs rotating over 3 sets): cfg2 decode+NMS
B=64, cfg3 assign/encode B=256 (and the fused assign+loss), cfg5 dense-anchor stress B=8.  Prints one JSON line per config."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import object_detector_b200 as odk
from object_detector_b200 import synth
from oracle.restated import VOC_PATTERNS
dev = torch.device('cuda', 0)
ROT = 3


def timed(graphs, steps=60):
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


def put(arrs, reps=1):
    t = tuple(torch.from_numpy(x).to(dev) for x in arrs)
    return tuple(x.repeat(reps, *([1] * (x.dim() - 1))) for x in t) if reps > 1 else t


pb = odk.ObjectDetector.load_voc((640, 640)).pb
N, C = len(pb), 20
# cfg2
heads = [put(synth.predictions(1234 + r, 64, N, C)) for r in range(ROT)]
d = torch.empty((64, 200, 6), device=dev)
us = timed([cap(lambda h=h: pb.decode_nms(*h, out=d)) for h in heads])
us_s = timed([cap(lambda h=h: pb.decode_nms(*h, out=d, nms_all_threshold=0.5)) for h in heads])
print(json.dumps({'config': 'cfg2 decode+NMS B=64 C=20', 'us': round(us, 1), 'img_per_s': round(64 / us * 1e6), 'strict_us': round(us_s, 1),
                  'read_gbs': round(64 * N * 4 * (1 + C + 4) / us / 1e3)}), flush=True)
# cfg3
gts = [put(synth.gt_batch(7 + r, 256, 64, C)) for r in range(ROT)]
ys = [torch.empty((256, N, 27), device=dev) for _ in range(ROT)]
us = timed([cap(lambda g=g, y=y: pb.encode_truth_batch(*g, y_true=y)) for g, y in zip(gts, ys)])
preds = [torch.rand((256, N, 27), device=dev) * 0.98 + 0.01 for _ in range(ROT)]
us_f = timed([cap(lambda g=g, p=p: pb.assign_loss_batch(*g, p)) for g, p in zip(gts, preds)])
us_l = timed([cap(lambda y=y, p=p: pb.loss_terms(y, p)) for y, p in zip(ys, preds)])
print(json.dumps({'config': 'cfg3 assign+encode B=256 G<=64', 'us': round(us, 1), 'img_per_s': round(256 / us * 1e6),
                  'write_gbs': round(256 * N * 27 * 4 / us / 1e3), 'assign_loss_fused_us': round(us_f, 1), 'loss_dense_us': round(us_l, 1),
                  'loss_dense_gbs': round(2 * 256 * N * 27 * 4 / us_l / 1e3)}), flush=True)
del ys, preds, gts, heads
torch.cuda.empty_cache()
# cfg5
sp = odk.PriorBoxes((1120, 1120), [140, 70, 35], 80)
sp.from_dict({'map_sizes': [140, 70, 35], 'pb_size_patterns': VOC_PATTERNS})
NS, CS = len(sp), 80
heads = [put(synth.predictions(55 + r, 8, NS, CS)) for r in range(ROT)]
d = torch.empty((8, 200, 6), device=dev)
us_p = timed([cap(lambda h=h: sp.decode_nms(*h, out=d)) for h in heads], steps=30)
gts = [put(synth.gt_batch(9 + r, 8, 512, CS, fixed_g=512, crowded_every=0, size=(0.01, 0.3))) for r in range(ROT)]
y = torch.empty((8, NS, 87), device=dev)
st = sp.encode_truth_batch(*gts[0], y_true=y)['status'].cpu().numpy().tolist()
us_t = timed([cap(lambda g=g: sp.encode_truth_batch(*g, y_true=y)) for g in gts], steps=30)
print(json.dumps({'config': 'cfg5 stress N=205800 C=80 G=512 B=8', 'predict_us': round(us_p, 1), 'predict_img_per_s': round(8 / us_p * 1e6),
                  'predict_read_gbs': round(8 * NS * 4 * (1 + CS + 4) / us_p / 1e3), 'train_us': round(us_t, 1),
                  'train_img_per_s': round(8 / us_t * 1e6), 'train_write_gbs': round(8 * NS * 87 * 4 / us_t / 1e3), 'status': st}), flush=True)
# evaluation kernels at dataset size: VOC07-test (4952 images, 20 classes) and COCO-val-sized (5000 images, 80 classes)
del heads, gts, y
torch.cuda.empty_cache()
from object_detector_b200 import coco, voc  # noqa: E402
for name, n_img, ncls in (('voc07-test-sized', 4952, 20), ('coco-val-sized', 5000, 80)):
    g_list, p_list = synth.coco_eval_set(2017, n_img, ncls, n_fp=60)
    y_true = [odk.ml.ObjectsAnnotation('.', w, h, c, b, areas=a, crowdeds=cr) for (w, h, c, b, a, cr) in g_list]
    y_pred = [odk.ml.ObjectsPrediction(c, s, b) for (c, s, b) in p_list]
    dets = voc.pack_predictions(y_pred, dev)
    gv, gc = voc.pack_ground_truth(y_true, dev), coco.pack_ground_truth(y_true, dev)
    out = {'config': f'eval {name}: {n_img} images x {dets.shape[1]} detections, {ncls} classes'}
    for label, fn in (('voc_ap_us', lambda: voc.evaluate_detections(dets, gv, ncls)), ('coco_ap_us', lambda: coco.evaluate_detections(dets, gc, ncls))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[label] = round(e0.elapsed_time(e1) / 10 * 1e3, 1)
    print(json.dumps(out), flush=True)
