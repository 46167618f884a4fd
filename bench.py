"""bench.py — images/sec of the detection hot path (assign+encode + decode+NMS) at 640x640.

    python bench.py --gpus N --steps K --warmup W            # this build (CUDA, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port)

One "step" = one pass of the hot path over one batch of synthetic input per GPU:
  train side   odk_assign_encode  on B=64 images, <=64 GT each (BASELINE config 3 shapes)
  predict side odk_decode_nms     on B=64 images, N=16800 anchors, C=20 (BASELINE config 2)
`value`  = images/s with inputs resident in HBM (images = B per step per GPU, all ranks summed).
`e2e`    = the same through the host-buffer API: pinned host inputs -> H2D -> kernels -> D2H of
           y_true and of the (B,200,6) detections, every step.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'images/sec assign+decode+NMS @640x640'
B, N, C, GMAX, TOPK = 64, 16800, 20, 64, 200
ROT = 8  # distinct input sets rotated through so every step reads data that is not L2-resident
LANES, GROUP = 2, 4  # pipeline.StepRing: GROUP consecutive slots per CUDA graph (a branch each), graphs alternating on LANES streams
WORKLOAD = 'cfg2+cfg3: Darknet53 640x640 VOC anchors (N=16800, C=20), batch 64: assign+encode (<=64 GT/img) + decode + per-class NMS (top-k 200)'

# algorithmic HBM bytes per image
BYTES_DECODE_NMS = 4 * N * (1 + C + 4) + 4800          # 1 684 800: read head outputs, write (200,6)  (whole fused op)
BYTES_SCORE_EMIT = 4 * N * (1 + C)                     # 1 411 200: what score_emit_kernel itself must read (locs are decoded lazily by NMS)
BYTES_ASSIGN_ENCODE = 20 * GMAX + 4 * N * (3 + C + 4)  # 1 815 680: read GT, write y_true


def load_peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured'
    except Exception:  # noqa: BLE001
        return 6650.0, 'fallback'


NCU_NAMES = {'score_emit_kernel': 'decode_score_kernel<256, 0', 'encode_rows_kernel': 'encode_rows_kernel<4'}  # ncu name prefixes


def load_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture
    (profiles/ncu_traffic.json, written by scripts/ncu_summary.py), or None."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
            kernels = json.load(f)['kernels']
        prefix = NCU_NAMES.get(kernel, kernel)
        k = next(v for name, v in kernels.items() if name.startswith(prefix))
        return k['dram_read_bytes'] + k['dram_write_bytes']
    except Exception:  # noqa: BLE001
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            'hw_slowdown': getattr(nv, 'nvmlClocksEventReasonHwSlowdown', 0x8),
            'hw_thermal_slowdown': getattr(nv, 'nvmlClocksEventReasonHwThermalSlowdown', 0x40),
            'sw_thermal_slowdown': getattr(nv, 'nvmlClocksEventReasonSwThermalSlowdown', 0x20),
            'sw_power_cap': getattr(nv, 'nvmlClocksEventReasonSwPowerCap', 0x4),
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {'sm_mhz': med, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons), 'samples': len(self.samples)}


# ------------------------------------------------------------------------------------------ CPU (oracle) legs
def _cpu_worker(args):
    """One image through the reference algorithm (oracle port): encode_truth + predict tail."""
    seed, = args
    from object_detector_b200 import synth
    from oracle import restated as R
    t = _cpu_worker.tables
    boxes, classes, counts = synth.gt_batch(1000 + seed, 1, GMAX, C, crowded_every=0)
    objs, clfs, locs = synth.predictions(2000 + seed, 1, N, C)
    t0 = time.perf_counter()
    R.encode_truth(t, [(classes[0, :counts[0]], boxes[0, :counts[0]])])
    R.predict_tail(t, objs, clfs, locs, inner=R.numpy_nms)
    return time.perf_counter() - t0


def _cpu_init():
    os.environ.setdefault('OMP_NUM_THREADS', '1')
    from oracle import restated as R
    _cpu_worker.tables = R.AnchorTables((640, 640), R.VOC_MAP_SIZES, R.VOC_PATTERNS, C)


def cpu_images_per_sec(n_images, procs):
    """Times the oracle port on `n_images` synthetic images with `procs` worker processes."""
    import multiprocessing as mp
    if procs <= 1:
        _cpu_init()
        t0 = time.perf_counter()
        for s in range(n_images):
            _cpu_worker((s,))
        return n_images / (time.perf_counter() - t0)
    ctx = mp.get_context('fork')
    with ctx.Pool(procs, initializer=_cpu_init) as pool:
        pool.map(_cpu_worker, [(s,) for s in range(procs)])  # warm the workers (imports, tables)
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(s,) for s in range(n_images)], chunksize=1)
        return n_images / (time.perf_counter() - t0)


def run_reference(args):
    """--impl reference: the reference's CPU algorithm for the same workload on all host cores.
    The unmodified reference cannot travel to the GPU box (Python tree mounted only in the build
    container), so this is the oracle port (cpu_baseline.kind = "port"); rank 0 only."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    per_step = procs  # one image per worker per step: a bounded sample of the B=64 batch
    import multiprocessing as mp
    ctx = mp.get_context('fork')
    with ctx.Pool(procs, initializer=_cpu_init) as pool:
        for _ in range(max(1, min(args.warmup, 1))):
            pool.map(_cpu_worker, [(s,) for s in range(per_step)], chunksize=1)
        t0 = time.perf_counter()
        for k in range(args.steps):
            pool.map(_cpu_worker, [(k * per_step + s,) for s in range(per_step)], chunksize=1)
        dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    sample = f'{per_step} images/step x {args.steps} steps of the batch-64 workload, {procs} worker processes'
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'images/s', 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': WORKLOAD, 'sample': sample},
        'cpu_baseline': {'value': value, 'unit': 'images/s', 'cores': procs, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': 'images/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ CUDA arm
def run_cuda(args):
    global ROT
    ROT = max(1, args.slots)
    import torch
    import torch.distributed as dist
    import object_detector_b200 as odk
    from object_detector_b200 import _lib, pipeline, synth

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    # native libraries (NCCL's version banner) write to fd 1: keep stdout clean for the single JSON line
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    assert world == args.gpus or world == 1, f'WORLD_SIZE={world} but --gpus {args.gpus}'
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    od = odk.ObjectDetector.load_voc((640, 640))
    pb = od.pb
    # ---- synthetic inputs (seeded per rank), ROT distinct sets resident in HBM
    host_sets = []
    for r in range(ROT):
        gtb, gtc, gtn = synth.gt_batch(7 + 100 * rank + r, B, GMAX, C)
        objs, clfs, locs = synth.predictions(1234 + 100 * rank + r, B, N, C)
        host_sets.append((gtb, gtc, gtn, objs, clfs, locs))
    dev_sets = [tuple(torch.from_numpy(x).to(dev) for x in s) for s in host_sets]
    y_trues = [torch.empty((B, N, 3 + C + 4), dtype=torch.float32, device=dev) for _ in range(ROT)]
    detss = [torch.empty((B, TOPK, 6), dtype=torch.float32, device=dev) for _ in range(ROT)]
    y_true, dets = y_trues[0], detss[0]

    def step_eager(i):
        gtb, gtc, gtn, objs, clfs, locs = dev_sets[i % ROT]
        pb.encode_truth_batch(gtb, gtc, gtn, y_true=y_true)
        pb.decode_nms(objs, clfs, locs, top_k=TOPK, out=dets)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for i in range(3):
        step_eager(i)  # uploads the anchor tables, sets kernel attributes
    torch.cuda.synchronize()
    # one CUDA graph per input set: the train side and the predict side of a step are independent and run on
    # parallel branches; replaying removes the per-launch host cost of the five small kernels
    # and `lanes` batches in flight: the latency-bound tail of one batch runs beside the HBM-bound head of the next
    group = max(1, min(args.group, ROT))
    while ROT % group:
        group -= 1
    lanes = max(1, min(args.lanes, ROT // group))
    while (ROT // group) % lanes:
        lanes -= 1
    slots = [(s[:3], s[3:], y_trues[k], detss[k]) for k, s in enumerate(dev_sets)]
    ring = pipeline.StepRing(pb, slots, lanes=lanes, group=group, top_k=TOPK, use_graph=not args.no_graph)
    graphs = ring.group_graphs

    def timed_ring(steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ring.fork()
        ring.run(steps)
        ring.join()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    timed_ring(max(args.warmup, 3))
    sampler = ClockSampler(local)
    sampler.start()
    ms_total = timed_ring(args.steps)
    clocks = sampler.stop()
    value = B * world * args.steps / (ms_total * 1e-3)
    # the same K steps with one batch in flight (latency of a step, and the figure earlier rounds reported)
    one = pipeline.StepRing(pb, slots, lanes=1, group=1, top_k=TOPK, use_graph=not args.no_graph) if lanes * group > 1 else ring
    ring_saved, ring = ring, one
    timed_ring(3)
    ms_serial = timed_ring(args.steps)
    ring = ring_saved

    # ---- per-kernel device time over the same K steps (second pass, library event hooks on)
    _lib.profile_enable(True)
    timed(step_eager, args.steps)  # eager launches on one stream: the library brackets every kernel with events
    _lib.profile_enable(False)
    prof = _lib.profile_read()
    launches_per_step = sum(c for _, c in prof.values()) // max(args.steps, 1)
    peak, peak_src = load_peaks()
    alg_bytes = {'score_emit_kernel': B * BYTES_SCORE_EMIT, 'encode_rows_kernel': B * BYTES_ASSIGN_ENCODE}
    kernels = {k: {'ms_per_launch': ms / c, 'launches': c, 'share': ms / max(1e-9, sum(m for m, _ in prof.values()))}
               for k, (ms, c) in prof.items()}
    # the roofline line is for the dominant HBM-streaming kernel (largest device time among the two
    # that carry the algorithmic bytes); the NMS kernels move <1% of the bytes and are latency-bound
    dom = max(alg_bytes, key=lambda k: prof.get(k, (0.0, 1))[0])
    dms, dcnt = prof[dom]
    achieved = alg_bytes[dom] / (dms / dcnt * 1e-3) / 1e9
    roofline = {'bound': 'hbm', 'kernel': dom, 'achieved': achieved, 'peak': peak, 'peak_source': peak_src, 'unit': 'GB/s',
                'frac': achieved / peak, 'traffic': load_traffic(dom), 'algorithmic_bytes_per_launch': alg_bytes[dom],
                'kernels': kernels}
    for k, nbytes in alg_bytes.items():
        if k in prof:
            kernels[k]['achieved_gbs'] = nbytes / (prof[k][0] / prof[k][1] * 1e-3) / 1e9
            kernels[k]['frac'] = kernels[k]['achieved_gbs'] / peak
    if all(k in prof for k in alg_bytes):
        # encode_rows leaves about half of its 116 MB dirty in the 126 MB L2 when it ends; that write-back lands on
        # whichever HBM-bound kernel runs next (here score_emit), so the two are also reported as one pair
        pair_ms = sum(prof[k][0] / prof[k][1] for k in alg_bytes)
        pair_gbs = sum(alg_bytes.values()) / (pair_ms * 1e-3) / 1e9
        roofline['streaming_pair'] = {'kernels': sorted(alg_bytes), 'algorithmic_bytes': sum(alg_bytes.values()),
                                      'ms': pair_ms, 'achieved_gbs': pair_gbs, 'frac': pair_gbs / peak}

    # ---- end to end through host buffers: pinned inputs -> H2D -> kernels -> D2H of y_true and detections, every step,
    #      through the public HostPipeline (copy-in / compute / copy-out streams, double buffered)
    pin_in = [tuple(torch.from_numpy(x).pin_memory() for x in s) for s in host_sets]
    pipe = pipeline.HostPipeline(pb, B, GMAX, top_k=TOPK, use_graph=not args.no_graph)
    h2d, d2h = pipe.h2d_bytes, pipe.d2h_bytes
    sink = [0.0]

    def e2e_step(i):
        k = pipe.submit(*pin_in[i % ROT])
        if i >= 1:  # consume the previous batch's host result while this one is in flight
            y, d = pipe.result(k ^ 1)
            sink[0] += float(d[0, 0, 1])

    def e2e_timed(steps):
        barrier()
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            e2e_step(i)
        for st in (pipe.s_in, pipe.s_run, pipe.s_out):
            cur.wait_stream(st)
        e1.record()
        barrier()
        dt = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return float(dt.item())

    e2e_steps = max(3, min(args.steps, 50))
    e2e_timed(4)
    e2e_ms = e2e_timed(e2e_steps)
    e2e_value = B * world * e2e_steps / (e2e_ms * 1e-3)

    # ---- CPU baseline (oracle port, 1 core, bounded sample) on rank 0 at N=1
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_img = 24
        v = cpu_images_per_sec(n_img, 1)
        cpu = {'value': v, 'unit': 'images/s', 'cores': 1, 'kind': 'port',
               'sample': f'{n_img} images of the batch-64 workload (encode_truth G<=64 + decode + per-class NumPy NMS), single process'}

    if rank == 0:
        line = {
            'metric': METRIC, 'value': value, 'unit': 'images/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms_total / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'batch_per_gpu': B, 'anchors': N, 'classes': C, 'gt_max': GMAX, 'top_k': TOPK,
                       'l2': f'inputs rotate over {ROT} distinct sets ({ROT * (h2d + B * N * 27 * 4) >> 20} MB touched) > 126 MB L2',
                       'launch': ('eager' if graphs is None else 'CUDA graph per input set, train and predict sides on parallel branches')
                       + f'; {group} consecutive batches per graph (a branch each), graphs alternating on {lanes} streams'
                       f' = up to {lanes * group} batches in flight (each slot has its own inputs, y_true, detections, workspaces)',
                       'batches_in_flight': lanes * group, 'batches_per_graph': group, 'lanes': lanes, 'ms_per_step_one_in_flight': ms_serial / args.steps,
                       'parallelism': f'images sharded over {world} GPU(s), no data-path collective'},
            'clocks': clocks, 'e2e': {'value': e2e_value, 'unit': 'images/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                                      'steps': e2e_steps},
            'gpu_launches': launches_per_step * args.steps, 'gpu_launches_per_step': launches_per_step,
            'roofline': roofline, 'cpu_baseline': cpu,
        }
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='cuda', choices=['cuda', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-graph', action='store_true', help='eager launches instead of CUDA-graph replay')
    ap.add_argument('--slots', type=int, default=ROT, help='distinct resident batches (inputs + outputs) rotated through')
    ap.add_argument('--lanes', type=int, default=LANES, help='streams the step graphs alternate on')
    ap.add_argument('--group', type=int, default=GROUP, help='consecutive batches captured into one CUDA graph (a branch each)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == '__main__':
    main()
