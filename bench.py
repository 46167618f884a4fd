"""bench.py — images/sec of the detection hot path (assign+encode + decode+NMS) at 640x640.

    python bench.py --gpus N --steps K --warmup W            # this build (CUDA, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref), else the oracle port

One "step" = one pass of the hot path over one batch of synthetic input per GPU:
  train side   odk_assign_encode  on B=64 images, <=64 GT each (BASELINE config 3 shapes)
  predict side odk_decode_nms     on B=64 images, N=16800 anchors, C=20 (BASELINE config 2)
`value`  = images/s with inputs resident in HBM (images = B per step per GPU, all ranks summed).
`e2e`    = the same through the host-buffer API: pinned host inputs -> H2D -> kernels -> D2H of
           y_true and of the (B,200,6) detections, every step; `e2e.copy_only_img_s` is the ceiling the
           same copies reach with no kernels in between, `e2e_compact` / `e2e_fused` the other two result
           contracts of pipeline.HostPipeline.
`configs` = BASELINE configs 1, 3 (B=256), 5 (stress) and 4 (sharded eval + NCCL all-gather + VOC AP) as
           sub-results with device time, images/s and HBM fraction; `variants` = the headline step with clustered
           head outputs and/or strict_nms.
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
SETS = 8            # distinct input sets rotated through so every step reads data that is not L2-resident
SLOTS = 20          # resident batches (each with its own outputs and workspaces); slot k reads input set k % SETS
# pipeline.StepRing: GROUP consecutive slots = a group (a graph branch each), CHAIN consecutive groups per CUDA graph, graphs
# alternating on LANES streams.  2 x 5 x 2: 10 batches in flight, one host launch per 10 batches.  Chosen on 8 ranks sharing a
# host (profiles/ring_probe_r02x_n8.log): every extra graph boundary on a lane costs there (2 x 4 x 1: 43.7 us per step at
# K = 20, 2 x 5 x 2: 41.7-42.8), while one GPU is indifferent to the shape (41.3-41.7 us, profiles/ring_probe_r02x_n1.log).
LANES, GROUP, CHAIN = 2, 5, 2
WORKLOAD = 'cfg2+cfg3: Darknet53 640x640 VOC anchors (N=16800, C=20), batch 64: assign+encode (<=64 GT/img) + decode + per-class NMS (top-k 200)'
EVAL_IMAGES = 4952  # config 4

# algorithmic HBM bytes per image
BYTES_DECODE_NMS = 4 * N * (1 + C + 4) + 4800          # 1 684 800: read head outputs, write (200,6)  (whole fused op)
BYTES_ASSIGN_ENCODE = 20 * GMAX + 4 * N * (3 + C + 4)  # 1 815 680: read GT, write y_true


def static_config(world):
    """The workload description: identical in both arms (the reference arm prints the same dict)."""
    return {'workload': WORKLOAD, 'batch_per_gpu': B, 'anchors': N, 'classes': C, 'gt_max': GMAX, 'top_k': TOPK,
            'l2': f'every step uses a different input set ({SETS} distinct sets, {SETS * B * N * 25 * 4 >> 20} MB of head outputs) > 126 MB L2',
            'parallelism': f'images sharded over {world} GPU(s), no data-path collective in the step; config 4 adds the NCCL all-gather'}


def load_peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            return float(json.load(f)['hbm_gbs']), 'measured'
    except Exception:  # noqa: BLE001
        return 6650.0, 'fallback'


NCU_NAMES = {'score_emit_kernel': ('decode_score_kernel<256, 0',), 'encode_rows_kernel': ('encode_rows_kernel<4', 'encode_rows_persistent_kernel<4')}


def load_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture
    (profiles/ncu_traffic.json, written by scripts/ncu_summary.py), or None."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'ncu_traffic.json')) as f:
            kernels = json.load(f)['kernels']
        prefixes = NCU_NAMES.get(kernel, (kernel,))
        k = next(v for name, v in kernels.items() if name.startswith(prefixes))
        return k['dram_read_bytes'] + k['dram_write_bytes']
    except Exception:  # noqa: BLE001
        return None


def pin_affinity(local, world_local, enable=True):
    """One core set per rank: the CPUs NVML reports as local to the rank's GPU, split evenly among the ranks that share
    them (else an even split of the process's CPU set).  Must run before CUDA is initialised so that pinned host
    allocations are first touched from the rank's own cores."""
    avail = sorted(os.sched_getaffinity(0))
    info = {'pinned': False, 'cores_before': len(avail)}
    if not enable or world_local <= 1 or len(avail) < 2 * world_local:
        info['cores'] = len(avail)
        return info
    ideal = {}
    try:
        import pynvml
        pynvml.nvmlInit()
        words = (max(avail) // 64) + 1
        for g in range(world_local):
            h = pynvml.nvmlDeviceGetHandleByIndex(g)
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1]
            ideal[g] = tuple(c for c in cpus if c in set(avail))
    except Exception:  # noqa: BLE001
        ideal = {}
    if len(ideal) != world_local or any(len(v) == 0 for v in ideal.values()):
        ideal = {g: tuple(avail) for g in range(world_local)}
    sharers = [g for g in range(world_local) if ideal[g] == ideal[local]]
    mine = list(ideal[local])
    per = len(mine) // len(sharers)
    if per < 1:
        info['cores'] = len(avail)
        return info
    k = sharers.index(local)
    chosen = mine[k * per:(k + 1) * per]
    try:
        os.sched_setaffinity(0, chosen)
        info.update({'pinned': True, 'cores': len(chosen), 'first_core': chosen[0], 'last_core': chosen[-1],
                     'numa_local': ideal[local] != tuple(avail)})
    except Exception as e:  # noqa: BLE001
        info.update({'cores': len(avail), 'error': str(e)})
    return info


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML every 2 ms from before the warm-up until the end of the soak
    that follows the timed region; samples taken while ``load`` is set count as "under load"."""

    def __init__(self, index, period_s=0.002):
        super().__init__(daemon=True)
        self.period_s = period_s
        self.index, self.samples, self.loaded, self.reasons, self.max_mhz = index, [], [], set(), None
        self._stop_evt, self.load = threading.Event(), threading.Event()
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
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                self.samples.append(mhz)
                if self.load.is_set():
                    self.loaded.append(mhz)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period_s)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        src = self.loaded if self.loaded else self.samples
        med = float(np.median(src)) if src else None
        return {'sm_mhz': med, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons), 'samples': len(src),
                'samples_total': len(self.samples),
                'window': 'warm-up + timed region + a 0.25 s soak of the same step (samples under load only)'}


# ------------------------------------------------------------------------------------------ CPU legs
def _cpu_init_port():
    os.environ.setdefault('OMP_NUM_THREADS', '1')
    from oracle import restated as R
    _cpu_worker.tables = R.AnchorTables((640, 640), R.VOC_MAP_SIZES, R.VOC_PATTERNS, C)
    _cpu_worker.kind = 'port'


def _cpu_init_reference():
    """The unmodified reference modules (oracle/_ref, compiled from /root/reference by oracle/make_ref.py)."""
    os.environ.setdefault('OMP_NUM_THREADS', '1')
    import logging
    from oracle import ref_loader, restated as R
    tk = ref_loader.load()
    logging.disable(logging.CRITICAL)  # the reference logs a table summary per PriorBoxes
    pb = tk.dl.od_pb.PriorBoxes((640, 640), list(R.VOC_MAP_SIZES), C)
    pb.from_dict({'map_sizes': list(R.VOC_MAP_SIZES), 'pb_size_patterns': R.VOC_PATTERNS})
    _cpu_worker.tk, _cpu_worker.pb = tk, pb
    _cpu_worker.tables = R.AnchorTables((640, 640), R.VOC_MAP_SIZES, R.VOC_PATTERNS, C)
    _cpu_worker.kind = 'reference'


def _cpu_worker(args):
    """One image through the reference algorithm: encode_truth + predict tail (decode, score, per-class NMS, top-k).
    kind 'reference': PriorBoxes.encode_truth, PriorBoxes.decode_locs(xp=np) and ml.non_maximum_suppression are the
    unmodified reference code; the arg-max / objectness product and the per-class loop of the NMS layer are TensorFlow
    graph ops in the reference (od_net.py:177-190, layers.py:925-970) and come from the oracle's restatement.
    kind 'port': everything from oracle/restated.py."""
    seed, = args
    from object_detector_b200 import synth
    from oracle import restated as R
    t = _cpu_worker.tables
    boxes, classes, counts = synth.gt_batch(1000 + seed, 1, GMAX, C, crowded_every=0)
    objs, clfs, locs = synth.predictions(2000 + seed, 1, N, C)
    g = int(counts[0])
    t0 = time.perf_counter()
    if _cpu_worker.kind == 'reference':
        tk, pb = _cpu_worker.tk, _cpu_worker.pb
        pb.encode_truth([tk.ml.ObjectsAnnotation('.', 640, 640, classes[0, :g], boxes[0, :g])])
        cls, conf = R.score_anchors(t, objs, clfs)
        dec = pb.decode_locs(locs[0], xp=np).astype(np.float32)
        R.nms_layer_image(cls[0], conf[0], dec, C, TOPK, 0.01, 0.45, inner=tk.ml.non_maximum_suppression)
    else:
        R.encode_truth(t, [(classes[0, :g], boxes[0, :g])])
        R.predict_tail(t, objs, clfs, locs, inner=R.numpy_nms)
    return time.perf_counter() - t0


def cpu_images_per_sec(n_images, procs, init=_cpu_init_port):
    """Times the CPU path on `n_images` synthetic images with `procs` worker processes."""
    import multiprocessing as mp
    if procs <= 1:
        init()
        t0 = time.perf_counter()
        for s in range(n_images):
            _cpu_worker((s,))
        return n_images / (time.perf_counter() - t0)
    ctx = mp.get_context('fork')
    with ctx.Pool(procs, initializer=init) as pool:
        pool.map(_cpu_worker, [(s,) for s in range(procs)])  # warm the workers (imports, tables)
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(s,) for s in range(n_images)], chunksize=1)
        return n_images / (time.perf_counter() - t0)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores (rank 0 only).
    oracle/_ref (compiled from the unmodified reference by oracle/make_ref.py) when present, else the oracle port."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from oracle import ref_loader
    have_ref = ref_loader.available()
    init, kind = (_cpu_init_reference, 'reference') if have_ref else (_cpu_init_port, 'port')
    cores = len(os.sched_getaffinity(0))
    procs = max(1, min(cores, 64))
    per_step = procs  # one image per worker per step: a bounded sample of the B=64 batch
    import multiprocessing as mp
    ctx = mp.get_context('fork')
    with ctx.Pool(procs, initializer=init) as pool:
        for _ in range(max(1, min(args.warmup, 1))):
            pool.map(_cpu_worker, [(s,) for s in range(per_step)], chunksize=1)
        t0 = time.perf_counter()
        for k in range(args.steps):
            pool.map(_cpu_worker, [(k * per_step + s,) for s in range(per_step)], chunksize=1)
        dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    sample = (f'{per_step} images/step x {args.steps} steps of the batch-{B} workload, {procs} worker processes; '
              + ('unmodified reference modules (' + ref_loader.source() + ': ' + str(ref_loader.REFERENCE_ROOT) + ')' if have_ref else 'oracle port (oracle/_ref absent)'))
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'images/s', 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': static_config(args.gpus), 'cores': procs, 'images_per_step': per_step,
        'cpu_baseline': {'value': value, 'unit': 'images/s', 'cores': procs, 'kind': kind, 'sample': sample},
        'e2e': {'value': value, 'unit': 'images/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ CUDA arm
def run_cuda(args):
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    world_local = int(os.environ.get('LOCAL_WORLD_SIZE', str(world)))
    affinity = pin_affinity(local, world_local, enable=args.affinity and not args.no_affinity)  # before CUDA init / pinned allocations

    import torch
    import torch.distributed as dist
    import object_detector_b200 as odk
    from object_detector_b200 import _lib, pipeline, synth

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

    def max_over_ranks(x):
        t = torch.tensor([float(x)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    peak, peak_src = load_peaks()
    use_graph = not args.no_graph
    od = odk.ObjectDetector.load_voc((640, 640))
    pb = od.pb
    spans = pb.span_stats(dev)
    # what score_emit must read: every objectness value, and the class rows of the anchors that can become candidates
    # (pb_mask = 1); the rows of masked anchors are not on the algorithmic path (conf = x * 0)
    bytes_score_emit = 4 * N + 4 * C * spans['valid'] if os.environ.get('ODK_EMIT_SPANS', '1') != '0' else 4 * N * (1 + C)

    # ---- synthetic inputs (seeded per rank), SETS distinct sets resident in HBM
    n_sets, n_slots = max(1, args.sets), max(1, args.slots)
    host_sets = []
    for r in range(n_sets):
        gtb, gtc, gtn = synth.gt_batch(7 + 100 * rank + r, B, GMAX, C)
        objs, clfs, locs = synth.predictions(1234 + 100 * rank + r, B, N, C)
        host_sets.append((gtb, gtc, gtn, objs, clfs, locs))
    dev_sets = [tuple(torch.from_numpy(x).to(dev) for x in s) for s in host_sets]
    y_trues = [torch.empty((B, N, 3 + C + 4), dtype=torch.float32, device=dev) for _ in range(n_slots)]
    detss = [torch.empty((B, TOPK, 6), dtype=torch.float32, device=dev) for _ in range(n_slots)]

    def step_eager(i, sets=dev_sets, **kw):
        gtb, gtc, gtn, objs, clfs, locs = sets[i % len(sets)]
        pb.encode_truth_batch(gtb, gtc, gtn, y_true=y_trues[0])
        pb.decode_nms(objs, clfs, locs, top_k=TOPK, out=detss[0], **kw)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    for i in range(3):
        step_eager(i)  # uploads the anchor tables, sets kernel attributes
    torch.cuda.synchronize()

    group = max(1, min(args.group, n_slots))
    while n_slots % group:
        group -= 1
    chain = max(1, args.chain)
    while n_slots % (group * chain):
        chain -= 1
    lanes = max(1, min(args.lanes, n_slots // (group * chain)))
    while (n_slots // (group * chain)) % lanes:
        lanes -= 1

    def host_calls(steps):
        """The launches StepRing.run makes for `steps` steps from step 0."""
        span, i, calls = group * chain, 0, 0
        while i < steps:
            i += span if (chain > 1 and i % span == 0 and i + span <= steps) else (group if (use_graph and i % group == 0 and i + group <= steps) else 1)
            calls += 1
        return calls

    def make_slots(sets):
        return [(sets[k % len(sets)][:3], sets[k % len(sets)][3:], y_trues[k], detss[k]) for k in range(n_slots)]

    def timed_ring(ring, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ring.fork()
        ring.run(steps)
        ring.join()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    sampler = ClockSampler(local, args.clock_period_ms * 1e-3)
    sampler.start()
    ring = pipeline.StepRing(pb, make_slots(dev_sets), lanes=lanes, group=group, top_k=TOPK, use_graph=use_graph, chain=chain)
    sampler.load.set()
    timed_ring(ring, max(args.warmup, 3))
    timed_ring(ring, n_slots)  # untimed as well: one pass over every slot, so that every graph of the ring has been launched once
    # the K steps launched group by group (K / group host calls inside the timed region): the headline.  The region is
    # ~1 ms at the driver's K = 20 and the figure is the max over ranks, so one descheduled launch thread on one of 8 ranks
    # doubles it (seen once in 12 runs: profiles/bench_r02o_n4.json); three regions of exactly K steps are timed back to
    # back and the MEDIAN is reported (all three are listed under schedule.timed_regions_ms) ...
    regions = [timed_ring(ring, args.steps) for _ in range(3)]
    ms_groups = float(np.median(regions))
    # ... or (--run-graph) as ONE graph of the whole K-step run (StepRing.capture_run: same kernels, lanes and order, one
    # host call).  Measured: the GPU executes the one large graph SLOWER than the group graphs launched back to back
    # (45.0 vs 40.8 us/step at K = 200), so it is an option for hosts that cannot keep up, not the default.
    run_graph = use_graph and args.run_graph and args.steps <= 2000
    if run_graph:
        ring.capture_run(args.steps)
        timed_ring(ring, args.steps)
    ms_total = timed_ring(ring, args.steps) if run_graph else ms_groups
    value = B * world * args.steps / (ms_total * 1e-3)
    # soak: the same step for ~0.25 s so that the clock record has a few hundred samples under this workload's load
    soak_steps = int(min(20000, max(200, 0.25 / max(ms_total / args.steps * 1e-3, 1e-6))))
    soak_steps -= soak_steps % (lanes * group * chain)
    ms_soak = timed_ring(ring, max(soak_steps, lanes * group * chain))
    sampler.load.clear()
    clocks = sampler.stop()
    # the same K steps with one batch in flight (latency of a step, and the figure earlier rounds reported)
    one = pipeline.StepRing(pb, make_slots(dev_sets), lanes=1, group=1, top_k=TOPK, use_graph=use_graph) if lanes * group > 1 else ring
    timed_ring(one, 3)
    ms_serial = timed_ring(one, args.steps)

    # ---- per-kernel device time over the same K steps (second pass, library event hooks on)
    def profiled(fn, steps):
        _lib.profile_enable(True)
        timed(fn, steps)  # eager launches on one stream: the library brackets every kernel with events
        _lib.profile_enable(False)
        return _lib.profile_read()

    prof = profiled(step_eager, args.steps)
    launches_per_step = sum(c for _, c in prof.values()) // max(args.steps, 1)
    alg_bytes = {'score_emit_kernel': B * bytes_score_emit, 'encode_rows_kernel': B * BYTES_ASSIGN_ENCODE}
    kernels = {k: {'ms_per_launch': ms / c, 'launches': c, 'share': ms / max(1e-9, sum(m for m, _ in prof.values()))}
               for k, (ms, c) in prof.items()}
    # the roofline line is for the dominant HBM-streaming kernel (largest device time among the two
    # that carry the algorithmic bytes); the NMS kernels move <1% of the bytes and are latency-bound
    dom = max(alg_bytes, key=lambda k: prof.get(k, (0.0, 1))[0])
    dms, dcnt = prof[dom]
    achieved = alg_bytes[dom] / (dms / dcnt * 1e-3) / 1e9
    roofline = {'bound': 'hbm', 'kernel': dom, 'achieved': achieved, 'peak': peak, 'peak_source': peak_src, 'unit': 'GB/s',
                'frac': achieved / peak, 'traffic': load_traffic(dom), 'algorithmic_bytes_per_launch': alg_bytes[dom],
                'timing': 'CUDA events around every launch of an eager pass over the same K steps; with the library\'s profiling switch on, odk_assign_encode runs its two '
                          'chains one after the other, so every kernel is timed alone (includes ~3 us of launch gap per kernel)',
                'kernels': kernels}
    for k, nbytes in alg_bytes.items():
        if k in prof:
            kernels[k]['achieved_gbs'] = nbytes / (prof[k][0] / prof[k][1] * 1e-3) / 1e9
            kernels[k]['frac'] = kernels[k]['achieved_gbs'] / peak
            kernels[k]['algorithmic_bytes'] = nbytes
            kernels[k]['traffic'] = load_traffic(k)
    if all(k in prof for k in alg_bytes):
        # encode_rows leaves part of its 116 MB dirty in the 126 MB L2 when it ends; that write-back lands on whichever
        # HBM-bound kernel runs next (here score_emit), so the two are also reported as one pair
        pair_ms = sum(prof[k][0] / prof[k][1] for k in alg_bytes)
        pair_gbs = sum(alg_bytes.values()) / (pair_ms * 1e-3) / 1e9
        roofline['streaming_pair'] = {'kernels': sorted(alg_bytes), 'algorithmic_bytes': sum(alg_bytes.values()),
                                      'ms': pair_ms, 'achieved_gbs': pair_gbs, 'frac': pair_gbs / peak}
    step_bytes = sum(alg_bytes.values())
    roofline['step'] = {'algorithmic_bytes': step_bytes, 'hbm_floor_us': step_bytes / (peak * 1e9) * 1e6,
                        'achieved_gbs': step_bytes / (ms_total / args.steps * 1e-3) / 1e9,
                        'frac': step_bytes / (ms_total / args.steps * 1e-3) / 1e9 / peak}
    roofline['emit_spans'] = spans

    extras = {}

    def leg(name, fn):
        """Optional sub-result: a failure is recorded, never fatal for the headline line."""
        t0 = time.perf_counter()
        try:
            extras[name] = fn()
        except Exception as e:  # noqa: BLE001
            extras[name] = {'error': f'{type(e).__name__}: {e}'[:300]}
        if isinstance(extras[name], dict):
            extras[name]['wall_s'] = round(time.perf_counter() - t0, 2)

    # ---- variants of the headline step: clustered head outputs (hundreds of near-duplicates around a few objects: many
    #      more select / sort / sweep rounds per image than the unclustered config-2 generator) and strict_nms
    def variants():
        out = {'headline': 'value is the config-2 primary generator (unclustered head outputs, per-class NMS, strict_nms off) as BASELINE '
                           'config 2 defines it; the variants below change only what their name says'}
        n_var = max(2, min(4, n_sets))
        clustered = [tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7 + 100 * rank + r, B, GMAX, C)
                           + synth.predictions(4321 + 100 * rank + r, B, N, C, clustered=True)) for r in range(n_var)]
        for name, sets, strict in (('clustered', clustered, None), ('strict_nms', dev_sets, 0.5), ('clustered_strict_nms', clustered, 0.5)):
            rg = pipeline.StepRing(pb, make_slots(sets), lanes=lanes, group=group, top_k=TOPK, use_graph=use_graph, nms_all_threshold=strict, chain=chain)
            timed_ring(rg, max(args.warmup, 3))
            timed_ring(rg, n_slots)
            if run_graph:
                rg.capture_run(args.steps)
                timed_ring(rg, args.steps)
            ms = timed_ring(rg, args.steps)
            pr = profiled(lambda i, s=sets, st=strict: step_eager(i, s, nms_all_threshold=st), min(args.steps, len(sets)))
            nms_key = 'nms_image_kernel_strict' if strict else 'nms_image_kernel'
            out[name] = {'value': B * world * args.steps / (ms * 1e-3), 'ms_per_step': ms / args.steps,
                         'nms_ms_per_launch': pr[nms_key][0] / pr[nms_key][1] if nms_key in pr else None,
                         'input_sets': len(sets)}
            del rg
        return out

    if not args.quick:
        leg('variants', variants)

    # ---- the other BASELINE configs
    def time_graph(fn, reps, flush=None):
        """Device us per call of `fn` replayed from a CUDA graph: `reps` back-to-back replays between one event pair, or —
        with a flush buffer — each replay timed on its own after a write larger than L2 (median)."""
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay()
        torch.cuda.synchronize()
        if flush is None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps * 1e3
        ts = []
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        return float(np.median(ts))

    def entry(us, images, nbytes, note=None):
        gbs = nbytes / (us * 1e-6) / 1e9
        e = {'us': us, 'img_s': images / (us * 1e-6), 'algorithmic_bytes': nbytes, 'gbs': gbs, 'frac': gbs / peak}
        if note:
            e['note'] = note
        return e

    def device_head(seed, batch, n, c):
        """Config-2 head outputs generated on the device (same distributions as synth.predictions): for the shapes whose
        host generation would take longer than the whole bench."""
        gen = torch.Generator(device=dev)
        gen.manual_seed(seed)
        objs = torch.sigmoid(torch.randn((batch, n, 1), generator=gen, device=dev) * 2.5 - 5.0)
        clfs = torch.softmax(3.0 * torch.randn((batch, n, c), generator=gen, device=dev), dim=-1)
        locs = torch.randn((batch, n, 4), generator=gen, device=dev)
        return objs, clfs, locs

    def cfg1():
        od1 = odk.ObjectDetector.load_voc((320, 320))
        gt = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(0, 1, 7, C, crowded_every=0))
        head = tuple(torch.from_numpy(x).to(dev) for x in synth.predictions(1234, 1, N, C))
        y, d = torch.empty((1, N, 27), device=dev), torch.empty((1, TOPK, 6), device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

        g = pipeline.capture_step(od1.pb, gt, head, y, d, top_k=TOPK)
        ts = []
        for _ in range(20):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        g_count = int(gt[2][0])
        nbytes = 20 * g_count + 4 * N * 27 + 4 * N * 25 + 4800
        return dict(entry(float(np.median(ts)), 1, nbytes, 'latency-bound: one image, both sides on parallel graph branches, L2 flushed before every replay'),
                    workload='cfg1: 320x320 VOC anchors, batch 1, assign+encode + decode + per-class NMS', gt=g_count)

    def cfg3():
        b3 = 256
        gt = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(7, b3, GMAX, C))
        y = torch.empty((b3, N, 27), dtype=torch.float32, device=dev)
        us = time_graph(lambda: pb.encode_truth_batch(*gt, y_true=y), 10)
        st = pb.encode_truth_batch(*gt, y_true=y)['status']
        _lib.profile_enable(True)
        for _ in range(3):
            pb.encode_truth_batch(*gt, y_true=y)
        torch.cuda.synchronize()
        _lib.profile_enable(False)
        per_kernel = {k: round(ms / c * 1e3, 1) for k, (ms, c) in _lib.profile_read().items()}
        return dict(entry(us, b3, b3 * BYTES_ASSIGN_ENCODE), kernel_us_eager=per_kernel, workload='cfg3: assign+encode 640x640, batch 256, <=64 GT/img (every 4th image crowded)',
                    status_ok=bool((st == 0).all().item()))

    def cfg5():
        n5, c5, b5, g5 = 205800, 80, 8, 512
        pb5 = odk.PriorBoxes((1120, 1120), [140, 70, 35], c5)
        pb5.from_dict({'map_sizes': [140, 70, 35], 'pb_size_patterns': odk.od._VOC_JSON_DATA['pb_size_patterns']})
        assert len(pb5) == n5
        gt = tuple(torch.from_numpy(x).to(dev) for x in synth.gt_batch(5, b5, g5, c5, fixed_g=g5, crowded_every=0, centre=(0.05, 0.95), size=(0.01, 0.3)))
        head = device_head(55, b5, n5, c5)
        y = torch.empty((b5, n5, 3 + c5 + 4), dtype=torch.float32, device=dev)
        d = torch.empty((b5, TOPK, 6), dtype=torch.float32, device=dev)
        us_t = time_graph(lambda: pb5.encode_truth_batch(*gt, y_true=y), 5)
        us_p = time_graph(lambda: pb5.decode_nms(*head, top_k=TOPK, out=d), 5)
        st = pb5.encode_truth_batch(*gt, y_true=y)['status']
        kept = int((d[:, :, 1] > 0).sum().item())
        w5 = 3 + c5 + 4
        _lib.profile_enable(True)
        for _ in range(3):
            pb5.encode_truth_batch(*gt, y_true=y)
            pb5.decode_nms(*head, top_k=TOPK, out=d)
        torch.cuda.synchronize()
        _lib.profile_enable(False)
        per_kernel = {k: round(ms / c * 1e3, 1) for k, (ms, c) in _lib.profile_read().items()}
        return {'workload': 'cfg5: dense-anchor stress, N=205800 anchors x 512 GT, C=80, batch 8, NMS IoU 0.45, top-k 200 (head outputs generated on the device)',
                'assign_encode': dict(entry(us_t, b5, b5 * (20 * g5 + 4 * n5 * w5)), status_ok=bool((st == 0).all().item())),
                'decode_nms': dict(entry(us_p, b5, b5 * (4 * n5 * (1 + c5 + 4) + 4800)), kept=kept),
                'kernel_us_eager': per_kernel}

    def configs():
        out = {}
        for name, fn in (('cfg1', cfg1), ('cfg3', cfg3), ('cfg5', cfg5)):
            try:
                out[name] = fn()
            except Exception as e:  # noqa: BLE001
                out[name] = {'error': f'{type(e).__name__}: {e}'[:300]}
            torch.cuda.empty_cache()
        return out

    if not args.quick:
        leg('configs', configs)

    # ---- config 4: VOC07-test-sized eval sharded over the ranks: decode+NMS -> NCCL all-gather -> VOC AP
    def cfg4():
        from object_detector_b200 import dist as odist, voc
        n_img = args.eval_images
        gts, preds = synth.eval_set(2007, n_img, C)
        y_true = [odk.ml.ObjectsAnnotation('.', w, h, cl, b, d) for (w, h, cl, b, d) in gts]
        plan = synth.plan_eval_head(pb, preds)                      # host: which anchor carries which planted detection
        lo, hi, per = odist.shard_range(n_img, rank, world)

        def head_for(a, b):                                         # resident head outputs of images [a, b), rank-independent
            return synth.eval_head_outputs(pb, plan, a, b, dev, seed=2007)
        objs, clfs, locs = head_for(lo, hi)
        buf = odist.new_gather_buffer(n_img, world, TOPK, dev)
        mine = odist.local_slab(buf, rank, world)[:hi - lo]
        gt = voc.pack_ground_truth(y_true, dev)
        batch = 256
        odist.decode_nms_shard(pb, objs, clfs, locs, mine, batch=batch, top_k=TOPK)  # warm-up (workspaces, attributes)
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        odist.decode_nms_shard(pb, objs, clfs, locs, mine, batch=batch, top_k=TOPK)
        ev[1].record()
        odist.all_gather_detections(buf, rank, world)               # in place: the slab the kernel wrote is the send buffer
        ev[2].record()
        ap, ap07 = voc.evaluate_detections(buf[:n_img], gt, C, 0.01)
        ev[3].record()
        barrier()
        ms = [max_over_ranks(ev[i].elapsed_time(ev[i + 1])) for i in range(3)]
        ap, ap07 = ap.cpu().numpy(), ap07.cpu().numpy()
        res = {'workload': f'cfg4: {n_img} synthetic VOC07-test-sized images, {world} GPU(s): decode+NMS on the resident shard ({hi - lo} images on rank 0, '
                           f'batches of {batch}) -> in-place NCCL all-gather of the (n,200,6) slabs -> VOC AP on every rank',
               'images': n_img, 'images_per_rank': per, 'ms_decode_nms': ms[0], 'ms_allgather': ms[1], 'ms_ap': ms[2], 'ms_total': sum(ms),
               'img_s': n_img / (sum(ms) * 1e-3), 'allgather_bytes': int(buf.numel() * 4), 'allgather_bytes_sent_per_rank': int(mine.numel() * 4) if world > 1 else 0,
               'decode_nms_gbs_per_gpu': (hi - lo) * BYTES_DECODE_NMS / (ms[0] * 1e-3) / 1e9,
               'mAP': float(np.nanmean(ap)), 'mAP_VOC': float(np.nanmean(ap07)), 'timing': 'CUDA events on the launching stream, max over ranks per phase'}
        if world > 1:
            res['allgather_busbw_gbs'] = buf.numel() * 4 * (world - 1) / world / (ms[1] * 1e-3) / 1e9
        del objs, clfs, locs
        # single-rank check on rank 0: the whole set through one GPU, chunk by chunk; detections and AP must be identical
        same = True
        if world > 1:
            if rank == 0:
                full = odist.new_gather_buffer(n_img, 1, TOPK, dev)
                for a in range(0, n_img, 512):
                    e = min(n_img, a + 512)
                    o, c_, l_ = head_for(a, e)
                    odist.decode_nms_shard(pb, o, c_, l_, full[a:e], batch=batch, top_k=TOPK)
                ap1, ap071 = (t.cpu().numpy() for t in voc.evaluate_detections(full[:n_img], gt, C, 0.01))
                same = bool(torch.equal(full[:n_img], buf[:n_img])) and np.array_equal(ap1, ap, equal_nan=True) and np.array_equal(ap071, ap07, equal_nan=True)
            flag = torch.tensor([1.0 if same else 0.0], device=dev)
            dist.broadcast(flag, 0)
            same = bool(flag.item() == 1.0)
        res['matches_single_rank'] = same
        return res

    if not args.quick and (world > 1 or args.eval_at_one):
        leg('cfg4', cfg4)
        torch.cuda.empty_cache()
    if 'configs' in extras and 'cfg4' in extras and isinstance(extras['configs'], dict):
        extras['configs']['cfg4'] = extras.pop('cfg4')

    # ---- end to end through host buffers: pinned inputs -> H2D -> kernels -> D2H of the results, every step,
    #      through the public HostPipeline (copy-in / compute / copy-out streams, double buffered)
    pin_in = [] if args.no_e2e else [tuple(torch.from_numpy(x).pin_memory() for x in s) for s in host_sets]
    e2e_steps = max(3, min(args.steps, 50))

    def e2e_run(pipe, n_inputs, steps):
        sink = [0.0]

        def one_step(i):
            k = pipe.submit(*pin_in[i % n_sets][:n_inputs])
            if i >= 1:  # consume the previous batch's host result while this one is in flight
                _, d = pipe.result(k ^ 1)
                sink[0] += float(d[0, 0, 1])

        def run(n):
            barrier()
            cur = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(n):
                one_step(i)
            for st in (pipe.s_in, pipe.s_run, pipe.s_out):
                cur.wait_stream(st)
            e1.record()
            barrier()
            return max_over_ranks(e0.elapsed_time(e1))
        run(4)
        ms = run(steps)
        return B * world * steps / (ms * 1e-3)

    if args.no_e2e:
        e2e = None
    pipe = None if args.no_e2e else pipeline.HostPipeline(pb, B, GMAX, top_k=TOPK, use_graph=use_graph)
    h2d, d2h = (0, 0) if args.no_e2e else (pipe.h2d_bytes, pipe.d2h_bytes)
    e2e_value = 0.0 if args.no_e2e else e2e_run(pipe, 6, e2e_steps)
    e2e = None if args.no_e2e else {'value': e2e_value, 'unit': 'images/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h, 'steps': e2e_steps,
           'contract': 'dense: the reference\'s own result arrays — y_true (B,N,27) f32 of encode_truth and the (B,200,6) detections — to pinned host memory',
           'h2d_gbs': h2d * e2e_value / (B * world) / 1e9, 'd2h_gbs': d2h * e2e_value / (B * world) / 1e9}
    del pipe

    def e2e_ceiling():
        p0 = pipeline.HostPipeline(pb, B, GMAX, top_k=TOPK, use_graph=False, kernels=False)
        v = e2e_run(p0, 6, e2e_steps)
        return v
    try:
        if args.no_e2e:
            raise RuntimeError('skipped')
        ceiling = e2e_ceiling()
        e2e['copy_only_img_s'] = ceiling
        e2e['of_copy_only'] = e2e_value / ceiling
        e2e['limiter'] = ('the host<->device link: the same ' + f'{(h2d + d2h) >> 20} MB per step through the same three streams with no kernels reach '
                          f'{ceiling:.0f} img/s ({d2h * ceiling / (B * world) / 1e9:.1f} GB/s D2H + {h2d * ceiling / (B * world) / 1e9:.1f} GB/s H2D per GPU)')
    except Exception as e:  # noqa: BLE001
        if e2e is not None:
            e2e['copy_only_error'] = str(e)[:200]
    torch.cuda.empty_cache()

    def e2e_other(mode):
        resident = None
        if mode == 'fused':
            gen = torch.Generator(device=dev)
            gen.manual_seed(99)
            y_pred = torch.rand((B, N, 3 + C + 4), generator=gen, device=dev) * 0.98 + 0.01  # the network's raw output, resident
            resident = {'y_pred': y_pred, 'head': dev_sets[0][3:]}
        p1 = pipeline.HostPipeline(pb, B, GMAX, top_k=TOPK, use_graph=use_graph, mode=mode, resident=resident)
        n_in = 3 if mode == 'fused' else 6
        v = e2e_run(p1, n_in, e2e_steps)
        out = {'value': v, 'unit': 'images/s', 'h2d_bytes_per_step': p1.h2d_bytes, 'd2h_bytes_per_step': p1.d2h_bytes, 'steps': e2e_steps}
        if mode == 'compact':
            p0 = pipeline.HostPipeline(pb, B, GMAX, top_k=TOPK, use_graph=False, mode=mode, kernels=False)
            out['copy_only_img_s'] = e2e_run(p0, n_in, e2e_steps)
            out['contract'] = ('same host inputs; the train-side result leaves the device as pb_valids bits + the <= G positive rows + indices '
                               '(odk_assign_encode_compact; od_pb.expand_truth rebuilds the dense y_true on the host), detections as in the dense contract')
        else:
            out['contract'] = ('the network lives on the same GPU (y_pred and the head outputs are resident): only the ground truth goes in, the consumer of '
                               'the assignment is odk_assign_loss_grad (y_true never exists), and only the (B,8) loss terms + the (B,200,6) detections come out')
        return out
    if not args.quick and not args.no_e2e:
        leg('e2e_compact', lambda: e2e_other('compact'))
        leg('e2e_fused', lambda: e2e_other('fused'))

    # ---- CPU baseline (oracle port, 1 core, bounded sample) on rank 0 at N=1
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_img = 24
        v = cpu_images_per_sec(n_img, 1)
        cpu = {'value': v, 'unit': 'images/s', 'cores': 1, 'kind': 'port',
               'sample': f'{n_img} images of the batch-64 workload (encode_truth G<=64 + decode + per-class NumPy NMS), single process; '
                         'the --impl reference arm times the unmodified reference modules on all cores'}

    if rank == 0:
        cfg = static_config(world)
        line = {
            'metric': METRIC, 'value': value, 'unit': 'images/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
            'ms_per_step': ms_total / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f32', 'data': 'synthetic', 'config': cfg,
            'schedule': {'launch': ('eager' if not use_graph else 'CUDA graphs, train and predict sides of a batch on parallel branches')
                         + f'; {group} consecutive batches per group (a branch each), {chain} consecutive group(s) per graph, graphs alternating on {lanes} streams'
                         + f' = up to {lanes * group} batches in flight (each slot has its own y_true, detections, workspaces)'
                         + ('; the K timed steps are one graph of the whole run (one host call)' if run_graph else f'; one graph launch per {group * chain} batches'),
                         'batches_in_flight': lanes * group, 'batches_per_graph': group * chain, 'groups_chained_per_graph': chain, 'lanes': lanes, 'slots': n_slots, 'input_sets': n_sets,
                         'host_calls_in_timed_region': 1 if run_graph else host_calls(args.steps),
                         'ms_per_step_group_launches': ms_groups / args.steps, 'timed_regions_ms': regions,
                         'headline_region': 'median of three timed regions of exactly K steps each',
                         'ms_per_step_one_in_flight': ms_serial / args.steps, 'ms_per_step_soak': ms_soak / max(soak_steps, lanes * group * chain),
                         'soak_steps': max(soak_steps, lanes * group * chain), 'affinity': affinity},
            'clocks': clocks, 'e2e': e2e,
            'gpu_launches': launches_per_step * args.steps, 'gpu_launches_per_step': launches_per_step,
            'roofline': roofline, 'cpu_baseline': cpu,
        }
        line.update(extras)
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
    ap.add_argument('--run-graph', action='store_true', help='time ONE graph of the whole K-step run (StepRing.capture_run) instead of K/group graph launches '
                    '(measured slower: 45.0 vs 40.8 us/step at K = 200, profiles/sweep_r02p.log)')
    ap.add_argument('--affinity', action='store_true', help='pin the rank to its own share of the cores (off by default: measured on 8 ranks, the launches '
                    'of a region take 188 us from a 4-core share and 113 us unpinned, profiles/ring_probe_r02x_n8*.log)')
    ap.add_argument('--no-affinity', action='store_true', help='(default) do not pin the rank to a core set')
    ap.add_argument('--clock-period-ms', type=float, default=2.0, help='NVML sampling period of the in-run clock record')
    ap.add_argument('--no-e2e', action='store_true', help='skip the host-buffer legs (knob sweeps)')
    ap.add_argument('--quick', action='store_true', help='headline + roofline + dense e2e only (no variants / configs / other e2e contracts)')
    ap.add_argument('--eval-at-one', action='store_true', help='run the config-4 leg on a single GPU too (no collective)')
    ap.add_argument('--eval-images', type=int, default=EVAL_IMAGES)
    ap.add_argument('--sets', type=int, default=SETS, help='distinct resident input sets rotated through')
    ap.add_argument('--slots', type=int, default=SLOTS, help='resident batches (outputs + workspaces) in the ring')
    ap.add_argument('--lanes', type=int, default=LANES, help='streams the step graphs alternate on')
    ap.add_argument('--chain', type=int, default=CHAIN, help='consecutive groups chained inside one CUDA graph (one host launch per group*chain batches)')
    ap.add_argument('--group', type=int, default=GROUP, help='consecutive batches captured into one CUDA graph (a branch each)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == '__main__':
    main()
