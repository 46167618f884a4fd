"""K-step timed regions of pipeline.StepRing shapes: device time (CUDA events, max over ranks) beside the HOST time the rank
spends issuing the launches of the region (max over ranks) — to tell a GPU-bound window from a launch-bound one.

    python scripts/gpu_ring_probe.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/gpu_ring_probe.py                                      # 8 ranks sharing one host

Every shape is "lanes,group,chain,slots[,nb]" (see StepRing; nb = chain_barrier=False); "+run" appended = also ONE graph of the whole K-step run.
Rank 0 prints one JSON line per shape.  The timed region is bench.py's (barrier + synchronize on both sides, events on
the current stream, fork / run / join), with the same inputs (8 resident sets) and per-rank core pinning.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench  # noqa: E402  (constants, pin_affinity)

DEFAULT_SHAPES = '2,4,1,8+run;2,4,2,16;2,5,1,10;2,5,2,20;2,10,1,20;2,2,1,8;4,4,1,16;1,4,1,4;1,8,1,8'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--regions', type=int, default=7)
    ap.add_argument('--shapes', default=DEFAULT_SHAPES)
    ap.add_argument('--no-affinity', action='store_true')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    world_local = int(os.environ.get('LOCAL_WORLD_SIZE', str(world)))
    affinity = bench.pin_affinity(local, world_local, enable=not args.no_affinity)

    import torch
    import torch.distributed as dist
    import object_detector_b200 as odk
    from object_detector_b200 import pipeline, synth

    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(xs):
        t = torch.tensor([float(x) for x in xs], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    B, N, C, GMAX, TOPK = bench.B, bench.N, bench.C, bench.GMAX, bench.TOPK
    pb = odk.ObjectDetector.load_voc((640, 640)).pb
    dev_sets = []
    for r in range(bench.SETS):
        gt = synth.gt_batch(7 + 100 * rank + r, B, GMAX, C)
        head = synth.predictions(1234 + 100 * rank + r, B, N, C)
        dev_sets.append(tuple(torch.from_numpy(x).to(dev) for x in gt + head))

    shapes = []
    for tok in args.shapes.split(';'):
        tok = tok.strip()
        if not tok:
            continue
        with_run = tok.endswith('+run')
        tok = tok.replace('+run', '')
        barrier_in_chain = not tok.endswith(',nb')  # ",nb" = no barrier between the groups of a chain
        lanes, group, chain, slots = (int(v) for v in tok.replace(',nb', '').split(','))
        assert slots % (lanes * group * chain) == 0, tok
        shapes.append((lanes, group, chain, slots, with_run, barrier_in_chain))
    max_slots = max(s[3] for s in shapes)
    y_trues = [torch.empty((B, N, 3 + C + 4), dtype=torch.float32, device=dev) for _ in range(max_slots)]
    detss = [torch.empty((B, TOPK, 6), dtype=torch.float32, device=dev) for _ in range(max_slots)]
    for i in range(3):  # anchor upload, kernel attributes
        s = dev_sets[i]
        pb.encode_truth_batch(*s[:3], y_true=y_trues[0])
        pb.decode_nms(*s[3:], top_k=TOPK, out=detss[0])
    torch.cuda.synchronize()

    def region(ring, steps):
        """-> (device ms, host ms spent issuing) of one timed region, each the max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        ring.fork()
        ring.run(steps)
        ring.join()
        e1.record()
        t1 = time.perf_counter()
        barrier()
        return max_over_ranks([e0.elapsed_time(e1), (t1 - t0) * 1e3])

    def measure(ring, steps, regions):
        for _ in range(2):
            region(ring, steps)
        rs = [region(ring, steps) for _ in range(regions)]
        devs, hosts = [r[0] for r in rs], [r[1] for r in rs]
        return {'us_per_step_median': float(np.median(devs)) / steps * 1e3, 'us_per_step_min': min(devs) / steps * 1e3,
                'us_per_step_max': max(devs) / steps * 1e3, 'host_issue_us_median': float(np.median(hosts)) * 1e3,
                'region_us_median': float(np.median(devs)) * 1e3}

    if rank == 0:
        print(json.dumps({'world': world, 'steps': args.steps, 'regions': args.regions, 'affinity': affinity,
                          'max_connections': os.environ.get('CUDA_DEVICE_MAX_CONNECTIONS')}), flush=True)
    for lanes, group, chain, slots, with_run, chain_barrier in shapes:
        sl = [(dev_sets[k % len(dev_sets)][:3], dev_sets[k % len(dev_sets)][3:], y_trues[k], detss[k]) for k in range(slots)]
        out = {'lanes': lanes, 'group': group, 'chain': chain, 'slots': slots, 'chain_barrier': chain_barrier}
        try:
            ring = pipeline.StepRing(pb, sl, lanes=lanes, group=group, top_k=TOPK, chain=chain, chain_barrier=chain_barrier)
            span, i, calls = group * chain, 0, 0
            while i < args.steps:  # the launches StepRing.run makes for this region
                i += span if (chain > 1 and i % span == 0 and i + span <= args.steps) else (
                    group if (i % group == 0 and i + group <= args.steps) else 1)
                calls += 1
            out['host_calls'] = calls
            out['k'] = measure(ring, args.steps, args.regions)
            long_steps = args.steps * 10
            out['k10'] = measure(ring, long_steps, 3)
            if with_run:
                ring.capture_run(args.steps)
                out['k_run_graph'] = measure(ring, args.steps, args.regions)
            del ring
        except Exception as e:  # noqa: BLE001
            out['error'] = f'{type(e).__name__}: {e}'[:300]
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        if rank == 0:
            print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
