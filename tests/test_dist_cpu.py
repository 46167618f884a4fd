"""World-size-2 gloo test of the evaluation-path sharding / all-gather plumbing (no GPU)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from object_detector_b200 import dist as odist


def test_shard_ranges_cover_everything():
    for n in (0, 1, 7, 64, 4952):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                lo, hi, per = odist.shard_range(n, r, world)
                assert hi - lo <= per and per == (n + world - 1) // world
                seen.extend(range(lo, hi))
            assert seen == list(range(n))
    assert odist.shard_range(4952, 7, 8) == (4333, 4952, 619)


def _worker(rank, world, port, n_items, top_k, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        buf = odist.new_gather_buffer(n_items, world, top_k, 'cpu')
        lo, hi, per = odist.shard_range(n_items, rank, world)
        mine = odist.local_slab(buf, rank, world)
        for i in range(lo, hi):  # stand-in for the NMS kernel writing image i's slab
            mine[i - lo] = float(i)
        odist.all_gather_detections(buf, rank, world)
        ok = all(bool((buf[i] == float(i)).all()) for i in range(n_items))
        pad = buf[n_items:]
        ok = ok and bool((pad[:, :, 0] == -1).all()) and bool((pad[:, :, 1:] == 0).all())
        out[rank] = ok
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.timeout(120)
def test_all_gather_world2_gloo():
    world, n_items = 2, 7  # odd count: the last rank's shard has one padding slab
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), n_items, 5, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}
