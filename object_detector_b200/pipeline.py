"""Stream / graph plumbing around the two halves of the hot path (no arithmetic here).

The train side (``odk_assign_encode``) and the predict side (``odk_decode_nms``) of one batch are
independent, their latency-bound kernels (greedy walk, per-image NMS) occupy a fraction of the SMs, and
all launches are tiny — so a step is captured once into a CUDA graph with the two sides on parallel
branches and replayed (``capture_step``).  ``HostPipeline`` adds what a caller with HOST buffers needs:
pinned staging, separate copy-in / compute / copy-out streams and double buffering, so that the
host->device copy of batch i+1, the kernels of batch i and the device->host copy of batch i-1 overlap
(PCIe is full duplex).  The reference has no counterpart: its generator threads call ``encode_truth`` per
sample on the CPU (pytoolkit/generator.py:160-166) and ``predict`` runs the whole graph per batch.
"""
import torch

from . import _dev


def _enqueue_step(pb, gt, head, y_true, dets, side, kw):
    """Both sides of one batch on the current stream; the predict side on ``side`` (forked / joined) if given."""
    main = torch.cuda.current_stream(y_true.device)
    if side is not None:
        side.wait_stream(main)
        with torch.cuda.stream(side):
            pb.decode_nms(*head, out=dets, **kw)
        pb.encode_truth_batch(*gt, y_true=y_true)
        main.wait_stream(side)
    else:
        pb.encode_truth_batch(*gt, y_true=y_true)
        pb.decode_nms(*head, out=dets, **kw)


def capture_steps(pb, slots, top_k=200, conf_threshold=0.01, nms_threshold=0.45, two_streams=True):
    """Captures the steps of all ``slots`` = [(gt, head, y_true, dets)] into ONE CUDA graph, every slot on its own
    branch (and within a slot the train and the predict side on parallel branches when ``two_streams``): one host
    launch per ``len(slots)`` batches.  All tensors must be CUDA tensors that stay alive and keep their addresses."""
    dev = slots[0][2].device
    kw = dict(top_k=top_k, conf_threshold=conf_threshold, nms_threshold=nms_threshold)
    with torch.cuda.device(dev):
        pb._handle(dev)  # the anchor tables are uploaded (cudaMalloc) on first use: not allowed while capturing
        branch = [torch.cuda.Stream(device=dev) for _ in slots[1:]]
        side = [torch.cuda.Stream(device=dev) if two_streams else None for _ in slots]
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.device(dev), torch.cuda.graph(graph):
        main = torch.cuda.current_stream(dev)
        for k, (gt, head, y, d) in enumerate(slots):
            if k == 0:
                continue
            branch[k - 1].wait_stream(main)
            with torch.cuda.stream(branch[k - 1]):
                _enqueue_step(pb, gt, head, y, d, side[k], kw)
        _enqueue_step(pb, *slots[0], side[0], kw)
        for st in branch:
            main.wait_stream(st)
    return graph


def capture_step(pb, gt, head, y_true, dets, top_k=200, conf_threshold=0.01, nms_threshold=0.45, two_streams=True):
    """Captures ``encode_truth_batch(*gt, y_true=y_true)`` and ``decode_nms(*head, out=dets)`` into one CUDA
    graph (parallel branches when ``two_streams``).  All tensors must be CUDA tensors that stay alive and
    keep their addresses; replay with ``graph.replay()`` on any stream."""
    return capture_steps(pb, [(gt, head, y_true, dets)], top_k, conf_threshold, nms_threshold, two_streams)


class StepRing:
    """Device-resident batches, several in flight.

    The six kernels of a step leave most of the GPU idle most of the time (two of them are HBM-bound for ~20 us each,
    the other four are latency-bound chains on 64-128 SMs), and consecutive batches are independent — so consecutive
    steps run concurrently: the tail of batch i (per-image NMS, greedy walk, positives) beside the head of batch i+1
    (score/emit, candidates).  Slots (= inputs + outputs + workspaces of one batch) are tied into groups of ``group``
    consecutive slots; a group is ONE CUDA graph with a branch per slot (one host launch per ``group`` batches — with
    8 ranks on one host the launch rate of single-step graphs is what limits), and groups are replayed round-robin on
    ``lanes`` streams.  A slot is only ever touched by work queued on its group's lane, so reuse is ordered by that
    stream.

    ``slots``: list of ``(gt, head, y_true, dets)`` with ``gt = (boxes, classes, count)``, ``head = (objs, clfs, locs)``
    CUDA tensors that keep their addresses; ``len(slots)`` must be a multiple of ``lanes * group``.
    """

    def __init__(self, pb, slots, lanes=2, group=4, top_k=200, conf_threshold=0.01, nms_threshold=0.45, use_graph=True):
        assert len(slots) >= 1 and lanes >= 1 and group >= 1 and len(slots) % (lanes * group) == 0, \
            'len(slots) must be a multiple of lanes * group'
        self.pb, self.slots, self.lanes, self.group = pb, slots, lanes, group
        self.dev = slots[0][2].device
        self._kw = dict(top_k=top_k, conf_threshold=conf_threshold, nms_threshold=nms_threshold)
        self.n_groups = len(slots) // group
        with torch.cuda.device(self.dev):
            self.streams = [torch.cuda.Stream(device=self.dev) for _ in range(lanes)]
            self._ev = torch.cuda.Event()
        self.use_graph = use_graph
        self.group_graphs = ([capture_steps(pb, slots[j * group:(j + 1) * group], **self._kw) for j in range(self.n_groups)]
                             if use_graph else None)
        # single-step graphs for steps outside whole groups; captured up front so that no capture (milliseconds) ever
        # happens between a caller's timing events
        self._slot_graphs = ({k: capture_step(pb, *slots[k], **self._kw) for k in range(len(slots))}
                             if (use_graph and group > 1) else {})

    def fork(self):
        """Orders every lane after the work already queued on the current stream."""
        cur = torch.cuda.current_stream(self.dev)
        self._ev.record(cur)
        for st in self.streams:
            st.wait_event(self._ev)

    def _lane_of_slot(self, k):
        return self.streams[(k // self.group) % self.lanes]

    def submit(self, i):
        """Queues the single step ``i`` (slot ``i % len(slots)``) on its lane; returns the slot index."""
        k = i % len(self.slots)
        gt, head, y, d = self.slots[k]
        with torch.cuda.device(self.dev), torch.cuda.stream(self._lane_of_slot(k)):
            if self.use_graph:
                (self.group_graphs[k] if self.group == 1 else self._slot_graphs[k]).replay()
            else:
                _enqueue_step(self.pb, gt, head, y, d, None, self._kw)
        return k

    def run(self, steps, start=0):
        """Queues steps ``start .. start+steps-1``: whole groups as one graph launch each, the rest step by step."""
        i, end = start, start + steps
        while i < end:
            if self.use_graph and i % self.group == 0 and i + self.group <= end:
                j = (i // self.group) % self.n_groups
                with torch.cuda.device(self.dev), torch.cuda.stream(self.streams[j % self.lanes]):
                    self.group_graphs[j].replay()
                i += self.group
            else:
                self.submit(i)
                i += 1

    def join(self):
        """Orders the current stream after everything queued on the lanes."""
        cur = torch.cuda.current_stream(self.dev)
        for st in self.streams:
            cur.wait_stream(st)


class HostPipeline:
    """Host-buffer front end for fixed batch shapes.

    ``submit(gt_boxes, gt_classes, gt_count, objs, clfs, locs)`` takes pinned host tensors (or NumPy
    arrays, which are staged through an internal pinned buffer), returns a slot id immediately;
    ``result(slot)`` waits for that batch only and returns pinned host views ``(y_true, dets)`` that stay
    valid until the slot is submitted again (``depth`` batches later).
    """

    def __init__(self, pb, batch, gmax, top_k=200, conf_threshold=0.01, nms_threshold=0.45, depth=2, device=None,
                 use_graph=True):
        self.pb, self.batch, self.gmax, self.top_k, self.depth = pb, batch, gmax, top_k, depth
        self.dev = _dev.device(device)
        n, c = len(pb.pb_locs), pb.num_classes
        f32, i32 = torch.float32, torch.int32
        shapes = [((batch, gmax, 4), f32), ((batch, gmax), i32), ((batch,), i32),
                  ((batch, n, 1), f32), ((batch, n, c), f32), ((batch, n, 4), f32)]
        with torch.cuda.device(self.dev):
            self.s_in, self.s_run, self.s_out = (torch.cuda.Stream(device=self.dev) for _ in range(3))
            self.d_in = [[torch.empty(s, dtype=t, device=self.dev) for s, t in shapes] for _ in range(depth)]
            self.h_in = [[torch.empty(s, dtype=t).pin_memory() for s, t in shapes] for _ in range(depth)]
            self.d_y = [torch.empty((batch, n, 3 + c + 4), dtype=f32, device=self.dev) for _ in range(depth)]
            self.d_det = [torch.empty((batch, top_k, 6), dtype=f32, device=self.dev) for _ in range(depth)]
            self.h_y = [torch.empty((batch, n, 3 + c + 4), dtype=f32).pin_memory() for _ in range(depth)]
            self.h_det = [torch.empty((batch, top_k, 6), dtype=f32).pin_memory() for _ in range(depth)]
            self.ev_in = [torch.cuda.Event() for _ in range(depth)]
            self.ev_run = [torch.cuda.Event() for _ in range(depth)]
            self.ev_out = [torch.cuda.Event() for _ in range(depth)]
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.d_in[0])
        self.d2h_bytes = self.h_y[0].numel() * 4 + self.h_det[0].numel() * 4
        self._kw = dict(top_k=top_k, conf_threshold=conf_threshold, nms_threshold=nms_threshold)
        self._count = 0
        self.graphs = None
        if use_graph:
            with torch.cuda.device(self.dev):
                for t in self.d_in[0]:
                    t.zero_()
                self.d_in[0][2].fill_(0)  # gt_count = 0: a valid empty batch for the eager warm-up call
                self._eager(0)            # uploads the anchor tables, sets kernel attributes
                torch.cuda.synchronize(self.dev)
                self.graphs = [capture_step(pb, self.d_in[k][:3], self.d_in[k][3:], self.d_y[k], self.d_det[k], **self._kw)
                               for k in range(depth)]

    def _eager(self, k):
        self.pb.encode_truth_batch(*self.d_in[k][:3], y_true=self.d_y[k])
        self.pb.decode_nms(*self.d_in[k][3:], out=self.d_det[k], **self._kw)

    def submit(self, *host_inputs):
        k = self._count % self.depth
        self._count += 1
        src = []
        for i, x in enumerate(host_inputs):
            if isinstance(x, torch.Tensor) and x.is_pinned():
                src.append(x)
            else:  # pageable memory: stage through the slot's pinned buffer (safe: the slot's last H2D has completed)
                self.ev_in[k].synchronize()
                self.h_in[k][i].copy_(torch.as_tensor(x).reshape(self.h_in[k][i].shape))
                src.append(self.h_in[k][i])
        with torch.cuda.device(self.dev):
            self.s_in.wait_event(self.ev_run[k])   # the kernels that last read this slot's inputs are done
            with torch.cuda.stream(self.s_in):
                for d, h in zip(self.d_in[k], src):
                    d.copy_(h.reshape(d.shape), non_blocking=True)
                self.ev_in[k].record(self.s_in)
            self.s_run.wait_event(self.ev_in[k])
            self.s_run.wait_event(self.ev_out[k])  # the slot's previous outputs have left the device
            with torch.cuda.stream(self.s_run):
                if self.graphs is not None:
                    self.graphs[k].replay()
                else:
                    self._eager(k)
                self.ev_run[k].record(self.s_run)
            self.s_out.wait_event(self.ev_run[k])
            with torch.cuda.stream(self.s_out):
                self.h_y[k].copy_(self.d_y[k], non_blocking=True)
                self.h_det[k].copy_(self.d_det[k], non_blocking=True)
                self.ev_out[k].record(self.s_out)
        return k

    def result(self, slot):
        self.ev_out[slot].synchronize()
        return self.h_y[slot], self.h_det[slot]

    def synchronize(self):
        for s in (self.s_in, self.s_run, self.s_out):
            s.synchronize()
