"""The batched boundary in front of ``odk_assign_encode``.

The reference's data pipeline calls ``encode_truth`` once per SAMPLE, from up to ``min(batch, 3*cores)``
joblib threads (pytoolkit/generator.py:158-166, hook installed at dl/od.py:169 as
``ProcessOutput(lambda y: pb.encode_truth([y])[0])``).  One GPU launch per sample is latency-bound, so this
module offers two drop-ins that keep the per-sample call shape but feed the kernel whole batches:

* ``BatchCoalescer`` — concurrent per-sample calls are queued, a worker thread gathers what arrived within a
  short window (or ``max_batch`` items) and runs ONE batched call; each caller gets its own row back.
* ``encode_truth_hook(od)`` — the function to hand to ``generator.ProcessOutput`` in place of the reference's
  lambda; ``collate_truth`` is the same for pipelines that already hold the whole batch of annotations.
"""
import queue
import threading
import time


class _Failed:
    """Per-item failure marker a batched function may return in place of a result: only that caller raises."""

    def __init__(self, error):
        self.error = error


class BatchCoalescer:
    """Turns a batched function ``fn_batch(list_of_items) -> sequence_of_results`` into a thread-safe
    per-item callable.  Items submitted by different threads within ``max_wait_ms`` of the first one (at most
    ``max_batch``) are served by one ``fn_batch`` call.  An exception raised by ``fn_batch`` is re-raised in
    every caller of that batch; a ``_Failed`` entry in its result fails only the caller of that item."""

    _STOP = object()

    def __init__(self, fn_batch, max_batch=64, max_wait_ms=2.0):
        self.fn_batch, self.max_batch, self.max_wait = fn_batch, int(max_batch), max_wait_ms / 1e3
        self.batches = 0  # how many batched calls were made (for tests / diagnostics)
        self.items = 0
        self._q = queue.Queue()
        self._worker = threading.Thread(target=self._run, name='odk-coalescer', daemon=True)
        self._worker.start()

    def __call__(self, item):
        slot = {'done': threading.Event(), 'item': item, 'result': None, 'error': None}
        self._q.put(slot)
        slot['done'].wait()
        if slot['error'] is not None:
            raise slot['error']
        return slot['result']

    def close(self):
        self._q.put(self._STOP)
        self._worker.join(timeout=5)

    def _run(self):
        while True:
            first = self._q.get()
            if first is self._STOP:
                return
            batch, stop = [first], False
            deadline = time.monotonic() + self.max_wait
            while len(batch) < self.max_batch:
                timeout = deadline - time.monotonic()
                try:
                    nxt = self._q.get(timeout=timeout) if timeout > 0 else self._q.get_nowait()
                except queue.Empty:
                    break
                if nxt is self._STOP:
                    stop = True
                    break
                batch.append(nxt)
            try:
                results = self.fn_batch([s['item'] for s in batch])
                if len(results) != len(batch):
                    raise RuntimeError(f'batched function returned {len(results)} results for {len(batch)} items')
                for s, r in zip(batch, results):
                    if isinstance(r, _Failed):
                        s['error'] = r.error
                    else:
                        s['result'] = r
            except BaseException as e:  # noqa: BLE001 - delivered to the callers
                for s in batch:
                    s['error'] = e
            self.batches += 1
            self.items += len(batch)
            for s in batch:
                s['done'].set()
            if stop:
                return


def collate_truth(pb, annotations):
    """Whole batch of annotations -> np.float32 (B, N, 3+C+4) in one launch (dl/od_pb.py:299-313)."""
    return pb.encode_truth(list(annotations))


def encode_truth_hook(od, max_batch=64, max_wait_ms=2.0, device=None):
    """Replacement for the reference's ``lambda y: pb.encode_truth([y])[0]`` (dl/od.py:169): same per-sample
    call shape, but concurrent calls from the generator's worker threads share one ``odk_assign_encode``
    launch.  Returns the callable; ``hook.coalescer.close()`` stops the worker thread."""
    import torch
    # the worker thread has its own "current device" (device 0 by default): pin it to the device that is current
    # where the hook is built — in a one-process-per-GPU job that is the rank's GPU
    dev = torch.cuda.current_device() if device is None else torch.device(device).index

    def run(ys):
        with torch.cuda.device(dev):
            try:
                return list(od.pb.encode_truth(ys))
            except AssertionError:
                if len(ys) == 1:
                    raise
            # a bad sample fails only its own call, as in the reference (od_pb.py:304,332,338 raise per sample)
            out = []
            for y in ys:
                try:
                    out.append(od.pb.encode_truth([y])[0])
                except AssertionError as e:
                    out.append(_Failed(e))
            return out
    co = BatchCoalescer(run, max_batch=max_batch, max_wait_ms=max_wait_ms)

    def hook(y):
        return co(y)
    hook.coalescer = co
    return hook
