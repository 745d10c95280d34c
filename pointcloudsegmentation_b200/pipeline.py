"""Look-ahead block preparation: the role of the reference's provider threads (provider.py:42-140 feeds
prepared blocks to the towers through a queue while the previous step trains).

`model.prepare` — the point-only part of a step: host->device copy of the block batch, two-level voxel
pooling, the nine radius searches, the transpose CSRs — reads no parameter, so it can run one batch ahead.
It contains host syncs (every search returns its edge count to size the index tensors); issued from the
training thread they stall the enqueueing of the step's ~1300 launches, and the step is within 10 % of
being launch-bound.  Here it runs on its own CUDA stream from its own host thread; the training thread only
waits on an event.  ctypes and the torch bindings release the GIL inside their calls, so the two threads
overlap.

    ahead = PrepareAhead(model, num_blocks)
    fut = ahead.submit(lambda: {k: v.to(dev, non_blocking=True) for k, v in host_batch.items()})
    ...                      # train on the previous batch
    prep = ahead.take(fut)   # tensors are safe to use on the current stream
    loss, _, _ = model(prepared=prep)
"""
from __future__ import annotations

import queue
import threading
from concurrent.futures import Future

import torch


def _tensors(x):
    if torch.is_tensor(x):
        yield x
    elif isinstance(x, dict):
        for y in x.values():
            yield from _tensors(y)
    elif isinstance(x, (list, tuple)):
        for y in x:
            yield from _tensors(y)


class PrepareAhead:
    def __init__(self, model, num_blocks, device=None, prepare_fn=None):
        """prepare_fn(batch) -> prepared (default: model.prepare of the S3DIS-style batch dict)."""
        self.model = model
        self.num_blocks = num_blocks
        self.prepare_fn = prepare_fn
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self._jobs: queue.Queue = queue.Queue()
        self._thread = threading.Thread(target=self._run, name="pcs-prepare", daemon=True)
        self._thread.start()

    def _run(self):
        torch.cuda.set_device(self.device)
        while True:
            job = self._jobs.get()
            if job is None:
                return
            fut, batch_fn = job
            try:
                with torch.cuda.stream(self.stream):
                    b = batch_fn()  # resident tensors, or the H2D copies of a pinned host batch
                    if self.prepare_fn is not None:
                        prep = self.prepare_fn(b)
                    else:
                        prep = self.model.prepare(b["xyzs"], b["feats"], b["labels"], b.get("block_ids"), self.num_blocks)
                    done = torch.cuda.Event()
                    done.record(self.stream)
                fut.set_result((b, prep, done))
            except BaseException as e:  # noqa: BLE001 - handed to the training thread
                fut.set_exception(e)

    def submit(self, batch_fn) -> Future:
        """batch_fn() -> {"xyzs", "feats", "labels", "block_ids"} of CUDA tensors; called on the look-ahead stream."""
        fut: Future = Future()
        self._jobs.put((fut, batch_fn))
        return fut

    def take(self, fut: Future):
        """Block until the job has been ENQUEUED (not executed), make the current stream wait for it on the
        device, and hand the prepared tuple over to the current stream's allocator bookkeeping."""
        b, prep, done = fut.result()
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(done)
        for t in list(_tensors(prep)) + list(_tensors(b)):
            t.record_stream(cur)
        return prep

    def close(self):
        self._jobs.put(None)
        self._thread.join()
