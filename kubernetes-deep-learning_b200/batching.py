"""Server-side dynamic batching: concurrent single-image requests are coalesced into one forward call.

The reference never turns on TF-Serving's batching (tf-serving.dockerfile has no ``--enable_batching``), so every request runs alone;
on a B200 a batch-1 forward leaves >95 % of the GPU idle, so the replacement offers the knob TF-Serving has, with the same parameter
names as its ``batching_parameters`` file (max_batch_size, batch_timeout_micros, max_enqueued_batches).  Host-side only: the batch is
assembled from Python objects and handed to one `run_batch` call; nothing here touches CUDA.

Semantics
* a request of `size` images joins the open batch if it fits, otherwise the open batch is closed and run first;
* an open batch is run when it reaches `max_batch_size` images or `batch_timeout_micros` after its first request arrived;
* a request larger than `max_batch_size` is run on its own (TF-Serving would reject it; the unbatched server accepts any size, and
  the gateway only ever sends 1);
* when `max_enqueued_batches` worth of images is already waiting the request fails with `QueueFull` (gRPC UNAVAILABLE upstream);
* an exception raised by `run_batch` is re-raised in every request of that batch.
Results come back in request order within a batch, so a request always receives the rows computed from its own images.
"""
import threading
import time
from collections import deque
from dataclasses import dataclass
from typing import Any, Callable, List, Optional, Sequence


class QueueFull(RuntimeError):
    pass


class BatcherClosed(RuntimeError):
    pass


@dataclass
class BatchingParameters:
    max_batch_size: int = 32            # images per forward call (configs[4] of BASELINE.json uses batches up to 32)
    batch_timeout_micros: int = 1000    # how long the first request of a batch may wait for company
    max_enqueued_batches: int = 64      # back-pressure: at most this many full batches of images may be waiting

    def __post_init__(self):
        if self.max_batch_size < 1:
            raise ValueError("max_batch_size must be >= 1")
        if self.batch_timeout_micros < 0:
            raise ValueError("batch_timeout_micros must be >= 0")
        if self.max_enqueued_batches < 1:
            raise ValueError("max_enqueued_batches must be >= 1")


class _Task:
    __slots__ = ("item", "size", "done", "result", "error", "t_enq")

    def __init__(self, item, size):
        self.item, self.size = item, size
        self.done = threading.Event()
        self.result = None
        self.error: Optional[BaseException] = None
        self.t_enq = time.monotonic()


class DynamicBatcher:
    """`submit(item, size)` blocks until the batch containing `item` has run and returns that item's result.

    run_batch(items: list) -> sequence of len(items) results; it is always called from the single scheduler thread, so a
    non-re-entrant model needs no further locking."""

    def __init__(self, run_batch: Callable[[List[Any]], Sequence[Any]], params: Optional[BatchingParameters] = None):
        self._run = run_batch
        self.params = params or BatchingParameters()
        self._q: deque = deque()
        self._queued_images = 0
        self._cv = threading.Condition()
        self._closed = False
        self.batches_run = 0
        self.images_run = 0
        self.batch_sizes: deque = deque(maxlen=1024)     # recent batch sizes in images (observability, tests)
        self._thread = threading.Thread(target=self._loop, name="kdl-batcher", daemon=True)
        self._thread.start()

    # -- client side -------------------------------------------------------------------------------------------------------------
    def submit(self, item, size: int = 1, timeout: Optional[float] = None):
        if size < 1:
            raise ValueError("size must be >= 1")
        t = _Task(item, size)
        with self._cv:
            if self._closed:
                raise BatcherClosed("batcher is closed")
            if self._queued_images + size > self.params.max_enqueued_batches * self.params.max_batch_size and self._q:
                raise QueueFull(f"{self._queued_images} images already enqueued "
                                f"(max_enqueued_batches={self.params.max_enqueued_batches} x max_batch_size={self.params.max_batch_size})")
            self._q.append(t)
            self._queued_images += size
            self._cv.notify_all()
        if not t.done.wait(timeout):
            raise TimeoutError(f"no result within {timeout} s")   # the batch still runs; its result for this task is dropped
        if t.error is not None:
            raise t.error
        return t.result

    def close(self):
        with self._cv:
            self._closed = True
            self._cv.notify_all()
        self._thread.join(timeout=30)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- scheduler thread --------------------------------------------------------------------------------------------------------
    def _take_batch(self) -> Optional[List[_Task]]:
        p = self.params
        with self._cv:
            while not self._q:
                if self._closed:
                    return None
                self._cv.wait()
            deadline = self._q[0].t_enq + p.batch_timeout_micros * 1e-6
            while True:
                n = 0
                full = False
                for t in self._q:
                    if n and n + t.size > p.max_batch_size:
                        full = True       # the next request does not fit: close the batch in front of it
                        break
                    n += t.size
                    if n >= p.max_batch_size:
                        full = True
                        break
                left = deadline - time.monotonic()
                if full or left <= 0 or self._closed:
                    break
                self._cv.wait(left)
            batch, n = [], 0
            while self._q and (not batch or n + self._q[0].size <= p.max_batch_size):
                t = self._q.popleft()
                batch.append(t)
                n += t.size
            self._queued_images -= n
            return batch

    def _loop(self):
        while True:
            batch = self._take_batch()
            if batch is None:
                return
            try:
                results = self._run([t.item for t in batch])
                if len(results) != len(batch):
                    raise RuntimeError(f"run_batch returned {len(results)} results for {len(batch)} requests")
                for t, r in zip(batch, results):
                    t.result = r
            except BaseException as e:   # noqa: BLE001 - delivered to every waiting request
                for t in batch:
                    t.error = e
            n = sum(t.size for t in batch)
            self.batches_run += 1
            self.images_run += n
            self.batch_sizes.append(n)
            for t in batch:
                t.done.set()
