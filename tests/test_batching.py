"""Host-side dynamic batcher (batching.py): routing, limits, time-out flush, errors.  CPU only."""
import threading
import time

import numpy as np
import pytest

import kdl_b200  # noqa: F401
from kdl_b200.batching import BatcherClosed, BatchingParameters, DynamicBatcher, QueueFull


def _run_threads(n, fn):
    out, errs = [None] * n, [None] * n

    def work(i):
        try:
            out[i] = fn(i)
        except BaseException as e:   # noqa: BLE001
            errs[i] = e
    ts = [threading.Thread(target=work, args=(i,)) for i in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(30)
    return out, errs


def test_results_are_routed_to_their_requests_and_batches_respect_the_limit():
    seen = []

    def run(items):
        seen.append(len(items))
        time.sleep(0.002)
        return [i * 10 for i in items]
    with DynamicBatcher(run, BatchingParameters(max_batch_size=8, batch_timeout_micros=20000)) as b:
        out, errs = _run_threads(50, lambda i: b.submit(i))
    assert errs == [None] * 50
    assert out == [i * 10 for i in range(50)]
    assert max(seen) <= 8 and sum(seen) == 50
    assert len(seen) < 50          # some coalescing happened
    assert b.images_run == 50 and b.batches_run == len(seen)


def test_single_request_is_flushed_after_the_timeout():
    with DynamicBatcher(lambda items: items, BatchingParameters(max_batch_size=32, batch_timeout_micros=30000)) as b:
        t0 = time.monotonic()
        assert b.submit("x") == "x"
        dt = time.monotonic() - t0
    assert 0.02 < dt < 2.0
    assert list(b.batch_sizes) == [1]


def test_full_batch_runs_without_waiting_for_the_timeout():
    with DynamicBatcher(lambda items: items, BatchingParameters(max_batch_size=4, batch_timeout_micros=5_000_000)) as b:
        t0 = time.monotonic()
        out, errs = _run_threads(4, lambda i: b.submit(i))
        dt = time.monotonic() - t0
    assert errs == [None] * 4 and sorted(out) == [0, 1, 2, 3]
    assert dt < 2.0


def test_sized_requests_never_overflow_a_batch_and_oversize_runs_alone():
    sizes = []

    def run(items):
        sizes.append(sum(a.shape[0] for a in items))
        return [a.sum(axis=1) for a in items]
    with DynamicBatcher(run, BatchingParameters(max_batch_size=8, batch_timeout_micros=20000)) as b:
        arrs = [np.full((n, 3), n, np.float32) for n in (3, 3, 3, 5, 8, 1, 20, 2)]
        out, errs = _run_threads(len(arrs), lambda i: b.submit(arrs[i], size=arrs[i].shape[0]))
    assert errs == [None] * len(arrs)
    for a, o in zip(arrs, out):
        np.testing.assert_array_equal(o, a.sum(axis=1))
    assert sum(sizes) == sum(a.shape[0] for a in arrs)
    assert all(s <= 8 or s == 20 for s in sizes)


def test_run_batch_exception_reaches_every_request_of_the_batch_and_the_batcher_survives():
    calls = []

    def run(items):
        calls.append(list(items))
        if "bad" in items:
            raise ValueError("boom")
        return items
    with DynamicBatcher(run, BatchingParameters(max_batch_size=2, batch_timeout_micros=200000)) as b:
        out, errs = _run_threads(2, lambda i: b.submit("bad" if i == 0 else "good"))
        assert all(isinstance(e, ValueError) for e in errs)
        assert b.submit("later") == "later"


def test_wrong_result_count_is_an_error():
    with DynamicBatcher(lambda items: [], BatchingParameters(max_batch_size=2, batch_timeout_micros=0)) as b:
        with pytest.raises(RuntimeError, match="returned 0 results"):
            b.submit(1)


def test_queue_full_and_closed():
    gate = threading.Event()

    def run(items):
        gate.wait(10)
        return items
    b = DynamicBatcher(run, BatchingParameters(max_batch_size=1, batch_timeout_micros=0, max_enqueued_batches=2))
    out, errs = [], []
    ts = [threading.Thread(target=lambda: out.append(b.submit(0))) for _ in range(3)]   # 1 running + 2 waiting
    for t in ts:
        t.start()
    deadline = time.monotonic() + 5
    while time.monotonic() < deadline:
        with b._cv:
            if b._queued_images == 2:
                break
        time.sleep(0.005)
    with pytest.raises(QueueFull):
        b.submit(0)
    gate.set()
    for t in ts:
        t.join(10)
    assert out == [0, 0, 0]
    b.close()
    with pytest.raises(BatcherClosed):
        b.submit(0)


def test_parameters_are_validated():
    for kw in ({"max_batch_size": 0}, {"batch_timeout_micros": -1}, {"max_enqueued_batches": 0}):
        with pytest.raises(ValueError):
            BatchingParameters(**kw)
    with DynamicBatcher(lambda items: items) as b:
        with pytest.raises(ValueError):
            b.submit(0, size=0)
