"""Host-side logic of the multi-GPU (replicas-only) path, on CPU: sharding and the world_size-2 gloo gather of logits."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import kdl_b200  # noqa: F401
from kdl_b200 import replicas


def test_shard_ranges_cover_everything_once():
    for n in (0, 1, 2, 5, 64, 1023, 1024):
        for world in (1, 2, 3, 4, 8):
            rs = replicas.shard_ranges(n, world)
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in rs]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        replicas.shard_range(4, 2, 2)


def _fake_predict(images):
    # deterministic "logits" that identify the image: row = image id * 10 + class
    return np.array([[float(im[0, 0, 0]) * 10 + c for c in range(10)] for im in images], np.float32).reshape(-1, 10)


def _worker(rank, world, port, n_images, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        images = [np.full((2, 2, 3), i, np.uint8) for i in range(n_images)]
        out = replicas.run_sharded(images, _fake_predict)
        if rank == 0:
            q.put(out)
        else:
            assert out is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("n_images", [5, 8, 1])
def test_world2_gloo_gather_preserves_order(n_images):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    want = _fake_predict([np.full((2, 2, 3), i, np.uint8) for i in range(n_images)])
    np.testing.assert_array_equal(out, want)


def test_single_process_passthrough():
    x = np.arange(30, dtype=np.float32).reshape(3, 10)
    np.testing.assert_array_equal(replicas.gather_logits(x), x)
