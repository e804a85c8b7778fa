"""Multi-GPU = independent replicas.  The reference scales by adding Kubernetes replicas behind a Service
(replicas: 1 in both deployment yamls :8, no shared state); here one model replica lives on each GPU of the box, the batch is
split on the host, and the only exchange is the host-side concatenation of (B,10) logits.  No collective on the data path.

Two ways to run it:
  * `ReplicaSet`        -- one process, one host thread + one `engine.Model` per GPU (ctypes releases the GIL in the C call);
  * process per GPU     -- `shard_range` / `gather_logits` under torch.distributed (torchrun); NCCL or gloo is used ONLY to
                           bring the tiny logit blocks to rank 0.
"""
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split: the first n_items % world ranks get one extra item.  Returns [start, end)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_ranges(n_items: int, world: int) -> List[Tuple[int, int]]:
    return [shard_range(n_items, r, world) for r in range(world)]


def gather_logits(local: np.ndarray, group=None, dst: int = 0) -> Optional[np.ndarray]:
    """Concatenate per-rank (b_r, 10) float32 logit blocks in rank order on `dst` (None elsewhere).
    Blocks may have different b_r (including 0)."""
    import torch
    import torch.distributed as dist

    if not dist.is_initialized():
        return np.ascontiguousarray(local, np.float32)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    use_cuda = dist.get_backend(group) == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    local = np.ascontiguousarray(local, np.float32).reshape(-1, 10)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[rank] = local.shape[0]
    dist.all_reduce(counts, group=group)
    cmax = int(counts.max().item())
    pad = torch.zeros((max(cmax, 1), 10), dtype=torch.float32, device=dev)
    if local.shape[0]:
        pad[:local.shape[0]] = torch.from_numpy(local).to(dev)
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    if rank != dst:
        return None
    parts = [bufs[r][:int(counts[r].item())].cpu().numpy() for r in range(world)]
    return np.concatenate(parts, 0) if parts else np.empty((0, 10), np.float32)


def run_sharded(images: Sequence, predict: Callable[[Sequence], np.ndarray], group=None) -> Optional[np.ndarray]:
    """Process-per-GPU data path: every rank calls this with the SAME image list; each predicts its shard with its local
    replica and rank 0 gets the logits in the original order."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(len(images), rank, world)
    local = predict(images[lo:hi]) if hi > lo else np.empty((0, 10), np.float32)
    return gather_logits(local, group)


class ReplicaSet:
    """One `engine.Model` per GPU inside one process.  `predict_uint8(images)` splits the batch across the replicas, runs
    them concurrently on their own host threads and returns the logits in input order."""

    def __init__(self, weights, devices: Optional[Sequence[int]] = None, **model_kwargs):
        import torch
        from .engine import Model

        if not torch.cuda.is_available():
            raise RuntimeError("ReplicaSet needs CUDA devices; there is no CPU fallback")
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise ValueError("no devices")
        self.devices = list(devices)
        self.models = [Model(weights, device=d, **model_kwargs) for d in self.devices]
        self.pool = ThreadPoolExecutor(max_workers=len(self.models))

    def predict_uint8(self, images: Sequence[np.ndarray]) -> np.ndarray:
        ranges = shard_ranges(len(images), len(self.models))
        futs = [self.pool.submit(m.predict_uint8, images[lo:hi]) if hi > lo else None
                for m, (lo, hi) in zip(self.models, ranges)]
        parts = [f.result() for f in futs if f is not None]
        return np.concatenate(parts, 0) if parts else np.empty((0, 10), np.float32)

    def close(self):
        self.pool.shutdown(wait=True)
        for m in self.models:
            m.close()
