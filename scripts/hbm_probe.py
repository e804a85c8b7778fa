"""Read / write / copy bandwidth of this GPU with torch ops (context for the roofline fractions of write-heavy kernels)."""
import torch
n = 1 << 30  # 1 Gi bf16 = 2 GiB
a = torch.empty(n, dtype=torch.bfloat16, device="cuda"); b = torch.empty_like(a)
def t(fn, bytes_, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True); best = 1e9
    for _ in range(it):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return bytes_ / best / 1e6
print("write-only fill_   GB/s", round(t(lambda: a.fill_(1.0), 2 * n)))
print("copy (r+w)         GB/s", round(t(lambda: b.copy_(a), 4 * n)))
print("read-only sum      GB/s", round(t(lambda: a.view(torch.int16).max(), 2 * n)))
c = torch.empty(n // 2, dtype=torch.bfloat16, device="cuda")
print("read 2 : write 1 (add out) GB/s", round(t(lambda: torch.add(a[: n // 2], b[: n // 2], out=c), 3 * n)))
