"""Pinned device-to-host bandwidth of the box with 1, 2 and 4 concurrent streams (403 MB, the fp32 prediction of one
bench step): the ceiling of bench.py's `e2e`.  [B200 r01] 56.7 / 55.5 / 55.0 GB/s: one copy engine saturates the link."""
import torch, time
n = 403 * 1024 * 1024 // 4
src = torch.empty(n, device="cuda"); dst = torch.empty(n).pin_memory()
def run(k, reps=8):
    streams = [torch.cuda.Stream() for _ in range(k)]
    chunks = [(i * n // k, (i + 1) * n // k) for i in range(k)]
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        for s, (a, b) in zip(streams, chunks):
            with torch.cuda.stream(s):
                dst[a:b].copy_(src[a:b], non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    return reps * n * 4 / dt / 1e9
for k in (1, 2, 4):
    run(k, 2); print(k, "streams: %.1f GB/s" % run(k))
