"""D2H bandwidth of one 8.3-MB frame to pinned memory: one cudaMemcpyAsync vs the frame split over k streams (k copy engines)."""
import sys, time
import torch
n = 1920 * 1080 * 4
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
host = torch.empty(n, dtype=torch.uint8).pin_memory()
for k in (1, 2, 3, 4):
    streams = [torch.cuda.Stream() for _ in range(k)]
    chunk = (n // k + 255) // 256 * 256
    def copy_once():
        for j, st in enumerate(streams):
            lo, hi = j * chunk, min(n, (j + 1) * chunk)
            with torch.cuda.stream(st):
                host[lo:hi].copy_(dev[lo:hi], non_blocking=True)
    for _ in range(20): copy_once()
    torch.cuda.synchronize()
    reps = 300
    t0 = time.perf_counter()
    for _ in range(reps): copy_once()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    print(f"{k} stream(s): {1e6 * dt:.1f} us / frame, {n / dt / 1e9:.1f} GB/s")
