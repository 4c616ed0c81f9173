#!/usr/bin/env python
"""What a plain device-to-device copy reaches between two GPUs of the box (the practical
ceiling for the gather of the packed slabs): one process, tensor.copy_ across devices."""
import torch, sys
n = torch.cuda.device_count()
size = 2 << 30
src = [torch.empty(size, dtype=torch.uint8, device="cuda:%d" % i) for i in range(1, n)]
dst = [torch.empty(size, dtype=torch.uint8, device="cuda:0") for _ in range(1, n)]
def run(k):
    """k senders into device 0 concurrently, each on its own stream of the source device"""
    streams = [torch.cuda.Stream(device="cuda:%d" % (i + 1)) for i in range(k)]
    best = 0.0
    for rep in range(4):
        for d in range(n):
            torch.cuda.synchronize(d)
        ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
        import time
        t0 = time.perf_counter()
        for i in range(k):
            with torch.cuda.stream(streams[i]):
                dst[i].copy_(src[i], non_blocking=True)
        for d in range(n):
            torch.cuda.synchronize(d)
        t = time.perf_counter() - t0
        best = max(best, k * size / t / 1e9)
    return best
for k in range(1, n):
    print("%d sender(s) into one GPU: %.1f GB/s (wall clock, best of 4, %d GiB each)" % (k, run(k), size >> 30), flush=True)
