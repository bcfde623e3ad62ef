"""H2D bandwidth of one pinned block, freshly written by the CPU or not (staging-path sanity check)."""
import time, torch
for mb in (4, 15, 64, 128):
    n = mb << 20
    src = torch.empty(n, dtype=torch.uint8).random_(0, 255)
    pin = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    dev = torch.empty(n, dtype=torch.uint8, device="cuda")
    pin.copy_(src); dev.copy_(pin, non_blocking=True); torch.cuda.synchronize()
    for fresh in (False, True):
        ts = []
        for _ in range(5):
            if fresh:
                pin.copy_(src)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            dev.copy_(pin, non_blocking=True); torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        t = sorted(ts)[2]
        print("%4d MB %s: %.3f ms  %.1f GB/s" % (mb, "fresh" if fresh else "cold ", t * 1e3, n / t / 1e9))
