"""Host<->device copy latencies that bound the single-frame call (development aid)."""
import time
import numpy as np, torch
a = torch.from_numpy(np.random.randint(0, 255, (480, 752), dtype=np.uint8))
p = a.pin_memory()
d = torch.empty_like(a, device="cuda")
r = torch.empty(73 * 1024, dtype=torch.uint8, device="cuda")
rh = torch.empty(73 * 1024, dtype=torch.uint8).pin_memory()
def t(fn, n=300):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    return (time.perf_counter() - t0) / n * 1e6
print("pageable 361 KB H2D + sync: %.1f us" % t(lambda: (d.copy_(a), torch.cuda.synchronize())))
print("pinned   361 KB H2D + sync: %.1f us" % t(lambda: (d.copy_(p, non_blocking=True), torch.cuda.synchronize())))
print("host memcpy 361 KB        : %.1f us" % t(lambda: p.copy_(a)))
print("pinned 73 KB D2H + sync   : %.1f us" % t(lambda: (rh.copy_(r, non_blocking=True), torch.cuda.synchronize())))
print("empty sync                : %.1f us" % t(lambda: torch.cuda.synchronize()))
