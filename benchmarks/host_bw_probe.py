import torch, time
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h2.copy_(d2, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5): fn()
    torch.cuda.synchronize()
    print(name, 5 * n / (time.perf_counter() - t0) / 1e9, "GB/s")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
print("both directions, each", 5 * n / (time.perf_counter() - t0) / 1e9, "GB/s")
import numpy as np
a = np.empty(n // 8); b = h.numpy().view(np.float64)
t0 = time.perf_counter(); np.copyto(b, a); print("host memcpy 1 thread", n / (time.perf_counter() - t0) / 1e9, "GB/s (first touch)")
t0 = time.perf_counter(); np.copyto(b, a); print("host memcpy 1 thread", n / (time.perf_counter() - t0) / 1e9, "GB/s")
import os; print("cpus", os.cpu_count())
