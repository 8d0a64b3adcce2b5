"""Pinned host<->device copy bandwidth of the box (what bounds gnc_classify_host): python tools/pcie_peak.py

Measured on the round-1 B200 box: H2D 55.6 GB/s alone, 54.6 GB/s with a D2H running beside it; the e2e line of bench.py
moves 2.11 GB up and 0.32 GB down per step in 42.3-43.0 ms = 50 GB/s up, i.e. ~90 % of this."""
import torch, time
n = 2_113_000_000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(320_000_000, dtype=torch.uint8, pin_memory=True)
d2 = torch.empty(320_000_000, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); d.copy_(h, non_blocking=True); e1.record(); torch.cuda.synchronize()
print("H2D alone GB/s", n / e0.elapsed_time(e1) / 1e6)
torch.cuda.synchronize()
t0 = time.perf_counter()
with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print("H2D + D2H concurrently: ms", dt * 1e3, "H2D GB/s", n / dt / 1e9)
