"""GPU box: page-locked D2H / H2D copy rates, the ceiling of bench.py's e2e line (74 MB D2H + 17 MB H2D per 1 Mi-game step)."""
import time, torch
n = 74 << 20
d = torch.empty(n, dtype=torch.uint8, device="cuda"); h = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(17 << 20, dtype=torch.uint8, device="cuda"); h2 = torch.empty(17 << 20, dtype=torch.uint8).pin_memory()
s2 = torch.cuda.Stream()
for name, both in (("D2H 74 MiB alone", False), ("D2H 74 MiB with a concurrent 17 MiB H2D", True)):
    for _ in range(3): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20):
        h.copy_(d, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): d2.copy_(h2, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
    print(f"{name}: {dt * 1e3:.3f} ms  {n / dt / 1e9:.1f} GB/s")
