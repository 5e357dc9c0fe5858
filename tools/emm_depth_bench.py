import sys, os
sys.path.insert(0, '.')
import torch
from ewn_gym_b200.engine import Engine
e = Engine(5, 3)
for n in (4096, 65536):
    roots = e.reset(n, seed=3)
    for s in range(1, 5):
        e.step(roots, None, seed=3, step=s, agent_random=True, autoreset=True, want=())
    for depth in (4, 5, 6):
        e.expectiminimax(roots, depth); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5): e.expectiminimax(roots, depth)
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"roots={n} depth={depth} {ms:9.3f} ms {n/ms*1e3:.3e} roots/s", flush=True)
