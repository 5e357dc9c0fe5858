"""GPU measurement of one Monte-Carlo decision (K3): ewn_mc_predict against the five-launch sequence it replaces, at the per-rank
playout share of 1, 2, 4 and 8 GPUs (1024 / world playouts per root action, 4096 mid-game roots).  python tools/mc_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200 import _lib
if os.environ.get("EWN_LIB_VARIANT"):
    _lib.LIB_PATH = os.path.abspath(os.environ["EWN_LIB_VARIANT"])
from ewn_gym_b200.engine import Engine
e = Engine(5, 3)
SEED, n = 9487, 4096
roots = e.reset(n, seed=SEED + 1)
for s in range(1, 5):
    e.step(roots, None, seed=SEED + 1, step=s, agent_random=True, autoreset=True, want=())
wins = torch.zeros(n, 6, dtype=torch.int32, device="cuda")
def old(it, c):
    wins.zero_(); e.mc_root_wins(roots, seed=it, first=0, count=c, wins=wins); return e.mc_best_action(roots, wins)
def new(it, c):
    return e.mc_predict(roots, seed=it, first=0, count=c, wins=wins)[0]
for c in (1024, 512, 256, 128):
    row = []
    for name, fn in (("5-launch", old), ("mc_predict", new)):
        fn(0, 8); torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for it in range(8): fn(it, c)
            b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / 8)
        row.append(f"{name} {best * 1e3:8.1f} us")
    print(f"{os.environ.get('EWN_LABEL', 'product'):12s} playouts/child {c:5d}: " + "   ".join(row), flush=True)
