"""T env steps per launch at the headline batch (ewn_rollout without a network, k_rollout_env): 1 Mi 5x5 games x T steps with pooled
uniform actions, against T launches of the VecEnv step kernel on the same games.  us per env-step-of-the-batch and fraction of the HBM peak
(79 B per env step: action 16 R, reward 4, done 1, board 50, dice 8 W -- the games stay in registers)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200.engine import Engine
e = Engine(5, 3)
dev = torch.device("cuda", 0)
n, SEED = int(os.environ.get("EWN_N", 1 << 20)), 9487
peak = 6548.5
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
g = torch.Generator(device=dev); g.manual_seed(7)
for T in (4, 16, 32):
    acts = torch.stack([torch.randint(0, 2, (T, n), generator=g, device=dev), torch.randint(0, 3, (T, n), generator=g, device=dev)], 2).contiguous()
    st = e.reset(n, seed=SEED)
    bufs = {"board": torch.empty((T + 1, n, 5, 5), dtype=torch.int16, device=dev), "dice": torch.empty((T + 1, n), dtype=torch.int64, device=dev),
            "actions": acts, "rewards": torch.empty((T, n), device=dev), "dones": torch.empty((T, n), dtype=torch.uint8, device=dev)}
    ctr = torch.zeros(1, dtype=torch.int64, device=dev)
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    roll = e.rollout_plan(None, st, T, bufs, seed=SEED, step_dev=ctr, stats=stats)
    def one():
        roll(1); e.counter_add(ctr, T)
    for _ in range(3): one()
    torch.cuda.synchronize()
    reps = max(4, 64 // T)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): one()
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) / reps / T * 1e3
    byt = 79 + (50 + 8) / T                                      # + the extra observation slot per launch
    print(f"n={n} T={T:3d}: {us:7.2f} us per step of the batch   {n / us * 1e6:.3e} env steps/s   {n * byt / us / 1e3:7.1f} GB/s = {n * byt / us / 1e3 / peak:.3f} of the HBM peak "
          f"({byt:.1f} B per env step)   finished per step {int(stats[3]) / ((3 + reps) * T):.0f}", flush=True)
    del bufs, acts
