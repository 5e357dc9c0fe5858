"""GPU measurement of K4 (expectiminimax): depth-4 roots/s of the fused on-chip kernel (k_emm_fused) against the record + replay
path (k_emm_expand + k_emm_replay), CUDA events on the launching stream, roots from random mid-game play.
  python tools/emm_bench.py [--ncu]     (--ncu: one call of each path at each size, nothing timed -- for ncu captures)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200.engine import Engine

SEED = 9487
ncu = "--ncu" in sys.argv
dev = torch.device("cuda", 0)
for N in (5, 7):
    e = Engine(N, 3)
    for n in (65536, 1 << 20):
        roots = e.reset(n, seed=SEED + 2)
        for s in range(1, 5):
            e.step(roots, None, seed=SEED + 2, step=s, agent_random=True, autoreset=True, want=())
        out = (torch.empty(n, dtype=torch.float64, device=dev), torch.empty(n, 2, dtype=torch.int8, device=dev))
        res = {}
        for depth in ((4,) if ncu else (3, 4)):
            for name, rp in (("fused", False), ("replay", True)):
                e.expectiminimax(roots, depth, replay_path=rp, out=out)          # warm-up (scratch allocation, tables)
                torch.cuda.synchronize()
                if ncu:
                    continue
                reps = 20 if n <= 65536 else 5
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(reps):
                    v, act = e.expectiminimax(roots, depth, replay_path=rp, out=out)
                b.record(); torch.cuda.synchronize()
                ms = a.elapsed_time(b) / reps
                res[(depth, name)] = (v.clone(), act.clone())
                print(f"N={N} roots={n:8d} depth={depth} {name:6s} {ms:8.4f} ms  {n / ms * 1e3:.3e} roots/s", flush=True)
            if not ncu:
                same = torch.equal(res[(depth, "fused")][0].view(torch.int64), res[(depth, "replay")][0].view(torch.int64)) and \
                    torch.equal(res[(depth, "fused")][1], res[(depth, "replay")][1])
                print(f"N={N} roots={n} depth={depth} fused == replay (value bits and actions): {same}", flush=True)
        del roots
