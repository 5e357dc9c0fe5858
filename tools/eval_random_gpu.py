#!/usr/bin/env python
"""eval_random.py of the reference (random agent vs random opponent, N seeded games, win rate + CI) with all games played
at once on the GPU: one fused k_step launch per agent step, in-kernel random agent and opponent (Philox), no auto-reset.

  python tools/eval_random_gpu.py --board_size 5 --cube_layer 3 --eval_num 10000 [--seed 0]

Same flags as eval_random.py:54-80 where they apply; unknown flags are ignored like the reference's **kwargs.
The reference prints `proportion_confint(count=winrate, ...)`, which feeds a rate where a count is expected
(eval_random.py:106); the interval printed here is the Wilson interval of wins / games.
"""
import argparse
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def wilson(k, n, z=1.959963984540054):
    p = k / n
    d = 1 + z * z / n
    c = p + z * z / (2 * n)
    h = z * math.sqrt(p * (1 - p) / n + z * z / (4 * n * n))
    return (c - h) / d, (c + h) / d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--board_size", "-bs", type=int, default=5)
    ap.add_argument("--cube_layer", "-cl", type=int, default=3)
    ap.add_argument("--eval_num", "-n", type=int, default=10000)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--numpy-exact", action="store_true",
                    help="replay the reference's own random stream (np.random.seed(seed) per game, seeds 0..eval_num-1): the same games, "
                         "dice and results as eval_random.py, from the seeds alone")
    args, _ = ap.parse_known_args()

    import torch
    from ewn_gym_b200.engine import Engine
    e = Engine(args.board_size, args.cube_layer)
    n = args.eval_num
    if args.numpy_exact:
        seeds = torch.arange(n, dtype=torch.int64, device="cuda").to(torch.int32)
        e.numpy_random_games(seeds[:32]); torch.cuda.synchronize()
        t0 = time.perf_counter()
        res, steps, _, _ = e.numpy_random_games(seeds)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        wins, agent_steps = int((res > 0).sum()), int(steps.sum())
        lo, hi = wilson(wins, n)
        print(f"[numpy-exact] games {n} (seeds 0..{n - 1})  wins {wins}  losses {n - wins}  win rate {wins / n:.4f}  95% CI [{lo:.4f}, {hi:.4f}]")
        print(f"[numpy-exact] agent steps {agent_steps} ({agent_steps / n:.2f} per game)  1 kernel launch  wall {dt:.4f} s  {n / dt:.0f} games/s")
        return
    t0 = time.perf_counter()
    state = e.reset(n, seed=args.seed)
    stats = torch.zeros(4, dtype=torch.int64, device=state.device)
    alive = torch.ones(n, dtype=torch.bool, device=state.device)
    steps = 0
    agent_steps = 0
    while bool(alive.any()):
        steps += 1
        agent_steps += int(alive.sum())
        # finished games keep reporting their result (check_win is sticky, no auto-reset): count each game once
        out = e.step(state, None, seed=args.seed, step=steps, agent_random=True, autoreset=False, want=("terminated", "code"))
        done_now = alive & (out["terminated"] != 0)
        code = out["code"][done_now]
        stats[0] += int((code == 2).sum()); stats[1] += int((code == 4).sum()); stats[3] += int(done_now.sum())
        alive &= ~done_now
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    wins, losses, _, fin = stats.tolist()
    lo, hi = wilson(wins, n)
    print(f"games {n}  wins {wins}  losses {losses}  win rate {wins / n:.4f}  95% CI [{lo:.4f}, {hi:.4f}]")
    print(f"agent steps {agent_steps} ({agent_steps / n:.2f} per game)  kernel launches {steps}  wall {dt:.3f} s  {n / dt:.0f} games/s")


if __name__ == "__main__":
    main()
