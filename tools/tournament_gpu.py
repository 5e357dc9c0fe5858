"""eval_pairs.py:3-34 on the GPU: policy x opponent win-rate matrix (1,024 games per pair by default) plus the expectiminimax
depth sweep of the reference's assets/p16.png / p16-2.png (agent = ExpectiMinimaxAgent(depth d) vs a random opponent).

    python tools/tournament_gpu.py [--num 1024] [--board_size 5] [--max_depth 5] [--num_simulations 10] [--json out.json]
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200.engine import Engine
from ewn_gym_b200 import arena as A

ap = argparse.ArgumentParser()
ap.add_argument("--num", type=int, default=1024)
ap.add_argument("--board_size", type=int, default=5)
ap.add_argument("--cube_layer", type=int, default=3)
ap.add_argument("--max_depth", type=int, default=5)
ap.add_argument("--num_simulations", type=int, default=10)
ap.add_argument("--json", default=None)
args = ap.parse_args()

e = Engine(args.board_size, args.cube_layer)
pols = {"random": A.RandomPolicy(seed=11), "minimax": A.MinimaxPolicy(args.max_depth), "mcts": A.MonteCarloPolicy(args.num_simulations, 5, seed=7)}
A.play_matches(e, pols["random"], pols["random"], 64, seed=0)          # warm-up
res = {"config": vars(args), "pairs": {}, "depth_sweep": {}}
print(f"== policy x opponent, {args.num} games per pair, {args.board_size}x{args.board_size}, agent = TOP_LEFT (rows)")
for an, apol in pols.items():
    for on, opol in pols.items():
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = A.play_matches(e, apol, opol, args.num, seed=hash((an, on)) & 0xffff)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        lo, hi = A.wilson(r["agent_wins"], r["n"])
        print(f"{apol.name:>22s} vs {opol.name:<22s} win rate {r['agent_wins'] / r['n']:.3f}  [{lo:.3f}, {hi:.3f}]  {r['plies']:3d} plies  {dt:7.3f} s")
        res["pairs"][f"{an} vs {on}"] = {"win_rate": r["agent_wins"] / r["n"], "ci95": [lo, hi], "seconds": dt, "unfinished": r["unfinished"]}
print("== ExpectiMinimaxAgent(depth d, hybrid) vs random, cf. assets/p16-2.png (~0.62 / 0.72 / 0.83 / 0.85 / 0.54) and p16.png (50 .. 10,800 s)")
for d in range(1, args.max_depth + 1):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r = A.play_matches(e, A.MinimaxPolicy(d), pols["random"], args.num, seed=100 + d)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    lo, hi = A.wilson(r["agent_wins"], r["n"])
    print(f"depth {d}: win rate {r['agent_wins'] / r['n']:.3f}  [{lo:.3f}, {hi:.3f}]  {dt:7.3f} s for {args.num} episodes")
    res["depth_sweep"][str(d)] = {"win_rate": r["agent_wins"] / r["n"], "ci95": [lo, hi], "seconds": dt}
if args.json:
    json.dump(res, open(args.json, "w"), indent=1)
