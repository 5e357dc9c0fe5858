"""TEST / MEASUREMENT INFRASTRUCTURE ONLY -- times the UNMODIFIED reference's own Python code for the hot path on the host cores.

What runs (SURVEY 8d "CPU baseline"), one worker process per core over disjoint seed ranges:
  * the evaluation() loop of eval_random.py:17-51 (RandomAgent vs the env's random opponent)        -> env steps / s
  * MctsAgent.simulate (classical_policies/mcts.py:21-45) from mid-game positions                   -> playouts / s
  * ExpectiMinimaxAgent(max_depth=4).predict (classical_policies/minimax.py:89-93), 'hybrid'        -> roots / s
The reference is imported through oracle/ref_harness.py (stubbed gymnasium / stable_baselines3 / pygame, no game logic in the
stubs) from EWN_REFERENCE_ROOT.  `python -m oracle.time_reference [seconds] [out.json]` writes the numbers as JSON; bench.py calls
`measure()` when a reference tree is present on the machine it runs on and otherwise quotes the committed container measurement.
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import sys
import time


def _worker(args):
    kind, seed0, seconds, N, L = args
    import numpy as np
    from oracle.ref_harness import load_reference
    R = load_reference()
    t_end = time.perf_counter() + seconds
    units, t0 = 0, time.perf_counter()
    if kind == "steps":                                       # eval_random.py:17-51
        env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
        agent = R.RandomAgent(env)
        seed = seed0
        while time.perf_counter() < t_end:
            obs, _ = env.reset(seed=seed)
            done = False
            while not done:
                action, _ = agent.predict(obs, deterministic=True)
                obs, reward, done, trunc, info = env.step(action)
                units += 1
            seed += 1
    elif kind == "playouts":                                  # mcts.py:21-45 on positions two random steps into a game
        import copy
        import random
        random.seed(seed0)
        ag = R.MctsAgent(cube_layer=L, board_size=N, num_simulations=20, num_env_copies=1)
        env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
        rnd = R.RandomAgent(env)
        seed = seed0
        while time.perf_counter() < t_end:
            obs, _ = env.reset(seed=seed)
            for _ in range(2):
                a, _ = rnd.predict(obs)
                obs, r, done, _, _ = env.step(a)
                if done:
                    break
            if done:
                seed += 1
                continue
            ag.restore_env_with_obs({"board": np.array(obs["board"], np.int16), "dice_roll": int(obs["dice_roll"])})
            la = ag.env.get_legal_actions(R.Player.TOP_LEFT)
            ag.env.make_simulated_action(R.Player.TOP_LEFT, la[0])
            ag.simulate(copy.deepcopy(ag.env))
            units += 20
            seed += 1
    else:                                                     # minimax.py:89-93, depth 4
        ag = R.ExpectiMinimaxAgent(max_depth=4, cube_layer=L, board_size=N, heuristic="hybrid")
        env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
        rnd = R.RandomAgent(env)
        seed = seed0
        while time.perf_counter() < t_end:
            obs, _ = env.reset(seed=seed)
            done = False
            while not done and time.perf_counter() < t_end:
                ag.predict({"board": np.array(obs["board"], np.int16), "dice_roll": int(obs["dice_roll"])})
                units += 1
                a, _ = rnd.predict(obs)
                obs, r, done, _, _ = env.step(a)
            seed += 1
    return units, time.perf_counter() - t0


def measure(seconds: float = 10.0, procs: int | None = None, N: int = 5, L: int = 3) -> dict:
    """-> {"steps_per_s", "playouts_per_s", "roots_per_s", "cores", ...}: aggregate over `procs` worker processes"""
    procs = procs or os.cpu_count() or 1
    out = {"cores": procs, "seconds_per_item": seconds, "board_size": N, "cube_layer": L,
           "what": "unmodified reference (jchen8tw/ewn-gym) through oracle/ref_harness.py, one process per core"}
    ctx = mp.get_context("fork")
    with ctx.Pool(procs) as pool:
        for kind, key in (("steps", "steps_per_s"), ("playouts", "playouts_per_s"), ("roots", "roots_per_s")):
            res = pool.map(_worker, [(kind, 1_000_000 + 100_000 * i, seconds, N, L) for i in range(procs)])
            out[key] = float(sum(u / t for u, t in res))
            out[key.replace("_per_s", "_per_s_per_core")] = out[key] / procs
    return out


if __name__ == "__main__":
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
    r = measure(secs)
    r["where"] = os.environ.get("EWN_WHERE", "build container")
    txt = json.dumps(r, indent=1)
    print(txt)
    if len(sys.argv) > 2:
        with open(sys.argv[2], "w") as f:
            f.write(txt + "\n")
