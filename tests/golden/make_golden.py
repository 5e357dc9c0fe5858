#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py            # all fixtures, ~10 min on 8 cores
    python tests/golden/make_golden.py games5 mm5 # a subset

Every fixture is the output of reference code imported through oracle/ref_harness.py (stubbed gymnasium /
stable_baselines3 / pygame, no game logic in the stubs).  Dice rolls and opponent answers are *recorded* by
wrapping ``env.roll_dice`` / ``env.opponent_action`` so the exact trajectories can be replayed elsewhere.
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.ref_harness import load_reference  # noqa: E402

R = load_reference()
NPROC = min(8, os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------------ helpers
class Recorder:
    """Wraps env.roll_dice and env.opponent_action to log what happened inside one step()."""

    def __init__(self, env):
        self.env = env
        self.rolls, self.opp = [], []
        orig_roll, orig_opp = env.roll_dice, env.opponent_action

        def roll():
            orig_roll()
            self.rolls.append(int(env.dice_roll))

        def opp():
            a = orig_opp()
            self.opp.append((int(a[0]), int(a[1])))
            return a

        env.roll_dice, env.opponent_action = roll, opp

    def clear(self):
        self.rolls.clear()
        self.opp.clear()


def legal_mask(env, player):
    m = 0
    for a in env.get_legal_actions(player):
        m |= 1 << (a[0] * 3 + a[1])
    return m


def play_random_games(args):
    """eval_random.py:17-51 evaluation() loop for a seed range, with full per-step logging."""
    N, L, seeds = args
    env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
    rec = Recorder(env)
    agent = R.RandomAgent(env)
    steps = []   # rows: game, dice_agent, a0, a1, dice_opp, o0, o1, dice_next, term, trunc, legal_mask_before
    rewards, boards, game_rewards, game_steps, init_dice = [], [], [], [], []
    for seed in seeds:
        rec.clear()
        obs, _ = env.reset(seed=seed)
        init_dice.append(int(obs["dice_roll"]))
        done, n, reward = False, 0, 0
        while not done:
            rec.clear()
            dice_agent = int(env.dice_roll)
            lm = legal_mask(env, env.current_player)
            action, _ = agent.predict(obs, deterministic=True)
            obs, reward, done, trunc, info = env.step(action)
            d_opp = rec.rolls[0] if len(rec.rolls) > 0 else 0
            d_next = rec.rolls[1] if len(rec.rolls) > 1 else 0
            o = rec.opp[0] if rec.opp else (-1, -1)
            steps.append((seed, dice_agent, int(action[0]), int(action[1]), d_opp, o[0], o[1], d_next, int(done), int(trunc), lm))
            rewards.append(float(reward))
            boards.append(obs["board"].astype(np.int8).copy())
            n += 1
        game_rewards.append(float(reward))
        game_steps.append(n)
    return (np.array(steps, np.int32), np.array(rewards, np.float64), np.array(boards, np.int8),
            np.array(game_rewards), np.array(game_steps, np.int32), np.array(init_dice, np.int32))


def pmap(fn, jobs):
    with mp.Pool(NPROC) as pool:
        return pool.map(fn, jobs)


def split(seq, parts):
    seq = list(seq)
    k = (len(seq) + parts - 1) // parts
    return [seq[i:i + k] for i in range(0, len(seq), k)]


def gen_games(name, N, L, n_games):
    t = time.time()
    out = pmap(play_random_games, [(N, L, s) for s in split(range(n_games), NPROC * 4)])
    steps = np.concatenate([o[0] for o in out])
    np.savez_compressed(os.path.join(HERE, name),
                        N=N, L=L, steps=steps.astype(np.int8 if False else np.int32),
                        rewards=np.concatenate([o[1] for o in out]).astype(np.int8),
                        boards=np.concatenate([o[2] for o in out]),
                        game_rewards=np.concatenate([o[3] for o in out]).astype(np.int8),
                        game_steps=np.concatenate([o[4] for o in out]),
                        init_dice=np.concatenate([o[5] for o in out]).astype(np.int8),
                        columns="game,dice_agent,a_larger,a_dir,dice_opp,o_larger,o_dir,dice_next,terminated,truncated,legal_mask_before")
    gr = np.concatenate([o[3] for o in out])
    print(f"{name}: {n_games} games, {len(steps)} steps, wins {(gr > 0).sum()}, {time.time() - t:.1f}s")


# ------------------------------------------------------------------------------------------------ state sampler
def sample_states(args):
    """Observations at the agent's (TOP_LEFT) turn from random self-play: game `seed`, after a uniformly drawn
    number of env steps in 0..8 (or the last non-terminal obs); plus the terminal obs of every 8th game."""
    N, L, seeds = args
    env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
    agent = R.RandomAgent(env)
    boards, dice = [], []
    for seed in seeds:
        obs, _ = env.reset(seed=seed)
        k = int(np.random.randint(0, 9))
        done = False
        want_terminal = (seed % 8 == 7)
        i = 0
        while True:
            if (i == k and not want_terminal) or done:
                break
            prev = (obs["board"].copy(), int(obs["dice_roll"]))
            action, _ = agent.predict(obs)
            obs, r, done, _, _ = env.step(action)
            i += 1
            if done and not want_terminal:
                obs = {"board": prev[0], "dice_roll": prev[1]}
                break
        boards.append(np.array(obs["board"], np.int8).copy())
        dice.append(int(obs["dice_roll"]))
    return np.array(boards, np.int8), np.array(dice, np.int8)


def get_states(N, L, n):
    out = pmap(sample_states, [(N, L, s) for s in split(range(n), NPROC * 2)])
    return np.concatenate([o[0] for o in out]), np.concatenate([o[1] for o in out])


def run_minimax(args):
    N, L, depth, heuristic, boards, dice = args
    ag = R.ExpectiMinimaxAgent(max_depth=depth, cube_layer=L, board_size=N, heuristic=heuristic)
    vals, acts = [], []
    for b, d in zip(boards, dice):
        obs = {"board": b.astype(np.int16), "dice_roll": int(d)}
        ag.restore_env_with_obs(obs)
        v, a = ag.expectiminimax(depth, ag.env.agent_player, None, -float("inf"), float("inf"))
        a2, _ = ag.predict(obs)
        assert (a is None and a2 is None) or list(a) == list(a2)
        vals.append(float(v))
        acts.append((-1, -1) if a is None else (int(a[0]), int(a[1])))
    return np.array(vals, np.float64), np.array(acts, np.int8)


def gen_minimax(name, N, L, n_states, depth_counts, heur_depth=3, heur_count=300):
    t = time.time()
    boards, dice = get_states(N, L, n_states)
    out = dict(N=N, L=L, boards=boards, dice=dice)
    for depth, cnt in depth_counts.items():
        jobs = [(N, L, depth, "hybrid", boards[i:j], dice[i:j]) for i, j in
                [(c[0], c[-1] + 1) for c in split(range(cnt), NPROC * 4)]]
        res = pmap(run_minimax, jobs)
        out[f"value_d{depth}"] = np.concatenate([r[0] for r in res]).view(np.uint64)
        out[f"action_d{depth}"] = np.concatenate([r[1] for r in res])
        print(f"  {name} depth {depth}: {cnt} states, {time.time() - t:.1f}s")
    for h in ("min_dist", "two_min_dist", "attk"):
        jobs = [(N, L, heur_depth, h, boards[i:j], dice[i:j]) for i, j in
                [(c[0], c[-1] + 1) for c in split(range(heur_count), NPROC * 2)]]
        res = pmap(run_minimax, jobs)
        out[f"value_{h}_d{heur_depth}"] = np.concatenate([r[0] for r in res]).view(np.uint64)
        out[f"action_{h}_d{heur_depth}"] = np.concatenate([r[1] for r in res])
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(f"{name}: done {time.time() - t:.1f}s")


# ------------------------------------------------------------------------------------------------ evaluate + plies
def gen_rules(name, N, L, n_states):
    """All heuristics on sampled states; every (dice, action) ply for both colours on those states."""
    t = time.time()
    boards, dice = get_states(N, L, n_states)
    menv = R.MinimaxEnv(board_size=N, cube_layer=L)
    ag = R.ExpectiMinimaxAgent(max_depth=1, cube_layer=L, board_size=N)
    ag.env = menv
    evals = {h: [] for h in ("hybrid", "min_dist", "two_min_dist", "attk")}
    wins = []
    ply_rows, ply_boards, legal_rows = [], [], []
    for si, b in enumerate(boards):
        obs = {"board": b.astype(np.int16), "dice_roll": int(dice[si])}
        ag.restore_env_with_obs(obs)
        win = bool(menv.check_win())
        wins.append(win)
        for h in evals:
            evals[h].append(float(menv.evaluate(heuristic=h)))
        if win or si >= 1500:
            continue
        for player in (R.Player.TOP_LEFT, R.Player.BOTTOM_RIGHT):
            for d in range(1, menv.cube_num + 1):            # envs/ewn.py:91: the die has cube_num faces
                ag.restore_env_with_obs({"board": b.astype(np.int16), "dice_roll": d})
                has_cube = (menv.board > 0).any() if player == R.Player.TOP_LEFT else (menv.board < 0).any()
                if not has_cube:
                    continue
                la = menv.get_legal_actions(player)
                m = 0
                for a in la:
                    m |= 1 << (a[0] * 3 + a[1])
                flat = [x for a in la for x in a] + [-1] * (12 - 2 * len(la))
                legal_rows.append([si, player.value, d, m, len(la)] + flat)
                for larger in (0, 1):
                    for direction in (0, 1, 2):
                        ag.restore_env_with_obs({"board": b.astype(np.int16), "dice_roll": d})
                        idx = menv.find_cube_to_move(larger == 1, player)
                        valid = menv.execute_move(idx, direction)
                        cube = idx + 1 if idx >= 0 else -idx
                        ply_rows.append((si, player.value, d, larger, direction, int(valid), int(menv.check_win()), cube))
                        ply_boards.append(menv.board.astype(np.int8).copy())
    np.savez_compressed(os.path.join(HERE, name), N=N, L=L, boards=boards, dice=dice, win=np.array(wins, np.uint8),
                        **{f"eval_{h}": np.array(v, np.float64).view(np.uint64) for h, v in evals.items()},
                        ply_rows=np.array(ply_rows, np.int16), ply_boards=np.array(ply_boards, np.int8),
                        legal_rows=np.array(legal_rows, np.int16),
                        ply_columns="state,player,dice,larger,dir,valid,win_after,moved_cube",
                        legal_columns="state,player,dice,mask,count,list[6][2]")
    print(f"{name}: {len(boards)} states, {len(ply_rows)} plies, {time.time() - t:.1f}s")


# ------------------------------------------------------------------------------------------------ shaped env
def run_shaped(args):
    N, L, env_seed, n_steps, tol, act_seed = args
    env = R.MiniMaxHeuristicEnv(board_size=N, cube_layer=L, seed=env_seed, illegal_move_reward=-1.0, illegal_move_tolerance=tol)
    rec = Recorder(env)
    rng = np.random.RandomState(act_seed)  # private stream for the actions; does not touch the global one
    rows, rewards, boards = [], [], []
    init_board, init_dice, init_prev = env.board.astype(np.int8).copy(), int(env.dice_roll), float(env.prev_score)
    for t in range(n_steps):
        rec.clear()
        # legal-biased but sometimes illegal actions, so both the tolerance path and long games are exercised
        if rng.rand() < 0.85:
            la = env.get_legal_actions(env.current_player)
            a = np.array(la[rng.randint(0, len(la))])
        else:
            a = np.array([rng.randint(0, 2), rng.randint(0, 3)])
        dice_agent = int(env.dice_roll)
        obs, reward, term, trunc, info = env.step(a)
        d_opp = rec.rolls[0] if len(rec.rolls) > 0 else 0
        d_next = rec.rolls[1] if len(rec.rolls) > 1 else 0
        o = rec.opp[0] if rec.opp else (-1, -1)
        boards.append(obs["board"].astype(np.int8).copy())
        rewards.append(float(reward))
        reset_dice = 0
        if term:
            rec.clear()
            obs, _ = env.reset()  # SB3 worker behaviour: reset(seed=None)
            reset_dice = rec.rolls[0]
        rows.append((dice_agent, int(a[0]), int(a[1]), d_opp, o[0], o[1], d_next, int(term), int(trunc), reset_dice,
                     int(env.illegal_move_tolerance)))
    return (np.array(rows, np.int32), np.array(rewards, np.float64), np.array(boards, np.int8), init_board, init_dice, init_prev)


def gen_shaped(name, N, L, n_envs, n_steps, tol):
    t = time.time()
    out = pmap(run_shaped, [(N, L, 1000 + i, n_steps, tol, 77 + i) for i in range(n_envs)])
    np.savez_compressed(os.path.join(HERE, name), N=N, L=L, tolerance0=tol, goal_reward_effective=1.0,
                        rows=np.stack([o[0] for o in out]), rewards=np.stack([o[1] for o in out]).view(np.uint64),
                        boards=np.stack([o[2] for o in out]), init_board=np.stack([o[3] for o in out]),
                        init_dice=np.array([o[4] for o in out], np.int32), init_prev=np.array([o[5] for o in out]).view(np.uint64),
                        columns="dice_agent,a_larger,a_dir,dice_opp,o_larger,o_dir,dice_next,terminated,truncated,reset_dice,tolerance_after")
    print(f"{name}: {n_envs} envs x {n_steps} steps, {time.time() - t:.1f}s")


# ------------------------------------------------------------------------------------------------ MC statistics
def run_sim(args):
    N, L, board, n_play, seed = args
    import random
    random.seed(seed)
    ag = R.MctsAgent(cube_layer=L, board_size=N, num_simulations=n_play, num_env_copies=1)
    ag.restore_env_with_obs({"board": board.astype(np.int16), "dice_roll": 1})
    return ag.simulate(ag.env)


def gen_mc(name, N, L, n_states, n_play):
    """TL-win counts of MctsAgent.simulate (mcts.py:21-45) from sampled positions with BOTTOM_RIGHT to move."""
    t = time.time()
    boards, dice = get_states(N, L, n_states * 3)
    menv = R.MinimaxEnv(board_size=N, cube_layer=L)
    ag = R.ExpectiMinimaxAgent(max_depth=1, cube_layer=L, board_size=N)
    ag.env = menv
    roots = []
    for b, d in zip(boards, dice):  # one TL move from each sampled obs gives a BR-to-move position
        ag.restore_env_with_obs({"board": b.astype(np.int16), "dice_roll": int(d)})
        if menv.check_win():
            continue
        la = menv.get_legal_actions(R.Player.TOP_LEFT)
        menv.make_simulated_action(R.Player.TOP_LEFT, la[len(roots) % len(la)])
        roots.append(menv.board.astype(np.int8).copy())
        if len(roots) == n_states:
            break
    parts = 8
    jobs = [(N, L, b, n_play // parts, 1234 + i * parts + j) for i, b in enumerate(roots) for j in range(parts)]
    res = np.array(pmap(run_sim, jobs)).reshape(len(roots), parts).sum(1)
    np.savez_compressed(os.path.join(HERE, name), N=N, L=L, boards=np.array(roots, np.int8), wins=res.astype(np.int64),
                        playouts=(n_play // parts) * parts)
    print(f"{name}: {len(roots)} states x {n_play} playouts, {time.time() - t:.1f}s")


def run_simulate(args):
    """MinimaxEnv.simulate (minimax_ewn.py:215-238) called `calls` times in a row on one position with a FRESH env (current_player =
    TOP_LEFT: the first playout is opened by BOTTOM_RIGHT; within the calls the never-restored toggle runs as in the reference)."""
    N, L, board, calls, seed = args
    import random
    random.seed(seed)
    env = R.MinimaxEnv(board_size=N, cube_layer=L)
    ag = R.ExpectiMinimaxAgent(max_depth=1, cube_layer=L, board_size=N)
    ag.env = env
    ag.restore_env_with_obs({"board": board.astype(np.int16), "dice_roll": 1})
    tot = 0
    first = None
    for c in range(calls):
        v = env.simulate()
        if first is None:
            first = v
        tot += int(round(v * env.num_simulations))
    return tot, int(round(first * env.num_simulations))


def gen_sim(name, N, L, n_states, calls):
    """'sim_winrate' statistics of the unmodified reference: per sampled position the TOP_LEFT wins over calls x 100 playouts (one env,
    consecutive calls) and of the first call alone."""
    t = time.time()
    boards, dice = get_states(N, L, n_states * 2)
    menv = R.MinimaxEnv(board_size=N, cube_layer=L)
    ag = R.ExpectiMinimaxAgent(max_depth=1, cube_layer=L, board_size=N)
    ag.env = menv
    keep = []
    for b, d in zip(boards, dice):
        ag.restore_env_with_obs({"board": b.astype(np.int16), "dice_roll": int(d)})
        if not menv.check_win():
            keep.append(b)
        if len(keep) == n_states:
            break
    res = pmap(run_simulate, [(N, L, b, calls, 4321 + i) for i, b in enumerate(keep)])
    np.savez_compressed(os.path.join(HERE, name), N=N, L=L, boards=np.array(keep, np.int8), wins=np.array([r[0] for r in res], np.int64),
                        wins_first_call=np.array([r[1] for r in res], np.int64), playouts=calls * 100, num_simulations=100)
    print(f"{name}: {len(keep)} states x {calls} calls x 100 playouts, {time.time() - t:.1f}s")


def run_sim_search(args):
    N, L, depth, boards, dice, seed = args
    import random
    random.seed(seed)
    ag = R.ExpectiMinimaxAgent(max_depth=depth, cube_layer=L, board_size=N, heuristic="sim_winrate")
    vals = []
    for b, d in zip(boards, dice):
        obs = {"board": b.astype(np.int16), "dice_roll": int(d)}
        ag.restore_env_with_obs(obs)
        v, a = ag.expectiminimax(depth, ag.env.agent_player, None, -float("inf"), float("inf"))
        vals.append(float(v))
    return np.array(vals)


def gen_sim_search(name, N, L, n_states, depth, reps):
    """root values of ExpectiMinimaxAgent(heuristic='sim_winrate') at `depth`, `reps` independent runs per state (they are random)"""
    t = time.time()
    boards, dice = get_states(N, L, n_states)
    res = pmap(run_sim_search, [(N, L, depth, boards, dice, 99 + r) for r in range(reps)])
    np.savez_compressed(os.path.join(HERE, name), N=N, L=L, depth=depth, boards=boards, dice=dice, values=np.stack(res))
    print(f"{name}: {n_states} states x {reps} runs at depth {depth}, {time.time() - t:.1f}s")


def count_games(args):
    N, L, seeds = args
    env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L)
    agent = R.RandomAgent(env)
    wins = steps = 0
    for seed in seeds:
        obs, _ = env.reset(seed=seed)
        done = False
        while not done:
            a, _ = agent.predict(obs)
            obs, r, done, _, _ = env.step(a)
            steps += 1
        wins += r > 0
    return wins, steps


def gen_stats(name):
    t = time.time()
    out = {}
    for N, n in ((5, 100000), (7, 20000)):
        res = pmap(count_games, [(N, 3, s) for s in split(range(100000, 100000 + n), NPROC * 4)])
        out[f"{N}x{N}"] = dict(games=n, first_seed=100000, wins=int(sum(r[0] for r in res)), agent_steps=int(sum(r[1] for r in res)))
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(out, f, indent=1)
    print(f"{name}: {out} {time.time() - t:.1f}s")


def count_policy_games(args):
    """eval_minimax.py / eval_random.py evaluation(): agent policy (TOP_LEFT) vs the env's opponent policy, reward of the last step"""
    N, L, seeds, agent_kind, agent_kw, opp_kind, opp_kw = args
    env = R.EinsteinWuerfeltNichtEnv(board_size=N, cube_layer=L, opponent_policy=getattr(R.ClassicalPolicy, opp_kind), **opp_kw)
    agent = R.RandomAgent(env) if agent_kind == "random" else R.ExpectiMinimaxAgent(cube_layer=L, board_size=N, **agent_kw)
    wins = 0
    for seed in seeds:
        obs, _ = env.reset(seed=seed)
        done = False
        while not done:
            a, _ = agent.predict(obs)
            obs, r, done, _, _ = env.step(a)
        wins += r > 0
    return wins


def gen_pair_stats(name):
    """agent-vs-opponent win counts of the unmodified reference for the tournament driver's statistical parity test"""
    t = time.time()
    out = {}
    cases = [("minimax_d1_vs_random", "minimax", dict(max_depth=1), "random", {}, 1600),
             ("minimax_d2_vs_random", "minimax", dict(max_depth=2), "random", {}, 1200),
             ("minimax_d3_vs_random", "minimax", dict(max_depth=3), "random", {}, 800),
             ("random_vs_minimax_d1", "random", {}, "minimax", dict(max_depth=1), 1600),
             ("random_vs_minimax_d2", "random", {}, "minimax", dict(max_depth=2), 1200),
             ("minimax_d4_vs_random", "minimax", dict(max_depth=4), "random", {}, 800),
             ("minimax_d5_vs_random", "minimax", dict(max_depth=5), "random", {}, 480),
             ("random_vs_minimax_d3", "random", {}, "minimax", dict(max_depth=3), 800),
             ("minimax_d2_vs_minimax_d1", "minimax", dict(max_depth=2), "minimax", dict(max_depth=1), 800),
             ("minimax_d3_min_dist_vs_random", "minimax", dict(max_depth=3, heuristic="min_dist"), "random", {}, 800)]
    for key, ak, akw, ok, okw, n in cases:
        res = pmap(count_policy_games, [(5, 3, s, ak, akw, ok, okw) for s in split(range(200000, 200000 + n), NPROC * 4)])
        out[key] = dict(games=n, first_seed=200000, agent_wins=int(sum(res)))
        print(key, out[key], f"{time.time() - t:.1f}s", flush=True)
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(out, f, indent=1)


TARGETS = {
    "games5": lambda: gen_games("random_games_5x5.npz", 5, 3, 10000),
    "games7": lambda: gen_games("random_games_7x7.npz", 7, 3, 1000),
    "games4": lambda: gen_games("random_games_4x4_L2.npz", 4, 2, 500),
    # cube_layer >= 4 (envs/ewn.py:47,51: 10 / 15 / 28 cubes per side, die with as many faces): the generic state codec
    "games7L4": lambda: gen_games("random_games_7x7_L4.npz", 7, 4, 400),
    "games9L7": lambda: gen_games("random_games_9x9_L7.npz", 9, 7, 60),
    "rules6L4": lambda: gen_rules("rules_6x6_L4.npz", 6, 4, 300),
    "rules7L5": lambda: gen_rules("rules_7x7_L5.npz", 7, 5, 150),
    "shaped7L4": lambda: gen_shaped("shaped_7x7_L4.npz", 7, 4, 16, 96, 10),
    "sim5": lambda: gen_sim("sim_winrate_5x5.npz", 5, 3, 40, 30),
    "simsearch5": lambda: gen_sim_search("sim_search_5x5.npz", 5, 3, 24, 1, 16),
    "mm5": lambda: gen_minimax("minimax_5x5.npz", 5, 3, 2000, {1: 2000, 2: 2000, 3: 2000, 4: 2000, 5: 320}),
    "mm7": lambda: gen_minimax("minimax_7x7.npz", 7, 3, 600, {1: 600, 2: 600, 3: 600, 4: 600, 5: 96}, heur_count=100),
    "rules5": lambda: gen_rules("rules_5x5.npz", 5, 3, 4000),
    "rules7": lambda: gen_rules("rules_7x7.npz", 7, 3, 2000),
    "shaped5": lambda: gen_shaped("shaped_5x5.npz", 5, 3, 32, 128, 10),
    "shaped7": lambda: gen_shaped("shaped_7x7.npz", 7, 3, 64, 128, 10),
    "mc5": lambda: gen_mc("mc_sim_5x5.npz", 5, 3, 48, 4000),
    "mc7": lambda: gen_mc("mc_sim_7x7.npz", 7, 3, 16, 2000),
    "stats": lambda: gen_stats("random_vs_random_stats.json"),
    "pairs": lambda: gen_pair_stats("policy_pairs_stats.json"),
}

if __name__ == "__main__":
    which = sys.argv[1:] or list(TARGETS)
    for w in which:
        TARGETS[w]()
