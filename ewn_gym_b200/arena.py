"""GPU tournament driver: many games of policy-vs-policy played in lock step on packed states.

Semantics are those of the reference's evaluation loops (`eval_random.py:17-51`, `eval_minimax.py`, `eval_mcts.py`,
driven pairwise by `eval_pairs.py:3-34`): the agent is TOP_LEFT and moves first, `env.step` plays the agent's ply, rolls
the die (`envs/ewn.py:458`), asks the opponent policy for its move on the flipped board `np.rot90(-board, 2)`
(`:289-296`), plays it and rolls again (`:483`); an episode is won when `check_win` fires after the agent's ply and lost when it
fires after the opponent's.  Every policy reasons as TOP_LEFT (`minimax.py:92`, `mcts.py:49`), which is why BOTTOM_RIGHT is
served through `Engine.flip`.

Everything stays on the device: one tensor of packed states, dice drawn with `torch.randint` on the device (numpy for the
host-simulator test backend), policies are batched engine calls:

  RandomPolicy            ewn_random_action            uniform over the ordered legal list (random_policy.py:11-15)
  MinimaxPolicy(d, h)     ewn_expectiminimax           bit-exact ExpectiMinimaxAgent (minimax.py:19-93)
  MonteCarloPolicy(k)     ewn_mc_root_wins + best      flat Monte-Carlo "MCTS" (mcts.py:47-84), k playouts per root action
  UctPolicy(r, b)         ewn_uct_search               UCB1 at the root, r rounds of b playouts (ucth_temp.py's intent)

Finished games stay in the batch (their state is frozen) so that all calls keep a fixed shape.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np

from .engine import Engine, TOP_LEFT, BOTTOM_RIGHT


class BatchPolicy:
    """action(engine, state, ply) -> int32[n, 2] for packed states with TOP_LEFT to move (device array of the engine)"""
    name = "policy"

    def action(self, engine: Engine, state, ply: int):
        raise NotImplementedError


class RandomPolicy(BatchPolicy):
    def __init__(self, seed: int = 0):
        self.seed, self.name = int(seed), "random"

    def action(self, engine, state, ply):
        a = engine.random_action(state, self.seed, ply)                      # int64[n, 2]
        return engine.xp.asarray(a, "i32")


class MinimaxPolicy(BatchPolicy):
    def __init__(self, max_depth: int, heuristic: str = "hybrid", pruned: bool = True):
        self.max_depth, self.heuristic, self.pruned = int(max_depth), heuristic, bool(pruned)
        self.name = f"minimax(d={max_depth},{heuristic})"

    def action(self, engine, state, ply):
        _, a = engine.expectiminimax(state, self.max_depth, self.heuristic, self.pruned, seed=ply)   # int8[n,2]; (-1,-1) on finished roots
        return engine.xp.asarray(a, "i32")


class MonteCarloPolicy(BatchPolicy):
    """num_simulations * num_env_copies playouts per root action, first arg-max (mcts.py:47-69)"""

    def __init__(self, num_simulations: int = 10, num_env_copies: int = 5, seed: int = 0, exchange=None):
        """exchange: a peer.PeerExchange sized for the batch -> every rank of the node plays 1/world of each root action's playouts
        and the counters are merged by one kernel over NVLink peer memory (all ranks hold the same games and get the same moves)"""
        self.playouts, self.seed, self.exchange = int(num_simulations) * int(num_env_copies), int(seed), exchange
        self.name = f"mcts({self.playouts})"

    def action(self, engine, state, ply):
        n = int(state.shape[0])
        rid = engine.xp.asarray((np.arange(n, dtype=np.int64) + ply * n).astype(np.uint32), "u32")   # fresh playout streams per ply
        if self.exchange is not None:
            from .sharding import mc_predict_peers
            a, _ = mc_predict_peers(engine, self.exchange, state, self.seed, self.playouts, root_ids=rid)
            return engine.xp.asarray(a, "i32")
        a, _, _ = engine.mc_predict(state, self.seed, 0, self.playouts, root_ids=rid)
        return engine.xp.asarray(a, "i32")


class UctPolicy(BatchPolicy):
    """UCB1 over the root's legal moves with batched random playouts (Engine.uct_search): `rounds` x `batch` playouts per decision,
    spent where the upper confidence bound is best instead of evenly as MonteCarloPolicy does"""

    def __init__(self, rounds: int = 16, batch: int = 32, c: float = 1.41, seed: int = 0):
        self.rounds, self.batch, self.c, self.seed = int(rounds), int(batch), float(c), int(seed)
        self.name = f"uct({rounds}x{batch})"

    def action(self, engine, state, ply):
        n = int(state.shape[0])
        rid = engine.xp.asarray((np.arange(n, dtype=np.int64) + ply * n).astype(np.uint32), "u32")     # fresh playout streams per ply
        a, _, _ = engine.uct_search(state, self.seed, self.rounds, self.batch, self.c, root_ids=rid)
        return engine.xp.asarray(a, "i32")


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class _Ops:
    """the few elementwise operations the driver needs, on torch (device) or numpy (host-simulator tests) arrays"""

    def __init__(self, engine: Engine, seed: int):
        self.e = engine
        self.torch = getattr(engine.xp, "torch", None)
        if self.torch is not None:
            self.gen = self.torch.Generator(device=engine.xp.device)
            self.gen.manual_seed(int(seed))
        else:
            self.rng = np.random.default_rng(int(seed))

    def dice(self, n, faces):
        if self.torch is not None:
            return self.torch.randint(1, faces + 1, (n,), generator=self.gen, device=self.e.xp.device, dtype=self.torch.int32)
        return self.rng.integers(1, faces + 1, n).astype(np.int32)

    def zeros(self, n, dt):
        return self.e.xp.zeros((n,), dt)

    def where_rows(self, mask, a, b):
        """rows of `a` where mask else rows of `b` (mask: bool[n]; a, b: [n] or [n, k])"""
        m = mask if a.ndim == 1 else mask[:, None]
        return self.torch.where(m, a, b) if self.torch is not None else np.where(m, a, b)

    def to_bool(self, x):
        return x.bool() if self.torch is not None else x.astype(bool)

    def host(self, x):
        return x.detach().cpu().numpy() if self.torch is not None else np.asarray(x)


def play_matches(engine: Engine, agent: BatchPolicy, opponent: BatchPolicy, n_games: int, seed: int = 0, max_plies: int = 400):
    """n_games of agent (TOP_LEFT, moves first) against opponent.  -> dict(agent_wins, opponent_wins, unfinished, plies,
    winner int8[n] (1 agent, 2 opponent, 0 unfinished))"""
    e, ops = engine, _Ops(engine, seed)
    n = int(n_games)
    state = e.reset_dice(ops.dice(n, e.cube_num))                              # setup_game + first roll (envs/ewn.py:94-108)
    winner = ops.zeros(n, "i8")
    done = ops.to_bool(ops.zeros(n, "u8"))
    plies = 0
    for ply in range(max_plies):
        side = TOP_LEFT if ply % 2 == 0 else BOTTOM_RIGHT
        pol = agent if side == TOP_LEFT else opponent
        view = state if side == TOP_LEFT else e.flip(state)                     # policies reason as TOP_LEFT (:289-296)
        act = pol.action(e, view, ply)
        before = state.clone() if ops.torch is not None else state.copy()
        valid, win, _ = e.ply(state, act, side)                                 # find_cube_to_move + execute_move + check_win
        won = ops.to_bool(win) & ops.to_bool(valid) & ~done
        # an invalid move (never produced by these policies on a live game) would end the episode against the mover (:445-448, :470-473)
        state = ops.where_rows(done, before, state)
        if ops.torch is not None:
            winner = ops.torch.where(won, ops.torch.full_like(winner, 1 if side == TOP_LEFT else 2), winner)
        else:
            winner = np.where(won, np.int8(1 if side == TOP_LEFT else 2), winner)
        done = done | won
        plies += 1
        e.set_dice(state, ops.dice(n, e.cube_num), None)                         # roll_dice for the next mover (:458, :483)
        if plies % 4 == 0 and bool(done.all()):                                  # the loop's one synchronisation point
            _check_exchanges(agent, opponent)
            break
    w = ops.host(winner)
    _check_exchanges(agent, opponent)
    return {"agent_wins": int((w == 1).sum()), "opponent_wins": int((w == 2).sum()), "unfinished": int((w == 0).sum()),
            "plies": plies, "winner": w, "n": n}


def _check_exchanges(*policies):
    """a peer-memory merge whose wait timed out leaves (-1, -1) actions and an error word behind: raise instead of playing on"""
    for p in policies:
        ex = getattr(p, "exchange", None)
        if ex is not None:
            ex.check()


class NumpyRandom(BatchPolicy):
    """marker policy for play_matches_numpy: the reference's RandomAgent drawing from the game's own numpy stream"""
    name = "random(numpy stream)"


def play_matches_numpy(engine: Engine, agent: BatchPolicy, opponent: BatchPolicy, seeds, max_plies: int = 400):
    """The reference's evaluation loop (eval_random.py / eval_minimax.py:17-51) for the given seeds with the reference's OWN random
    stream: game i seeds numpy's global MT19937 with seeds[i] in env.reset (envs/ewn.py:488-494) and then draws, in this order, the
    first die (:108), per step the agent's pick if it is a RandomAgent, the die at :458, the opponent's pick if it is a RandomAgent,
    the die at :483.  A deterministic policy (MinimaxPolicy) on the other seat draws nothing, so whole games -- and therefore win
    counts -- are reproduced exactly.  `agent` / `opponent`: NumpyRandom() or any deterministic BatchPolicy."""
    e, ops = engine, _Ops(engine, 0)
    sd = np.asarray(seeds)
    n = len(sd)
    raw, cur = e.numpy_stream(sd.astype(np.uint32))
    state = e.reset_dice(np.ones(n, np.int32))
    e.numpy_draw(state, raw, cur, roll_dice=True)                                # first roll of env.reset
    winner = ops.zeros(n, "i8")
    done = ops.to_bool(ops.zeros(n, "u8"))
    steps = ops.zeros(n, "i32")
    plies = 0
    for ply in range(max_plies):
        side = TOP_LEFT if ply % 2 == 0 else BOTTOM_RIGHT
        pol = agent if side == TOP_LEFT else opponent
        if isinstance(pol, NumpyRandom):
            act = e.numpy_draw(state, raw, cur, roll_dice=False, pick_player=side)     # RandomAgent reads the env directly
        else:
            act = pol.action(e, state if side == TOP_LEFT else e.flip(state), ply)
        before = state.clone() if ops.torch is not None else state.copy()
        valid, win, _ = e.ply(state, act, side)
        won = ops.to_bool(win) & ops.to_bool(valid) & ~done
        state = ops.where_rows(done, before, state)
        if side == TOP_LEFT:
            steps = steps + (~done).to(steps.dtype) if ops.torch is not None else steps + (~done).astype(steps.dtype)
        if ops.torch is not None:
            winner = ops.torch.where(won, ops.torch.full_like(winner, 1 if side == TOP_LEFT else 2), winner)
        else:
            winner = np.where(won, np.int8(1 if side == TOP_LEFT else 2), winner)
        done = done | won
        plies += 1
        e.numpy_draw(state, raw, cur, roll_dice=True)                            # :458 / :483 (finished games: irrelevant draws)
        if plies % 4 == 0 and bool(done.all()):
            break
    w = ops.host(winner)
    return {"agent_wins": int((w == 1).sum()), "opponent_wins": int((w == 2).sum()), "unfinished": int((w == 0).sum()),
            "plies": plies, "winner": w, "agent_steps": ops.host(steps), "n": n}


def wilson(k: int, n: int, z: float = 1.959964):
    """Wilson score interval (statsmodels proportion_confint(method='wilson') of the eval_*.py scripts)"""
    if n == 0:
        return 0.0, 1.0
    p = k / n
    d = 1 + z * z / n
    c = (p + z * z / (2 * n)) / d
    h = z * math.sqrt(p * (1 - p) / n + z * z / (4 * n * n)) / d
    return c - h, c + h


def tournament(engine: Engine, policies: dict, n_games: int = 1024, seed: int = 0):
    """eval_pairs.py:3-34: every policy against every policy (rows = agent = TOP_LEFT).  -> {(agent, opponent): result}"""
    out = {}
    for i, (an, ap) in enumerate(policies.items()):
        for j, (on, op) in enumerate(policies.items()):
            r = play_matches(engine, ap, op, n_games, seed=seed + 1000 * i + j)
            r["win_rate"] = r["agent_wins"] / max(1, r["n"])
            r["ci95"] = wilson(r["agent_wins"], r["n"])
            out[(an, on)] = r
    return out
