"""The classical_policies surface of the reference on top of the GPU engine.

  PolicyBase.predict(obs, **kwargs) -> (action, None)                       classical_policies/base.py:6-9
  RandomAgent(env)                                                          classical_policies/random_policy.py:6-15
  ExpectiMinimaxAgent(max_depth, cube_layer, board_size, heuristic, **kw)   classical_policies/minimax.py:9-93
  MctsAgent(cube_layer, board_size, num_simulations, num_env_copies, **kw)  classical_policies/mcts.py:10-106

``get_action`` is an alias of ``predict`` (BASELINE.json's north_star uses that name; the reference has only predict).
Every agent also has ``predict_batch`` taking ``board int16[n,N,N]`` and ``dice_roll [n]`` -- the form in which the
engine is actually fast.  Agents reason as TOP_LEFT on the observation they are given (minimax.py:92, mcts.py:49).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .engine import Engine


class PolicyBase:
    def predict(self, obs, **kwargs) -> Tuple[Optional[np.ndarray], None]:
        raise NotImplementedError

    def get_action(self, obs, **kwargs):
        return self.predict(obs, **kwargs)


class RandomAgent(PolicyBase):
    """Uniform over ``env.get_legal_actions(env.current_player)`` drawn from the process-global numpy RNG, like the
    reference: it shares one MT19937 stream with the env's dice, the observation is ignored."""

    def __init__(self, env):
        self.env = env

    def predict(self, obs=None, **kwargs):
        legal_moves = self.env.get_legal_actions(self.env.current_player)
        action_idx = np.random.randint(0, len(legal_moves))
        return np.array(legal_moves[action_idx]), None


def _obs_arrays(obs, N):
    board = np.asarray(obs["board"], dtype=np.int16).reshape(-1, N, N)
    dice = np.asarray(obs["dice_roll"], dtype=np.int32).reshape(-1)
    return board, dice


class ExpectiMinimaxAgent(PolicyBase):
    """Expectiminimax with the reference's alpha/beta hand-through at chance nodes, bit-exact values
    (kernels k_emm_expand / k_emm_replay)."""

    def __init__(self, max_depth: int, cube_layer: int, board_size: int, heuristic: str = "hybrid", engine: Engine = None,
                 pruned: bool = True, **kwargs):
        self.max_depth, self.heuristic, self.pruned = int(max_depth), heuristic, bool(pruned)
        self.engine = engine if engine is not None else Engine(board_size, cube_layer)
        self.board_size = board_size

    def predict_batch(self, board, dice_roll):
        """-> (actions int8[n,2] with (-1,-1) for finished positions, values float64[n]) as host arrays"""
        e = self.engine
        st = e.encode(board, dice_roll)
        self._calls = getattr(self, "_calls", 0) + 1          # 'sim_winrate': fresh playout streams for every call
        v, a = e.expectiminimax(st, self.max_depth, self.heuristic, self.pruned, seed=self._calls)
        return _host(a), _host(v)

    def expectiminimax_value(self, obs):
        board, dice = _obs_arrays(obs, self.board_size)
        a, v = self.predict_batch(board, dice)
        return float(v[0]), (None if a[0, 0] < 0 else [int(a[0, 0]), int(a[0, 1])])

    def predict(self, obs, **kwargs):
        _, action = self.expectiminimax_value(obs)
        return action, None                      # a python list, like minimax.py:39,93 (None when the game is over)


class MctsAgent(PolicyBase):
    """Flat Monte-Carlo over the root's legal actions (the reference's "MCTS"): num_simulations * num_env_copies random
    playouts per action, first arg-max of the TOP_LEFT win counts (mcts.py:47-69).  Playouts are keyed
    (seed, call number, action, playout index) in Philox, so a prediction is reproducible and can be sharded."""

    def __init__(self, cube_layer: int, board_size: int, num_simulations: int = 10, num_env_copies: int = 5, engine: Engine = None,
                 seed: int = 0, **kwargs):
        self.num_simulations, self.num_env_copies = int(num_simulations), int(num_env_copies)
        self.engine = engine if engine is not None else Engine(board_size, cube_layer)
        self.board_size, self.seed, self._calls = board_size, int(seed), 0

    @property
    def playouts_per_action(self):
        return self.num_simulations * self.num_env_copies

    def win_counts(self, board, dice_roll, first=0, count=None):
        e = self.engine
        n = len(board)
        st = e.encode(board, dice_roll)
        rid = (np.arange(n, dtype=np.int64) + self._calls).astype(np.uint32)
        self._calls += n
        wins, nch = e.mc_root_wins(st, self.seed, first, self.playouts_per_action if count is None else count, root_ids=rid)
        return st, wins, nch

    def predict_batch(self, board, dice_roll):
        st, wins, _ = self.win_counts(board, dice_roll)
        return _host(self.engine.mc_best_action(st, wins)), _host(wins)

    def predict(self, obs, **kwargs):
        board, dice = _obs_arrays(obs, self.board_size)
        a, _ = self.predict_batch(board, dice)
        return (None if a[0, 0] < 0 else np.array([int(a[0, 0]), int(a[0, 1])])), None


class UctAgent(PolicyBase):
    """ClassicalPolicy.uct (constants/policy.py:8; the reference's ucth_temp.py:6-89 never got past a sketch): UCB1 over the root's legal
    moves, `iterations` random playouts per decision in rounds of `batch`, final move = best win rate."""

    def __init__(self, cube_layer: int, board_size: int, iterations: int = 512, batch: int = 32, exploration_constant: float = 1.41,
                 engine: Engine = None, seed: int = 0, **kwargs):
        self.engine = engine if engine is not None else Engine(board_size, cube_layer)
        self.board_size, self.seed, self._calls = board_size, int(seed), 0
        self.batch, self.rounds, self.c = int(batch), max(1, int(iterations) // int(batch)), float(exploration_constant)

    def predict_batch(self, board, dice_roll):
        e = self.engine
        st = e.encode(board, dice_roll)
        n = len(board)
        rid = (np.arange(n, dtype=np.int64) + self._calls).astype(np.uint32)
        self._calls += n
        a, w, v = e.uct_search(st, self.seed, self.rounds, self.batch, self.c, root_ids=rid)
        return _host(a), (_host(w), _host(v))

    def predict(self, obs, **kwargs):
        board, dice = _obs_arrays(obs, self.board_size)
        a, _ = self.predict_batch(board, dice)
        return (None if a[0, 0] < 0 else np.array([int(a[0, 0]), int(a[0, 1])])), None


def _host(x):
    return x if isinstance(x, np.ndarray) else x.detach().cpu().numpy()
