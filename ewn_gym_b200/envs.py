"""The reference's single-environment gym surface as a batch-of-one view on the GPU engine.

  EinsteinWuerfeltNichtEnv   envs/ewn.py:18-494          (alias EinsteinWurfeltNichtEnv, the spelling BASELINE.json uses)
  MinimaxEnv                 envs/minimax_ewn.py:11-238
  MiniMaxHeuristicEnv        envs/training_ewn.py:12-99

Same constructor keywords, attributes (board, cube_pos, dice_roll, current_player, agent_player, cube_num, cube_layer,
reward, action_space, observation_space, history, opponent_policy) and public helper methods as the reference.  Every
rule decision is made by the engine kernels through the C ABI (ewn_ply / ewn_legal / ewn_check_win / ewn_set_dice /
ewn_evaluate / ewn_playouts on a batch of one game): this module holds no rules logic, it only mirrors the device
state into `self.board`.  Like the reference, dice and the random agent draw from the process-global numpy RNG in the
reference's order, so `reset(seed)` + the same actions reproduce the reference's games exactly.

A batch of one is latency-bound (a few kernel launches per ply); the fast surfaces are EWNVecEnv and Engine.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .constants import ClassicalPolicy, Player
from .engine import Engine, MESSAGES, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_OPP_ILLEGAL, CODE_AGENT_LOST
from .spaces import Box, Dict, Discrete, MultiDiscrete


def _np(x):
    return x if isinstance(x, np.ndarray) else x.detach().cpu().numpy()


class EinsteinWuerfeltNichtEnv:
    metadata = {"render_modes": ["human", "rgb_array", "ansi"], "render_fps": 30}

    def __init__(self, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.,
                 agent_player: Player = Player.TOP_LEFT, render_mode: Optional[str] = None,
                 opponent_policy=ClassicalPolicy.random, **policy_kwargs):
        assert cube_layer < board_size - 1                                    # envs/ewn.py:47
        self._engine: Engine = policy_kwargs.pop("engine", None) or Engine(board_size, cube_layer)
        self.board_size = int(board_size)
        self.board: np.ndarray = np.zeros((board_size, board_size), dtype=np.int16)
        self.cube_num: int = cube_layer * (cube_layer + 1) // 2
        self.cube_layer: int = cube_layer
        self.dice_roll: int = 1
        self.action_space = MultiDiscrete([2, 3])
        self.observation_space = Dict({
            "board": Box(low=-self.cube_num, high=self.cube_num, shape=(board_size, board_size), dtype=np.int16),
            "dice_roll": Discrete(self.cube_num + 1, start=1),                # the +1 is the reference's SB3 work-around (:66-68)
        })
        self.current_player = Player.TOP_LEFT
        self.agent_player: Player = agent_player
        self.reward = reward
        assert opponent_policy is not None
        self._state = None
        self.history: list = []
        self.load_opponent_policy(opponent_policy, **policy_kwargs)
        self.reset(seed=seed)
        self.render_mode = render_mode
        self.history = []

    # ------------------------------------------------------------------ device state <-> mirror
    def _pull(self):
        """refresh self.board (in place: observations alias it, envs/ewn.py:485) from the device state"""
        b, _ = self._engine.decode(self._state)
        self.board[...] = _np(b)[0]

    def _push_turn(self, state=None):
        """write dice_roll / current_player into the packed state"""
        st = self._state if state is None else state
        self._engine.set_dice(st, [int(self.dice_roll)], [self.current_player.value])
        return st

    def _clone(self):
        return self._state.copy() if isinstance(self._state, np.ndarray) else self._state.clone()

    @property
    def cube_pos(self):
        """the reference's masked piece list (envs/ewn.py:53-58): index k-1 <-> cube +k, index -k <-> cube -k"""
        n = self.cube_num
        pos = np.ma.zeros((2 * n,), dtype=[("x", int), ("y", int)])
        pos[:] = np.ma.masked
        for x, y in zip(*np.nonzero(self.board)):
            v = int(self.board[x, y])
            pos[v - 1 if v > 0 else v] = (x, y)
        return pos

    def _obs(self):
        return {"board": self.board, "dice_roll": self.dice_roll}

    # ------------------------------------------------------------------ envs/ewn.py:90-142
    def roll_dice(self):
        self.dice_roll = int(np.random.randint(1, self.cube_num + 1))        # same global stream as the reference (:91)

    def setup_game(self):
        self.roll_dice()
        self._state = self._engine.reset_dice([self.dice_roll])
        self._pull()
        if self.agent_player == Player.BOTTOM_RIGHT:                          # :111-129 the opponent opens
            self.current_player = Player.TOP_LEFT
            a = self.opponent_action()
            valid, win = self._ply(a, Player.TOP_LEFT)
            if win:
                return self._obs(), -self.reward, True, {"message": "You lost!"}
            self.switch_player()
            self.roll_dice()

    def check_win(self) -> bool:
        return bool(_np(self._engine.check_win(self._state))[0])

    def switch_player(self):
        self.current_player = Player.get_opponent(self.current_player)

    # ------------------------------------------------------------------ one ply on the device
    def _ply(self, action, player: Optional[Player] = None, state=None, pull=True):
        """find_cube_to_move + execute_move + check_win (envs/ewn.py:439-451) -> (valid, win)"""
        player = self.current_player if player is None else player
        st = self._state if state is None else state
        self._engine.set_dice(st, [int(self.dice_roll)], [player.value])
        a = [[int(action[0] == 1), int(action[1]) if 0 <= int(action[1]) <= 2 else 3]]
        valid, win, _ = self._engine.ply(st, a, player.value)
        if pull and state is None:
            self._pull()
        return bool(_np(valid)[0]), bool(_np(win)[0])

    # ------------------------------------------------------------------ envs/ewn.py:144-263 (cube selection, move)
    def _index_of(self, cube_number: int, player: Player) -> int:
        return cube_number - 1 if player == Player.TOP_LEFT else -cube_number

    def _number_of(self, cube_pos_index: int):
        """piece-list index -> (cube number 1..cube_num, owner)"""
        if cube_pos_index >= 0:
            if cube_pos_index < self.cube_num:
                return cube_pos_index + 1, Player.TOP_LEFT
            return 2 * self.cube_num - cube_pos_index, Player.BOTTOM_RIGHT    # positive spelling of a BR index
        return -cube_pos_index, Player.BOTTOM_RIGHT

    def find_near_cube(self, cube_pos_index: int, chose_larger: bool, player: Optional[Player] = None):
        """nearest alive larger / smaller own cube of `cube_pos_index`, or None (envs/ewn.py:144-176)"""
        if player is None:
            player = self.current_player
        k, _ = self._number_of(cube_pos_index)
        b = self.board.copy()
        b[b == (k if player == Player.TOP_LEFT else -k)] = 0                 # the scan starts next to k, k itself is not a candidate
        st = self._engine.encode(b[None], [k], [player.value])
        _, lst, cnt = self._engine.legal(st, player.value)
        lst, cnt = _np(lst)[0], int(_np(cnt)[0])
        flags = {int(lst[j, 0]) for j in range(cnt)}
        # with cube k gone the list is [1,d].. for the larger neighbour then [0,d].. for the smaller one (:338-375)
        if (1 if chose_larger else 0) not in flags:
            return None
        _, _, moved = self._engine.ply(st, [[int(chose_larger), 0]], player.value)
        return self._index_of(int(_np(moved)[0]), player)

    def find_cube_to_move(self, chose_larger: bool, player: Optional[Player] = None) -> int:
        """the dice cube if alive, else the requested neighbour, else the other one (envs/ewn.py:178-215)"""
        if player is None:
            player = self.current_player
        st = self._push_turn(self._clone())
        self._engine.set_dice(st, [int(self.dice_roll)], [player.value])
        _, _, moved = self._engine.ply(st, [[int(bool(chose_larger)), 0]], player.value)
        m = int(_np(moved)[0])
        assert m > 0                                                          # :214
        return self._index_of(m, player)

    def update_position(self, x: int, y: int, direction: int, cube: int):
        """envs/ewn.py:301-319 (pure index arithmetic, kept for API compatibility)"""
        s = 1 if cube > 0 else -1
        if direction == 0:
            y += s
        elif direction == 1:
            x += s
        elif direction == 2:
            x += s; y += s
        return x, y

    def is_within_board(self, x: int, y: int):
        return 0 <= x < self.board.shape[0] and 0 <= y < self.board.shape[1]

    def execute_move(self, cube_index: int, action, dry_run: bool = False) -> bool:
        """move one given cube in direction `action`; False and no change when it leaves the board (envs/ewn.py:217-263)"""
        k, owner = self._number_of(cube_index)
        assert (self.board == (k if owner == Player.TOP_LEFT else -k)).any()  # :227 the cube must exist
        st = self._clone() if dry_run else self._state
        self._engine.set_dice(st, [k], [owner.value])                         # cube k alive => it is the one that moves
        valid, _, _ = self._engine.ply(st, [[0, int(action)]], owner.value)
        if not dry_run:
            self._push_turn()
            self._pull()
        return bool(_np(valid)[0])

    def get_cube_legal_directions(self, cube_pos_index: int):
        k, owner = self._number_of(cube_pos_index)
        st = self._clone()
        self._engine.set_dice(st, [k], [owner.value])
        _, lst, cnt = self._engine.legal(st, owner.value)
        lst, cnt = _np(lst)[0], int(_np(cnt)[0])
        return [int(lst[j, 1]) for j in range(cnt)]

    def get_legal_actions(self, player: Player) -> List[List[int]]:
        """ordered [larger, direction] list (envs/ewn.py:338-375) from kernel k_legal"""
        self._engine.set_dice(self._state, [int(self.dice_roll)], None)
        _, lst, cnt = self._engine.legal(self._state, player.value)
        lst, cnt = _np(lst)[0], int(_np(cnt)[0])
        return [[int(lst[j, 0]), int(lst[j, 1])] for j in range(cnt)]

    # ------------------------------------------------------------------ envs/ewn.py:377-434
    def make_simulated_action(self, player: Player, action):
        self.history.append((self._clone(), self.dice_roll))
        self._ply(action, player)

    def undo_simulated_action(self):
        if not self.history:
            return
        self._state, _ = self.history.pop()
        self._pull()

    # ------------------------------------------------------------------ opponent (envs/ewn.py:265-296)
    def load_opponent_policy(self, opponent_policy, **policy_kwargs):
        from . import policies as P
        if isinstance(opponent_policy, str):
            opponent_policy = ClassicalPolicy.from_string(opponent_policy)
        kw = dict(policy_kwargs)
        if opponent_policy == ClassicalPolicy.random:
            self.opponent_policy = P.RandomAgent(self)
        elif opponent_policy == ClassicalPolicy.minimax:
            self.opponent_policy = P.ExpectiMinimaxAgent(cube_layer=self.cube_layer, board_size=self.board.shape[0],
                                                         engine=self._engine, **kw)
        elif opponent_policy == ClassicalPolicy.mcts:
            self.opponent_policy = P.MctsAgent(cube_layer=self.cube_layer, board_size=self.board.shape[0], engine=self._engine, **kw)
        elif opponent_policy == ClassicalPolicy.alpha_zero:
            raise NotImplementedError("AlphaZero opponents need the reference's checkpoints (out of scope, DESIGN.md 9)")
        else:
            assert isinstance(opponent_policy, str)
            try:
                from stable_baselines3 import A2C
            except ImportError as e:
                raise ImportError("loading an SB3 checkpoint as opponent needs stable_baselines3") from e
            self.opponent_policy = A2C.load(opponent_policy)

    def opponent_action(self):
        # policies reason as TOP_LEFT: hand them the board rotated by 180 degrees and negated (envs/ewn.py:289-296)
        action, _ = self.opponent_policy.predict({"board": np.rot90(-self.board, 2), "dice_roll": self.dice_roll}, deterministic=True)
        return action

    # ------------------------------------------------------------------ envs/ewn.py:436-494
    def _agent_invalid(self):
        return self._obs(), -self.reward, True, True, {"message": MESSAGES[CODE_AGENT_ILLEGAL]}

    def step(self, action):
        valid, win = self._ply(action)
        if not valid:
            return self._agent_invalid()
        if win:
            return self._obs(), self.reward, True, False, {"message": MESSAGES[CODE_AGENT_WON]}
        self.switch_player()
        self.roll_dice()
        opp = self.opponent_action()
        valid, win = self._ply(opp)
        if not valid:
            return self._obs(), 0, True, True, {"message": MESSAGES[CODE_OPP_ILLEGAL]}
        if win:
            return self._obs(), -self.reward, True, False, {"message": MESSAGES[CODE_AGENT_LOST]}
        self.switch_player()
        self.roll_dice()
        return self._step_tail()

    def _step_tail(self):
        return self._obs(), 0, False, False, {}

    def reset(self, seed: Optional[int] = None):
        self.current_player = Player.TOP_LEFT
        np.random.seed(seed)
        self.action_space.seed(seed)
        self.setup_game()
        return self._obs(), {}

    def render(self):
        if self.render_mode == "ansi":
            return "\n".join(" ".join(f"{int(v):3d}" for v in row) for row in self.board) + f"\ndice {self.dice_roll}\n"
        return None          # pygame drawing (envs/ewn.py:496-569) is out of scope: no graphics stack on this box

    def close(self):
        pass


EinsteinWurfeltNichtEnv = EinsteinWuerfeltNichtEnv


class MinimaxEnv(EinsteinWuerfeltNichtEnv):
    """Search simulator (envs/minimax_ewn.py:11-238): set_dice_roll, evaluate(heuristic), simulate."""

    def __init__(self, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.,
                 agent_player: Player = Player.TOP_LEFT, render_mode: Optional[str] = None,
                 opponent_policy=ClassicalPolicy.random, **policy_kwargs):
        # the reference forwards board_size and cube_layer ONLY (minimax_ewn.py:17-19): seed 9487, reward 1.0, TOP_LEFT agent and
        # a random opponent whatever the caller asked for
        kw = {"engine": policy_kwargs["engine"]} if "engine" in policy_kwargs else {}
        super().__init__(board_size=board_size, cube_layer=cube_layer, **kw)
        self.num_simulations = 100
        self._sim_calls = 0

    def set_dice_roll(self, roll: int):
        self.dice_roll = roll

    def evaluate(self, heuristic="hybrid") -> float:
        if heuristic == "sim_winrate":
            return self.simulate()
        h = heuristic if heuristic in ("min_dist", "two_min_dist", "attk") else "hybrid"
        return float(_np(self._engine.evaluate(self._state, h))[0])

    def evaluate_min_dist(self):
        return self.evaluate("min_dist")

    def evaluate_two_min_dist(self):
        return self.evaluate("two_min_dist")

    def evaluate_attack(self):
        return self.evaluate("attk")

    def simulate(self) -> float:
        """num_simulations random playouts from the current position, fraction TOP_LEFT wins (minimax_ewn.py:215-238), on the engine's
        'sim_winrate' kernels: the opener of every playout follows the reference's never-restored current_player toggle within the
        call (Engine.sim_winrate); the call's first playout is opened by the opponent of current_player, as in the reference.
        (The reference also leaves current_player toggled by the number of plies it played; this class leaves it alone.)"""
        n = self.num_simulations
        if self.check_win():                                           # the playout loop never runs: win_count / n is 1.0 or 0.0 (:231-238)
            return 1.0 if (self.board[-1, -1] > 0 or not np.any(self.board < 0)) else 0.0
        first = Player.get_opponent(self.current_player)
        st = self._clone()
        if first == Player.TOP_LEFT:
            # the kernels open a position's first playout with BOTTOM_RIGHT: hand them the mirrored position (envs/ewn.py:289-296) and
            # read the result from the other side: TOP_LEFT wins = n - mirrored TOP_LEFT wins
            st = self._engine.encode(np.rot90(-self.board, 2)[None].copy(), [1], [Player.BOTTOM_RIGHT.value])
        v = float(_np(self._engine.sim_winrate(st, n, seed=int(np.random.randint(0, 2 ** 31 - 1)), unit0=self._sim_calls))[0])
        self._sim_calls += 1
        return v if first == Player.BOTTOM_RIGHT else (n - round(v * n)) / n


class MiniMaxHeuristicEnv(MinimaxEnv):
    """Reward-shaping env train.py registers as EWN-v0 (envs/training_ewn.py:12-99): an illegal move costs
    illegal_move_reward until the lifetime tolerance budget is spent; otherwise reward = evaluate() - previous evaluate()."""

    def __init__(self, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, goal_reward: float = 10.,
                 agent_player: Player = Player.TOP_LEFT, render_mode: Optional[str] = None,
                 opponent_policy=ClassicalPolicy.random, illegal_move_reward: float = -1.0, illegal_move_tolerance: int = 10,
                 **policy_kwargs):
        super().__init__(board_size=board_size, cube_layer=cube_layer, seed=seed, reward=goal_reward, agent_player=agent_player,
                         render_mode=render_mode, opponent_policy=opponent_policy, **policy_kwargs)
        self.prev_score: float = self.evaluate()                              # never reset afterwards (training_ewn.py:35-38)
        self.illegal_move_reward: float = illegal_move_reward
        self.illegal_move_tolerance: int = illegal_move_tolerance

    def _agent_invalid(self):                                                  # training_ewn.py:48-56
        self.illegal_move_tolerance -= 1
        if self.illegal_move_tolerance <= 0:
            return super()._agent_invalid()
        return self._obs(), self.illegal_move_reward, False, False, {
            "message": f"Invalid move for player! Tolerance left {self.illegal_move_tolerance}."}

    def _step_tail(self):                                                      # training_ewn.py:93-99
        self._push_turn()
        score = self.evaluate()
        reward = score - self.prev_score
        self.prev_score = score
        return self._obs(), reward, False, False, {}
