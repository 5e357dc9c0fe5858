"""EWNDeviceVecEnv -- the VecEnv contract with everything resident in HBM ("zero-copy torch obs").

`EWNVecEnv` (vec_env.py) serves numpy users: actions come from host memory and observations land in host memory, so its
throughput is the PCIe link's.  A policy that lives on the same GPU -- an SB3 `MultiInputPolicy` moved to CUDA,
the arena's batched policies -- needs none of that traffic: this class hands out the CUDA tensors the step kernel writes and
takes the action tensor the policy produced.  Same semantics, same Philox schedule: for equal `(seed, env_id0)` and equal actions
the two classes produce identical observations, rewards, dones and terminal observations step by step
(`tests/test_vec_env.py::test_device_env_equals_host_env`).

    env = EWNDeviceVecEnv(1 << 20)
    obs = env.reset()                              # {"board": int16[n,N,N] cuda, "dice_roll": int64[n] cuda}
    obs, rew, done, infos = env.step(actions)      # actions: int64[n,2] cuda tensor (or anything torch.as_tensor takes)

Contract details (envs/ewn.py:436-494, SB3 VecEnv):
  * finished games are reset inside the same call; `obs` then holds the new episode's first observation and the terminal
    observation is appended, compacted, to `infos.terminal_rows()` -> (env index int32[k], board int16[k,N,N], dice int64[k]);
  * `infos[i]` builds the SB3 dict (`terminal_observation`, `TimeLimit.truncated`, `message`) on demand (one device read);
  * returned tensors are owned by the env and overwritten by the next `step`;
  * `shaped=True` = `MiniMaxHeuristicEnv` (envs/training_ewn.py:40-99), `stats` = int64[4] {wins, losses, illegal endings, episodes}.
One kernel launch per step, enqueued on torch's current stream; nothing synchronises unless the caller reads a value.
"""
from __future__ import annotations

from collections.abc import Sequence

import numpy as np

from .engine import Engine, MESSAGES, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_AGENT_LOST
from .spaces import Box, Dict, Discrete, MultiDiscrete


class DeviceInfos(Sequence):
    """`infos` of a device step.  Stays on the GPU until somebody asks for an entry."""

    def __init__(self, env, count, index, board, dice, done, trunc, reward):
        self._env, self._count, self._index, self._board, self._dice = env, count, index, board, dice
        self._done, self._trunc, self._reward = done, trunc, reward
        self._host = None

    def __len__(self):
        return self._env.num_envs

    def terminal_count(self) -> int:
        """number of games that ended in this step (synchronises)"""
        return int(self._count.item()) if self._count is not None else int(self._done.sum().item())

    def terminal_rows(self):
        """(env index int32[k], board int16[k,N,N], dice int64[k]) of the games that ended, in no particular order"""
        if self._count is None:
            raise RuntimeError("the env was created with terminal_observation=False")
        k = self.terminal_count()
        return self._index[:k], self._board[:k], self._dice[:k]

    def _pull(self):
        if self._host is None:
            done, trunc, rew = (x.cpu().numpy() for x in (self._done, self._trunc, self._reward))
            rows = {}
            if self._count is not None:
                idx, b, d = (x.cpu().numpy() for x in self.terminal_rows())
                rows = {int(i): k for k, i in enumerate(idx)}
                self._tb, self._td = b, d
            self._host = (done, trunc, rew, rows)
        return self._host

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        done, trunc, rew, rows = self._pull()
        if not done[i]:
            return {}
        r = float(rew[i])
        if trunc[i]:
            code = CODE_AGENT_ILLEGAL if r < 0 else 3
        else:
            code = CODE_AGENT_WON if r > 0 else CODE_AGENT_LOST
        info = {"message": MESSAGES[code], "TimeLimit.truncated": False}
        if i in rows:
            k = rows[i]
            info["terminal_observation"] = {"board": self._tb[k].copy(), "dice_roll": int(self._td[k])}
        return info


class EWNDeviceVecEnv:
    metadata = {"render_modes": []}

    def __init__(self, num_envs: int, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.0,
                 shaped: bool = False, illegal_move_reward: float = -1.0, illegal_move_tolerance: int = 10, device=None,
                 env_id0: int = 0, terminal_observation: bool = True, engine: Engine = None, **_ignored):
        assert cube_layer < board_size - 1
        self.e = engine if engine is not None else Engine(board_size, cube_layer, device=device)
        xp = self.e.xp
        self.torch = xp.torch
        self.num_envs, self.board_size, self.cube_layer = int(num_envs), int(board_size), int(cube_layer)
        self.cube_num = cube_layer * (cube_layer + 1) // 2
        self.shaped = bool(shaped)
        self.reward = 1.0 if shaped else float(reward)             # MinimaxEnv drops goal_reward (envs/minimax_ewn.py:17-19)
        self.illegal_move_reward, self.illegal_move_tolerance = float(illegal_move_reward), int(illegal_move_tolerance)
        self.seed_value, self.env_id0 = int(seed), int(env_id0)
        self.action_space = MultiDiscrete([2, 3])
        self.observation_space = Dict({
            "board": Box(low=-self.cube_num, high=self.cube_num, shape=(board_size, board_size), dtype=np.int16),
            "dice_roll": Discrete(self.cube_num + 1, start=1),
        })
        n, N = self.num_envs, self.board_size
        self.state = None
        self._t = 0
        self.stats = xp.zeros((4,), "i64")
        self._action = xp.zeros((n, 2), "i64")
        self._out = {"reward_f32": xp.empty((n,), "f32"), "terminated": xp.empty((n,), "u8"), "truncated": xp.empty((n,), "u8"),
                     "obs_board": xp.empty((n, N, N), "i16"), "obs_dice": xp.empty((n,), "i64")}
        self._want_term = bool(terminal_observation)
        if self._want_term:
            self._out["term_board"] = xp.empty((n, N, N), "i16")
            self._out["term_dice"] = xp.empty((n,), "i64")
            self._cnt = xp.zeros((2,), "i32")
            self._tidx = xp.empty((n,), "i32")
        if self.shaped:
            self._tol = xp.empty((n,), "i32")
            self._prev = xp.empty((n,), "f64")
        self._plans = None
        self._pending = False
        self.render_mode = None
        self.reset_infos = [{} for _ in range(min(n, 1))]

    # ------------------------------------------------------------------ VecEnv API
    def _obs(self):
        return {"board": self._out["obs_board"], "dice_roll": self._out["obs_dice"]}

    def reset(self):
        e, n = self.e, self.num_envs
        fresh = e.reset(n, seed=self.seed_value, step=self._t, env_id0=self.env_id0)
        if self.state is None:
            self.state = fresh
            if self.shaped:                                         # training_ewn.py:35-38: set once, at construction
                self._tol.fill_(self.illegal_move_tolerance)
                self._prev.copy_(e.evaluate(self.state))
            self._bind()
        else:
            self.state.copy_(fresh)                                 # the pre-bound launches keep pointing at self.state
        if self._want_term:
            self._cnt.zero_()                                       # a reset shifts the step parity: no counter may carry a count over
        b, d = e.decode(self.state)
        self._out["obs_board"].copy_(b)
        self._out["obs_dice"].copy_(d)
        self._t += 1
        return self._obs()

    def _bind(self):
        kw = dict(seed=self.seed_value, step=0, env_id0=self.env_id0, reward=self.reward, autoreset=True, shaped=self.shaped,
                  illegal_move_reward=self.illegal_move_reward, stats=self.stats, want=tuple(self._out), out=self._out)
        if self.shaped:
            kw.update(tolerance=self._tol, prev_score=self._prev)
        if self._want_term:                                         # two append counters alternate; a launch clears the other one
            self._plans = [self.e.step_plan(self.state, self._action, term_compact=(self._cnt[b:b + 1], self._tidx, self._cnt[1 - b:2 - b]), **kw)
                           for b in (0, 1)]
        else:
            self._plans = [self.e.step_plan(self.state, self._action, **kw)] * 2

    def step_async(self, actions):
        if self.state is None:
            raise RuntimeError("reset() first")
        a = self._action
        if actions is not a:
            a.copy_(self.torch.as_tensor(actions, device=a.device).reshape(self.num_envs, 2))
        self._pending = True

    def step_wait(self):
        assert self._pending, "step_async first"
        self._pending = False
        b = self._t & 1
        self._plans[b].stream = self.e.xp.stream()                 # torch's current stream, not the one reset() saw
        self._plans[b].launch(self._t)
        self._t += 1
        o = self._out
        done = o["terminated"].view(self.torch.bool)
        cnt = self._cnt[b:b + 1] if self._want_term else None
        infos = DeviceInfos(self, cnt, self._tidx if self._want_term else None, o.get("term_board"), o.get("term_dice"), done, o["truncated"],
                            o["reward_f32"])
        return self._obs(), o["reward_f32"], done, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    @property
    def action_buffer(self):
        """the int64[n,2] CUDA tensor the step kernel reads: a policy that writes its actions here skips one copy"""
        return self._action

    def seed(self, seed=None):
        return [None if seed is None else seed + i for i in range(min(self.num_envs, 1))]

    def close(self):
        self._plans = None

    def get_attr(self, name, indices=None):
        return [getattr(self, name)] * len(self._indices(indices))

    def set_attr(self, name, value, indices=None):
        setattr(self, name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        raise NotImplementedError("games live in one packed device array; there is no per-env Python object")

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * len(self._indices(indices))

    def get_images(self):
        return [None] * self.num_envs

    def render(self, mode=None):
        return None

    def _indices(self, indices):
        if indices is None:
            return range(self.num_envs)
        return [indices] if isinstance(indices, int) else list(indices)
