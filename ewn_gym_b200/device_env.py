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
  * `infos[i]` builds the SB3 dict (`terminal_observation`, `TimeLimit.truncated`, `message`) on demand (one device read); the
    message comes from the result code the kernel stored for the finished game, not from the reward's sign;
  * the tensors a `step` returns stay valid across the NEXT `step` (two output sets alternate): SB3's `collect_rollouts` adds
    `_last_obs` / `_last_episode_starts` to its buffer only after it has called `env.step()` again;
  * `seed(s)` (train.py:135): from the next `reset()` on, game i draws from the Philox stream (key s, counter env_id0 + i);
  * `opponent_policy` (envs/ewn.py:265-296): None / "random" = the in-kernel random opponent (one launch per step); "minimax",
    "mcts" or any `arena.BatchPolicy` = the opponent's move comes from that policy on the flipped board, between the two plies;
  * `shaped=True` = `MiniMaxHeuristicEnv` (envs/training_ewn.py:40-99), `stats` = int64[4] {wins, losses, illegal endings, episodes}.
Everything is enqueued on torch's current stream; nothing synchronises unless the caller reads a value.
"""
from __future__ import annotations

from collections.abc import Sequence

import numpy as np

from .constants import ClassicalPolicy
from .engine import Engine, MESSAGES, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_AGENT_LOST, CODE_OPP_ILLEGAL, BOTTOM_RIGHT, TOP_LEFT
from .spaces import Box, Dict, Discrete, MultiDiscrete


class EmptyInfos(Sequence):
    """`reset_infos` of a VecEnv with n envs -- n empty dicts -- without building a million of them"""

    def __init__(self, n):
        self._n = int(n)

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [{} for _ in range(*i.indices(self._n))]
        if not -self._n <= i < self._n:
            raise IndexError(i)
        return {}


def seed_list(seed, n):
    """what VecEnv.seed returns: the seed of every sub-environment, seed + i (None: unseeded)"""
    if seed is None:
        return [None] * n
    r = range(int(seed), int(seed) + n)
    return list(r) if n <= (1 << 16) else r              # a range is a Sequence too; a million-entry list helps nobody


def code_from_flags(reward, truncated):
    """Fallback when no result code was stored (terminal_observation=False): envs/ewn.py:445-478 by elimination on
    (reward sign, truncated).  Ambiguous only for a goal reward of 0; envs with terminal observations never come here."""
    if truncated:
        return CODE_AGENT_ILLEGAL if reward < 0 else CODE_OPP_ILLEGAL
    return CODE_AGENT_WON if reward > 0 else CODE_AGENT_LOST


def resolve_opponent(opponent_policy, policy_kwargs):
    """envs/ewn.py:265-287 load_opponent_policy for a batch: -> None (in-kernel random opponent) or an arena.BatchPolicy"""
    from . import arena
    if opponent_policy is None or isinstance(opponent_policy, arena.BatchPolicy):
        return opponent_policy
    if isinstance(opponent_policy, str):
        opponent_policy = ClassicalPolicy.from_string(opponent_policy)
    kw = dict(policy_kwargs or {})
    if opponent_policy == ClassicalPolicy.random:
        return None
    if opponent_policy == ClassicalPolicy.minimax:
        return arena.MinimaxPolicy(int(kw.get("max_depth", 3)), kw.get("heuristic", "hybrid"))
    if opponent_policy == ClassicalPolicy.mcts:
        return arena.MonteCarloPolicy(int(kw.get("num_simulations", 10)), int(kw.get("num_env_copies", 5)), seed=int(kw.get("seed", 0)))
    if opponent_policy == ClassicalPolicy.uct:
        return arena.UctPolicy(int(kw.get("rounds", 16)), int(kw.get("batch", 32)), float(kw.get("exploration_constant", 1.41)), seed=int(kw.get("seed", 0)))
    raise NotImplementedError(f"opponent_policy {opponent_policy!r}: AlphaZero / SB3-checkpoint opponents are out of scope (DESIGN.md 9)")


def _host(x):
    return x.cpu().numpy() if hasattr(x, "cpu") else np.asarray(x)


class DeviceInfos(Sequence):
    """`infos` of a device step.  Stays on the GPU until somebody asks for an entry."""

    def __init__(self, env, count, index, board, dice, code, done, trunc, reward):
        self._env, self._count, self._index, self._board, self._dice, self._code = env, count, index, board, dice, code
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

    def terminal_codes(self):
        """EWN_CODE_* uint8[k] of the rows of terminal_rows()"""
        if self._count is None:
            raise RuntimeError("the env was created with terminal_observation=False")
        return self._code[: self.terminal_count()]

    def _pull(self):
        if self._host is None:
            done, trunc, rew = (_host(x) for x in (self._done, self._trunc, self._reward))
            rows = {}
            if self._count is not None:
                idx, b, d = (_host(x) for x in self.terminal_rows())
                rows = {int(i): k for k, i in enumerate(idx)}
                self._tb, self._td, self._tc = b, d, _host(self.terminal_codes())
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
        if i in rows:
            k = rows[i]
            return {"message": MESSAGES[int(self._tc[k])], "TimeLimit.truncated": False,
                    "terminal_observation": {"board": self._tb[k].copy(), "dice_roll": int(self._td[k])}}
        return {"message": MESSAGES[code_from_flags(float(rew[i]), bool(trunc[i]))], "TimeLimit.truncated": False}


class EWNDeviceVecEnv:
    metadata = {"render_modes": []}

    def __init__(self, num_envs: int, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.0,
                 shaped: bool = False, illegal_move_reward: float = -1.0, illegal_move_tolerance: int = 10, device=None,
                 env_id0: int = 0, terminal_observation: bool = True, engine: Engine = None, opponent_policy=None, **policy_kwargs):
        assert cube_layer < board_size - 1
        self.e = engine if engine is not None else Engine(board_size, cube_layer, device=device)
        xp = self.e.xp
        self.torch = getattr(xp, "torch", None)
        self.num_envs, self.board_size, self.cube_layer = int(num_envs), int(board_size), int(cube_layer)
        self.cube_num = cube_layer * (cube_layer + 1) // 2
        self.shaped = bool(shaped)
        self.reward = 1.0 if shaped else float(reward)             # MinimaxEnv drops goal_reward (envs/minimax_ewn.py:17-19)
        self.illegal_move_reward, self.illegal_move_tolerance = float(illegal_move_reward), int(illegal_move_tolerance)
        self.seed_value, self.env_id0 = int(seed), int(env_id0)
        self._next_seed = None
        self.opponent = resolve_opponent(opponent_policy, policy_kwargs)
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
        self._want_term = bool(terminal_observation)
        # two output sets alternate by step parity: what step t returned is still intact after step t + 1
        self._outs = []
        for _ in (0, 1):
            o = {"reward_f32": xp.empty((n,), "f32"), "terminated": xp.empty((n,), "u8"), "truncated": xp.empty((n,), "u8"),
                 "obs_board": xp.empty((n, N, N), "i16"), "obs_dice": xp.empty((n,), "i64")}
            if self._want_term:
                o["term_board"], o["term_dice"], o["term_code"] = xp.empty((n, N, N), "i16"), xp.empty((n,), "i64"), xp.empty((n,), "u8")
            self._outs.append(o)
        if self._want_term:
            self._cnt = xp.zeros((2,), "i32")
            self._tidx = [xp.empty((n,), "i32"), xp.empty((n,), "i32")]
        if self.shaped:
            self._tol = xp.empty((n,), "i32")
            self._prev = xp.empty((n,), "f64")
        if self.opponent is not None:
            self._br = xp.asarray(np.full(n, BOTTOM_RIGHT, np.int32), "i32")
        self._plans = None
        self._pending = False
        self._last = 0                                              # output set of the latest reset / step
        self.render_mode = None
        self.reset_infos = EmptyInfos(n)

    # ------------------------------------------------------------------ VecEnv API
    def _obs(self, b):
        return {"board": self._outs[b]["obs_board"], "dice_roll": self._outs[b]["obs_dice"]}

    def _copy_into(self, dst, src):
        if self.torch is not None:
            dst.copy_(src)
        else:
            np.copyto(dst, src)

    def reset(self):
        e, n = self.e, self.num_envs
        if self._next_seed is not None:                             # VecEnv.seed(): takes effect here (train.py:135-136)
            self.seed_value, self._next_seed = int(self._next_seed), None
            self._plans = None
        fresh = e.reset(n, seed=self.seed_value, step=self._t, env_id0=self.env_id0)
        if self.state is None:
            self.state = fresh
            if self.shaped:                                         # training_ewn.py:35-38: set once, at construction
                self._tol[...] = self.illegal_move_tolerance
                self._copy_into(self._prev, e.evaluate(self.state))
        else:
            self._copy_into(self.state, fresh)                      # the pre-bound launches keep pointing at self.state
        if self._plans is None and self.opponent is None:
            self._bind()
        if self._want_term:
            self._cnt[...] = 0                                      # a reset shifts the step parity: no counter may carry a count over
        b = self._t & 1
        bd, dc = e.decode(self.state)
        self._copy_into(self._outs[b]["obs_board"], bd)
        self._copy_into(self._outs[b]["obs_dice"], dc)
        self._t += 1
        self._last = b
        return self._obs(b)

    def _step_kw(self, b):
        kw = dict(seed=self.seed_value, step=0, env_id0=self.env_id0, reward=self.reward, autoreset=True, shaped=self.shaped,
                  illegal_move_reward=self.illegal_move_reward, stats=self.stats, want=tuple(self._outs[b]), out=self._outs[b])
        if self.shaped:
            kw.update(tolerance=self._tol, prev_score=self._prev)
        if self._want_term:                                         # two append counters alternate; a launch clears the other one
            kw["term_compact"] = (self._cnt[b:b + 1], self._tidx[b], self._cnt[1 - b:2 - b])
        return kw

    def _bind(self):
        self._plans = [self.e.step_plan(self.state, self._action, **self._step_kw(b)) for b in (0, 1)]

    def bind_step(self, action, out, step=0, step_dev=None):
        """A pre-bound step of THIS env's games (same seed, env ids, reward shaping, statistics) that reads `action` (int64[n,2]) and
        writes the output tensors in `out` (names of ewn_step_args) -- e.g. slices of a rollout buffer -- instead of the env's own
        two output sets; no terminal observations.  The caller launches it with ``plan.launch(step)``; with `step_dev` (int64[1] on the
        device) the kernel adds that counter to the step index, so a CUDA graph of such launches can be replayed.  After driving the
        games this way call ``advance(k)`` with the number of steps taken."""
        if self.opponent is not None:
            raise RuntimeError("bind_step: the in-kernel random opponent only")
        kw = self._step_kw(0)
        kw.pop("term_compact", None)
        kw.update(want=tuple(out), out=dict(out), step=int(step))
        if step_dev is not None:
            kw["step_dev"] = step_dev
        return self.e.step_plan(self.state, action, **kw)

    def advance(self, steps: int, obs=None):
        """account for `steps` steps taken through bind_step plans; `obs` = (board, dice) after the last of them becomes the env's
        latest observation"""
        self._t += int(steps)
        if obs is not None:
            self._copy_into(self._outs[self._last]["obs_board"], obs[0])
            self._copy_into(self._outs[self._last]["obs_dice"], obs[1])

    def rollout(self, actions=None, n_steps=None, agent_random=False):
        """T env steps of every game in ONE launch (ewn_rollout without a network): `actions` int64[T, n, 2] on the device (e.g. pooled
        or pre-sampled), or `agent_random=True` for the in-kernel uniform-over-the-legal-list agent (n_steps = T; the moves it played
        come back in the result).  Same games, dice and statistics as T `step` calls.  -> dict with `board` int16[T + 1, n, N, N] and
        `dice_roll` int64[T + 1, n] (slot t = the observation step t acted on, slot T = the one after the last step), `actions`,
        `rewards` f32[T, n], `dones` bool[T, n].  In-kernel random opponent and six-cube layouts only; no terminal observations."""
        if self.opponent is not None or self.torch is None:
            raise RuntimeError("rollout: the device env with its in-kernel random opponent only")
        if self.state is None:
            self.reset()
        th, n, N = self.torch, self.num_envs, self.board_size
        dev = self.state.device
        if agent_random:
            T = int(n_steps)
            acts = th.empty((T, n, 2), dtype=th.int64, device=dev)
        else:
            acts = th.as_tensor(actions, device=dev).to(th.int64).reshape(-1, n, 2).contiguous()
            T = int(acts.shape[0])
        bufs = {"board": th.empty((T + 1, n, N, N), dtype=th.int16, device=dev), "dice": th.empty((T + 1, n), dtype=th.int64, device=dev),
                "actions": acts, "rewards": th.empty((T, n), device=dev), "dones": th.empty((T, n), dtype=th.uint8, device=dev)}
        launch = self.e.rollout_plan(None, self.state, T, bufs, seed=self.seed_value, env_id0=self.env_id0, reward=self.reward, shaped=self.shaped,
                                     illegal_move_reward=self.illegal_move_reward, tolerance=getattr(self, "_tol", None),
                                     prev_score=getattr(self, "_prev", None), stats=self.stats, agent_random=agent_random)
        launch(self._t)
        self.advance(T, obs=(bufs["board"][T], bufs["dice"][T]))
        return {"board": bufs["board"], "dice_roll": bufs["dice"], "actions": acts, "rewards": bufs["rewards"], "dones": bufs["dones"].view(th.bool)}

    @property
    def step_index(self) -> int:
        """the step counter the NEXT step will use (the Philox counter of its dice)"""
        return self._t

    def step_async(self, actions):
        if self.state is None:
            raise RuntimeError("reset() first")
        a = self._action
        if actions is not a:
            if self.torch is not None:
                a.copy_(self.torch.as_tensor(actions, device=a.device).reshape(self.num_envs, 2))
            else:
                np.copyto(a, np.asarray(actions).reshape(self.num_envs, 2), casting="unsafe")
        self._pending = True

    def _step_external(self, b):
        """envs/ewn.py:436-486 with a policy as opponent: its answer is computed between the two plies on the flipped board
        (:289-296) and handed to the step kernel in replay mode together with the very dice the fused step would have drawn."""
        e, n, t = self.e, self.num_envs, self._t
        d_opp, d_next = e.step_dice(n, self.seed_value, t, self.env_id0)
        peek = self.state.clone() if self.torch is not None else self.state.copy()
        e.ply(peek, e.xp.asarray(self._action, "i32"), TOP_LEFT)     # the agent's ply on a copy (:439-442)
        e.set_dice(peek, d_opp, self._br)                           # :457-458 switch_player + roll_dice
        opp = self.opponent.action(e, e.flip(peek), t)              # :289-296; ignored by the kernel where the game ended at the agent's ply
        kw = self._step_kw(b)
        kw["step"] = t
        e.step(self.state, self._action, inject=(d_opp, opp, d_next), **kw)

    def step_wait(self):
        assert self._pending, "step_async first"
        self._pending = False
        b = self._t & 1
        if self.opponent is None:
            self._plans[b].stream = self.e.xp.stream()             # torch's current stream, not the one reset() saw
            self._plans[b].launch(self._t)
        else:
            self._step_external(b)
        self._t += 1
        self._last = b
        o = self._outs[b]
        done = o["terminated"].view(self.torch.bool) if self.torch is not None else o["terminated"].view(np.bool_)
        T = self._want_term
        infos = DeviceInfos(self, self._cnt[b:b + 1] if T else None, self._tidx[b] if T else None, o.get("term_board"), o.get("term_dice"),
                            o.get("term_code"), done, o["truncated"], o["reward_f32"])
        return self._obs(b), o["reward_f32"], done, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    @property
    def action_buffer(self):
        """the int64[n,2] CUDA tensor the step kernel reads: a policy that writes its actions here skips one copy"""
        return self._action

    def seed(self, seed=None):
        """SB3 VecEnv.seed (train.py:135): sub-environment i is seeded with seed + i at the next reset().  Here every game already
        owns a Philox stream (key, env_id0 + i), so the key becomes `seed` at the next reset(); returns the per-env seeds."""
        if seed is not None:
            self._next_seed = int(seed)
        return seed_list(seed, self.num_envs)

    def close(self):
        self._plans = None

    def get_attr(self, name, indices=None):
        return [getattr(self, name)] * len(self._indices(indices))

    def set_attr(self, name, value, indices=None):
        setattr(self, name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        """Per-env Python objects do not exist (the games are one packed array); the methods of the reference env that read a game
        are served from the batch: get_legal_actions / check_win / evaluate -> one result per selected env."""
        idx = list(self._indices(indices))
        e = self.e
        sel = self.state[idx] if self.torch is None else self.state[self.torch.as_tensor(idx, device=self.state.device)]
        host = (lambda x: x.cpu().numpy()) if self.torch is not None else np.asarray
        if method_name == "check_win":
            return [bool(v) for v in host(e.check_win(sel))]
        if method_name == "evaluate":
            return [float(v) for v in host(e.evaluate(sel, kwargs.get("heuristic", args[0] if args else "hybrid")))]
        if method_name == "get_legal_actions":
            player = args[0] if args else kwargs.get("player", TOP_LEFT)
            player = getattr(player, "value", player)
            _, lst, cnt = e.legal(sel, int(player))
            lst, cnt = host(lst), host(cnt)
            return [[list(map(int, a)) for a in lst[i, : cnt[i]]] for i in range(len(idx))]
        raise NotImplementedError(f"env_method({method_name!r}): games live in one packed device array; available: "
                                  "check_win, evaluate, get_legal_actions")

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
