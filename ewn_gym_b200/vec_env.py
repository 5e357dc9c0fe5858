"""EWNVecEnv -- n concurrent EinStein-wuerfelt-nicht games behind Stable-Baselines3's ``VecEnv`` contract.

Replaces ``SubprocVecEnv([make_env] * num_envs)`` of the reference (train.py:126-136): one fused kernel launch per
``step`` instead of one OS process and one pipe round trip per env.  SB3 is duck-typed (it is not installed in this
image): ``num_envs / observation_space / action_space / reset / step_async / step_wait / step / seed / close /
get_attr / set_attr / env_method / env_is_wrapped / get_images / render / reset_infos``.

Contract details reproduced:
  * ``reset() -> obs`` with ``obs = {"board": int16[n,N,N], "dice_roll": int64[n]}``
  * ``step(actions int64[n,2]) -> (obs, rewards float32[n], dones bool[n], infos)``; finished games are reset in the
    same call, the returned obs is the new episode's and ``infos[i]["terminal_observation"]`` holds the last one,
    ``infos[i]["TimeLimit.truncated"] = truncated and not terminated`` (always False here: envs/ewn.py sets
    terminated on every truncation), ``infos[i]["message"]`` as envs/ewn.py:448-478.
  * ``shaped=True`` gives ``MiniMaxHeuristicEnv`` semantics (envs/training_ewn.py:40-99), the env train.py registers.

  * ``seed(s)`` (train.py:135): from the next ``reset()`` on, game i draws from the Philox stream (key s, counter env_id0 + i);
    returns the per-env seeds ``s + i``.
  * ``opponent_policy`` (envs/ewn.py:265-296): default / "random" = the in-kernel random opponent; "minimax" / "mcts" / an
    ``arena.BatchPolicy`` = that policy answers on the flipped board between the two plies (served by ``EWNDeviceVecEnv``).

Host arrays are views of page-locked buffers owned by the env.  Two sets alternate, so what ``step`` t returned is still intact
after ``step`` t + 1 (SB3's ``collect_rollouts`` stores ``_last_obs`` / ``_last_episode_starts`` only after the next ``env.step``);
it is overwritten by ``step`` t + 2.  When stable_baselines3 is importable the class derives from its ``VecEnv``.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Sequence

import numpy as np

from . import _lib
from .device_env import EmptyInfos, code_from_flags, resolve_opponent, seed_list
from .engine import MESSAGES
from .spaces import Box, Dict, Discrete, MultiDiscrete

try:                                               # SB3's _wrap_env leaves a VecEnv instance alone and wraps anything else in DummyVecEnv
    from stable_baselines3.common.vec_env import VecEnv as _VecEnvBase
except Exception:                                  # noqa: BLE001 - not installed in this image: the duck type stands alone
    _VecEnvBase = object


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_to_device(lib, device: int, sysfs: str = "/sys"):
    """Pin the calling thread to the CPUs of the NUMA node the GPU is attached to, so that the page-locked buffers allocated next
    (first touch) and the thread that drives `step` sit on the GPU's side of the socket interconnect.  With one process per GPU
    on a two-socket host this keeps the 91 MB per step of PCIe traffic off the inter-socket link.
    -> {"pci", "node", "cpus"} or None when the platform does not say (no sysfs entry, node -1, single node)."""
    import os
    buf = C.create_string_buffer(32)
    if lib.ewn_device_pci_bus_id(int(device), buf, 32) != 0:
        return None
    pci = buf.value.decode().lower()
    try:
        node = int(open(f"{sysfs}/bus/pci/devices/{pci}/numa_node").read())
        if node < 0:
            return None
        cpus = _parse_cpulist(open(f"{sysfs}/devices/system/node/node{node}/cpulist").read())
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return {"pci": pci, "node": node, "cpus": len(allowed), "changed": False}
        os.sched_setaffinity(0, cpus)
        return {"pci": pci, "node": node, "cpus": len(cpus), "changed": True}
    except (OSError, ValueError):
        return None


class _Pinned:
    """A page-locked host buffer exposed as a numpy array."""

    def __init__(self, lib, shape, dtype):
        self.lib = lib
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        _lib.check(lib.ewn_host_alloc(C.byref(p), n), "ewn_host_alloc", lib)
        self.ptr = p.value
        self.array = np.frombuffer((C.c_char * max(n, 1)).from_address(self.ptr), dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.lib.ewn_host_free(self.ptr)
            self.ptr = None


class LazyInfos(Sequence):
    """``infos`` of VecEnv.step_wait without building n dicts per step: dict i is made when it is asked for."""

    def __init__(self, env, done, trunc, reward, term_board, term_dice, term_index, term_code, term_count):
        self._env, self._done, self._trunc, self._reward = env, done, trunc, reward
        self._tb, self._td, self._ti, self._tcode, self._tc = term_board, term_dice, term_index, term_code, term_count
        self._order = None

    def __len__(self):
        return len(self._done)

    def _row(self, i):
        """row of env i in the compacted terminal-observation list"""
        if self._order is None:
            idx = self._ti[: self._tc]
            self._order = np.argsort(idx, kind="stable")
            self._sorted = idx[self._order]
        k = int(np.searchsorted(self._sorted, i))
        assert k < self._tc and self._sorted[k] == i
        return int(self._order[k])

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        if not self._done[i]:
            return {}
        if self._tb is None:                       # no terminal rows were requested: the code by elimination (device_env.code_from_flags)
            return {"message": MESSAGES[code_from_flags(float(self._reward[i]), bool(self._trunc[i]))], "TimeLimit.truncated": False}
        k = self._row(i)                           # the kernel stored how the game ended next to its terminal observation
        return {"message": MESSAGES[int(self._tcode[k])], "TimeLimit.truncated": False,
                "terminal_observation": {"board": self._tb[k].copy(), "dice_roll": int(self._td[k])}}


class EWNVecEnv(_VecEnvBase):
    """SB3-VecEnv-compatible batched environment on one GPU (host-buffer interface)."""

    metadata = {"render_modes": []}

    def __init__(self, num_envs: int, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.0,
                 shaped: bool = False, illegal_move_reward: float = -1.0, illegal_move_tolerance: int = 10, device: int = 0,
                 env_id0: int = 0, terminal_observation: bool = True, chunks: int = 0, staged: bool = False, numa_bind: bool = False,
                 opponent_policy=None, **policy_kwargs):
        """numa_bind: restrict this process to the CPUs next to the GPU before the host buffers are allocated (opt-in: it
        changes the caller's CPU affinity; meant for one-process-per-GPU launches, see bind_host_to_device)"""
        assert cube_layer < board_size - 1
        self.lib = _lib.lib()
        self._device = int(device)
        self.numa = bind_host_to_device(self.lib, device) if numa_bind else None
        self.num_envs, self.board_size, self.cube_layer = int(num_envs), int(board_size), int(cube_layer)
        self.cube_num = cube_layer * (cube_layer + 1) // 2
        # MinimaxEnv drops every ctor argument but board_size / cube_layer (envs/minimax_ewn.py:17-19), so the env that
        # train.py registers always has reward 1.0 whatever goal_reward says
        self.reward = 1.0 if shaped else float(reward)
        self.shaped = bool(shaped)
        act_space, obs_space = MultiDiscrete([2, 3]), Dict({
            "board": Box(low=-self.cube_num, high=self.cube_num, shape=(board_size, board_size), dtype=np.int16),
            "dice_roll": Discrete(self.cube_num + 1, start=1),
        })
        if _VecEnvBase is not object:              # SB3 validates gymnasium spaces: use the real ones next to the real base class
            import gymnasium.spaces as gs
            act_space = gs.MultiDiscrete([2, 3])
            obs_space = gs.Dict({"board": gs.Box(low=-self.cube_num, high=self.cube_num, shape=(board_size, board_size), dtype=np.int16),
                                 "dice_roll": gs.Discrete(self.cube_num + 1, start=1)})
            _VecEnvBase.__init__(self, int(num_envs), obs_space, act_space)
        self.action_space, self.observation_space = act_space, obs_space
        self._t = 0
        self._want_term = bool(terminal_observation)
        self.render_mode = None
        self.reset_infos = EmptyInfos(self.num_envs)
        self._dev = None
        if resolve_opponent(opponent_policy, policy_kwargs) is not None:
            # a policy as opponent (envs/ewn.py:265-296): the step is a sequence of device calls, served by the device-resident class;
            # this class then only moves its results into host arrays
            from .device_env import EWNDeviceVecEnv
            import torch
            torch.cuda.set_device(device)
            self._dev = EWNDeviceVecEnv(num_envs, board_size, cube_layer, seed=seed, reward=reward, shaped=shaped,
                                        illegal_move_reward=illegal_move_reward, illegal_move_tolerance=illegal_move_tolerance,
                                        device=torch.device("cuda", device), env_id0=env_id0, terminal_observation=terminal_observation,
                                        opponent_policy=opponent_policy, **policy_kwargs)
            self._h, self._b = None, {}
            return
        cfg = _lib.VecConfig(self.num_envs, board_size, cube_layer, self.reward, float(illegal_move_reward), int(seed), int(env_id0),
                             int(self.shaped), int(illegal_move_tolerance), int(chunks), int(device), 1 if staged else 0, 0)
        h = C.c_void_p()
        _lib.check(self.lib.ewn_vec_create(C.byref(cfg), C.byref(h)), "ewn_vec_create", self.lib)
        self._h = h
        n, N = self.num_envs, board_size
        mk = lambda shape, dt: _Pinned(self.lib, shape, dt)
        self._b = dict(action=mk((n, 2), np.int64))
        for p in (0, 1):                           # two output sets alternate by step parity (see the module docstring)
            self._b.update({f"board{p}": mk((n, N, N), np.int16), f"dice{p}": mk((n,), np.int64), f"reward{p}": mk((n,), np.float32),
                            f"done{p}": mk((n,), np.uint8), f"trunc{p}": mk((n,), np.uint8)})
            if self._want_term:
                self._b.update({f"tboard{p}": mk((n, N, N), np.int16), f"tdice{p}": mk((n,), np.int64), f"tindex{p}": mk((n,), np.int32),
                                f"tcode{p}": mk((n,), np.uint8), f"tcount{p}": mk((16,), np.uint32)})
        self._last = 0

    # ------------------------------------------------------------------ VecEnv API
    def _obs(self, p):
        return {"board": self._b[f"board{p}"].array, "dice_roll": self._b[f"dice{p}"].array}

    def _to_host(self, obs):
        return {k: v.cpu().numpy() for k, v in obs.items()}

    def reset(self):
        if self._dev is not None:
            return self._to_host(self._dev.reset())
        p = self._t & 1
        _lib.check(self.lib.ewn_vec_reset(self._h, self._b[f"board{p}"].ptr, self._b[f"dice{p}"].ptr), "ewn_vec_reset", self.lib)
        self._t += 1
        self._last = p
        return self._obs(p)

    def step_async(self, actions):
        if self._dev is not None:
            self._dev.step_async(np.asarray(actions))
            return
        a = self._b["action"].array
        if actions is not a:
            np.copyto(a, np.asarray(actions).reshape(self.num_envs, 2), casting="unsafe")

    def step_wait(self):
        if self._dev is not None:
            obs, rew, done, infos = self._dev.step_wait()
            return self._to_host(obs), rew.cpu().numpy(), done.cpu().numpy(), infos
        b, p = self._b, self._t & 1
        T = self._want_term
        g = lambda name: b[name + str(p)]
        _lib.check(self.lib.ewn_vec_step2(self._h, b["action"].ptr, g("board").ptr, g("dice").ptr, g("reward").ptr, g("done").ptr,
                                          g("trunc").ptr, g("tboard").ptr if T else None, g("tdice").ptr if T else None,
                                          g("tindex").ptr if T else None, g("tcode").ptr if T else None, g("tcount").ptr if T else None),
                   "ewn_vec_step", self.lib)
        self._t += 1
        self._last = p
        done = g("done").array.view(np.bool_)
        if T:
            infos = LazyInfos(self, done, g("trunc").array, g("reward").array, g("tboard").array, g("tdice").array, g("tindex").array,
                              g("tcode").array, int(g("tcount").array[0]))
        else:
            infos = LazyInfos(self, done, g("trunc").array, g("reward").array, None, None, None, None, 0)
        return self._obs(p), g("reward").array, done, infos

    @property
    def terminal_count(self):
        """finished games of the latest step (rows of the compacted terminal observations)"""
        return int(self._b[f"tcount{self._last}"].array[0]) if self._want_term and self._dev is None else None

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    @property
    def action_buffer(self):
        """The page-locked int64[n,2] array step() reads from: write actions here to skip one host copy."""
        return self._b["action"].array if self._dev is None else None

    def seed(self, seed=None):
        """SB3 VecEnv.seed (train.py:135): sub-environment i is seeded with seed + i at the next reset().  Every game already owns a
        Philox stream (key, env_id0 + i); the key becomes `seed` at the next reset().  Returns the per-env seeds."""
        if self._dev is not None:
            return self._dev.seed(seed)
        if seed is not None:
            _lib.check(self.lib.ewn_vec_set_seed(self._h, int(seed)), "ewn_vec_set_seed", self.lib)
        return seed_list(seed, self.num_envs)

    def close(self):
        if getattr(self, "_dev", None) is not None:
            self._dev.close()
            self._dev = None
        if getattr(self, "_h", None):
            self.lib.ewn_vec_destroy(self._h)
            self._h = None
            for p in self._b.values():
                p.free()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_attr(self, name, indices=None):
        return [getattr(self, name)] * len(self._indices(indices))

    def set_attr(self, name, value, indices=None):
        setattr(self, name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        """the reference env's read-only methods, served for the selected games from their current observation (the agent is to
        move): check_win / evaluate(heuristic) / get_legal_actions(player)"""
        if self._dev is not None:
            return self._dev.env_method(method_name, *args, indices=indices, **kwargs)
        import torch
        from .device_env import EWNDeviceVecEnv
        from .engine import Engine
        idx = list(self._indices(indices))
        obs = self._obs(self._last)
        view = EWNDeviceVecEnv.__new__(EWNDeviceVecEnv)            # batch-of-len(idx) view: only what env_method reads
        view.e = Engine(self.board_size, self.cube_layer, device=torch.device("cuda", self._device))
        view.torch, view.num_envs = torch, len(idx)
        view.state = view.e.encode(obs["board"][idx], obs["dice_roll"][idx].astype(np.int32))
        return EWNDeviceVecEnv.env_method(view, method_name, *args, **kwargs)

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
