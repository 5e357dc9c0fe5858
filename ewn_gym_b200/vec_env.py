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

Host arrays are views of page-locked buffers owned by the env; they are overwritten by the next ``step``.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Sequence

import numpy as np

from . import _lib
from .engine import MESSAGES, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_AGENT_LOST
from .spaces import Box, Dict, Discrete, MultiDiscrete


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

    def __init__(self, env, done, trunc, reward, term_board, term_dice, term_index, term_count):
        self._env, self._done, self._trunc, self._reward = env, done, trunc, reward
        self._tb, self._td, self._ti, self._tc = term_board, term_dice, term_index, term_count
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
        r = float(self._reward[i])
        if self._trunc[i]:
            code = CODE_AGENT_ILLEGAL if r < 0 else 3
        else:
            code = CODE_AGENT_WON if r > 0 else CODE_AGENT_LOST
        info = {"message": MESSAGES[code], "TimeLimit.truncated": False}
        if self._tb is not None:
            k = self._row(i)
            info["terminal_observation"] = {"board": self._tb[k].copy(), "dice_roll": int(self._td[k])}
        return info


class EWNVecEnv:
    """SB3-VecEnv-compatible batched environment on one GPU (host-buffer interface)."""

    metadata = {"render_modes": []}

    def __init__(self, num_envs: int, board_size: int = 5, cube_layer: int = 3, seed: int = 9487, reward: float = 1.0,
                 shaped: bool = False, illegal_move_reward: float = -1.0, illegal_move_tolerance: int = 10, device: int = 0,
                 env_id0: int = 0, terminal_observation: bool = True, chunks: int = 0, staged: bool = False, numa_bind: bool = False,
                 **_ignored):
        """numa_bind: restrict this process to the CPUs next to the GPU before the host buffers are allocated (opt-in: it
        changes the caller's CPU affinity; meant for one-process-per-GPU launches, see bind_host_to_device)"""
        assert cube_layer < board_size - 1
        self.lib = _lib.lib()
        self.numa = bind_host_to_device(self.lib, device) if numa_bind else None
        self.num_envs, self.board_size, self.cube_layer = int(num_envs), int(board_size), int(cube_layer)
        self.cube_num = cube_layer * (cube_layer + 1) // 2
        # MinimaxEnv drops every ctor argument but board_size / cube_layer (envs/minimax_ewn.py:17-19), so the env that
        # train.py registers always has reward 1.0 whatever goal_reward says
        self.reward = 1.0 if shaped else float(reward)
        self.shaped = bool(shaped)
        self.action_space = MultiDiscrete([2, 3])
        self.observation_space = Dict({
            "board": Box(low=-self.cube_num, high=self.cube_num, shape=(board_size, board_size), dtype=np.int16),
            "dice_roll": Discrete(self.cube_num + 1, start=1),
        })
        cfg = _lib.VecConfig(self.num_envs, board_size, cube_layer, self.reward, float(illegal_move_reward), int(seed), int(env_id0),
                             int(self.shaped), int(illegal_move_tolerance), int(chunks), int(device), 1 if staged else 0, 0)
        h = C.c_void_p()
        _lib.check(self.lib.ewn_vec_create(C.byref(cfg), C.byref(h)), "ewn_vec_create", self.lib)
        self._h = h
        n, N = self.num_envs, board_size
        mk = lambda shape, dt: _Pinned(self.lib, shape, dt)
        self._b = dict(action=mk((n, 2), np.int64), board=mk((n, N, N), np.int16), dice=mk((n,), np.int64), reward=mk((n,), np.float32),
                       done=mk((n,), np.uint8), trunc=mk((n,), np.uint8))
        self._want_term = bool(terminal_observation)
        if self._want_term:
            self._b["tboard"] = mk((n, N, N), np.int16)
            self._b["tdice"] = mk((n,), np.int64)
            self._b["tindex"] = mk((n,), np.int32)
            self._b["tcount"] = mk((16,), np.uint32)
        self._actions = None
        self.reset_infos = [{} for _ in range(min(n, 1))]
        self.render_mode = None

    # ------------------------------------------------------------------ VecEnv API
    def _obs(self):
        return {"board": self._b["board"].array, "dice_roll": self._b["dice"].array}

    def reset(self):
        _lib.check(self.lib.ewn_vec_reset(self._h, self._b["board"].ptr, self._b["dice"].ptr), "ewn_vec_reset", self.lib)
        return self._obs()

    def step_async(self, actions):
        a = self._b["action"].array
        if actions is not a:
            np.copyto(a, np.asarray(actions).reshape(self.num_envs, 2), casting="unsafe")

    def step_wait(self):
        b = self._b
        T = self._want_term
        _lib.check(self.lib.ewn_vec_step(self._h, b["action"].ptr, b["board"].ptr, b["dice"].ptr, b["reward"].ptr, b["done"].ptr,
                                         b["trunc"].ptr, b["tboard"].ptr if T else None, b["tdice"].ptr if T else None,
                                         b["tindex"].ptr if T else None, b["tcount"].ptr if T else None), "ewn_vec_step", self.lib)
        done = b["done"].array.view(np.bool_)
        if T:
            infos = LazyInfos(self, done, b["trunc"].array, b["reward"].array, b["tboard"].array, b["tdice"].array, b["tindex"].array,
                              int(b["tcount"].array[0]))
        else:
            infos = LazyInfos(self, done, b["trunc"].array, b["reward"].array, None, None, None, 0)
        return self._obs(), b["reward"].array, done, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    @property
    def action_buffer(self):
        """The page-locked int64[n,2] array step() reads from: write actions here to skip one host copy."""
        return self._b["action"].array

    def seed(self, seed=None):
        return [None if seed is None else seed + i for i in range(min(self.num_envs, 1))]

    def close(self):
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
