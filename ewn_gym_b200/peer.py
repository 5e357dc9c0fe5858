"""Root-statistics merge across the GPUs of one node without a collective library.

The flat Monte-Carlo policy of the reference (`classical_policies/mcts.py:47-69`) sums the wins of all its env copies per root
action and takes `np.argmax`.  With the playout index range of every (root, action) split over the ranks -- the Philox streams
are keyed on global ids, so the sum does not depend on the rank count -- that sum is an all-reduce of ``u32[n, 6]`` followed by
the arg-max.  `PeerExchange` maps one small buffer of every rank into every other rank (CUDA IPC; loads travel over
NVLink 5 / NVSwitch) and `best_action` runs ONE kernel per rank that publishes its epoch, waits for the peers' epochs, reads all
counters with peer loads, sums them and picks the move (`ewn_mc_best_action_peers`, include/ewn_b200.h).

    ex = PeerExchange(engine, n_roots)            # collective: every rank of the group calls it
    w = ex.wins_buffer()                          # zeroed u32[n, 6] view of this rank's buffer for the next epoch
    engine.mc_root_wins(roots, seed, first=rank * per, count=per, wins=w)
    action, total = ex.best_action(roots)         # identical on every rank

Buffer layout per rank: ``[flags u32[32] | wins parity 0 | wins parity 1]``.  Two win buffers alternate by epoch: the wait of
epoch e+1 proves that every peer finished reading epoch e, so rewriting parity p at epoch e+2 is safe.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .engine import Engine

FLAG_WORDS, ERR_WORD, MAX_PEERS = 32, 16, 8


class _DevView:
    """`__cuda_array_interface__` holder so that torch can wrap memory this library allocated"""

    def __init__(self, ptr, shape, typestr="<i4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class PeerExchange:
    def __init__(self, engine: Engine, n_roots: int, group=None, timeout_ms: int = 2000, handles=None, rank=None, world=None):
        """`group`: a torch.distributed group (default: WORLD; no process group at all = a single rank).  `handles / rank /
        world` are for tests that play several ranks in one process (the list collects every rank's handle)."""
        self.e, self.lib, self.n = engine, engine.lib, int(n_roots)
        self.timeout_ms = int(timeout_ms)
        self.row_bytes = max(self.n, 1) * 6 * 4
        self.bytes = FLAG_WORDS * 4 + 2 * self.row_bytes
        self.epoch = 0
        self._own = C.c_void_p()
        self._opened = []
        h = (C.c_ubyte * 64)()

        def alloc():
            with engine.xp.guard():
                _lib.check(self.lib.ewn_peer_alloc(self.bytes, C.byref(self._own), h), "ewn_peer_alloc", self.lib)
            return bytes(h)

        if handles is not None:                                   # several ranks in one process (tests)
            self.rank, self.world = int(rank), int(world)
            handles.append((alloc(), self._own.value))
            self._all = None
        else:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                # every step below either succeeds on all ranks or raises on all ranks: a rank that failed alone would leave the
                # others waiting in the next collective
                self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
                try:
                    mine, err = alloc(), None
                except Exception as ex:                           # noqa: BLE001 - reported to every rank below
                    mine, err = None, ex
                every = [None] * self.world
                dist.all_gather_object(every, mine, group=group)
                if any(b is None for b in every):
                    self.close()
                    raise RuntimeError(f"peer buffer allocation failed on rank(s) {[r for r, b in enumerate(every) if b is None]}: {err}")
                try:
                    self._map([(b, None) for b in every])
                    ok, err = True, None
                except Exception as ex:                           # noqa: BLE001
                    ok, err = False, ex
                oks = [None] * self.world
                dist.all_gather_object(oks, ok, group=group)
                if not all(oks):
                    self.close()
                    raise RuntimeError(f"mapping the peers' buffers (CUDA IPC) failed on rank(s) {[r for r, o in enumerate(oks) if not o]}: {err}")
            else:
                self.rank, self.world = 0, 1
                self._map([(alloc(), None)])
        if self.world > MAX_PEERS:
            raise ValueError(f"at most {MAX_PEERS} ranks (one NVSwitch domain)")

    def connect(self, handles):
        """tests: map the buffers of the ranks that live in this process"""
        self._map(handles)

    def _map(self, every):
        base = []
        for r, (hb, ptr) in enumerate(every):
            if r == self.rank:
                base.append(self._own.value)
            elif ptr is not None:
                base.append(ptr)                                   # same process: the allocation is already addressable
            else:
                p = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(hb)
                with self.e.xp.guard():
                    _lib.check(self.lib.ewn_peer_open(buf, C.byref(p)), "ewn_peer_open", self.lib)
                self._opened.append(p.value)
                base.append(p.value)
        self._base = base
        self._flag_ptrs = (C.c_void_p * self.world)(*base)
        self._win_ptrs = [(C.c_void_p * self.world)(*[b + FLAG_WORDS * 4 + par * self.row_bytes for b in base]) for par in (0, 1)]
        xp = self.e.xp
        if getattr(xp, "torch", None) is not None:
            t = xp.torch
            self._views = [_DevView(self._own.value, (FLAG_WORDS,))] + \
                          [_DevView(self._own.value + FLAG_WORDS * 4 + par * self.row_bytes, (max(self.n, 1), 6)) for par in (0, 1)]
            self.flags = t.as_tensor(self._views[0], device=xp.device)
            self._wins = [t.as_tensor(v, device=xp.device)[: self.n] for v in self._views[1:]]
        else:                                                      # host simulator
            def arr(ptr, count, shape):
                return np.frombuffer((C.c_char * (count * 4)).from_address(ptr), dtype=np.uint32, count=count).reshape(shape)
            self.flags = arr(self._own.value, FLAG_WORDS, (FLAG_WORDS,))
            self._wins = [arr(self._own.value + FLAG_WORDS * 4 + par * self.row_bytes, max(self.n, 1) * 6, (max(self.n, 1), 6))[: self.n]
                          for par in (0, 1)]

    # ------------------------------------------------------------------
    def wins_buffer(self, zero: bool = True):
        """u32[n, 6] of the NEXT epoch, zeroed unless the producer overwrites it (Engine.mc_predict(wins=...) clears it in its own kernel)"""
        w = self._wins[(self.epoch + 1) & 1]
        if not zero:
            return w
        if hasattr(w, "zero_"):
            w.zero_()
        else:
            w[...] = 0
        return w

    def best_action(self, state, want_total: bool = True):
        """-> (action int8[n, 2], total u32[n, 6] or None): summed over all ranks, identical on all ranks"""
        e = self.e
        self.epoch += 1
        n = e._n(state)
        assert n == self.n
        a = e.xp.empty((n, 2), "i8")
        tot = e.xp.empty((n, 6), "u32") if want_total else None
        e._chk(self.lib.ewn_mc_best_action_peers(e.xp.ptr(state), self._win_ptrs[self.epoch & 1], self._flag_ptrs, self.rank, self.world,
                                                 self.epoch, self.timeout_ms, e.xp.ptr(tot), e.xp.ptr(a), n, e.N, e.L, e.xp.stream()),
               "ewn_mc_best_action_peers")
        return a, tot

    def check(self):
        """raise if a wait for the peers ever timed out (synchronises the device)"""
        f = self.flags
        err = int(f[ERR_WORD].item()) if hasattr(f, "item") and not isinstance(f, np.ndarray) else int(f[ERR_WORD])
        if err:
            raise RuntimeError(f"rank {self.rank}: peers did not publish epoch {err} within {self.timeout_ms} ms")

    def close(self):
        if getattr(self, "_own", None) is None:
            return
        xp = self.e.xp
        if getattr(xp, "torch", None) is not None:
            xp.torch.cuda.synchronize(xp.device)
        self.flags, self._wins, self._views = None, None, None
        with xp.guard():
            for p in self._opened:
                self.lib.ewn_peer_close(p)
            if self._own.value:
                self.lib.ewn_peer_free(self._own)
        self._opened, self._own = [], None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
