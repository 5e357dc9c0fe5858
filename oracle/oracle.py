"""TEST INFRASTRUCTURE ONLY -- ctypes/numpy front-end of the CPU oracle (oracle/ewn_oracle.c).

Importers allowed: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline leg and --impl reference).
The product package ewn_gym_b200 must never import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libewn_oracle.so")

H_HYBRID, H_MIN_DIST, H_TWO_MIN_DIST, H_ATTK = 0, 1, 2, 3
HEURISTICS = {"hybrid": 0, "min_dist": 1, "two_min_dist": 2, "attk": 3}
TOP_LEFT, BOTTOM_RIGHT = 1, 2
CODE_ONGOING, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_OPP_ILLEGAL, CODE_AGENT_LOST, CODE_ILLEGAL_TOLERATED = range(6)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ewn_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        cmd = ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fvisibility=hidden", "-o", _SO, src, "-lm"]
        subprocess.run(cmd, check=True, cwd=_HERE)
    return _SO


class StepCfg(C.Structure):
    _fields_ = [
        ("N", C.c_int32), ("L", C.c_int32), ("reward", C.c_double), ("agent_random", C.c_int32),
        ("autoreset", C.c_int32), ("shaped", C.c_int32), ("illegal_move_reward", C.c_double),
        ("seed", C.c_uint64), ("step", C.c_uint64), ("env_id0", C.c_uint64),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_random_games_range.restype = C.c_long
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _b16(board, N):
    b = np.ascontiguousarray(board, dtype=np.int16)
    return b.reshape(-1, N * N)


def philox(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32).copy()
    k = np.asarray(key, dtype=np.uint32).copy()
    out = np.zeros(4, dtype=np.uint32)
    lib().orc_philox(_p(c), _p(k), _p(out))
    return out


def init_boards(n, N=5, L=3):
    b = np.zeros((n, N * N), dtype=np.int16)
    lib().orc_init_boards(_p(b), C.c_long(n), N, L)
    return b.reshape(n, N, N)


def check_win(board, N=5, L=3):
    b = _b16(board, N)
    out = np.zeros(len(b), dtype=np.uint8)
    lib().orc_check_win(_p(b), C.c_long(len(b)), N, L, _p(out))
    return out


def legal(board, dice, player, N=5, L=3):
    """-> (mask u8[n] with bit larger*3+dir, ordered list int8[n,6,2] (-1 padded), count int32[n])."""
    b = _b16(board, N)
    n = len(b)
    d, pl = _i32(dice).reshape(n), _i32(np.broadcast_to(player, (n,)))
    mask = np.zeros(n, np.uint8)
    lst = np.zeros((n, 6, 2), np.int8)
    cnt = np.zeros(n, np.int32)
    lib().orc_legal(_p(b), _p(d), _p(pl), C.c_long(n), N, L, _p(mask), _p(lst), _p(cnt))
    return mask, lst, cnt


def ply(board, dice, player, larger, direction, N=5, L=3):
    """One ply (find_cube_to_move + execute_move + check_win). -> (board', valid, win, moved_cube)."""
    b = _b16(board, N).copy()
    n = len(b)
    d, pl = _i32(dice).reshape(n), _i32(np.broadcast_to(player, (n,)))
    la, di = _i32(larger).reshape(n), _i32(direction).reshape(n)
    valid, win, mc = np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.int32)
    lib().orc_ply(_p(b), _p(d), _p(pl), _p(la), _p(di), C.c_long(n), N, L, _p(valid), _p(win), _p(mc))
    return b.reshape(n, N, N), valid, win, mc


def step_replay(board, dice, action, dice_opp, opp_action, dice_next, N=5, L=3, reward=1.0, player=None):
    """envs/ewn.py:436-486 with injected dice / opponent answer. Returns dict of arrays."""
    b = _b16(board, N).copy()
    n = len(b)
    d = _i32(dice).reshape(n).copy()
    pl = np.full(n, TOP_LEFT, np.int32) if player is None else _i32(player).copy()
    a, oa = _i32(action).reshape(n, 2), _i32(opp_action).reshape(n, 2)
    do, dn = _i32(dice_opp).reshape(n), _i32(dice_next).reshape(n)
    rew, te, tr, code = np.zeros(n), np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.int32)
    lib().orc_step_replay(_p(b), _p(d), _p(pl), _p(a), _p(do), _p(oa), _p(dn), C.c_long(n), N, L, C.c_double(reward),
                          _p(rew), _p(te), _p(tr), _p(code))
    return dict(board=b.reshape(n, N, N), dice=d, player=pl, reward=rew, terminated=te, truncated=tr, code=code)


def evaluate(board, N=5, L=3, heuristic="hybrid"):
    b = _b16(board, N)
    out = np.zeros(len(b))
    lib().orc_eval(_p(b), C.c_long(len(b)), N, L, HEURISTICS[heuristic], _p(out))
    return out


def _chunks(n, parts):
    parts = max(1, min(parts, n))
    edges = np.linspace(0, n, parts + 1).astype(np.int64)
    return [(int(edges[i]), int(edges[i + 1])) for i in range(parts) if edges[i + 1] > edges[i]]


def expectiminimax(board, dice, depth, N=5, L=3, heuristic="hybrid", pruned=True, threads=1):
    """classical_policies/minimax.py:89-93. -> (value f64[n], action int8[n,2] (-1 = None), leaves int64[n])."""
    b = _b16(board, N)
    n = len(b)
    d = _i32(dice).reshape(n)
    val, act, leaves = np.zeros(n), np.zeros((n, 2), np.int8), np.zeros(n, np.int64)

    def run(lo, hi):
        lib().orc_expectiminimax(_p(b[lo:hi]), _p(d[lo:hi]), C.c_long(hi - lo), N, L, depth, HEURISTICS[heuristic], int(pruned),
                                 _p(val[lo:hi]), _p(act[lo:hi]), _p(leaves[lo:hi]))

    _parallel(run, n, threads)
    return val, act, leaves


def _parallel(run, n, threads):
    if threads <= 1 or n <= 1:
        run(0, n)
        return
    with ThreadPoolExecutor(threads) as ex:  # ctypes drops the GIL during the foreign call
        list(ex.map(lambda r: run(*r), _chunks(n, threads)))


def mc_root_wins(board, dice, root_ids, seed, first, count, N=5, L=3, threads=1):
    """Flat Monte-Carlo (mcts.py:47-69) with the engine's Philox schedule.
    -> (wins uint32[n,6] by list position, n_children int32[n], total plies)."""
    b = _b16(board, N)
    n = len(b)
    d = _i32(dice).reshape(n)
    rid = np.ascontiguousarray(root_ids, dtype=np.uint32).reshape(n)
    wins, nch = np.zeros((n, 6), np.uint32), np.zeros(n, np.int32)
    plies = []

    def run(lo, hi):
        tp = C.c_int64(0)
        lib().orc_mc_root_wins(_p(b[lo:hi]), _p(d[lo:hi]), _p(rid[lo:hi]), C.c_long(hi - lo), N, L, C.c_uint64(seed),
                               C.c_uint32(first), C.c_uint32(count), _p(wins[lo:hi]), _p(nch[lo:hi]), C.byref(tp))
        plies.append(tp.value)

    _parallel(run, n, threads)
    return wins, nch, int(sum(plies))


def playouts(board, root_ids, seed, first, count, N=5, L=3, threads=1):
    """Random playouts from given states, BOTTOM_RIGHT to move (mcts.py:24-41). -> (TL wins uint32[n], total plies)."""
    b = _b16(board, N)
    n = len(b)
    rid = np.ascontiguousarray(root_ids, dtype=np.uint32).reshape(n)
    wins = np.zeros(n, np.uint32)
    plies = []

    def run(lo, hi):
        tp = C.c_int64(0)
        lib().orc_playouts(_p(b[lo:hi]), _p(rid[lo:hi]), C.c_long(hi - lo), N, L, C.c_uint64(seed), C.c_uint32(first),
                           C.c_uint32(count), _p(wins[lo:hi]), C.byref(tp))
        plies.append(tp.value)

    _parallel(run, n, threads)
    return wins, int(sum(plies))


def reset_philox(n, seed, step, env_id0=0, N=5, L=3):
    b = np.zeros((n, N * N), np.int16)
    d = np.zeros(n, np.int32)
    lib().orc_reset_philox(_p(b), _p(d), C.c_long(n), N, L, C.c_uint64(seed), C.c_uint64(step), C.c_uint64(env_id0))
    return b.reshape(n, N, N), d


def step_philox(board, dice, action, seed, step, env_id0=0, N=5, L=3, reward=1.0, agent_random=False, autoreset=True,
                shaped=False, illegal_move_reward=-1.0, tolerance=None, prev_score=None, threads=1):
    """CPU twin of the engine's free-running step kernel. board/dice are updated copies in the result."""
    b = _b16(board, N).copy()
    n = len(b)
    d = _i32(dice).reshape(n).copy()
    a = None if action is None else _i32(action).reshape(n, 2)
    rew, te, tr, code = np.zeros(n), np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.int32)
    tb, td = np.zeros((n, N * N), np.int16), np.zeros(n, np.int32)
    tol = None if tolerance is None else _i32(tolerance).copy()
    ps = None if prev_score is None else np.ascontiguousarray(prev_score, dtype=np.float64).copy()
    aout = np.zeros((n, 2), np.int8)

    def run(lo, hi):
        cfg = StepCfg(N, L, reward, int(agent_random), int(autoreset), int(shaped), illegal_move_reward, seed, step, env_id0 + lo)
        lib().orc_step_philox(_p(b[lo:hi]), _p(d[lo:hi]), None if a is None else _p(a[lo:hi]), C.c_long(hi - lo), C.byref(cfg),
                              _p(rew[lo:hi]), _p(te[lo:hi]), _p(tr[lo:hi]), _p(code[lo:hi]), _p(tb[lo:hi]), _p(td[lo:hi]),
                              None if tol is None else _p(tol[lo:hi]), None if ps is None else _p(ps[lo:hi]), _p(aout[lo:hi]))

    _parallel(run, n, threads)
    return dict(board=b.reshape(n, N, N), dice=d, reward=rew, terminated=te, truncated=tr, code=code,
                term_board=tb.reshape(n, N, N), term_dice=td, tolerance=tol, prev_score=ps, agent_action=aout)


def random_games(n_games, seed, N=5, L=3, threads=1):
    """eval_random.py:17-51 loop with Philox dice. -> (wins, agent steps, plies)."""
    res = []

    def run(lo, hi):
        # games are keyed by env id, so ranges need an id offset: run each chunk as its own id range
        st, pl = C.c_int64(0), C.c_int64(0)
        w = lib().orc_random_games_range(C.c_long(lo), C.c_long(hi), N, L, C.c_uint64(seed), C.byref(st), C.byref(pl))
        res.append((w, st.value, pl.value))

    _parallel(run, n_games, threads)
    return tuple(int(sum(x[i] for x in res)) for i in range(3))
