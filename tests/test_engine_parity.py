"""Parity of the engine (C ABI) with (a) golden vectors recorded from the unmodified reference and (b) the CPU oracle.

Each test runs twice: on "hostsim" (the kernel bodies compiled for the host, CPU tier) and on "cuda" (the product
library on a B200, `-m gpu`).  Integer / index results must be bit-exact; fp64 values are compared by bit pattern.
"""
import numpy as np
import pytest

from oracle import oracle as O
from tests.backends import BACKENDS, make_engine, to_np
from tests.golden_util import game_traces, load

pytestmark = pytest.mark.parametrize("backend", BACKENDS)


def bits(x):
    return np.ascontiguousarray(to_np(x), dtype=np.float64).view(np.uint64)


# --------------------------------------------------------------------------------------------------- codec + rules
@pytest.mark.parametrize("name", ["rules_5x5.npz", "rules_7x7.npz", "rules_6x6_L4.npz", "rules_7x7_L5.npz"])
def test_rules_golden(backend, name):
    """restore_env_with_obs round trip, check_win, all four heuristics, every (die, action) ply, ordered legal lists."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    e = make_engine(backend, N, L)
    boards = z["boards"].astype(np.int16)
    st = e.encode(boards, z["dice"].astype(np.int32))
    b2, d2 = e.decode(st)
    assert np.array_equal(to_np(b2), boards) and np.array_equal(to_np(d2), z["dice"].astype(np.int64))
    assert np.array_equal(to_np(e.check_win(st)), z["win"])
    for h in ("hybrid", "min_dist", "two_min_dist", "attk"):
        assert np.array_equal(bits(e.evaluate(st, h)), z[f"eval_{h}"]), h
    pr = z["ply_rows"].astype(np.int32)          # state,player,dice,larger,dir,valid,win_after,moved_cube
    for player in (1, 2):
        rows = pr[pr[:, 1] == player]
        s = e.encode(boards[rows[:, 0]], rows[:, 2], np.full(len(rows), player, np.int32))
        valid, win, moved = e.ply(s, rows[:, 3:5], player=0)
        assert np.array_equal(to_np(valid), rows[:, 5]) and np.array_equal(to_np(win), rows[:, 6])
        assert np.array_equal(to_np(moved), rows[:, 7])
        nb, _, pl = e.decode(s, want_player=True)
        assert np.array_equal(to_np(nb), z["ply_boards"].astype(np.int16)[pr[:, 1] == player])
        # the side to move flips exactly when the move was valid
        assert np.array_equal(to_np(pl) != player, rows[:, 5] == 1)
    lr = z["legal_rows"].astype(np.int32)        # state,player,dice,mask,count,list[6][2]
    for player in (1, 2):
        rows = lr[lr[:, 1] == player]
        s = e.encode(boards[rows[:, 0]], rows[:, 2])
        mask, lst, cnt = e.legal(s, player=player)
        assert np.array_equal(to_np(mask), rows[:, 3]) and np.array_equal(to_np(cnt), rows[:, 4])
        assert np.array_equal(to_np(lst).reshape(len(rows), 12), rows[:, 5:17])


def test_initial_layout_and_reset(backend):
    for N, L in ((5, 3), (7, 3), (4, 2), (9, 3), (3, 1), (6, 2)):
        e = make_engine(backend, N, L)
        st = e.reset(1000, seed=77, step=3, env_id0=12345)
        b, d = e.decode(st)
        ob, od = O.reset_philox(1000, 77, 3, 12345, N, L)
        assert np.array_equal(to_np(b), ob) and np.array_equal(to_np(d), od)
        assert to_np(d).min() >= 1 and to_np(d).max() <= L * (L + 1) // 2
    e = make_engine(backend, 5, 3)
    b, _ = e.decode(e.reset_dice(np.array([5], np.int32)))
    assert to_np(b)[0].tolist() == [[1, 2, 4, 0, 0], [3, 5, 0, 0, 0], [6, 0, 0, 0, -6], [0, 0, 0, -5, -3], [0, 0, -4, -2, -1]]


# --------------------------------------------------------------------------------------------------- config 1: replay
@pytest.mark.parametrize("name", ["random_games_5x5.npz", "random_games_7x7.npz", "random_games_4x4_L2.npz", "random_games_7x7_L4.npz",
                                  "random_games_9x9_L7.npz"])
def test_replay_reference_games(backend, name):
    """Config 1: every env.step of the reference's eval_random loop, all games in one batch, dice/opponent injected.
    Checks legal mask before, board / reward / flags / dice after every step; 5x5: 5,215 wins over 108,978 steps."""
    g = game_traces(name)
    N, L, S, t = g["N"], g["L"], g["steps"], g["t"]
    e = make_engine(backend, N, L)
    n_games = len(g["first_row"])
    state = e.reset_dice(g["init_dice"])
    wins = 0
    for step in range(int(t.max()) + 1):
        rows = np.flatnonzero(t == step)
        gi = g["gidx"][rows]
        s = S[rows]
        sub = e.xp.asarray(to_np(state)[gi], "u64")
        b0, d0 = e.decode(sub)
        assert np.array_equal(to_np(d0), s[:, 1]), "agent dice"
        mask, _, _ = e.legal(sub, player=1, want_list=False)
        assert np.array_equal(to_np(mask), s[:, 10].astype(np.uint8)), "legal mask before step"
        out = e.step(sub, e.xp.asarray(s[:, 2:4], "i64"), inject=(s[:, 4], s[:, 5:7], s[:, 7]), autoreset=False,
                     want=("reward_f64", "terminated", "truncated", "code", "obs_board", "obs_dice"))
        assert np.array_equal(to_np(out["obs_board"]), g["boards"][rows]), f"board after step {step}"
        assert np.array_equal(to_np(out["reward_f64"]), g["rewards"][rows])
        assert np.array_equal(to_np(out["terminated"]), s[:, 8]) and np.array_equal(to_np(out["truncated"]), s[:, 9])
        ongoing = s[:, 8] == 0
        assert np.array_equal(to_np(out["obs_dice"])[ongoing], s[ongoing, 7])
        wins += int((to_np(out["code"]) == O.CODE_AGENT_WON).sum())
        full = to_np(state).copy()
        full[gi] = to_np(sub)
        state = e.xp.asarray(full, "u64")
    assert wins == int((g["game_rewards"] > 0).sum())
    if name == "random_games_5x5.npz":
        assert wins == 5215 and len(S) == 108978 and n_games == 10000


# --------------------------------------------------------------------------------------------------- config 4: search
@pytest.mark.parametrize("name", ["minimax_5x5.npz", "minimax_7x7.npz"])
def test_expectiminimax_golden_bits(backend, name):
    """Value bits and root action of classical_policies/minimax.py (pruned semantics), depth 1..5, all heuristics."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    e = make_engine(backend, N, L)
    boards, dice = z["boards"].astype(np.int16), z["dice"].astype(np.int32)
    for key in z.files:
        if not key.startswith("value_"):
            continue
        tail = key[len("value_"):]
        heuristic = "hybrid" if tail.startswith("d") and tail[1:].isdigit() else tail.rsplit("_d", 1)[0]
        depth = int(key.rsplit("d", 1)[1])
        want_v, want_a = z[key], z["action_" + tail]
        n = len(want_v)
        if backend == "hostsim" and depth >= 5:
            n = min(n, 48)                       # keep the CPU tier quick; the GPU tier runs everything
        st = e.encode(boards[:n], dice[:n])
        v, a = e.expectiminimax(st, depth, heuristic)
        assert np.array_equal(bits(v), want_v[:n]), key
        assert np.array_equal(to_np(a), want_a[:n]), key


@pytest.mark.parametrize("N", [5, 7])
def test_expectiminimax_fused_equals_replay_path(backend, N):
    """Depth 3 / 4: the fused on-chip kernel (closed-form prefix rule, six lanes per root) against the record + replay kernels that
    restate minimax.py:19-73 literally -- value bits and actions, every heuristic, pruned and unpruned, on positions from random play
    (ragged root count so that partly filled warps and idle lane groups are exercised)."""
    e = make_engine(backend, N, 3)
    n = 333 if backend == "hostsim" else 20011
    st = e.reset(n, seed=5)
    for s in range(1, 7):
        e.step(st, None, seed=5, step=s, agent_random=True, autoreset=(s % 3 != 0), want=())     # some roots are finished games
    for depth in (3, 4):
        for heuristic in ("hybrid", "min_dist", "two_min_dist", "attk"):
            for pruned in (True, False):
                v, a = e.expectiminimax(st, depth, heuristic, pruned=pruned)
                rv, ra = e.expectiminimax(st, depth, heuristic, pruned=pruned, replay_path=True)
                assert np.array_equal(bits(v), bits(rv)), (depth, heuristic, pruned)
                assert np.array_equal(to_np(a), to_np(ra)), (depth, heuristic, pruned)


def test_expectiminimax_vs_oracle_unpruned_and_deep(backend):
    """Oracle cross-check beyond the golden set: unpruned mode, depth 6, finished roots (action None)."""
    z = load("minimax_5x5.npz")
    boards, dice = z["boards"].astype(np.int16), z["dice"].astype(np.int32)
    e = make_engine(backend, 5, 3)
    for depth, n, pruned in ((3, 400, False), (4, 400, False), (6, 24 if backend == "hostsim" else 160, True)):
        st = e.encode(boards[:n], dice[:n])
        v, a = e.expectiminimax(st, depth, "hybrid", pruned=pruned)
        ov, oa, _ = O.expectiminimax(boards[:n], dice[:n], depth, 5, 3, "hybrid", pruned=pruned, threads=8)
        assert np.array_equal(bits(v), ov.view(np.uint64)) and np.array_equal(to_np(a), oa), (depth, pruned)
    fin = O.check_win(boards) == 1
    assert fin.sum() > 50
    st = e.encode(boards[fin], dice[fin])
    v, a = e.expectiminimax(st, 4)
    assert (to_np(a) == -1).all() and set(np.unique(to_np(v))) <= {10.0, -10.0}
    # SURVEY 8c probe values at the initial position, die 5
    import struct
    want = {1: "bfd5555555555554", 2: "bfd5555555555554", 3: "bfdbbbbbbbbbbbbb", 4: "bfdbbbbbbbbbbbbb", 5: "bfd6eeeeeeeeeeed"}
    st = e.reset_dice(np.array([5], np.int32))
    for d, hx in want.items():
        v, a = e.expectiminimax(st, d)
        assert struct.pack(">d", float(to_np(v)[0])).hex() == hx and to_np(a)[0].tolist() == [0, 2]


# --------------------------------------------------------------------------------------------------- config 2: free-running step
@pytest.mark.parametrize("N,L", [(5, 3), (7, 3), (4, 2), (7, 4), (9, 7)])
@pytest.mark.parametrize("agent_random", [True, False])
def test_step_philox_vs_oracle(backend, N, L, agent_random):
    """Config 2 on a 4,096-env subset: Philox dice, in-kernel random opponent, auto-reset, observation + terminal
    observation rows, 40 consecutive steps; supplied actions are uniform over [2]x[3] (so illegal moves occur)."""
    n, seed, env_id0 = 4096 + 37, 9487, 1 << 33          # ragged tail warp, env ids beyond 32 bits
    e = make_engine(backend, N, L)
    state = e.reset(n, seed=seed, step=0, env_id0=env_id0)
    ob, od = O.reset_philox(n, seed, 0, env_id0, N, L)
    rng = np.random.default_rng(5)
    wins = losses = illegal = 0
    stats = e.xp.zeros((4,), "i64")
    for t in range(1, 41):
        act = None if agent_random else np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
        prev_state = to_np(state).copy()
        out = e.step(state, None if act is None else e.xp.asarray(act, "i64"), seed=seed, step=t, env_id0=env_id0, agent_random=agent_random,
                     autoreset=True, stats=stats, want=("reward_f32", "reward_f64", "terminated", "truncated", "code", "action_out", "obs_board",
                                           "obs_dice", "term_board", "term_dice"))
        ref = O.step_philox(ob, od, act, seed, t, env_id0, N, L, agent_random=agent_random, autoreset=True, threads=4)
        assert np.array_equal(to_np(out["code"]), ref["code"]), t
        assert np.array_equal(to_np(out["obs_board"]), ref["board"]) and np.array_equal(to_np(out["obs_dice"]), ref["dice"])
        assert np.array_equal(to_np(out["reward_f64"]), ref["reward"]) and np.array_equal(to_np(out["reward_f32"]), ref["reward"].astype(np.float32))
        assert np.array_equal(to_np(out["terminated"]), ref["terminated"]) and np.array_equal(to_np(out["truncated"]), ref["truncated"])
        done = ref["terminated"] == 1
        assert np.array_equal(to_np(out["term_board"])[done], ref["term_board"][done])
        assert np.array_equal(to_np(out["term_dice"])[done], ref["term_dice"][done])
        ao = to_np(out["action_out"])
        if agent_random:
            assert np.array_equal(np.stack([ao // 3, ao % 3], 1), ref["agent_action"])
        b2, d2 = e.decode(state)                                     # the stored state is what the observation shows
        assert np.array_equal(to_np(b2), ref["board"]) and np.array_equal(to_np(d2), ref["dice"])
        if t % 4 == 0:                                               # compacted terminal observations: same rows, any order
            s3 = e.xp.asarray(prev_state, "u64")
            cnt, tidx = e.xp.zeros((1,), "i32"), e.xp.zeros((n,), "i32")
            o3 = e.step(s3, None if act is None else e.xp.asarray(act, "i64"), seed=seed, step=t, env_id0=env_id0, agent_random=agent_random,
                        autoreset=True, term_compact=(cnt, tidx), want=("term_board", "term_dice"))
            k = int(to_np(cnt)[0])
            assert k == int(done.sum())
            order = np.argsort(to_np(tidx)[:k])
            assert np.array_equal(to_np(tidx)[:k][order], np.flatnonzero(done))
            assert np.array_equal(to_np(o3["term_board"])[:k][order], ref["term_board"][done])
            assert np.array_equal(to_np(o3["term_dice"])[:k][order], ref["term_dice"][done])
            assert np.array_equal(to_np(s3), to_np(state))
        ob, od = ref["board"], ref["dice"]
        wins += int((ref["code"] == O.CODE_AGENT_WON).sum()); losses += int((ref["code"] == O.CODE_AGENT_LOST).sum())
        illegal += int((ref["code"] == O.CODE_AGENT_ILLEGAL).sum())
    thr = 300 if L <= 3 else 40                    # games on the big layouts last several times longer
    assert wins > thr and losses > thr
    assert to_np(stats).tolist()[:2] == [wins, losses] and to_np(stats)[2] == illegal and to_np(stats)[3] == wins + losses + illegal
    assert (illegal == 0) if agent_random else (illegal > thr)


def test_random_action_matches_in_kernel_agent(backend):
    """RandomAgent.predict as a separate call == the in-kernel random agent (same Philox word)."""
    e = make_engine(backend, 5, 3)
    n, seed = 3000, 11
    s1 = e.reset(n, seed=seed)
    s2 = e.xp.asarray(to_np(s1).copy(), "u64")
    for t in range(1, 25):
        a = e.random_action(s1, seed=seed, step=t)
        mask, _, _ = e.legal(s1, player=1, want_list=False)
        an = to_np(a)
        assert ((to_np(mask) >> (an[:, 0] * 3 + an[:, 1])) & 1).all(), "sampled action is legal"
        o1 = e.step(s1, a, seed=seed, step=t, want=("code",))
        o2 = e.step(s2, None, seed=seed, step=t, agent_random=True, want=("code", "action_out"))
        assert np.array_equal(to_np(o1["code"]), to_np(o2["code"])) and np.array_equal(to_np(s1), to_np(s2))
        assert np.array_equal(an[:, 0] * 3 + an[:, 1], to_np(o2["action_out"]))


def test_action_dtypes_and_out_of_range(backend):
    e = make_engine(backend, 5, 3)
    n = 512
    rng = np.random.default_rng(3)
    act = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1)
    res = []
    for dt, arr in (("i64", act.astype(np.int64)), ("i32", act.astype(np.int32)), ("u8", (act[:, 0] * 3 + act[:, 1]).astype(np.uint8))):
        st = e.reset(n, seed=4)
        out = e.step(st, e.xp.asarray(arr, dt), seed=4, step=1, want=("code",))
        res.append((to_np(st).copy(), to_np(out["code"]).copy()))
    assert all(np.array_equal(res[0][0], r[0]) and np.array_equal(res[0][1], r[1]) for r in res[1:])
    # outside MultiDiscrete([2,3]): the reference's execute_move matches no direction branch, leaves the cube in place and
    # reports a VALID move (envs/ewn.py:229-247); only `action[0] == 1` means larger (:439).  Same here, bit for bit.
    odd = np.array([[0, 3], [1, -1], [0, 7], [5, 1], [1, 2], [0, -9]], np.int64)
    st = e.reset(len(odd), seed=4)
    b0, d0 = (to_np(x) for x in e.decode(st))
    out = e.step(st, e.xp.asarray(odd, "i64"), seed=4, step=1, want=("code", "obs_board", "obs_dice"))
    ref = O.step_philox(b0, d0, odd, 4, 1)
    assert np.array_equal(to_np(out["code"]), ref["code"]) and np.array_equal(to_np(out["obs_board"]), ref["board"])
    assert np.array_equal(to_np(out["obs_dice"]), ref["dice"])
    assert O.CODE_AGENT_ILLEGAL not in to_np(out["code"])[[0, 1, 2, 5]].tolist()


# --------------------------------------------------------------------------------------------------- config 5: shaped env
@pytest.mark.parametrize("name", ["shaped_5x5.npz", "shaped_7x7.npz", "shaped_7x7_L4.npz"])
def test_shaped_env_traces(backend, name):
    """envs/training_ewn.py:40-99 recorded from the reference (tolerance budget, delta-evaluate reward, resets)."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    e = make_engine(backend, N, L)
    rows, rewards, boards = z["rows"], z["rewards"], z["boards"].astype(np.int16)
    n_env, n_steps = rows.shape[:2]
    state = e.encode(z["init_board"].astype(np.int16), z["init_dice"].astype(np.int32))
    prev = e.xp.asarray(z["init_prev"].view(np.float64).copy(), "f64")
    tol = e.xp.asarray(np.full(n_env, int(z["tolerance0"]), np.int32), "i32")
    assert np.array_equal(bits(e.evaluate(state)), z["init_prev"])
    for t in range(n_steps):
        r = rows[:, t]
        _, d0 = e.decode(state)
        assert np.array_equal(to_np(d0), r[:, 0])
        # after a terminal step the reference's next die comes from reset(): column 9
        nxt = np.where(r[:, 7] == 1, r[:, 9], r[:, 6]).astype(np.int32)
        out = e.step(state, e.xp.asarray(r[:, 1:3], "i64"), inject=(r[:, 3], r[:, 4:6], nxt), reward=float(z["goal_reward_effective"]),
                     shaped=True, illegal_move_reward=-1.0, tolerance=tol, prev_score=prev, autoreset=True,
                     want=("reward_f64", "terminated", "truncated", "term_board", "obs_board"))
        assert np.array_equal(bits(out["reward_f64"]), rewards[:, t]), f"reward step {t}"
        assert np.array_equal(to_np(out["terminated"]), r[:, 7]) and np.array_equal(to_np(out["truncated"]), r[:, 8])
        done = r[:, 7] == 1
        shown = np.where(done[:, None, None], to_np(out["term_board"]), to_np(out["obs_board"]))
        assert np.array_equal(shown, boards[:, t])
        assert np.array_equal(to_np(tol), r[:, 10])


# --------------------------------------------------------------------------------------------------- config 3: playouts
@pytest.mark.parametrize("N", [5, 7])
def test_playouts_exact_vs_oracle(backend, N):
    """Same Philox schedule => identical win counts per state, for any split of the playout index range."""
    z = load(f"mc_sim_{N}x{N}.npz")
    boards = z["boards"].astype(np.int16)
    n = len(boards)
    e = make_engine(backend, N, 3)
    st = e.encode(boards, None, np.full(n, 2, np.int32))
    rid = np.arange(100, 100 + n, dtype=np.uint32)
    count = 96 if backend == "hostsim" else 1000
    plies = e.xp.zeros((1,), "i64")
    w = e.playouts(st, seed=42, first=7, count=count, root_ids=rid, plies=plies)
    ow, oplies = O.playouts(boards, rid, 42, 7, count, N, 3, threads=8)
    assert np.array_equal(to_np(w).astype(np.uint32), ow) and int(to_np(plies)[0]) == oplies
    w2 = e.playouts(st, seed=42, first=7, count=count // 3, root_ids=rid)
    w2 = e.playouts(st, seed=42, first=7 + count // 3, count=count - count // 3, root_ids=rid, wins=w2)
    assert np.array_equal(to_np(w2), to_np(w)), "index-range split (the multi-GPU sharding) changes nothing"


def test_mc_root_wins_exact_vs_oracle(backend):
    z = load("minimax_5x5.npz")
    n = 64 if backend == "hostsim" else 600
    boards, dice = z["boards"][:n].astype(np.int16), z["dice"][:n].astype(np.int32)
    e = make_engine(backend, 5, 3)
    st = e.encode(boards, dice)
    rid = np.arange(n, dtype=np.uint32) * 3 + 1
    count = 40 if backend == "hostsim" else 256
    w, nch = e.mc_root_wins(st, seed=9, first=0, count=count, root_ids=rid)
    ow, onch, _ = O.mc_root_wins(boards, dice, rid, 9, 0, count, 5, 3, threads=8)
    fin = O.check_win(boards) == 1
    onch = np.where(fin, 0, onch)                       # finished roots have no children here (predict returns None)
    ow[fin] = 0
    assert np.array_equal(to_np(nch), onch) and np.array_equal(to_np(w).astype(np.uint32), ow)
    a = to_np(e.mc_best_action(st, w))
    _, lst, _ = O.legal(boards, dice, 1)
    for i in range(n):
        if fin[i]:
            assert a[i].tolist() == [-1, -1]
        else:
            assert a[i].tolist() == lst[i, int(np.argmax(ow[i, :onch[i]]))].tolist()     # np.argmax: first maximum (mcts.py:68)


@pytest.mark.parametrize("N", [5, 7])
def test_playout_win_rates_vs_reference(backend, N):
    """Statistical parity with MctsAgent.simulate (python `random`, not reproducible): per-state TL win rate of the
    engine must lie within a 4.5-sigma two-sample binomial bound of the reference's estimate (Bonferroni over <=48
    states: overall false-alarm probability < 1e-3)."""
    z = load(f"mc_sim_{N}x{N}.npz")
    boards, ref_w, ref_n = z["boards"].astype(np.int16), z["wins"].astype(np.float64), float(z["playouts"])
    n = len(boards)
    e = make_engine(backend, N, 3)
    st = e.encode(boards, None, np.full(n, 2, np.int32))
    count = 3000 if backend == "hostsim" else 100000
    w = to_np(e.playouts(st, seed=2024, first=0, count=count)).astype(np.float64)
    p, q = w / count, ref_w / ref_n
    pool = (w + ref_w) / (count + ref_n)
    se = np.sqrt(np.maximum(pool * (1 - pool), 1e-9) * (1 / count + 1 / ref_n))
    assert (np.abs(p - q) <= 4.5 * se + 1e-12).all(), np.c_[p, q, se][np.abs(p - q) > 4.5 * se]


def test_device_step_counter(backend):
    """ewn_step_args.step_dev: step = baked step + *step_dev, so a captured graph can be replayed with fresh Philox counters."""
    e = make_engine(backend, 5, 3)
    n = 2048
    rng = np.random.default_rng(11)
    act = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
    sa, sb = e.reset(n, seed=8), e.reset(n, seed=8)
    ctr = e.xp.asarray(np.zeros(1, np.int64), "i64")
    want = ("reward_f32", "terminated", "obs_board", "obs_dice")
    for t in range(1, 7):
        ra = e.step(sa, e.xp.asarray(act, "i64"), seed=8, step=t, want=want)
        rb = e.step(sb, e.xp.asarray(act, "i64"), seed=8, step=1 + (t - 1) % 3, step_dev=ctr, want=want)   # "graph" of 3 baked steps
        if t % 3 == 0:
            e.counter_add(ctr, 3)
        for k in want:
            assert np.array_equal(to_np(ra[k]), to_np(rb[k])), (t, k)
    assert int(to_np(ctr)[0]) == 6


@pytest.mark.parametrize("N,L", [(3, 1), (4, 1), (4, 2), (5, 1), (5, 2), (6, 3), (8, 3), (9, 3), (9, 1), (6, 2), (6, 4), (7, 4), (7, 5), (8, 6), (9, 7),
                                 (9, 4)])
def test_every_supported_board_size(backend, N, L):
    """3 <= board_size <= 9 with cube_layer < board_size - 1 (envs/ewn.py:47) up to the maximum 7 (28 cubes per side): the three codecs
    (one word up to 5x5, two words above, the generic byte-per-cube codec for cube_layer >= 4),
    the run-time-N kernel instantiations, VecEnv surface + generic surface, legal lists, and search / playouts where the reference
    allows them (six-face die: cube_layer 3)."""
    e = make_engine(backend, N, L)
    n, seed = 1500, 17
    st = e.reset(n, seed=seed)
    ob, od = O.reset_philox(n, seed, 0, 0, N, L)
    assert np.array_equal(to_np(e.decode(st)[0]), ob)
    rng = np.random.default_rng(N * 10 + L)
    for t in range(1, 13):
        act = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
        want = ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice") + (("code",) if t % 2 else ())
        out = e.step(st, e.xp.asarray(act, "i64"), seed=seed, step=t, want=want)
        ref = O.step_philox(ob, od, act, seed, t, 0, N, L)
        assert np.array_equal(to_np(out["obs_board"]), ref["board"]) and np.array_equal(to_np(out["obs_dice"]), ref["dice"]), t
        assert np.array_equal(to_np(out["terminated"]), ref["terminated"]) and np.array_equal(to_np(out["truncated"]), ref["truncated"])
        assert np.array_equal(to_np(out["reward_f32"]), ref["reward"].astype(np.float32))
        ob, od = ref["board"], ref["dice"]
        for player in (1, 2):
            mask, lst, cnt = (to_np(x) for x in e.legal(st, player))
            rm, rl, rc = O.legal(ob, od, player, N, L)
            assert np.array_equal(cnt, rc) and np.array_equal(lst, rl) and np.array_equal(mask, rm)
    if L == 3:
        m = 200
        v, a = e.expectiminimax(e.encode(ob[:m], od[:m]), 3)
        ov, oa, _ = O.expectiminimax(ob[:m], od[:m], 3, N=N, L=L)
        assert np.array_equal(to_np(v).view(np.uint64), ov.view(np.uint64)) and np.array_equal(to_np(a), oa)
        w, nch = e.mc_root_wins(e.encode(ob[:m], od[:m]), seed=3, first=0, count=16)
        ow, onch, _ = O.mc_root_wins(ob[:m], od[:m], np.arange(m, dtype=np.uint32), 3, 0, 16, N=N, L=L)
        assert np.array_equal(to_np(w).astype(np.uint32), ow) and np.array_equal(to_np(nch), onch)


@pytest.mark.parametrize("name", ["random_games_5x5.npz", "random_games_7x7.npz", "random_games_4x4_L2.npz", "random_games_7x7_L4.npz",
                                  "random_games_9x9_L7.npz"])
def test_eval_random_from_seeds_alone(backend, name):
    """ewn_numpy_random_games: the reference's eval_random.py games reproduced from the SEEDS (numpy legacy MT19937 + randint in
    the reference's consumption order) -- result, number of env.step calls, final board and the whole dice / action log (as a hash)
    of every recorded game; 5x5 seeds 0..9999 give the survey's 5,215 wins / 108,978 agent steps."""
    d = load(name)
    N, L = int(d["N"]), int(d["L"])
    e = make_engine(backend, N, L)
    n_games = len(d["game_steps"])
    res, steps, fin, tr = e.numpy_random_games(np.arange(n_games, dtype=np.uint32))
    res, steps, tr = to_np(res), to_np(steps), to_np(tr).view(np.uint64)
    assert np.array_equal(res, d["game_rewards"].astype(np.int8)) and np.array_equal(steps, d["game_steps"])
    last = np.cumsum(d["game_steps"]) - 1
    assert np.array_equal(to_np(e.decode(fin)[0]), d["boards"][last].astype(np.int16))
    rows = d["steps"]
    M = (1 << 64) - 1
    h = [0xcbf29ce484222325] * n_games
    for game, da, a0, a1, do, o0, o1, dn, term, trunc, _ in rows.tolist():
        x = h[game]
        vals = [da, a0, a1] + ([do, o0, o1] if do else []) + ([dn] if dn else [])
        for v in vals:
            x = ((x ^ (v & 0xff)) * 0x100000001b3) & M
        h[game] = x
    assert np.array_equal(tr, np.array(h, dtype=np.uint64))
    if name == "random_games_5x5.npz":
        assert int((res > 0).sum()) == 5215 and int(steps.sum()) == 108978


def test_eval_random_aggregate_counts_from_seeds(backend):
    """120,000 more reference games (tests/golden/random_vs_random_stats.json: seeds 100000.. on 5x5 and 7x7, played by the unmodified
    reference) reproduced from their seeds: exact win and agent-step totals."""
    import json, os
    ref = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "random_vs_random_stats.json")))
    for key, N in (("5x5", 5), ("7x7", 7)):
        r = ref[key]
        e = make_engine(backend, N, 3)
        seeds = np.arange(r["first_seed"], r["first_seed"] + r["games"], dtype=np.uint32)
        res, steps, _, _ = e.numpy_random_games(seeds)
        assert int((to_np(res) > 0).sum()) == r["wins"] and int(to_np(steps).sum()) == r["agent_steps"], key


def test_playout_jobs_beyond_2_31_games_are_sliced(backend):
    """k_playouts indexes games with 32 bits; the library slices a bigger job over the playout index range (ewn_kernels.cu
    launch_playouts).  A sliced call must equal the same range played as explicit pieces that each fit one launch."""
    if backend != "cuda":
        pytest.skip("2^31 playouts: GPU only")
    z = load("minimax_5x5.npz")
    e = make_engine(backend, 5, 3)
    # Monte-Carlo path: 4,096 roots -> the launch limit is (2^31 - 256) / 24,576 = 87,381 playouts per child
    n = 4096
    idx = np.arange(n) % len(z["boards"])
    st = e.encode(z["boards"][idx].astype(np.int16), z["dice"][idx].astype(np.int32))
    whole, nch = e.mc_root_wins(st, seed=5, first=3, count=100_000)
    parts, _ = e.mc_root_wins(st, seed=5, first=3, count=50_000)
    e.mc_root_wins(st, seed=5, first=50_003, count=50_000, wins=parts)
    w = to_np(whole).view(np.uint32)
    assert np.array_equal(w, to_np(parts).view(np.uint32))
    assert 0 < w.max() <= 100_000 and int(to_np(nch).sum()) > n
    # direct path (ewn_playouts): 40,000 states x 60,000 playouts = 2.4e9 games, limit 53,687 per launch
    m = 40_000
    idx = np.arange(m) % len(z["boards"])
    st = e.encode(z["boards"][idx].astype(np.int16), z["dice"][idx].astype(np.int32), np.full(m, 2, np.int32))
    whole = e.playouts(st, seed=6, first=0, count=60_000)
    parts = e.playouts(st, seed=6, first=0, count=30_000)
    parts = e.playouts(st, seed=6, first=30_000, count=30_000, wins=parts)
    assert np.array_equal(to_np(whole), to_np(parts))


def test_mc_predict_equals_root_wins_then_best(backend):
    """ewn_mc_predict (wins cleared in the expansion kernel, three chained launches, self-resetting pair counter) == zero + ewn_mc_root_wins +
    ewn_mc_best_action, repeatedly on the same workspace and with a dirty wins buffer"""
    e = make_engine(backend, 5, 3)
    n = 300 if backend == "hostsim" else 5000
    st = e.reset(n, seed=8)
    for s in range(1, 6):
        e.step(st, None, seed=8, step=s, agent_random=True, autoreset=(s != 5), want=())
    wins = e.xp.asarray(np.full((n, 6), 12345, np.uint32), "u32")
    for it in range(4):
        a, w, nch = e.mc_predict(st, seed=it, first=3, count=24, wins=wins)
        w0, nch0 = e.mc_root_wins(st, seed=it, first=3, count=24)
        assert np.array_equal(to_np(w), to_np(w0)) and np.array_equal(to_np(nch), to_np(nch0)), it
        assert np.array_equal(to_np(a), to_np(e.mc_best_action(st, w0)))
