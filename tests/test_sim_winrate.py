"""The 'sim_winrate' heuristic (envs/minimax_ewn.py:36-37, :215-238) as a leaf evaluation and as a search (eval_minimax.py --heuristic
sim_winrate): statistical parity with the unmodified reference's recorded win counts (CPython `random` there, Philox here), exact
agreement between the CUDA kernels and the serial host build of the same definition, and the chain semantics of simulate()."""
import numpy as np
import pytest

from tests.backends import BACKENDS, make_engine, to_np
from tests.golden_util import load


@pytest.mark.parametrize("backend", BACKENDS)
def test_leaf_win_rates_vs_reference(backend):
    """Per position: wins over 30 consecutive simulate() calls of the reference (3,000 playouts) against 3,000 playouts here (30 units).
    Pooled over the positions, and per position within 4.5 sigma of the two-sample difference."""
    z = load("sim_winrate_5x5.npz")
    boards, ref = z["boards"].astype(np.int16), z["wins"].astype(np.float64)
    n_ref = float(z["playouts"])
    e = make_engine(backend, 5, 3)
    m = len(boards) if backend == "cuda" else 12
    st = e.encode(boards[:m], np.ones(m, np.int32))
    calls = int(n_ref) // 100
    tot = np.zeros(m)
    for c in range(calls):
        v = to_np(e.sim_winrate(st, 100, seed=77 + c, unit0=c * m))
        assert ((v >= 0) & (v <= 1)).all() and np.allclose(v * 100, np.round(v * 100))
        tot += np.round(v * 100)
    p_ref, p = ref[:m] / n_ref, tot / n_ref
    pooled = (p_ref + p) / 2
    sigma = np.sqrt(np.maximum(pooled * (1 - pooled), 1e-4) * 2 / n_ref)
    assert (np.abs(p - p_ref) < 4.5 * sigma).all(), np.abs(p - p_ref) / sigma
    assert abs(p.mean() - p_ref.mean()) < 4.5 * np.sqrt(0.25 * 2 / (n_ref * m))


@pytest.mark.parametrize("backend", BACKENDS)
def test_simulate_is_a_chain_not_a_fixed_opener(backend):
    """simulate() never restores current_player (minimax_ewn.py:220): the opener of a playout is the loser of the previous one.  From the
    initial position the mixture differs measurably from 'BOTTOM_RIGHT always opens' (mcts.py-style playouts): the first mover has
    the edge, and TOP_LEFT opens whenever it has just lost."""
    e = make_engine(backend, 5, 3)
    reps = 600 if backend == "cuda" else 60
    st = e.reset_dice(np.ones(reps, np.int32))
    chain = to_np(e.sim_winrate(st, 100, seed=5)).mean()
    fixed = to_np(e.playouts(st, seed=5, first=0, count=100)).astype(np.float64).mean() / 100
    assert chain > fixed + 2.5 * np.sqrt(0.5 / (100 * reps)), (chain, fixed)


@pytest.mark.parametrize("backend", BACKENDS)
def test_search_with_sim_winrate_leaves(backend):
    """depth-1 root values against the reference's (16 independent runs per state there): the mean over runs must agree; plus the
    definitional properties at depth 1..4 -- values in [0, 1], legal first-best action, finished roots 1.0 / 0.0 with action None,
    reproducible in (seed, root index)."""
    z = load("sim_search_5x5.npz")
    boards, dice, ref = z["boards"].astype(np.int16), z["dice"].astype(np.int32), z["values"]
    e = make_engine(backend, 5, 3)
    m = len(boards) if backend == "cuda" else 6
    st = e.encode(boards[:m], dice[:m])
    fin = to_np(e.check_win(st)).astype(bool)
    runs = 16 if backend == "cuda" else 4
    vals = np.stack([to_np(e.expectiminimax(st, 1, "sim_winrate", seed=100 + r)[0]) for r in range(runs)])
    live = ~fin
    # a depth-1 value is the max over <= 6 children of a 100-playout estimate: sd of one run <= 0.05; compare the run means
    d = np.abs(vals.mean(0) - ref[:, :m].mean(0))[live]
    assert (d < 4.5 * 0.05 * np.sqrt(1 / runs + 1 / ref.shape[0])).all(), d
    _, lst, cnt = (to_np(x) for x in e.legal(st, 1))
    for depth in (1, 2, 3, 4) if backend == "cuda" else (1, 3):
        k = m if depth <= 2 else min(m, 4)
        v, a = (to_np(x) for x in e.expectiminimax(st[:k], depth, "sim_winrate", seed=9, num_simulations=40))
        v2, a2 = (to_np(x) for x in e.expectiminimax(st[:k], depth, "sim_winrate", seed=9, num_simulations=40))
        assert np.array_equal(v, v2) and np.array_equal(a, a2)
        assert ((v >= 0) & (v <= 1)).all()
        for i in range(k):
            if fin[i]:
                assert a[i].tolist() == [-1, -1] and v[i] in (0.0, 1.0)
            else:
                assert a[i].tolist() in lst[i, : cnt[i]].tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("N", [5, 7])
def test_cuda_kernels_equal_the_serial_definition(N):
    """k_sim_playouts + k_sim_values (one thread per playout, bit planes, chain) == sim_unit_value run serially by the host build, and
    the search built on them == the host build's, value bits and actions (same Philox keys)."""
    g, h = make_engine("cuda", N, 3), make_engine("hostsim", N, 3)
    n = 200
    sg = g.reset(n, seed=3)
    for s in range(1, 8):
        g.step(sg, None, seed=3, step=s, agent_random=True, autoreset=(s % 4 != 0), want=())
    sh = to_np(sg).copy()
    for nsim in (100, 33):
        assert np.array_equal(to_np(g.sim_winrate(sg, nsim, seed=11, unit0=5)), to_np(h.sim_winrate(sh, nsim, seed=11, unit0=5)))
    for depth, k in ((1, 60), (2, 30), (3, 6), (4, 3)):
        vg, ag = (to_np(x) for x in g.expectiminimax(sg[:k], depth, "sim_winrate", seed=2, num_simulations=20))
        vh, ah = (to_np(x) for x in h.expectiminimax(sh[:k], depth, "sim_winrate", seed=2, num_simulations=20))
        assert np.array_equal(vg.view(np.uint64), vh.view(np.uint64)) and np.array_equal(ag, ah), depth
