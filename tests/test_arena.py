"""Tournament driver: flipped-board opponents, lock-step games, win rates against the reference's recorded ones."""
import json
import os

import numpy as np
import pytest

from ewn_gym_b200 import arena as A
from tests.backends import BACKENDS, make_engine, to_np

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("N,L", [(5, 3), (7, 3), (4, 2)])
def test_flip_is_rot90_of_negated_board(backend, N, L):
    e = make_engine(backend, N, L)
    st = e.reset(700, seed=3)
    for t in range(1, 6):
        e.step(st, None, seed=3, step=t, agent_random=True, want=())
    b, d, p = (to_np(x) for x in e.decode(st, want_player=True))
    f = e.flip(st)
    fb, fd, fp = (to_np(x) for x in e.decode(f, want_player=True))
    assert np.array_equal(fb, np.rot90(-b, 2, axes=(1, 2))) and np.array_equal(fd, d) and np.array_equal(fp, 3 - p)   # envs/ewn.py:294
    assert np.array_equal(to_np(e.flip(f)), to_np(st))


@pytest.mark.parametrize("backend", BACKENDS)
def test_random_vs_random_matches_reference_win_rate(backend):
    """eval_random.py on the reference (tests/golden/random_vs_random_stats.json, recorded by make_golden.py).  The driver's
    rate must lie inside the 99.9 % Wilson interval around it (both are binomial samples of the same game)."""
    e = make_engine(backend, 5, 3)
    n = 20000 if backend == "cuda" else 6000
    r = A.play_matches(e, A.RandomPolicy(seed=1), A.RandomPolicy(seed=2), n, seed=5)
    assert r["unfinished"] == 0 and r["agent_wins"] + r["opponent_wins"] == n
    ref = json.load(open(os.path.join(GOLD, "random_vs_random_stats.json")))["5x5"]    # 52,978 wins in 100,000 reference games
    p_ref = ref["wins"] / ref["games"]
    se = (p_ref * (1 - p_ref) * (1 / n + 1 / ref["games"])) ** 0.5
    assert abs(r["agent_wins"] / n - p_ref) < 3.29 * se, (r["agent_wins"] / n, p_ref)


@pytest.mark.parametrize("backend", BACKENDS)
def test_minimax_and_monte_carlo_beat_random_from_both_seats(backend):
    e = make_engine(backend, 5, 3)
    n = 1024 if backend == "cuda" else 300
    mm, mc, rnd = A.MinimaxPolicy(2), A.MonteCarloPolicy(4, 4, seed=3), A.RandomPolicy(seed=9)
    as_agent = A.play_matches(e, mm, rnd, n, seed=1)
    as_opp = A.play_matches(e, rnd, mm, n, seed=2)
    assert as_agent["agent_wins"] / n > 0.62 and as_opp["opponent_wins"] / n > 0.58        # assets/p16-2.png: depth 2 wins ~0.72 as agent
    r = A.play_matches(e, mc, rnd, n, seed=3)
    assert r["agent_wins"] / n > 0.70 and r["unfinished"] == 0
    lo, hi = A.wilson(r["agent_wins"], n)
    assert lo < r["agent_wins"] / n < hi
