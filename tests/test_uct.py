"""UCT at the root on the playout kernel (the reference's ucth_temp.py:6-89 is an unfinished sketch of it): the selection /
back-up rule against an independent numpy loop over the ORACLE's playouts (same Philox keys -> identical counters), behaviour against
flat Monte-Carlo, and the agent / arena surfaces."""
import math

import numpy as np
import pytest

from oracle import oracle as O
from tests.backends import BACKENDS, make_engine, to_np


def _roots(e, n, seed, plies):
    st = e.reset(n, seed=seed)
    for s in range(1, plies + 1):
        e.step(st, None, seed=seed, step=s, agent_random=True, autoreset=(s <= plies - 4), want=())   # the last steps leave finished games in
    return st


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("N", [5, 7])
def test_uct_counters_equal_an_independent_loop_over_oracle_playouts(backend, N):
    e = make_engine(backend, N, 3)
    n, seed, rounds, batch, c = (48 if backend == "hostsim" else 160), 13, 9, 16, 1.41
    st = _roots(e, n, 21, 8 if N == 5 else 22)
    a, w, v = (to_np(x) for x in e.uct_search(st, seed, rounds, batch, c))
    board, dice = (to_np(x) for x in e.decode(st))
    fin = O.check_win(board, N, 3) == 1
    _, lst, cnt = O.legal(board, dice, 1, N, 3)
    assert fin.any() and (~fin).any()
    for i in range(n):
        if fin[i]:
            assert a[i].tolist() == [-1, -1] and not w[i].any() and not v[i].any()
            continue
        kids = []
        for j in range(cnt[i]):                                    # positions after each legal move (oracle ply), BOTTOM_RIGHT to move
            nb, valid, win, _ = O.ply(board[i][None], [dice[i]], [1], [lst[i, j, 0]], [lst[i, j, 1]], N, 3)
            assert valid[0]
            kids.append(nb[0])
        W, V = np.zeros(6, np.int64), np.zeros(6, np.int64)
        for r in range(rounds):
            score = [math.inf if V[j] == 0 else W[j] / V[j] + c * math.sqrt(math.log(r * batch) / V[j]) for j in range(cnt[i])]
            j = int(np.argmax(score))                               # first maximum
            got, _ = O.playouts(kids[j][None], [i * 6 + j], seed, r * batch, batch, N, 3)
            W[j] += int(got[0]); V[j] += batch
        assert W.tolist() == w[i].tolist() and V.tolist() == v[i].tolist(), i
        rate = [W[j] / V[j] if V[j] else 0.0 for j in range(cnt[i])]
        assert a[i].tolist() == lst[i, int(np.argmax(rate))].tolist()


@pytest.mark.parametrize("backend", BACKENDS)
def test_uct_spends_its_playouts_on_the_better_moves(backend):
    """against flat Monte-Carlo with the same budget: the move UCT returns has a (separately estimated) win rate close to the best
    move's, and the visit counts are ordered like the win rates"""
    e = make_engine(backend, 5, 3)
    n = 64 if backend == "hostsim" else 512
    st = _roots(e, n, 5, 4)
    live = to_np(e.check_win(st)) == 0
    a, w, v = (to_np(x) for x in e.uct_search(st, 3, 32, 16))
    flat, nch = (to_np(x) for x in e.mc_root_wins(st, seed=99, first=0, count=1024))
    rate = flat / 1024.0
    _, lst, cnt = (to_np(x) for x in e.legal(st, 1))
    regret, skew = [], []
    for i in np.flatnonzero(live):
        k = int(cnt[i])
        j = [lst[i, q].tolist() for q in range(k)].index(a[i].tolist())
        regret.append(rate[i, :k].max() - rate[i, j])
        assert v[i, :k].sum() == 32 * 16 and (v[i, :k] >= 16).all()          # every move is tried once
        if k >= 2 and rate[i, :k].max() - rate[i, :k].min() > 0.2:
            skew.append(v[i, int(np.argmax(rate[i, :k]))] > v[i, int(np.argmin(rate[i, :k]))])
    assert np.mean(regret) < 0.03 and np.mean(np.array(regret) < 0.1) > 0.9
    assert np.mean(skew) > 0.9


@pytest.mark.parametrize("backend", BACKENDS)
def test_uct_agent_and_arena_policy(backend):
    from ewn_gym_b200.arena import RandomPolicy, UctPolicy, play_matches
    from ewn_gym_b200.policies import UctAgent
    e = make_engine(backend, 5, 3)
    ag = UctAgent(cube_layer=3, board_size=5, iterations=128, batch=16, engine=e)
    b = O.init_boards(1, 5, 3)[0]
    act, _ = ag.predict({"board": b, "dice_roll": 2})
    assert act.tolist() in [[0, 0], [0, 1], [0, 2]]
    games = 48 if backend == "hostsim" else 256
    res = play_matches(e, UctPolicy(rounds=6, batch=8, seed=1), RandomPolicy(seed=2), games, seed=4)
    assert res["unfinished"] == 0 and res["agent_wins"] > 0.7 * games          # the reference's charts: MCTS(50) vs random 0.91-0.97
