"""BASELINE.json's full sizes through size-independent properties (the oracle cannot follow here in seconds):
  config 4  1 Mi depth-4 roots: two independent implementations of minimax.py:19-73 -- the fused on-chip kernel and the record + replay
            kernels (the latter pinned to the reference's value bits by tests/golden/minimax_*.npz) -- agree in every value bit and action;
            finished roots report evaluate() and no action; every reported action is in the root's legal list
  config 3  1024 playouts per leaf: the win counts of a decision do not depend on how the playout index range is cut (the 8-GPU split),
            nor on which entry point plays them, and stay inside [0, playouts]
(config 2 at 1 Mi games: tests/test_vec_env.py)"""
import pytest

from tests.backends import make_engine

pytestmark = pytest.mark.gpu


def _midgame(e, n, seed, plies=5):
    st = e.reset(n, seed=seed)
    for s in range(1, plies + 1):
        e.step(st, None, seed=seed, step=s, agent_random=True, autoreset=(s % 3 != 0), want=())
    return st


@pytest.mark.parametrize("N", [5, 7])
def test_depth4_search_at_one_million_roots(N):
    import torch
    e = make_engine("cuda", N, 3)
    n = 1 << 20
    st = _midgame(e, n, seed=31, plies=6)
    v, a = e.expectiminimax(st, 4, "hybrid")
    rv, ra = e.expectiminimax(st, 4, "hybrid", replay_path=True)
    assert torch.equal(v.view(torch.int64), rv.view(torch.int64)) and torch.equal(a, ra)
    over = e.check_win(st).bool()
    assert 0 < int(over.sum()) < n                                             # the sample holds finished and live roots
    assert torch.equal(v[over].view(torch.int64), e.evaluate(st, "hybrid")[over].view(torch.int64))     # minimax.py:22-23
    assert bool((a[over] == -1).all()) and bool((a[~over, 1] >= 0).all()) and bool((a[~over, 1] <= 2).all())
    # every reported action is in the root's legal list (envs/ewn.py:338-375)
    _, lst, cnt = e.legal(st, 0)
    k = torch.arange(6, device=a.device)[None, :]
    listed = ((lst == a[:, None, :]).all(2) & (k < cnt[:, None])).any(1)
    assert bool(listed[~over].all())


def test_1024_playouts_per_leaf_do_not_depend_on_the_split():
    import torch
    e = make_engine("cuda", 5, 3)
    n, P = 1024, 1024
    st = _midgame(e, n, seed=17, plies=4)
    a, whole, nch = e.mc_predict(st, seed=77, first=0, count=P)
    parts = torch.zeros_like(whole)
    for r in range(8):                                                          # eight ranks' shares of the playout index range
        w, _ = e.mc_root_wins(st, seed=77, first=r * (P // 8), count=P // 8)
        parts += w
    assert torch.equal(whole, parts)
    assert int(whole.min()) >= 0 and int(whole.max()) <= P
    k = torch.arange(6, device=whole.device)[None, :]
    assert bool((whole[k >= nch[:, None]] == 0).all())                          # no wins beyond the legal list
    assert torch.equal(a, e.mc_best_action(st, whole))
    # a different seed is a different sample of the same win rates: counts differ, rates agree within 6 sigma of a binomial difference
    _, other, _ = e.mc_predict(st, seed=78, first=0, count=P)
    assert not torch.equal(other, whole)
    p = (whole + other).double() / (2 * P)
    sigma = torch.sqrt(p * (1 - p) * 2 / P).clamp_min(1e-9)
    z = ((whole - other).double() / P).abs() / sigma
    assert float(z[(k < nch[:, None]) & (p > 0) & (p < 1)].max()) < 6.0
