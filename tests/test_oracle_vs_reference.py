"""Build container only (marker `reference`: needs /root/reference): the C oracle against the LIVE, unmodified reference -- fresh
seeds and configurations beyond the frozen fixtures under tests/golden/, so that the oracle is pinned on every CPU run here and
not only through recordings.  Skipped on machines without the reference (the GPU box)."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.reference


def _gold():
    from tests.golden import make_golden as G          # imports the reference through oracle/ref_harness.py
    return G


@pytest.mark.parametrize("N,L,seed0,games", [(5, 3, 50_000, 40), (6, 4, 60_000, 12), (8, 6, 70_000, 5), (4, 2, 80_000, 30)])
def test_random_games_replayed_on_the_oracle(N, L, seed0, games):
    """eval_random.py:17-51 on fresh seeds: every step's legal mask, board, reward and flags (dice and opponent answers injected)"""
    G = _gold()
    steps, rewards, boards, game_rewards, game_steps, init_dice = G.play_random_games((N, L, list(range(seed0, seed0 + games))))
    board = {s: O.init_boards(1, N, L)[0] for s in range(seed0, seed0 + games)}
    dice = {s: int(d) for s, d in zip(range(seed0, seed0 + games), init_dice)}
    for row, s in enumerate(steps):
        g = int(s[0])
        assert dice[g] == s[1]
        mask, _, _ = O.legal(board[g][None], [dice[g]], 1, N, L)
        assert mask[0] == s[10]
        out = O.step_replay(board[g][None], [dice[g]], s[None, 2:4], s[None, 4], s[None, 5:7], s[None, 7], N, L)
        assert np.array_equal(out["board"][0], boards[row]) and out["reward"][0] == rewards[row]
        assert out["terminated"][0] == s[8] and out["truncated"][0] == s[9]
        board[g], dice[g] = out["board"][0], int(out["dice"][0])


@pytest.mark.parametrize("depth,heuristic", [(2, "hybrid"), (3, "hybrid"), (4, "hybrid"), (3, "attk"), (3, "two_min_dist")])
def test_expectiminimax_on_fresh_states(depth, heuristic):
    """ExpectiMinimaxAgent (minimax.py:19-93) on freshly sampled 5x5 states: value bits and root action"""
    G = _gold()
    boards, dice = G.sample_states((5, 3, list(range(9000 + depth * 100, 9000 + depth * 100 + 24))))
    v_ref, a_ref = G.run_minimax((5, 3, depth, heuristic, boards, dice))
    v, a, _ = O.expectiminimax(boards.astype(np.int16), dice.astype(np.int32), depth, 5, 3, heuristic)
    assert np.array_equal(v.view(np.uint64), v_ref.view(np.uint64)) and np.array_equal(a, a_ref)
