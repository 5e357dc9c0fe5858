"""CPU: the C oracle (oracle/ewn_oracle.c) against golden vectors recorded from the unmodified reference."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests.golden_util import GOLDEN, game_traces, load


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    assert [hex(x) for x in O.philox([0] * 4, [0] * 2)] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in O.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in O.philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0])] == [
        "0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_initial_layout_matches_reference_probe():
    # reset(seed=0) board, envs/ewn.py:94-107
    b = O.init_boards(1, 5, 3)[0]
    assert b.tolist() == [[1, 2, 4, 0, 0], [3, 5, 0, 0, 0], [6, 0, 0, 0, -6], [0, 0, 0, -5, -3], [0, 0, -4, -2, -1]]
    for d in range(1, 7):
        for pl in (1, 2):
            m, lst, c = O.legal(b[None], [d], pl)
            assert c[0] == 3 and lst[0, :3].tolist() == [[0, 0], [0, 1], [0, 2]]


@pytest.mark.parametrize("name", ["random_games_5x5.npz", "random_games_7x7.npz", "random_games_4x4_L2.npz", "random_games_7x7_L4.npz",
                                  "random_games_9x9_L7.npz"])
def test_replay_reference_games(name):
    """Every env.step of the reference's eval_random loop: legal mask before, board/reward/flags after."""
    g = game_traces(name)
    N, L, S, t = g["N"], g["L"], g["steps"], g["t"]
    n_games = len(g["first_row"])
    board = O.init_boards(n_games, N, L)
    dice = g["init_dice"].copy()
    alive = np.arange(n_games)
    for step in range(int(t.max()) + 1):
        rows = np.flatnonzero(t == step)
        gi = g["gidx"][rows]
        assert np.array_equal(gi, alive[: len(gi)]) or set(gi) <= set(alive)
        s = S[rows]
        assert np.array_equal(dice[gi], s[:, 1]), "agent dice"
        mask, _, _ = O.legal(board[gi], dice[gi], 1, N, L)
        assert np.array_equal(mask, s[:, 10].astype(np.uint8)), "legal mask before step"
        out = O.step_replay(board[gi], dice[gi], s[:, 2:4], s[:, 4], s[:, 5:7], s[:, 7], N, L)
        assert np.array_equal(out["board"], g["boards"][rows]), f"board after step {step}"
        assert np.array_equal(out["reward"], g["rewards"][rows])
        assert np.array_equal(out["terminated"], s[:, 8]) and np.array_equal(out["truncated"], s[:, 9])
        board[gi] = out["board"]
        dice[gi] = out["dice"]
        alive = gi[out["terminated"] == 0]
    if name == "random_games_5x5.npz":
        assert (g["game_rewards"] > 0).sum() == 5215 and len(S) == 108978


@pytest.mark.parametrize("name", ["rules_5x5.npz", "rules_7x7.npz", "rules_6x6_L4.npz", "rules_7x7_L5.npz"])
def test_rules_plies_and_heuristics(name):
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    boards = z["boards"].astype(np.int16)
    assert np.array_equal(O.check_win(boards, N, L), z["win"])
    for h in ("hybrid", "min_dist", "two_min_dist", "attk"):
        assert np.array_equal(O.evaluate(boards, N, L, h).view(np.uint64), z[f"eval_{h}"]), h
    pr = z["ply_rows"]
    nb, valid, win, cube = O.ply(boards[pr[:, 0]], pr[:, 2], pr[:, 1], pr[:, 3], pr[:, 4], N, L)
    assert np.array_equal(valid, pr[:, 5]) and np.array_equal(win, pr[:, 6]) and np.array_equal(cube, pr[:, 7])
    assert np.array_equal(nb, z["ply_boards"].astype(np.int16))
    lr = z["legal_rows"]
    mask, lst, cnt = O.legal(boards[lr[:, 0]], lr[:, 2], lr[:, 1], N, L)
    assert np.array_equal(mask, lr[:, 3]) and np.array_equal(cnt, lr[:, 4])
    assert np.array_equal(lst.reshape(len(lr), 12), lr[:, 5:17])
    assert set(np.unique(cnt)) <= {1, 2, 3, 4, 6}


@pytest.mark.parametrize("name", ["minimax_5x5.npz", "minimax_7x7.npz"])
def test_expectiminimax_bits(name):
    """Value bits and root action of classical_policies/minimax.py (pruned semantics) for depth 1..5 + other heuristics."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    boards, dice = z["boards"].astype(np.int16), z["dice"].astype(np.int32)
    for key in z.files:
        if not key.startswith("value_"):
            continue
        parts = key[len("value_"):].rsplit("_d", 1) if "_d" in key[len("value_"):] else ["", key[len("value_d"):]]
        heuristic = "hybrid" if key.startswith("value_d") else parts[0]
        depth = int(key.rsplit("d", 1)[1])
        want_v, want_a = z[key], z["action_" + key[len("value_"):]]
        n = len(want_v)
        v, a, _ = O.expectiminimax(boards[:n], dice[:n], depth, N, L, heuristic, threads=8)
        assert np.array_equal(v.view(np.uint64), want_v), key
        assert np.array_equal(a, want_a), key


def test_expectiminimax_survey_probe_values():
    import struct
    b5, b7 = O.init_boards(1, 5), O.init_boards(1, 7)
    want5 = {1: "bfd5555555555554", 2: "bfd5555555555554", 3: "bfdbbbbbbbbbbbbb", 4: "bfdbbbbbbbbbbbbb", 5: "bfd6eeeeeeeeeeed"}
    want7 = {1: "bfe5555555555554", 2: "bfe5555555555554", 3: "bfe7ffffffffffff", 4: "bfe7ffffffffffff", 5: "bfe582d82d82d82c"}
    for d in range(1, 6):
        v, a, _ = O.expectiminimax(b5, [5], d, 5)
        assert struct.pack(">d", v[0]).hex() == want5[d] and a[0].tolist() == [0, 2]
        v, a, _ = O.expectiminimax(b7, [5], d, 7)
        assert struct.pack(">d", v[0]).hex() == want7[d] and a[0].tolist() == [0, 2]


@pytest.mark.parametrize("name", ["shaped_5x5.npz", "shaped_7x7.npz", "shaped_7x7_L4.npz"])
def test_shaped_env_traces(name):
    """envs/training_ewn.py:40-99 through the oracle's replay (tolerance budget, delta-evaluate reward)."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    rows, rewards, boards = z["rows"], z["rewards"].view(np.float64), z["boards"].astype(np.int16)
    n_env, n_steps = rows.shape[:2]
    board = z["init_board"].astype(np.int16)
    dice = z["init_dice"].astype(np.int32)
    prev = z["init_prev"].view(np.float64).copy()
    tol = np.full(n_env, int(z["tolerance0"]), np.int32)
    R = float(z["goal_reward_effective"])
    init = O.init_boards(1, N, L)[0]
    for t in range(n_steps):
        r = rows[:, t]
        assert np.array_equal(dice, r[:, 0])
        nb, valid, _, _ = O.ply(board, dice, 1, r[:, 1], r[:, 2], N, L)
        out = O.step_replay(board, dice, r[:, 1:3], r[:, 3], r[:, 4:6], r[:, 6], N, L, reward=R)
        exp_r = out["reward"].copy()
        term = out["terminated"].copy()
        new_board, new_dice = out["board"].copy(), out["dice"].copy()
        # training_ewn.py:48-56 tolerated illegal move: state unchanged, reward = illegal_move_reward
        ill = valid == 0
        tol[ill] -= 1
        tolerated = ill & (tol > 0)
        exp_r[tolerated] = -1.0
        term[tolerated] = 0
        new_board[tolerated] = board[tolerated]
        new_dice[tolerated] = dice[tolerated]
        ongoing = (out["code"] == O.CODE_ONGOING)
        score = O.evaluate(new_board, N, L, "hybrid")
        exp_r[ongoing] = score[ongoing] - prev[ongoing]          # training_ewn.py:94-96
        prev[ongoing] = score[ongoing]
        assert np.array_equal(exp_r.view(np.uint64), rewards[:, t].view(np.uint64)), f"reward step {t}"
        assert np.array_equal(term, r[:, 7]) and np.array_equal(new_board, boards[:, t])
        assert np.array_equal(tol, r[:, 10])
        done = term == 1
        new_board[done] = init
        new_dice[done] = r[done, 9]
        board, dice = new_board, new_dice


def test_random_vs_random_statistics():
    """Philox-driven random-vs-random win rate against a 100k-game run of the reference (different RNG => CI)."""
    with open(os.path.join(GOLDEN, "random_vs_random_stats.json")) as f:
        st = json.load(f)
    for N, n in ((5, 400000), (7, 100000)):
        ref = st[f"{N}x{N}"]
        p_ref = ref["wins"] / ref["games"]
        wins, steps, plies = O.random_games(n, seed=2024, N=N, threads=8)
        p = wins / n
        se = (p_ref * (1 - p_ref) * (1 / ref["games"] + 1 / n)) ** 0.5
        assert abs(p - p_ref) < 4.0 * se, (N, p, p_ref, se)      # 4 sigma two-sample bound
        ms_ref, ms = ref["agent_steps"] / ref["games"], steps / n
        assert abs(ms - ms_ref) < 0.05, (ms, ms_ref)
