"""The SB3 VecEnv contract of the batched env classes (train.py:126-136 is the reference's caller):
buffers that survive the next step, seed(), reset_infos, result codes that do not depend on the reward's sign, env_method, and the
opponent hook of envs/ewn.py:265-296 checked step for step against the single-env class.

EWNDeviceVecEnv holds no rules logic, so it runs on both backends (host build of the kernel bodies / CUDA)."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.backends import BACKENDS, make_engine, to_np

from ewn_gym_b200.device_env import EWNDeviceVecEnv
from ewn_gym_b200.envs import EinsteinWuerfeltNichtEnv


def _acts(rng, n):
    return np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)


@pytest.mark.parametrize("backend", BACKENDS)
def test_outputs_of_a_step_survive_the_next_step(backend):
    """SB3 collect_rollouts adds (_last_obs, _last_episode_starts) to its buffer AFTER the next env.step(): what step t returned
    must still be intact then."""
    e = make_engine(backend, 5, 3)
    n = 4096
    env = EWNDeviceVecEnv(n, 5, 3, seed=11, engine=e)
    rng = np.random.default_rng(0)
    obs0 = env.reset()
    keep0 = {k: to_np(v).copy() for k, v in obs0.items()}
    obs1, rew1, done1, infos1 = env.step(_acts(rng, n))
    assert all(np.array_equal(to_np(obs0[k]), keep0[k]) for k in keep0), "reset() observation was overwritten by the first step"
    keep1 = ({k: to_np(v).copy() for k, v in obs1.items()}, to_np(rew1).copy(), to_np(done1).copy())
    rows1 = [to_np(x).copy() for x in infos1.terminal_rows()]
    obs2, rew2, done2, infos2 = env.step(_acts(rng, n))
    assert all(np.array_equal(to_np(obs1[k]), keep1[0][k]) for k in keep1[0])
    assert np.array_equal(to_np(rew1), keep1[1]) and np.array_equal(to_np(done1), keep1[2])
    assert all(np.array_equal(to_np(a), b) for a, b in zip(infos1.terminal_rows(), rows1))
    assert not np.array_equal(to_np(obs2["board"]), keep1[0]["board"])           # and the new step really is a new one


@pytest.mark.parametrize("backend", BACKENDS)
def test_seed_and_reset_infos(backend):
    """train.py:135-136: env.seed(s); env.reset().  Sub-environment i is seeded with s + i -- here: keeps its own Philox stream under
    key s -- from the next reset on; seed() returns the per-env seeds; reset_infos has one entry per env."""
    N, L, n, id0 = 5, 3, 600, 40
    env = EWNDeviceVecEnv(n, N, L, seed=1, env_id0=id0, engine=make_engine(backend, N, L))
    assert len(env.reset_infos) == n and env.reset_infos[0] == {} and env.reset_infos[n - 1] == {}
    assert env.seed(None) == [None] * n
    env.reset()
    rng = np.random.default_rng(3)
    env.step(_acts(rng, n))
    assert list(env.seed(777)) == list(range(777, 777 + n))
    obs = env.reset()                                                            # third call of the env: RESET block of step index 2
    ob, od = O.reset_philox(n, 777, 2, id0, N, L)
    assert np.array_equal(to_np(obs["board"]), ob) and np.array_equal(to_np(obs["dice_roll"]), od)
    for t in range(3, 12):
        a = _acts(rng, n)
        obs, rew, done, infos = env.step(a)
        ref = O.step_philox(ob, od, a, 777, t, id0, N, L)
        assert np.array_equal(to_np(obs["board"]), ref["board"]) and np.array_equal(to_np(obs["dice_roll"]), ref["dice"])
        assert np.array_equal(to_np(done), ref["terminated"] == 1)
        ob, od = ref["board"], ref["dice"]


@pytest.mark.parametrize("backend", BACKENDS)
def test_messages_come_from_the_result_code(backend):
    """goal reward 0: every reward is 0, so (reward sign, truncated) cannot tell a win from a loss -- the stored code can"""
    N, L, n = 5, 3, 3000
    env = EWNDeviceVecEnv(n, N, L, seed=5, reward=0.0, engine=make_engine(backend, N, L))
    env.reset()
    ob, od = O.reset_philox(n, 5, 0, 0, N, L)
    rng = np.random.default_rng(1)
    want = {O.CODE_AGENT_ILLEGAL: "Invalid move for player! End the game.", O.CODE_AGENT_WON: "You won!", O.CODE_AGENT_LOST: "You lost!"}
    seen = set()
    for t in range(1, 25):
        # mostly legal moves so that games reach wins and losses
        a = _acts(rng, n)
        _, lst, cnt = O.legal(ob, od, 1, N, L)
        pick = lst[np.arange(n), rng.integers(0, 1 << 30, n) % cnt]
        legal_rows = rng.random(n) < 0.9
        a[legal_rows] = pick[legal_rows]
        obs, rew, done, infos = env.step(a)
        ref = O.step_philox(ob, od, a, 5, t, 0, N, L, reward=0.0)
        assert not to_np(rew).any()
        d = np.flatnonzero(to_np(done))
        codes = dict(zip(to_np(infos.terminal_rows()[0]).tolist(), to_np(infos.terminal_codes()).tolist()))
        assert all(codes[int(i)] == int(ref["code"][i]) for i in d)
        for i in d[:25]:
            assert infos[int(i)]["message"] == want[int(ref["code"][i])]
            seen.add(int(ref["code"][i]))
        ob, od = ref["board"], ref["dice"]
    assert seen == set(want)


@pytest.mark.parametrize("backend", BACKENDS)
def test_env_method_serves_the_read_only_methods(backend):
    N, L, n = 5, 3, 64
    env = EWNDeviceVecEnv(n, N, L, seed=2, engine=make_engine(backend, N, L))
    obs = env.reset()
    ob, od = to_np(obs["board"]), to_np(obs["dice_roll"])
    assert env.env_method("check_win") == [False] * n
    la = env.env_method("get_legal_actions", 1, indices=[0, 5, 63])
    _, lst, cnt = O.legal(ob[[0, 5, 63]], od[[0, 5, 63]], 1, N, L)
    assert la == [lst[k, : cnt[k]].tolist() for k in range(3)]
    ev = env.env_method("evaluate", heuristic="hybrid", indices=3)
    assert ev == [float(O.evaluate(ob[3:4], N, L, "hybrid")[0])]
    with pytest.raises(NotImplementedError):
        env.env_method("render")


class _Fed(EinsteinWuerfeltNichtEnv):
    """the single-env class with its dice taken from a feed (the vec env's Philox dice) instead of numpy's global stream"""
    dice_feed = None

    def roll_dice(self):
        if self.dice_feed:
            self.dice_roll = int(self.dice_feed.pop(0))
        else:
            super().roll_dice()


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("N,depth", [(5, 2), (7, 1)])
def test_minimax_opponent_equals_the_single_env_class(backend, N, depth):
    """envs/ewn.py:265-296,457-483: a VecEnv whose opponent is ExpectiMinimaxAgent(max_depth) against n single-env objects with the
    same opponent, step for step (same actions; the single envs are fed the dice the vec env draws)."""
    L, n, seed, id0 = 3, 40, 21, 1000
    e = make_engine(backend, N, L)
    env = EWNDeviceVecEnv(n, N, L, seed=seed, env_id0=id0, engine=e, opponent_policy="minimax", max_depth=depth)
    obs = env.reset()
    dice0 = to_np(obs["dice_roll"])
    singles = []
    for i in range(n):
        s = _Fed.__new__(_Fed)
        s.dice_feed = [int(dice0[i])]
        s.__init__(board_size=N, cube_layer=L, opponent_policy="minimax", max_depth=depth, engine=e)
        assert np.array_equal(s.board, to_np(obs["board"])[i]) and s.dice_roll == dice0[i]
        singles.append(s)
    rng = np.random.default_rng(9)
    finished = 0
    for t in range(1, 41):
        ob, od = to_np(obs["board"]), to_np(obs["dice_roll"])
        a = _acts(rng, n)
        _, lst, cnt = O.legal(ob, od, 1, N, L)
        legal_rows = rng.random(n) < 0.93                                         # a few illegal moves: the truncation path
        a[legal_rows] = lst[np.arange(n), rng.integers(0, 1 << 30, n) % cnt][legal_rows]
        d_opp, d_next = (to_np(x) for x in e.step_dice(n, seed, t, id0))
        obs, rew, done, infos = env.step(a)
        nb, nd, rew, done = to_np(obs["board"]), to_np(obs["dice_roll"]), to_np(rew), to_np(done)
        for i, s in enumerate(singles):
            s.dice_feed = [int(d_opp[i]), int(d_next[i])]
            o, r, term, trunc, info = s.step(a[i])
            assert r == rew[i] and term == bool(done[i]), (t, i)
            if term:
                ti = infos[i]
                assert np.array_equal(ti["terminal_observation"]["board"], o["board"]) and ti["message"] == info["message"]
                assert ti["terminal_observation"]["dice_roll"] == o["dice_roll"]
                s.dice_feed = [int(nd[i])]                                        # auto-reset: the new episode's first die
                o, _ = s.reset()
                finished += 1
            assert np.array_equal(o["board"], nb[i]) and o["dice_roll"] == nd[i], (t, i)
    assert finished > n // 2
