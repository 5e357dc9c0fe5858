"""GPU: EWNVecEnv (the SB3 VecEnv duck type over the host-buffer C ABI) against the CPU oracle, step by step."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,L,shaped", [(5, 3, False), (7, 3, True), (5, 3, True), (4, 2, False)])
def test_vec_env_matches_oracle(N, L, shaped):
    from ewn_gym_b200.vec_env import EWNVecEnv
    n, seed, env_id0 = 10000 + 123, 321, 7
    env = EWNVecEnv(n, N, L, seed=seed, shaped=shaped, illegal_move_reward=-1.0, illegal_move_tolerance=3, env_id0=env_id0, chunks=5)
    obs = env.reset()
    ob, od = O.reset_philox(n, seed, 0, env_id0, N, L)
    assert obs["board"].dtype == np.int16 and obs["dice_roll"].dtype == np.int64
    assert np.array_equal(obs["board"], ob) and np.array_equal(obs["dice_roll"], od)
    tol = np.full(n, 3, np.int32)
    prev = O.evaluate(ob, N, L, "hybrid")
    rng = np.random.default_rng(0)
    seen_done = 0
    for t in range(1, 31):
        act = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
        obs, rew, done, infos = env.step(act)
        ref = O.step_philox(ob, od, act, seed, t, env_id0, N, L, reward=1.0, autoreset=True, shaped=shaped, illegal_move_reward=-1.0,
                            tolerance=tol if shaped else None, prev_score=prev if shaped else None, threads=4)
        assert np.array_equal(obs["board"], ref["board"]) and np.array_equal(obs["dice_roll"], ref["dice"]), t
        assert rew.dtype == np.float32 and np.array_equal(rew, ref["reward"].astype(np.float32))
        assert done.dtype == np.bool_ and np.array_equal(done, ref["terminated"] == 1)
        assert len(infos) == n
        idx = np.flatnonzero(done)
        for i in list(idx[:40]) + list(idx[-5:]):
            info = infos[int(i)]
            assert np.array_equal(info["terminal_observation"]["board"], ref["term_board"][i])
            assert info["terminal_observation"]["dice_roll"] == ref["term_dice"][i]
            assert info["TimeLimit.truncated"] is False
            want = {O.CODE_AGENT_ILLEGAL: "Invalid move for player! End the game.", O.CODE_AGENT_WON: "You won!", O.CODE_AGENT_LOST: "You lost!"}[int(ref["code"][i])]
            assert info["message"] == want
        for i in np.flatnonzero(~done)[:5]:
            assert infos[int(i)] == {}
        seen_done += len(idx)
        ob, od = ref["board"], ref["dice"]
        if shaped:
            tol, prev = ref["tolerance"], ref["prev_score"]
    assert seen_done > n // 2
    env.close()


def test_vec_env_surface():
    from ewn_gym_b200.vec_env import EWNVecEnv
    env = EWNVecEnv(64, 5, 3, seed=1, terminal_observation=False)
    assert env.num_envs == 64 and env.action_space.nvec.tolist() == [2, 3]
    assert env.observation_space["board"].shape == (5, 5) and env.observation_space["dice_roll"].n == 7
    env.reset()
    env.step_async(np.zeros((64, 2), np.int64))
    obs, rew, done, infos = env.step_wait()
    assert obs["board"].shape == (64, 5, 5) and rew.shape == (64,) and done.shape == (64,)
    assert env.env_is_wrapped(object) == [False] * 64 and env.get_attr("board_size") == [5] * 64
    env.close()
