"""The torch-only A2C of ewn_gym_b200/train.py (the reference's train.py:35-63,126-155 drives SB3's A2C): rollout collection on the
batched env, n-step returns across episode boundaries, and a short training run.  Runs on the host build (CPU torch) and on CUDA."""
import numpy as np
import pytest
import torch as th

from oracle import oracle as O
from tests.backends import BACKENDS, make_engine, to_np

from ewn_gym_b200.device_env import EWNDeviceVecEnv
from ewn_gym_b200.train import A2C, evaluate, return_model


@pytest.mark.parametrize("backend", BACKENDS)
def test_collect_rollouts_buffer_and_returns(backend):
    """train.py:144-155 -> OnPolicyAlgorithm.collect_rollouts: the buffer holds, per step, the observation the action was computed
    from (not the one env.step returned), and the returns are the n-step returns cut at episode boundaries."""
    N, L, n, seed = 5, 3, 300, 4
    env = EWNDeviceVecEnv(n, N, L, seed=seed, shaped=True, illegal_move_tolerance=3, engine=make_engine(backend, N, L))
    model = A2C("MultiInputPolicy", env, n_steps=6, seed=1)
    ob, od = O.reset_philox(n, seed, 0, 0, N, L)
    tol, prev = np.full(n, 3, np.int32), O.evaluate(ob, N, L, "hybrid")
    t = 0
    for rollout in range(4):
        model.collect_rollouts()
        b = model.buf
        rew, dones = [], []
        for k in range(model.n_steps):
            assert np.array_equal(to_np(b["board"][k]), ob) and np.array_equal(to_np(b["dice"][k]), od), (rollout, k)
            t += 1
            ref = O.step_philox(ob, od, to_np(b["actions"][k]), seed, t, 0, N, L, shaped=True, illegal_move_reward=-1.0, tolerance=tol, prev_score=prev)
            assert np.array_equal(to_np(b["rewards"][k]), ref["reward"].astype(np.float32))
            rew.append(ref["reward"].astype(np.float32)); dones.append(ref["terminated"].astype(np.float32))
            ob, od, tol, prev = ref["board"], ref["dice"], ref["tolerance"], ref["prev_score"]
        assert np.array_equal(to_np(model._last_obs["board"]), ob)
        # n-step returns with gamma = 0.99, bootstrapped from the value of the last observation unless the episode ended
        values = to_np(b["values"])
        with th.no_grad():
            last = to_np(model.policy.predict_values(model._last_obs))
        ret = np.zeros_like(values)
        nxt = last
        for k in reversed(range(model.n_steps)):
            nxt = rew[k] + 0.99 * nxt * (1.0 - dones[k])
            ret[k] = nxt
        assert np.allclose(to_np(model.returns), ret, atol=1e-5)
        assert np.allclose(to_np(model.advantages), ret - values, atol=1e-5)
    assert model.num_timesteps == 4 * 6 * n


@pytest.mark.parametrize("backend", BACKENDS)
def test_a_short_training_run_learns_to_move_legally(backend):
    """A2C on MiniMaxHeuristicEnv semantics (an illegal move costs -1, train.py's env): after a few hundred updates the policy makes
    clearly fewer illegal moves than at the start; evaluate() plays greedy episodes against a minimax opponent."""
    N, L, n = 5, 3, 512
    e = make_engine(backend, N, L)
    env = EWNDeviceVecEnv(n, N, L, seed=11, shaped=True, illegal_move_tolerance=10, engine=e)
    env.seed(3); env.reset()
    model = return_model({"algorithm": "A2C", "n_steps": 5, "learning_rate": 7e-4, "model_seed": 0}, env)
    illegal = []
    for it in range(240):
        model.collect_rollouts()
        illegal.append(float((model.buf["rewards"] == -1.0).float().mean()))      # illegal_move_reward, never a shaped reward
        model.train()
        assert all(np.isfinite(v) for v in model.logger.values())
    early, late = np.mean(illegal[:20]), np.mean(illegal[-20:])
    assert late < 0.75 * early, (early, late)
    win_rate, avg = evaluate({"board_size": N, "cube_layer": L, "eval_episode_num": 64, "eval_opponent_policy": "minimax", "eval_max_depth": 1},
                             model, engine=e)
    assert 0.0 <= win_rate <= 1.0 and -1.0 <= avg <= 1.0
    a, _ = model.predict({"board": np.zeros((N, N), np.int16), "dice_roll": 3}, deterministic=True)     # train.py:87: one un-batched obs
    assert a.shape == (2,) and 0 <= a[0] < 2 and 0 <= a[1] < 3
