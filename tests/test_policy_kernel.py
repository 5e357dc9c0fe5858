"""K5 on the GPU: ewn_policy_forward (SB3 "MultiInputPolicy" forward + categorical draw in one launch) against the torch module it reads
its parameters from, ewn_gae against the tensor loop of OnPolicyAlgorithm, and the fused rollout collector (one CUDA graph per
rollout) against the generic one.  train.py:35-63 (MultiInputPolicy, Tanh), :144-155 (model.learn -> collect_rollouts)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 2e-5          # fp32 dot products of 56..64 terms summed in a different order than cuBLAS does, tanhf on both sides


def _module(N, L, seed=0):
    import torch as th
    from ewn_gym_b200.train import MultiInputActorCritic
    th.manual_seed(seed)
    m = MultiInputActorCritic(N, L * (L + 1) // 2).cuda()
    with th.no_grad():                               # SB3's init makes action_net tiny (gain 0.01): spread the logits for the test
        m.action_net.weight.mul_(60.0); m.action_net.bias.uniform_(-0.5, 0.5); m.value_net.bias.fill_(0.25)
    return m


def _obs(e, n, seed, steps=6):
    st = e.reset(n, seed=seed)
    for t in range(1, steps + 1):
        e.step(st, None, seed=seed, step=t, agent_random=True, autoreset=True, want=())
    b, d = e.decode(st)
    return b, d


@pytest.mark.parametrize("N,L,n", [(7, 3, 4096), (5, 3, 1000 + 7), (7, 3, 5), (6, 3, 333)])
def test_policy_forward_matches_the_torch_module(N, L, n):
    import torch as th
    from ewn_gym_b200.engine import Engine
    e = Engine(N, L)
    m = _module(N, L, seed=N + n)
    net = e.policy_net(m)
    board, dice = _obs(e, n, seed=11)
    obs = {"board": board, "dice_roll": dice}
    out = e.policy_forward(net, board, dice, seed=5, step=3, env_id0=100, want=("actions", "values", "logp", "logits"))
    with th.no_grad():
        f = m.features(obs)
        logits = m.action_net(m.pi(f))
        values = m.value_net(m.vf(f)).squeeze(1)
        v2, lp, _ = m.evaluate_actions(obs, out["actions"])
    assert th.allclose(out["logits"], logits, atol=TOL, rtol=0), float((out["logits"] - logits).abs().max())
    assert th.allclose(out["values"], values, atol=TOL, rtol=0) and th.allclose(out["values"], v2, atol=TOL, rtol=0)
    assert th.allclose(out["logp"], lp, atol=4 * TOL, rtol=0), float((out["logp"] - lp).abs().max())
    a = out["actions"]
    assert int(a[:, 0].min()) >= 0 and int(a[:, 0].max()) <= 1 and int(a[:, 1].min()) >= 0 and int(a[:, 1].max()) <= 2
    # the same call again draws the same actions (counter-based stream); another step index draws others
    again = e.policy_forward(net, board, dice, seed=5, step=3, env_id0=100)
    assert th.equal(again["actions"], a)
    if n >= 1000:
        other = e.policy_forward(net, board, dice, seed=5, step=4, env_id0=100)
        assert not th.equal(other["actions"], a)
    # deterministic = the arg-max of the probabilities, wherever the two best logits are not a numerical tie
    det = e.policy_forward(net, board, dice, deterministic=True)["actions"]
    top2 = logits[:, :2].sort(1).values
    clear0 = (top2[:, 1] - top2[:, 0]) > 1e-4
    top3 = logits[:, 2:].sort(1).values
    clear1 = (top3[:, 2] - top3[:, 1]) > 1e-4
    assert th.equal(det[clear0, 0], logits[:, :2].argmax(1)[clear0]) and th.equal(det[clear1, 1], logits[:, 2:].argmax(1)[clear1])
    # predict_values: no actions asked for
    pv = e.policy_forward(net, board, dice, want=("values",))
    assert set(pv) == {"values"} and th.equal(pv["values"], out["values"])


def test_sampled_actions_follow_the_categoricals():
    """one observation repeated 200,000 times (different env ids -> different Philox counters): the frequencies of the 2 x 3 actions
    against softmax(logits), 5 sigma per cell"""
    import torch as th
    from ewn_gym_b200.engine import Engine
    N, L, n = 7, 3, 200_000
    e = Engine(N, L)
    m = _module(N, L, seed=3)
    b1, d1 = _obs(e, 1, seed=2)
    board, dice = b1.expand(n, N, N).contiguous(), d1.expand(n).contiguous()
    out = e.policy_forward(e.policy_net(m), board, dice, seed=9, step=1, want=("actions", "logits"))
    p0 = th.softmax(out["logits"][0, :2].double(), 0).cpu().numpy()
    p1 = th.softmax(out["logits"][0, 2:].double(), 0).cpu().numpy()
    a = out["actions"].cpu().numpy()
    for i in range(2):
        for j in range(3):
            p = p0[i] * p1[j]
            k = int(((a[:, 0] == i) & (a[:, 1] == j)).sum())
            assert abs(k - n * p) <= 5 * np.sqrt(n * p * (1 - p)) + 1, (i, j, k, n * p)


def test_unsupported_width_is_refused():
    import torch as th
    from ewn_gym_b200.engine import Engine
    from ewn_gym_b200.train import MultiInputActorCritic
    e = Engine(5, 3)
    m = MultiInputActorCritic(5, 6, net_arch=(32, 32)).cuda()
    b, d = _obs(e, 8, seed=1)
    with pytest.raises(RuntimeError):
        e.policy_forward(e.policy_net(m), b, d)
    _ = th


@pytest.mark.parametrize("T,n,lam", [(128, 4096, 1.0), (5, 77, 0.95), (1, 3, 1.0)])
def test_gae_equals_the_tensor_loop(T, n, lam):
    import torch as th
    from ewn_gym_b200.engine import Engine
    e = Engine(5, 3)
    g = th.Generator(device="cuda"); g.manual_seed(T * n)
    rewards = th.randn(T, n, device="cuda", generator=g)
    values = th.randn(T, n, device="cuda", generator=g)
    dones = (th.rand(T, n, device="cuda", generator=g) < 0.1).to(th.uint8)
    last = th.randn(n, device="cuda", generator=g)
    gamma = 0.99
    adv, ret = e.gae(rewards, values, dones, last, gamma, lam)
    # OnPolicyAlgorithm / RolloutBuffer.compute_returns_and_advantage as ewn_gym_b200/train.py's generic collector writes it
    ref = th.zeros_like(rewards)
    last_gae = th.zeros(n, device="cuda")
    for t in reversed(range(T)):
        nonterminal = 1.0 - dones[t].float()
        next_values = last if t == T - 1 else values[t + 1]
        delta = rewards[t] + gamma * next_values * nonterminal - values[t]
        last_gae = delta + gamma * lam * nonterminal * last_gae
        ref[t] = last_gae
    assert th.equal(adv, ref) and th.equal(ret, ref + values)


@pytest.mark.parametrize("persistent", [False, True])
@pytest.mark.parametrize("N,L,shaped", [(7, 3, True), (5, 3, False)])
def test_fused_rollouts_are_consistent_and_continue_across_replays(N, L, shaped, persistent):
    """the graph-replayed collector: observations / rewards / dones of the buffer are what the env produces for the stored actions
    (replayed through a second env from the same seed), values and log-probs are the torch module's, the rollouts join up, stepping
    the env by hand in between is picked up, and a training step changes what the next replay computes (parameters read in place)"""
    import torch as th
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.train import A2C
    n, T, seed = 1000, 16, 21
    env = EWNDeviceVecEnv(n, N, L, seed=seed, shaped=shaped, illegal_move_tolerance=3)
    twin = EWNDeviceVecEnv(n, N, L, seed=seed, shaped=shaped, illegal_move_tolerance=3)
    model = A2C("MultiInputPolicy", env, n_steps=T, seed=2, persistent=persistent)
    assert model.fused and model.persistent == persistent
    obs = twin.reset()
    tb, td = obs["board"].clone(), obs["dice_roll"].clone()
    for rollout in range(4):
        if rollout == 2:                                          # somebody steps the env between two rollouts
            act = th.zeros((n, 2), dtype=th.int64, device="cuda")
            env.step(act)
            o2, _, _, _ = twin.step(act)
            tb, td = o2["board"].clone(), o2["dice_roll"].clone()
        model.collect_rollouts()
        b = model.buf
        for t in range(T):
            assert th.equal(b["board"][t], tb) and th.equal(b["dice"][t], td), (rollout, t)
            o2, r2, d2, _ = twin.step(b["actions"][t].clone())
            assert th.equal(b["rewards"][t], r2) and th.equal(b["dones"][t].bool(), d2), (rollout, t)
            tb, td = o2["board"].clone(), o2["dice_roll"].clone()
        assert th.equal(model._last_obs["board"], tb) and th.equal(env.state, twin.state)
        flat = {"board": b["board"].reshape(T * n, N, N), "dice_roll": b["dice"].reshape(T * n)}
        with th.no_grad():
            v, lp, _ = model.policy.evaluate_actions(flat, b["actions"].reshape(T * n, 2))
            last = model.policy.predict_values(model._last_obs)
        assert th.allclose(b["values"].reshape(-1), v, atol=TOL, rtol=0) and th.allclose(b["log_probs"].reshape(-1), lp, atol=4 * TOL, rtol=0)
        adv, ret = env.e.gae(b["rewards"], b["values"], b["dones"], model.fb["last_values"], model.gamma, model.gae_lambda)
        assert th.allclose(model.fb["last_values"], last, atol=TOL, rtol=0)
        assert th.equal(model.advantages, adv) and th.equal(model.returns, ret)
        assert model.logger["rollout/episodes"] == float(b["dones"].sum())
        model.train()                                             # RMSprop updates the parameters in place: the graph reads the new ones
    assert env.step_index == twin.step_index and model.num_timesteps == 4 * T * n


@pytest.mark.parametrize("N,L,shaped,n", [(7, 3, True, 4096), (5, 3, False, 1000 + 13), (6, 3, True, 77), (9, 3, False, 300)])
def test_one_launch_rollout_equals_two_launches_per_step(N, L, shaped, n):
    """ewn_rollout (the whole rollout in one launch, games in registers) against the graph of ewn_policy_forward + ewn_step launches:
    every buffer bit for bit (observations, actions, values, log-probs, rewards, dones, V(last), advantages, returns), the games and the
    env statistics afterwards, over several rollouts with training steps in between"""
    import torch as th
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.train import A2C
    T, seed = 12, 8
    envs = [EWNDeviceVecEnv(n, N, L, seed=seed, shaped=shaped, illegal_move_tolerance=2, env_id0=5000) for _ in (0, 1)]
    a = A2C("MultiInputPolicy", envs[0], n_steps=T, seed=4, persistent=False)
    b = A2C("MultiInputPolicy", envs[1], n_steps=T, seed=4, persistent=True)
    assert a.fused and b.fused and b.persistent and not a.persistent
    b.policy.load_state_dict(a.policy.state_dict())
    for rollout in range(4):
        a.collect_rollouts(); b.collect_rollouts()
        for k in ("board", "dice", "actions", "values", "log_probs", "rewards", "dones", "last_values", "adv", "ret"):
            assert th.equal(a.fb[k], b.fb[k]), (rollout, k)
        assert th.equal(envs[0].state, envs[1].state) and th.equal(envs[0].stats, envs[1].stats)
        assert a.logger["rollout/episodes"] == b.logger["rollout/episodes"]
        if shaped:
            assert th.equal(envs[0]._tol, envs[1]._tol) and th.equal(envs[0]._prev, envs[1]._prev)
        a.train()
        b.policy.load_state_dict(a.policy.state_dict())           # (autograd's reductions need not be bitwise repeatable: keep the twins in step)
    assert int(envs[0].stats[3]) > 0


@pytest.mark.parametrize("N,L,shaped,n", [(7, 3, True, 4096), (5, 3, False, 2000 + 5), (9, 3, True, 100)])
def test_rollout_without_a_policy_equals_step_calls(N, L, shaped, n):
    """ewn_rollout with net == NULL (actions read from a pre-filled tensor): T env steps of every game in one launch against T ewn_step
    calls -- observations of every slot, rewards, dones, the games, tolerance / previous score and the statistics"""
    import torch as th
    from ewn_gym_b200.engine import Engine
    e = Engine(N, L)
    T, seed, env_id0 = 16, 31, 700
    g = th.Generator(device="cuda"); g.manual_seed(3)
    acts = th.stack([th.randint(0, 2, (T, n), generator=g, device="cuda"), th.randint(-1, 4, (T, n), generator=g, device="cuda")], 2).contiguous()
    sa, sb = e.reset(n, seed=seed, env_id0=env_id0), e.reset(n, seed=seed, env_id0=env_id0)
    kw = {}
    if shaped:
        ta, tb = (th.full((n,), 2, dtype=th.int32, device="cuda") for _ in (0, 1))
        pa, pb = e.evaluate(sa, "hybrid"), e.evaluate(sb, "hybrid")
        kw = dict(shaped=True, illegal_move_reward=-1.0)
    stats_a, stats_b = th.zeros(4, dtype=th.int64, device="cuda"), th.zeros(4, dtype=th.int64, device="cuda")
    bufs = {"board": th.zeros((T + 1, n, N, N), dtype=th.int16, device="cuda"), "dice": th.zeros((T + 1, n), dtype=th.int64, device="cuda"),
            "actions": acts, "rewards": th.zeros((T, n), device="cuda"), "dones": th.zeros((T, n), dtype=th.uint8, device="cuda")}
    ctr = th.full((1,), 5, dtype=th.int64, device="cuda")
    roll = e.rollout_plan(None, sb, T, bufs, seed=seed, env_id0=env_id0, step_dev=ctr, stats=stats_b,
                          tolerance=tb if shaped else None, prev_score=pb if shaped else None, **kw)
    b0, d0 = e.decode(sa)
    roll(1)
    assert th.equal(bufs["board"][0], b0) and th.equal(bufs["dice"][0], d0)
    for t in range(T):
        r = e.step(sa, acts[t], seed=seed, step=6 + t, env_id0=env_id0, stats=stats_a, tolerance=ta if shaped else None,
                   prev_score=pa if shaped else None, want=("reward_f32", "terminated", "truncated", "obs_board", "obs_dice"), **kw)
        assert th.equal(bufs["board"][t + 1], r["obs_board"]) and th.equal(bufs["dice"][t + 1], r["obs_dice"]), t
        assert th.equal(bufs["rewards"][t], r["reward_f32"]) and th.equal(bufs["dones"][t], r["terminated"]), t
    assert th.equal(sa, sb) and th.equal(stats_a, stats_b) and (n < 1000 or int(stats_a[3]) > 0)
    if shaped:
        assert th.equal(ta, tb) and th.equal(pa, pb)


@pytest.mark.parametrize("persistent", [False, True])
def test_fused_collector_follows_env_seed_and_reset(persistent):
    """train.py:135-136: env.seed(s); env.reset() after the collector has bound its launches -- the next rollout must play the re-seeded
    games (the bound launches carry the Philox key: the collector rebinds)"""
    import torch as th
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.train import A2C
    n, T = 500, 6
    env = EWNDeviceVecEnv(n, 5, 3, seed=1)                             # (plain env: the shaped one carries tolerance / previous score over a reset)
    model = A2C("MultiInputPolicy", env, n_steps=T, seed=9, persistent=persistent)
    model.collect_rollouts()
    env.seed(42); first = env.reset()
    b0 = first["board"].clone()
    model.collect_rollouts()
    assert th.equal(model.buf["board"][0], b0)
    twin = EWNDeviceVecEnv(n, 5, 3, seed=42)
    twin._t = env.step_index - T - 1                                  # the same step counter as the re-seeded env had at its reset
    obs = twin.reset()
    assert th.equal(obs["board"], b0)
    for t in range(T):
        obs, r, d, _ = twin.step(model.buf["actions"][t].clone())
        assert th.equal(model.buf["rewards"][t], r) and th.equal(model.buf["dones"][t].bool(), d), t
    assert th.equal(model._last_obs["board"], obs["board"])


@pytest.mark.parametrize("N,n", [(5, 3000 + 1), (7, 1500)])
def test_rollout_with_the_in_kernel_random_agent_equals_step_calls(N, n):
    """ewn_rollout, net == NULL, agent_random: random-vs-random games (eval_random.py's loop) T steps per launch against T ewn_step calls with
    the in-kernel agent -- observations, rewards, dones, the moves played, the games and the statistics"""
    import torch as th
    from ewn_gym_b200.engine import Engine
    e = Engine(N, 3)
    T, seed = 24, 5
    sa, sb = e.reset(n, seed=seed), e.reset(n, seed=seed)
    stats_a, stats_b = th.zeros(4, dtype=th.int64, device="cuda"), th.zeros(4, dtype=th.int64, device="cuda")
    bufs = {"board": th.zeros((T + 1, n, N, N), dtype=th.int16, device="cuda"), "dice": th.zeros((T + 1, n), dtype=th.int64, device="cuda"),
            "actions": th.full((T, n, 2), -7, dtype=th.int64, device="cuda"), "rewards": th.zeros((T, n), device="cuda"),
            "dones": th.zeros((T, n), dtype=th.uint8, device="cuda")}
    e.rollout_plan(None, sb, T, bufs, seed=seed, stats=stats_b, agent_random=True)(1)
    for t in range(T):
        r = e.step(sa, None, seed=seed, step=1 + t, agent_random=True, stats=stats_a,
                   want=("reward_f32", "terminated", "truncated", "obs_board", "obs_dice", "action_out"), block_kernel=True)
        assert th.equal(bufs["board"][t + 1], r["obs_board"]) and th.equal(bufs["dice"][t + 1], r["obs_dice"]), t
        assert th.equal(bufs["rewards"][t], r["reward_f32"]) and th.equal(bufs["dones"][t], r["terminated"]), t
        assert th.equal(bufs["actions"][t, :, 0] * 3 + bufs["actions"][t, :, 1], r["action_out"].long()), t
    assert th.equal(sa, sb) and th.equal(stats_a, stats_b) and int(stats_a[0]) > 0 and int(stats_a[2]) == 0     # wins happen, illegal endings do not


def test_env_rollout_equals_env_steps():
    """EWNDeviceVecEnv.rollout (one launch) against the same env class stepped T times: pre-sampled actions and the in-kernel agent"""
    import torch as th
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    n, T = 777, 10
    a, b = EWNDeviceVecEnv(n, 7, 3, seed=6, shaped=True), EWNDeviceVecEnv(n, 7, 3, seed=6, shaped=True)
    oa, ob = a.reset(), b.reset()
    g = th.Generator(device="cuda"); g.manual_seed(1)
    acts = th.stack([th.randint(0, 2, (T, n), generator=g, device="cuda"), th.randint(0, 3, (T, n), generator=g, device="cuda")], 2)
    r = a.rollout(acts)
    assert th.equal(r["board"][0], ob["board"]) and th.equal(r["dice_roll"][0], ob["dice_roll"])
    for t in range(T):
        o, rew, done, _ = b.step(acts[t].clone())
        assert th.equal(r["board"][t + 1], o["board"]) and th.equal(r["rewards"][t], rew) and th.equal(r["dones"][t], done), t
    assert th.equal(a.state, b.state) and th.equal(a.stats, b.stats) and a.step_index == b.step_index
    o2, rew2, done2, _ = a.step(acts[0].clone())                      # the env carries on from the rollout's last observation
    o3, rew3, done3, _ = b.step(acts[0].clone())
    assert th.equal(o2["board"], o3["board"]) and th.equal(rew2, rew3)
    illegal_before = int(a.stats[2])
    rr = a.rollout(n_steps=5, agent_random=True)
    assert rr["actions"].shape == (5, n, 2) and int(rr["actions"].min()) >= 0 and int(rr["actions"][..., 1].max()) <= 2
    assert int(a.stats[2]) == illegal_before                         # a legal-list agent never ends a game on an illegal move


def test_generic_collector_still_runs_on_cuda():
    """fused=False keeps the torch loop (what a policy with another net_arch, or an env with a policy opponent, gets): same buffer
    contract; the learning run of tests/test_train.py covers the fused collector on this backend"""
    import torch as th
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.train import A2C
    env = EWNDeviceVecEnv(256, 5, 3, seed=5, shaped=True)
    model = A2C("MultiInputPolicy", env, n_steps=4, seed=3, fused=False)
    assert not model.fused
    model.collect_rollouts()
    first_last = model._last_obs["board"].clone()
    model.train()
    model.collect_rollouts()
    assert th.equal(model.buf["board"][0], first_last) and model.num_timesteps == 2 * 4 * 256
    wide = A2C("MultiInputPolicy", EWNDeviceVecEnv(64, 5, 3), n_steps=2, policy_kwargs={"net_arch": (32, 32)})
    assert not wide.fused
    wide.collect_rollouts(); wide.train()
