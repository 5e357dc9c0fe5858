"""GPU: EWNVecEnv (the SB3 VecEnv duck type over the host-buffer C ABI) against the CPU oracle, step by step."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,L,shaped,staged", [(5, 3, False, False), (5, 3, False, True), (7, 3, False, False), (7, 3, True, False),
                                               (5, 3, True, False), (4, 2, False, False), (7, 4, False, False), (6, 4, True, False),
                                               (9, 7, False, True)])
def test_vec_env_matches_oracle(N, L, shaped, staged):
    """staged = outputs through device staging arrays + D2H copies; default = the kernel writes the page-locked host buffers directly"""
    from ewn_gym_b200.vec_env import EWNVecEnv
    n, seed, env_id0 = 10000 + 123, 321, 7
    env = EWNVecEnv(n, N, L, seed=seed, shaped=shaped, illegal_move_reward=-1.0, illegal_move_tolerance=3, env_id0=env_id0, chunks=5,
                    staged=staged)
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
    assert seen_done > (n // 2 if L <= 3 else 200)            # games on the big layouts last much longer
    env.close()


def test_host_env_contract_buffers_seed_codes():
    """the VecEnv contract on the host-buffer class: outputs of step t intact after step t + 1, seed() effective at the next reset
    and returning seed + i per env, reset_infos of length num_envs, messages from the stored result code (goal reward 0)"""
    from ewn_gym_b200.vec_env import EWNVecEnv
    N, L, n, id0 = 5, 3, 5000, 9
    env = EWNVecEnv(n, N, L, seed=1, reward=0.0, env_id0=id0, chunks=3)
    assert len(env.reset_infos) == n and env.reset_infos[n - 1] == {}
    rng = np.random.default_rng(0)
    mk = lambda: np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
    obs0 = env.reset()
    keep0 = obs0["board"].copy()
    obs1, rew1, done1, infos1 = env.step(mk())
    assert np.array_equal(obs0["board"], keep0)
    keep1 = (obs1["board"].copy(), obs1["dice_roll"].copy(), done1.copy())
    env.step(mk())
    assert np.array_equal(obs1["board"], keep1[0]) and np.array_equal(obs1["dice_roll"], keep1[1]) and np.array_equal(done1, keep1[2])
    assert list(env.seed(4242)) == list(range(4242, 4242 + n))
    obs = env.reset()                                               # fourth call: RESET block of step index 3
    ob, od = O.reset_philox(n, 4242, 3, id0, N, L)
    assert np.array_equal(obs["board"], ob) and np.array_equal(obs["dice_roll"], od)
    want = {O.CODE_AGENT_ILLEGAL: "Invalid move for player! End the game.", O.CODE_AGENT_WON: "You won!", O.CODE_AGENT_LOST: "You lost!"}
    seen = set()
    for t in range(4, 30):
        a = mk()
        _, lst, cnt = O.legal(ob, od, 1, N, L)
        rows = rng.random(n) < 0.9
        a[rows] = lst[np.arange(n), rng.integers(0, 1 << 30, n) % cnt][rows]
        obs, rew, done, infos = env.step(a)
        ref = O.step_philox(ob, od, a, 4242, t, id0, N, L, reward=0.0)
        assert np.array_equal(obs["board"], ref["board"]) and not rew.any()
        for i in np.flatnonzero(done)[:30]:
            assert infos[int(i)]["message"] == want[int(ref["code"][i])]
            seen.add(int(ref["code"][i]))
        ob, od = ref["board"], ref["dice"]
    assert seen == set(want)
    la = env.env_method("get_legal_actions", 1, indices=[0, 7])
    _, lst, cnt = O.legal(ob[[0, 7]], od[[0, 7]], 1, N, L)
    assert la == [lst[k, : cnt[k]].tolist() for k in range(2)]
    env.close()


def test_host_env_with_minimax_opponent_equals_device_env():
    """opponent_policy on the host-buffer class is served by the device class: same games"""
    import torch
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.vec_env import EWNVecEnv
    n = 512
    kw = dict(board_size=5, cube_layer=3, seed=3, env_id0=5, opponent_policy="minimax", max_depth=2)
    host, devv = EWNVecEnv(n, **kw), EWNDeviceVecEnv(n, **kw)
    oh, od = host.reset(), devv.reset()
    assert np.array_equal(oh["board"], od["board"].cpu().numpy())
    rng = np.random.default_rng(2)
    for t in range(20):
        a = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
        oh, rh, dh, ih = host.step(a)
        od, rd, dd, idv = devv.step(torch.from_numpy(a).cuda())
        assert np.array_equal(oh["board"], od["board"].cpu().numpy()) and np.array_equal(rh, rd.cpu().numpy()) and np.array_equal(dh, dd.cpu().numpy())
    host.close(); devv.close()


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


@pytest.mark.parametrize("N,L,n", [(5, 3, 300_000 + 77), (7, 3, 260_000 + 5), (5, 3, 100), (4, 2, 4096)])
def test_persistent_step_kernel_vs_block_kernel_and_oracle(N, L, n):
    """k_vstep (persistent, prefetching, deferred terminal compaction) against k_step (one block per 128 games) on the
    VecEnv surface, every output incl. the compacted terminal observations, several units per block; a slice against the oracle."""
    import torch
    from ewn_gym_b200.engine import Engine
    e = Engine(N, L)
    seed, env_id0 = 99, 1000
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    want = ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice", "term_board", "term_dice")
    sa, sb = e.reset(n, seed=seed, env_id0=env_id0), e.reset(n, seed=seed, env_id0=env_id0)
    board0, dice0 = (x.cpu().numpy() for x in e.decode(sa))
    m = min(n, 3000)
    ob, od = board0[:m].copy(), dice0[:m].copy()
    stats_a, stats_b = torch.zeros(4, dtype=torch.int64, device="cuda"), torch.zeros(4, dtype=torch.int64, device="cuda")
    for t in range(1, 13):
        act = torch.stack([torch.randint(0, 2, (n,), generator=g, device="cuda"), torch.randint(-1, 4, (n,), generator=g, device="cuda")], 1).contiguous()
        ca, cb = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
        ia, ib = torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty(n, dtype=torch.int32, device="cuda")
        ra = e.step(sa, act, seed=seed, step=t, env_id0=env_id0, stats=stats_a, term_compact=(ca, ia), want=want)
        rb = e.step(sb, act, seed=seed, step=t, env_id0=env_id0, stats=stats_b, term_compact=(cb, ib), want=want, block_kernel=True)
        assert torch.equal(sa, sb), t
        for k in ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice"):
            assert torch.equal(ra[k], rb[k]), (t, k)
        F = int(ca.item())
        assert F == int(cb.item()) == int(ra["terminated"].sum().item())
        pa, pb = torch.argsort(ia[:F]), torch.argsort(ib[:F])           # row order is arbitrary: compare by env index
        assert torch.equal(ia[:F][pa], ib[:F][pb])
        assert torch.equal(ia[:F][pa].long(), torch.nonzero(ra["terminated"]).flatten())
        assert torch.equal(ra["term_board"][:F][pa], rb["term_board"][:F][pb]) and torch.equal(ra["term_dice"][:F][pa], rb["term_dice"][:F][pb])
        ref = O.step_philox(ob, od, act[:m].cpu().numpy(), seed, t, env_id0, N, L, reward=1.0, autoreset=True, threads=4)
        assert np.array_equal(ra["obs_board"][:m].cpu().numpy(), ref["board"]) and np.array_equal(ra["obs_dice"][:m].cpu().numpy(), ref["dice"])
        sel = ia[:F][pa] < m
        assert np.array_equal(ra["term_board"][:F][pa][sel].cpu().numpy(), ref["term_board"][ref["terminated"] == 1])
        ob, od = ref["board"], ref["dice"]
    assert torch.equal(stats_a, stats_b) and int(stats_a[3].item()) > 0


@pytest.mark.parametrize("N,L,n", [(5, 3, 200_000 + 33), (7, 3, 70_001)])
def test_persistent_kernel_with_in_kernel_random_agent(N, L, n):
    """k_vstep<AR>: the VecEnv surface with the agent = uniform over its legal list inside the kernel (no action array), against the
    block kernel and the oracle; result codes of the finished games (term_code) included"""
    import torch
    from ewn_gym_b200.engine import Engine
    e = Engine(N, L)
    seed, env_id0 = 77, 123456
    want = ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice", "term_board", "term_dice", "term_code")
    sa, sb = e.reset(n, seed=seed, env_id0=env_id0), e.reset(n, seed=seed, env_id0=env_id0)
    m = min(n, 3000)
    ob, od = (x.cpu().numpy()[:m].copy() for x in e.decode(sa))
    wins = 0
    for t in range(1, 30):
        ca, cb = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32, device="cuda")
        ia, ib = torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty(n, dtype=torch.int32, device="cuda")
        ra = e.step(sa, None, seed=seed, step=t, env_id0=env_id0, agent_random=True, term_compact=(ca, ia), want=want)
        rb = e.step(sb, None, seed=seed, step=t, env_id0=env_id0, agent_random=True, term_compact=(cb, ib), want=want, block_kernel=True)
        assert torch.equal(sa, sb), t
        for k in ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice"):
            assert torch.equal(ra[k], rb[k]), (t, k)
        F = int(ca.item())
        assert F == int(cb.item()) == int(ra["terminated"].sum().item())
        pa, pb = torch.argsort(ia[:F]), torch.argsort(ib[:F])
        assert torch.equal(ia[:F][pa], ib[:F][pb])
        for k in ("term_board", "term_dice", "term_code"):
            assert torch.equal(ra[k][:F][pa], rb[k][:F][pb]), (t, k)
        ref = O.step_philox(ob, od, None, seed, t, env_id0, N, L, reward=1.0, autoreset=True, agent_random=True, threads=4)
        assert np.array_equal(ra["obs_board"][:m].cpu().numpy(), ref["board"]) and np.array_equal(ra["obs_dice"][:m].cpu().numpy(), ref["dice"])
        sel = (ia[:F][pa] < m).cpu().numpy()
        assert np.array_equal(ra["term_code"][:F][pa].cpu().numpy()[sel], ref["code"][ref["terminated"] == 1])
        assert not ra["truncated"].any()                              # a legal-list agent never makes an illegal move
        wins += int((ref["code"] == O.CODE_AGENT_WON).sum())
        ob, od = ref["board"], ref["dice"]
    assert wins > 100


def test_full_size_properties_1mi_games():
    """BASELINE.json configs[1] at full size (1,048,576 games): properties that do not need the oracle -- every board holds each
    cube number at most once and only in range, finished games were reset to the initial layout, the compacted terminal list has
    exactly one row per finished game, the stats counters add up, and a second run from the same seed gives the same checksums;
    a 4,096-game slice is replayed on the oracle."""
    import torch
    from ewn_gym_b200.engine import Engine
    N, L, n, seed = 5, 3, 1 << 20, 9487
    e = Engine(N, L)
    g = torch.Generator(device="cuda")
    want = ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice", "term_board", "term_dice")
    sums = []
    for rep in range(2):
        g.manual_seed(1)
        st = e.reset(n, seed=seed)
        stats = torch.zeros(4, dtype=torch.int64, device="cuda")
        b0, d0 = e.decode(st)
        ob, od = b0[:4096].cpu().numpy(), d0[:4096].cpu().numpy()
        acc = torch.zeros((), dtype=torch.int64, device="cuda")
        fin_total = 0
        for t in range(1, 21):
            act = torch.stack([torch.randint(0, 2, (n,), generator=g, device="cuda"), torch.randint(0, 3, (n,), generator=g, device="cuda")], 1).contiguous()
            cnt, idx = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.empty(n, dtype=torch.int32, device="cuda")
            r = e.step(st, act, seed=seed, step=t, stats=stats, term_compact=(cnt, idx), want=want)
            F = int(cnt.item())
            fin = r["terminated"].bool()
            assert F == int(fin.sum().item())
            assert torch.equal(torch.sort(idx[:F].long()).values, torch.nonzero(fin).flatten())
            fin_total += F
            board = r["obs_board"].reshape(n, N * N).long()
            assert int(board.abs().max().item()) <= 6
            for k in (1, 3, 6):                                    # a cube number appears at most once per colour
                assert int((board == k).sum(1).max().item()) <= 1 and int((board == -k).sum(1).max().item()) <= 1
            assert torch.equal(r["obs_board"][fin], b0[:1].expand(F, N, N))                    # auto-reset: initial layout
            assert int((r["obs_dice"] < 1).sum().item()) == 0 and int((r["obs_dice"] > 6).sum().item()) == 0
            assert torch.equal(r["truncated"].bool() & ~fin, torch.zeros_like(fin))           # truncated implies terminated
            if F:
                tb = r["term_board"][:F].reshape(F, N * N).long()
                assert int((tb != 0).sum(1).min().item()) >= 1                                   # a terminal board is never empty
            acc = acc * 1000003 + (board * torch.arange(1, 26, device="cuda")).sum() + r["obs_dice"].sum() * 31 + (r["reward_f32"].double().sum() * 7).long()
            if rep == 0:
                ref = O.step_philox(ob, od, act[:4096].cpu().numpy(), seed, t, 0, N, L, reward=1.0, autoreset=True, threads=4)
                assert np.array_equal(r["obs_board"][:4096].cpu().numpy(), ref["board"]) and np.array_equal(r["reward_f32"][:4096].cpu().numpy(), ref["reward"].astype(np.float32))
                ob, od = ref["board"], ref["dice"]
        s = stats.cpu().tolist()
        assert s[3] == fin_total and s[0] + s[1] + s[2] == s[3] and s[0] > 0 and s[1] > 0
        sums.append((int(acc.item()), tuple(s), int(st.sum().item())))
    assert sums[0] == sums[1]



@pytest.mark.parametrize("N,L,shaped,term", [(5, 3, False, True), (5, 3, True, True), (7, 3, True, True), (6, 3, False, False), (7, 4, False, True),
                                             (8, 5, True, True)])
def test_device_env_equals_host_env(N, L, shaped, term):
    """EWNDeviceVecEnv (tensors stay in HBM) and EWNVecEnv (host buffers) are the same environment"""
    import torch
    from ewn_gym_b200.device_env import EWNDeviceVecEnv
    from ewn_gym_b200.vec_env import EWNVecEnv
    n = 3000
    kw = dict(board_size=N, cube_layer=L, seed=77, shaped=shaped, illegal_move_tolerance=3, env_id0=12345, terminal_observation=term)
    host, devv = EWNVecEnv(n, **kw), EWNDeviceVecEnv(n, **kw)
    try:
        oh, od = host.reset(), devv.reset()
        assert np.array_equal(oh["board"], od["board"].cpu().numpy()) and np.array_equal(oh["dice_roll"], od["dice_roll"].cpu().numpy())
        rng = np.random.default_rng(5)
        episodes = 0
        for t in range(60):
            a = np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64)
            oh, rh, dh, ih = host.step(a)
            od, rd, dd, idv = devv.step(torch.from_numpy(a).cuda())
            assert np.array_equal(oh["board"], od["board"].cpu().numpy()), t
            assert np.array_equal(oh["dice_roll"], od["dice_roll"].cpu().numpy())
            assert np.array_equal(rh, rd.cpu().numpy()) and np.array_equal(dh, dd.cpu().numpy())
            k = int(dh.sum())
            episodes += k
            assert idv.terminal_count() == k
            if term:
                idx, tb, td = (x.cpu().numpy() for x in idv.terminal_rows())
                assert sorted(idx.tolist()) == np.flatnonzero(dh).tolist()
                for i in np.flatnonzero(dh)[:5]:
                    a_, b_ = ih[int(i)], idv[int(i)]
                    assert a_["message"] == b_["message"] and a_["TimeLimit.truncated"] == b_["TimeLimit.truncated"]
                    assert np.array_equal(a_["terminal_observation"]["board"], b_["terminal_observation"]["board"])
                    assert a_["terminal_observation"]["dice_roll"] == b_["terminal_observation"]["dice_roll"]
            live = int(np.flatnonzero(~dh)[0])
            assert idv[live] == {} and ih[live] == {}
        st = devv.stats.cpu().numpy()
        assert st[3] == episodes > 0 and st[0] + st[1] + st[2] == st[3]
        oh, od = host.reset(), devv.reset()                        # a second reset draws fresh first dice, identically
        assert np.array_equal(oh["dice_roll"], od["dice_roll"].cpu().numpy())
        a = np.zeros((n, 2), np.int64)
        oh, rh, dh, _ = host.step(a)
        od, rd, dd, idv = devv.step(a)                            # numpy actions are accepted too
        assert np.array_equal(oh["board"], od["board"].cpu().numpy()) and np.array_equal(rh, rd.cpu().numpy())
        assert idv.terminal_count() == int(dh.sum())                # the append counters start from zero after a reset
        if term:
            assert host.terminal_count == int(dh.sum())
    finally:
        host.close(); devv.close()
