"""Root-statistics merge over peer memory (ewn_gym_b200/peer.py, ewn_mc_best_action_peers): the fused sum + arg-max must equal
all-reduce + ewn_mc_best_action, which in turn equals one rank playing every playout (mcts.py:59-68).

CPU tier: several ranks played in ONE process on the host build.  GPU tier: the same on one device, the time-out path, and
(with >= 2 GPUs) real processes exchanging CUDA IPC handles and reading each other's counters over NVLink."""
import os
import socket

import numpy as np
import pytest
import torch

from tests.backends import BACKENDS, make_engine, to_np

SEED = 77


def _roots(e, n):
    st = e.reset(n, seed=SEED + 1)
    for s in range(1, 7):
        e.step(st, None, seed=SEED + 1, step=s, agent_random=True, autoreset=True, want=())
    return st


def _set_flag(ex, src, epoch):
    ex.flags[src] = epoch


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_fused_merge_equals_sum_then_argmax(backend, world):
    from ewn_gym_b200.peer import PeerExchange
    from ewn_gym_b200.sharding import split_range
    e = make_engine(backend, 5, 3)
    n, playouts = 257, 24
    roots = _roots(e, n)
    handles = []
    exs = [PeerExchange(e, n, handles=handles, rank=r, world=world) for r in range(world)]
    for ex in exs:
        ex.connect(handles)
    try:
        for epoch in (1, 2, 3):                                   # three epochs: both win buffers and a reuse
            seed = SEED + epoch
            one, _ = e.mc_root_wins(roots, seed, 0, playouts)
            want = to_np(e.mc_best_action(roots, one))
            for r, ex in enumerate(exs):
                lo, hi = split_range(playouts, r, world)
                e.mc_root_wins(roots, seed, lo, hi - lo, wins=ex.wins_buffer())
            for ex in exs:                                        # one process plays the ranks in turn: nobody can wait for a later
                for s in range(world):                            # rank, so the publications are made up front
                    _set_flag(ex, s, epoch)
            for ex in exs:
                a, tot = ex.best_action(roots)
                assert np.array_equal(to_np(tot).view(np.uint32), to_np(one).view(np.uint32))
                assert np.array_equal(to_np(a), want)
                ex.check()
            for ex in exs:                                        # the kernel's own publication
                f = to_np(ex.flags).view(np.uint32)
                assert (f[:world] == epoch).all() and f[16] == 0
    finally:
        for ex in exs:
            ex.close()


@pytest.mark.parametrize("backend", BACKENDS)
def test_missing_peer_times_out_instead_of_hanging(backend):
    from ewn_gym_b200.peer import PeerExchange
    e = make_engine(backend, 5, 3)
    roots = _roots(e, 64)
    handles = []
    exs = [PeerExchange(e, 64, handles=handles, rank=r, world=2, timeout_ms=30) for r in range(2)]
    for ex in exs:
        ex.connect(handles)
    try:
        e.mc_root_wins(roots, SEED, 0, 4, wins=exs[0].wins_buffer())
        exs[0].best_action(roots)                                 # rank 1 never publishes epoch 1
        with pytest.raises(RuntimeError, match="did not publish epoch 1"):
            exs[0].check()
    finally:
        for ex in exs:
            ex.close()


def test_argument_checks():
    from ewn_gym_b200.peer import PeerExchange
    e = make_engine("hostsim", 5, 3)
    roots = _roots(e, 8)
    ex = PeerExchange(e, 8, handles=[], rank=0, world=1)
    ex.connect([(b"", ex._own.value)])
    try:
        import ctypes as C
        a = np.zeros((8, 2), np.int8)
        call = lambda rank, world, epoch: e.lib.ewn_mc_best_action_peers(roots.ctypes.data, ex._win_ptrs[0], ex._flag_ptrs, rank, world, epoch,
                                                                       0, None, a.ctypes.data, 8, 5, 3, None)
        assert call(0, 1, 0) == -1 and call(1, 1, 1) == -1 and call(0, 9, 1) == -1 and call(0, 0, 1) == -1
        assert call(0, 1, 1) == 0
    finally:
        ex.close()


# ---------------------------------------------------------------------------------------------- real processes, real NVLink
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from ewn_gym_b200.peer import PeerExchange
        from ewn_gym_b200.sharding import mc_predict_peers, mc_predict_sharded
        e = make_engine("cuda", 5, 3)
        n = 1000
        roots = _roots(e, n)                                       # same seed on every rank -> same roots
        ex = PeerExchange(e, n)
        out = []
        for it in range(5):
            a, tot = mc_predict_peers(e, ex, roots, SEED + it, 40)
            a2, w2 = mc_predict_sharded(e, roots, SEED + it, 40)
            out.append((to_np(a).copy(), to_np(tot).copy(), to_np(a2).copy(), to_np(w2).copy()))
        ex.check()
        # a whole arena with the Monte-Carlo agent sharded over the ranks: same games, same moves, same winners as one rank
        from ewn_gym_b200.arena import MonteCarloPolicy, RandomPolicy, play_matches
        ex2 = PeerExchange(e, 512)
        r = play_matches(e, MonteCarloPolicy(4, 5, seed=3, exchange=ex2), RandomPolicy(seed=5), 512, seed=11)
        ex2.check()
        dist.barrier()
        ex.close(); ex2.close()
        q.put((rank, out, r["winner"].copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 8])
def test_processes_merge_over_nvlink(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=300) for _ in procs), key=lambda x: x[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    e = make_engine("cuda", 5, 3)
    roots = _roots(e, 1000)
    from ewn_gym_b200.arena import MonteCarloPolicy, RandomPolicy, play_matches
    solo = play_matches(e, MonteCarloPolicy(4, 5, seed=3), RandomPolicy(seed=5), 512, seed=11)
    assert solo["unfinished"] == 0 and solo["agent_wins"] > 300                 # 20 playouts per action already beat random
    for rank, out, winner in res:
        assert np.array_equal(winner, solo["winner"])
    for it in range(5):
        one, _ = e.mc_root_wins(roots, SEED + it, 0, 40)
        want = to_np(e.mc_best_action(roots, one))
        for rank, out, _w in res:
            a, tot, a2, w2 = out[it]
            assert np.array_equal(tot, to_np(one)) and np.array_equal(w2, to_np(one))
            assert np.array_equal(a, want) and np.array_equal(a2, want)
