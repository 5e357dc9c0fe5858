"""The N > 1 path (ewn_gym_b200/sharding.py): world_size-2 gloo on CPU with the test-only host build of the kernel bodies,
and the same checks under NCCL on the GPU box (`-m gpu`, needs >= 2 GPUs).  Sharded results must be IDENTICAL to one rank."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.backends import make_engine, to_np

N_GAMES, N_ROOTS, PLAYOUTS, SEED = 4096, 96, 50, 4242


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _single(backend):
    """one-rank reference results"""
    from ewn_gym_b200.sharding import ShardedRollout, mc_predict_sharded
    e = make_engine(backend, 5, 3)
    ro = ShardedRollout(e, N_GAMES, SEED)
    for _ in range(12):
        ro.step(None, want=())
    stats = to_np(ro.reduce_stats()).copy()
    state = to_np(ro.state).copy()
    roots = e.xp.asarray(state[:N_ROOTS], "u64")
    act, wins = mc_predict_sharded(e, roots, SEED, PLAYOUTS)
    return stats, state, to_np(act).copy(), to_np(wins).copy()


def _worker(rank, world_size, port, backend, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    if backend == "cuda":
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world_size, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        from ewn_gym_b200.sharding import ShardedRollout, mc_predict_sharded, split_range
        e = make_engine(backend, 5, 3)
        ro = ShardedRollout(e, N_GAMES, SEED)
        assert (ro.lo, ro.hi) == split_range(N_GAMES, rank, world_size)
        for _ in range(12):
            ro.step(None, want=())
        stats = to_np(ro.reduce_stats()).copy()
        # every rank evaluates the same roots (the first N_ROOTS games of rank 0's slice, broadcast as packed states)
        roots_np = to_np(ro.state)[:N_ROOTS].copy() if rank == 0 else np.zeros(N_ROOTS, np.int64)
        t = torch.from_numpy(roots_np)
        if backend == "cuda":
            t = t.cuda()
        dist.broadcast(t, 0)
        roots = e.xp.asarray(t.cpu().numpy(), "u64")
        act, wins = mc_predict_sharded(e, roots, SEED, PLAYOUTS)
        q.put((rank, stats, to_np(ro.state).copy(), to_np(act).copy(), to_np(wins).copy()))
    finally:
        dist.destroy_process_group()


def _run(backend, world_size):
    ctx = mp.get_context("spawn")
    q, port = ctx.Queue(), _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, backend, q)) for r in range(world_size)]
    for p in procs:
        p.start()
    out = sorted((q.get(timeout=300) for _ in procs), key=lambda x: x[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return out


def _check(backend, world_size):
    stats1, state1, act1, wins1 = _single(backend)
    out = _run(backend, world_size)
    assert stats1[3] > 0 and stats1[0] + stats1[1] + stats1[2] == stats1[3]
    for rank, stats, state, act, wins in out:
        assert np.array_equal(stats, stats1)                       # all-reduced episode counters == single rank
        assert np.array_equal(wins, wins1) and np.array_equal(act, act1)   # merged root statistics and chosen actions
    assert np.array_equal(np.concatenate([o[2] for o in out]), state1)    # the shards ARE the single-rank games


def test_split_range():
    from ewn_gym_b200.sharding import split_range
    for total in (0, 1, 7, 1024, 1025):
        for ws in (1, 2, 3, 8):
            parts = [split_range(total, r, ws) for r in range(ws)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


def test_two_ranks_gloo_match_single_rank():
    _check("hostsim", 2)


@pytest.mark.gpu
def test_two_ranks_nccl_match_single_rank():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    _check("cuda", 2)
