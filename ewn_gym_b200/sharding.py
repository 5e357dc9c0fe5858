"""Multi-GPU sharding of the path: one process per GPU, independent games / playouts per rank, ONE all-reduce of counts.

Nothing on this path exchanges game state: games, roots and playouts are independent units, and Philox is
keyed on GLOBAL ids (env id, root id, playout index), so results do not depend on the number of ranks.

  * env step / rollouts : rank r owns games [r*n_local, (r+1)*n_local)  -> `env_id0 = r * n_local`;
                          after a rollout `all_reduce(SUM)` of stats int64[4] = {wins, losses, illegal endings, episodes}
  * flat Monte-Carlo    : every rank holds all roots and plays playout indices [first_r, first_r + count_r) of every
                          (root, action); `all_reduce(SUM)` of wins[n][6] is the root-statistics merge, followed by the same
                          first-arg-max on every rank (classical_policies/mcts.py:68)
  * expectiminimax      : roots are split contiguously; values / actions are gathered only if the caller wants them

The collective is `torch.distributed.all_reduce` (NCCL over NVLink on the GPU box, gloo in the CPU tests); messages are
at most a few hundred KB once per batch, i.e. latency-bound.  For the Monte-Carlo merge, where that latency is a third of
the step on 8 GPUs, `mc_predict_peers` replaces all-reduce + arg-max by one kernel over NVLink peer memory (peer.py).
"""
from __future__ import annotations

import numpy as np


def world():
    """(rank, world_size) of the default process group, (0, 1) when torch.distributed is not initialised"""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


def split_range(total: int, rank: int, world_size: int):
    """contiguous [lo, hi) share of `total` units for `rank`; the first `total % world_size` ranks get one more"""
    base, extra = divmod(int(total), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def all_reduce_sum(x, group=None):
    """In-place SUM all-reduce of a count tensor / array.  numpy arrays (CPU tests) go through a torch view."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return x
    if isinstance(x, np.ndarray):
        t = torch.from_numpy(x.view(np.int32) if x.dtype == np.uint32 else x)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return x
    dist.all_reduce(x, op=dist.ReduceOp.SUM, group=group)
    return x


def mc_root_wins_sharded(engine, state, seed: int, playouts_per_action: int, root_ids=None, first: int = 0, group=None):
    """Flat Monte-Carlo statistics of `state` (all ranks pass the same roots): rank r plays its share of the playout index
    range, the win counts are summed over ranks.  -> (wins[n][6], n_children[n]); identical on every rank and identical to a
    single-rank call with the same (seed, root_ids, first, playouts_per_action)."""
    rank, ws = world()
    lo, hi = split_range(playouts_per_action, rank, ws)
    wins, nch = engine.mc_root_wins(state, seed, first + lo, hi - lo, root_ids=root_ids)
    all_reduce_sum(wins, group)
    return wins, nch


def mc_predict_sharded(engine, state, seed: int, playouts_per_action: int, root_ids=None, group=None):
    """MctsAgent.predict on every root with the playouts sharded over the ranks -> (action int8[n][2], wins)"""
    wins, _ = mc_root_wins_sharded(engine, state, seed, playouts_per_action, root_ids, 0, group)
    return engine.mc_best_action(state, wins), wins


def mc_predict_peers(engine, exchange, state, seed: int, playouts_per_action: int, root_ids=None, first: int = 0, check: bool = False):
    """mc_predict_sharded without the collective library: the win counters are written straight into this rank's NVLink-visible
    buffer and ONE kernel per rank merges all ranks' counters with peer loads and takes the arg-max (ewn_gym_b200/peer.py).
    -> (action int8[n][2], wins u32[n][6]); identical on every rank and identical to mc_predict_sharded.
    check=True raises if the wait for the peers timed out (that costs a device synchronisation; arena.play_matches checks at its own)."""
    lo, hi = split_range(playouts_per_action, exchange.rank, exchange.world)
    w = exchange.wins_buffer(zero=False)                      # three launches per decision: expand (clears w), playouts, merge + arg-max
    engine.mc_predict(state, seed, first + lo, hi - lo, root_ids=root_ids, wins=w, want_action=False)
    out = exchange.best_action(state)
    if check:                                    # synchronises; a timed-out wait also shows as action (-1, -1) on every root
        exchange.check()
    return out


class ShardedRollout:
    """Rank-local slice of a global batch of `n_global` games stepped with a caller-supplied action function.

    env ids are global, so the concatenation of all ranks' games equals the single-process run with n_global games."""

    def __init__(self, engine, n_global: int, seed: int, shaped: bool = False, reward: float = 1.0, group=None):
        self.engine, self.seed, self.shaped, self.reward, self.group = engine, int(seed), bool(shaped), float(reward), group
        self.rank, self.world = world()
        self.lo, self.hi = split_range(n_global, self.rank, self.world)
        self.n = self.hi - self.lo
        self.state = engine.reset(self.n, seed=self.seed, step=0, env_id0=self.lo)
        self.stats = engine.xp.zeros((4,), "i64")
        self.t = 0
        if shaped:
            self.tolerance = engine.xp.asarray(np.full(self.n, 10, np.int32), "i32")
            self.prev_score = engine.evaluate(self.state)

    def step(self, action=None, want=("reward_f32", "terminated", "truncated"), **kw):
        self.t += 1
        extra = dict(tolerance=self.tolerance, prev_score=self.prev_score) if self.shaped else {}
        return self.engine.step(self.state, action, seed=self.seed, step=self.t, env_id0=self.lo, reward=self.reward,
                                agent_random=action is None, autoreset=True, shaped=self.shaped, stats=self.stats, want=want, **extra, **kw)

    def reduce_stats(self):
        """-> int64[4] {agent wins, agent losses, illegal endings, finished episodes} summed over all ranks"""
        return all_reduce_sum(self.stats, self.group)
