"""Batched engine calls over device-resident tensors: a thin, typed veneer over the C ABI (include/ewn_b200.h).

PyTorch is used for what it is good at here -- device memory, streams, torch.distributed -- and nothing else:
every operation below is one or two launches of the hand-written sm_100a kernels in csrc/ewn_kernels.cu.
Tensors that hold packed states are ``int64`` with shape ``[n]`` (board_size <= 5) or ``[n, 2]`` (<= 9).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

from . import _lib
from ._lib import StepArgs

HEURISTICS = {"hybrid": 0, "min_dist": 1, "two_min_dist": 2, "attk": 3, "sim_winrate": 4}
TOP_LEFT, BOTTOM_RIGHT = 1, 2
CODE_ONGOING, CODE_AGENT_ILLEGAL, CODE_AGENT_WON, CODE_OPP_ILLEGAL, CODE_AGENT_LOST, CODE_ILLEGAL_TOLERATED = range(6)
MESSAGES = {  # info["message"] strings of envs/ewn.py:448,454,473,478
    CODE_AGENT_ILLEGAL: "Invalid move for player! End the game.", CODE_AGENT_WON: "You won!", CODE_OPP_ILLEGAL: "Invalid move for opponent! End the game.",
    CODE_AGENT_LOST: "You lost!",
}


class TorchCuda:
    """Array provider: CUDA tensors on one device, work enqueued on torch's current stream."""

    def __init__(self, device=None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("ewn_gym_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.torch = torch
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else torch.device(device).index or 0)
        t = torch
        self._dt = {"i8": t.int8, "u8": t.uint8, "i16": t.int16, "i32": t.int32, "u32": t.int32, "i64": t.int64, "u64": t.int64,
                    "f32": t.float32, "f64": t.float64}

    def empty(self, shape, dt):
        return self.torch.empty(shape, dtype=self._dt[dt], device=self.device)

    def zeros(self, shape, dt):
        return self.torch.zeros(shape, dtype=self._dt[dt], device=self.device)

    def asarray(self, x, dt):
        t = self.torch
        if isinstance(x, t.Tensor):
            return x.to(device=self.device, dtype=self._dt[dt]).contiguous()
        import numpy as np
        a = np.ascontiguousarray(x)
        if a.dtype == np.uint32:
            a = a.view(np.int32)
        elif a.dtype == np.uint64:
            a = a.view(np.int64)
        return t.as_tensor(a).to(device=self.device, dtype=self._dt[dt]).contiguous()

    def ptr(self, a):
        return None if a is None else a.data_ptr()

    def stream(self):
        return self.torch.cuda.current_stream(self.device).cuda_stream

    def guard(self):
        return self.torch.cuda.device(self.device)


class Engine:
    """One (board_size, cube_layer) configuration bound to a library and an array provider."""

    def __init__(self, board_size: int = 5, cube_layer: int = 3, device=None, arrays=None, library=None):
        assert cube_layer < board_size - 1, "envs/ewn.py:47"
        self.N, self.L = int(board_size), int(cube_layer)
        self.cube_num = self.L * (self.L + 1) // 2
        self.lib = library if library is not None else _lib.lib()
        self.xp = arrays if arrays is not None else TorchCuda(device)
        self.words = self.lib.ewn_state_words_cfg(self.N, self.L)
        if self.words < 1:
            raise ValueError(f"board_size {board_size} not supported (3..9)")

    # ------------------------------------------------------------------ helpers
    def _chk(self, rc, what):
        _lib.check(rc, what, self.lib)

    def _state_shape(self, n):
        return (n,) if self.words == 1 else (n, self.words)

    def new_state(self, n):
        return self.xp.empty(self._state_shape(n), "u64")

    def _n(self, state):
        return int(state.shape[0])

    def _opt(self, x, dt):
        return None if x is None else self.xp.asarray(x, dt)

    # ------------------------------------------------------------------ codec
    def reset(self, n: int, seed: int = 0, step: int = 0, env_id0: int = 0, out=None):
        """setup_game + first die (envs/ewn.py:94-108) from the Philox RESET block."""
        st = self.new_state(n) if out is None else out
        self._chk(self.lib.ewn_reset(self.xp.ptr(st), n, self.N, self.L, seed, step, env_id0, self.xp.stream()), "ewn_reset")
        return st

    def reset_dice(self, dice, out=None):
        d = self.xp.asarray(dice, "i32")
        n = int(d.shape[0])
        st = self.new_state(n) if out is None else out
        self._chk(self.lib.ewn_reset_dice(self.xp.ptr(st), self.xp.ptr(d), n, self.N, self.L, self.xp.stream()), "ewn_reset_dice")
        return st

    def encode(self, board, dice=None, player=None):
        b = self.xp.asarray(board, "i16").reshape(-1, self.N, self.N)
        n = int(b.shape[0])
        d, p = self._opt(dice, "i32"), self._opt(player, "i32")
        if d is not None:
            d = d.reshape(n)
        if p is not None:
            p = p.reshape(-1)
            if int(p.shape[0]) == 1 and n != 1:
                p = self.xp.asarray([int(p[0])] * n, "i32")
        st = self.new_state(n)
        self._chk(self.lib.ewn_encode(self.xp.ptr(b), self.xp.ptr(d), self.xp.ptr(p), self.xp.ptr(st), n, self.N, self.L,
                                      self.xp.stream()), "ewn_encode")
        return st

    def decode(self, state, want_player=False):
        n = self._n(state)
        board, dice = self.xp.empty((n, self.N, self.N), "i16"), self.xp.empty((n,), "i64")
        player = self.xp.empty((n,), "i32") if want_player else None
        self._chk(self.lib.ewn_decode(self.xp.ptr(state), self.xp.ptr(board), self.xp.ptr(dice), self.xp.ptr(player), n, self.N, self.L,
                                      self.xp.stream()), "ewn_decode")
        return (board, dice, player) if want_player else (board, dice)

    def decode_onehot(self, state, want_dice=True):
        """-> (board f32[n,N,N,cube_num] with +1 / -1, dice f32[n,cube_num] one-hot): the NN-facing observation of
        alpha_zero.py:39-58 ewn_obs_one_hot"""
        n = self._n(state)
        board = self.xp.empty((n, self.N, self.N, self.cube_num), "f32")
        dice = self.xp.empty((n, self.cube_num), "f32") if want_dice else None
        self._chk(self.lib.ewn_decode_onehot(self.xp.ptr(state), self.xp.ptr(board), self.xp.ptr(dice), n, self.N, self.L, self.xp.stream()),
                  "ewn_decode_onehot")
        return board, dice

    # ------------------------------------------------------------------ rules
    def legal(self, state, player: int = 0, want_list=True):
        n = self._n(state)
        mask, cnt = self.xp.empty((n,), "u8"), self.xp.empty((n,), "i32")
        lst = self.xp.empty((n, 6, 2), "i8") if want_list else None
        self._chk(self.lib.ewn_legal(self.xp.ptr(state), player, self.xp.ptr(mask), self.xp.ptr(lst), self.xp.ptr(cnt), n, self.N, self.L,
                                     self.xp.stream()), "ewn_legal")
        return mask, lst, cnt

    def check_win(self, state):
        n = self._n(state)
        win = self.xp.empty((n,), "u8")
        self._chk(self.lib.ewn_check_win(self.xp.ptr(state), self.xp.ptr(win), n, self.N, self.L, self.xp.stream()), "ewn_check_win")
        return win

    def ply(self, state, action, player: int = 0):
        """In place. -> (valid u8[n], win u8[n], moved_cube i32[n])"""
        n = self._n(state)
        a = self.xp.asarray(action, "i32").reshape(n, 2)
        valid, win, moved = self.xp.empty((n,), "u8"), self.xp.empty((n,), "u8"), self.xp.empty((n,), "i32")
        self._chk(self.lib.ewn_ply(self.xp.ptr(state), player, self.xp.ptr(a), self.xp.ptr(valid), self.xp.ptr(win), self.xp.ptr(moved), n,
                                   self.N, self.L, self.xp.stream()), "ewn_ply")
        return valid, win, moved

    def set_dice(self, state, dice=None, player=None):
        n = self._n(state)
        d, p = self._opt(dice, "i32"), self._opt(player, "i32")
        self._chk(self.lib.ewn_set_dice(self.xp.ptr(state), self.xp.ptr(d), self.xp.ptr(p), n, self.N, self.L, self.xp.stream()),
                  "ewn_set_dice")
        return state

    def numpy_random_games(self, seeds):
        """eval_random.py's games for `seeds`, reproduced from the seeds alone (the reference's MT19937 / randint stream).
        -> (result i8[n] +1/-1, agent_steps i32[n], final_state, trace u64[n])"""
        sd = self.xp.asarray(seeds, "u32")
        n = int(sd.shape[0])
        res, steps = self.xp.empty((n,), "i8"), self.xp.empty((n,), "i32")
        fin, tr = self.new_state(n), self.xp.empty((n,), "u64")
        self._chk(self.lib.ewn_numpy_random_games(self.xp.ptr(sd), n, self.N, self.L, self.xp.ptr(res), self.xp.ptr(steps), self.xp.ptr(fin),
                                                  self.xp.ptr(tr), self.xp.stream()), "ewn_numpy_random_games")
        return res, steps, fin, tr

    def numpy_stream(self, seeds, K: int = 512):
        """first K outputs of numpy's legacy MT19937 per seed -> (raw u32[n,K], cursor i32[n] = 0)"""
        sd = self.xp.asarray(seeds, "u32")
        n = int(sd.shape[0])
        raw, cur = self.xp.empty((n, K), "u32"), self.xp.zeros((n,), "i32")
        self._chk(self.lib.ewn_numpy_stream(self.xp.ptr(sd), n, K, self.xp.ptr(raw), self.xp.stream()), "ewn_numpy_stream")
        return raw, cur

    def numpy_draw(self, state, raw, cursor, roll_dice=True, pick_player: int = 0):
        """the reference's next draws for every game: the die (in place) and / or a RandomAgent action int32[n,2] for `pick_player`"""
        n = self._n(state)
        act = self.xp.empty((n, 2), "i32") if pick_player else None
        self._chk(self.lib.ewn_numpy_draw(self.xp.ptr(state), self.xp.ptr(raw), self.xp.ptr(cursor), int(raw.shape[1]), n, self.N, self.L,
                                          int(bool(roll_dice)), int(pick_player), self.xp.ptr(act), self.xp.stream()), "ewn_numpy_draw")
        return act

    def flip(self, state, out=None):
        """the position as the other side sees it: np.rot90(-board, 2), side to move mirrored (envs/ewn.py:289-296)"""
        n = self._n(state)
        o = self.new_state(n) if out is None else out
        self._chk(self.lib.ewn_flip(self.xp.ptr(state), self.xp.ptr(o), n, self.N, self.L, self.xp.stream()), "ewn_flip")
        return o

    def random_action(self, state, seed: int, step: int, env_id0: int = 0, out=None):
        n = self._n(state)
        a = self.xp.empty((n, 2), "i64") if out is None else out
        self._chk(self.lib.ewn_random_action(self.xp.ptr(state), self.xp.ptr(a), n, self.N, self.L, seed, step, env_id0,
                                             self.xp.stream()), "ewn_random_action")
        return a

    # ------------------------------------------------------------------ env step
    def step(self, state, action=None, *, seed=0, step=0, env_id0=0, reward=1.0, agent_random=False, autoreset=True,
             shaped=False, illegal_move_reward=-1.0, tolerance=None, prev_score=None, inject=None, stats=None, term_compact=None,
             want=("reward_f32", "terminated", "truncated"), out: Optional[dict] = None, block_kernel=False, no_pdl=False, step_dev=None):
        """One env step for every game (envs/ewn.py:436-486), in place on `state`.

        action : int64[n,2] / int32[n,2] / uint8[n] tensor, or None with agent_random.
        inject : (dice_opp i32[n], opp_action i32[n,2], dice_next i32[n]) for replay mode.
        term_compact : (count int32[1], index int32[n]) -> terminal observations are appended compacted (see ewn_b200.h).
        stats  : int64[4] accumulator {agent wins, agent losses, illegal-move endings, finished episodes}.
        block_kernel : run the VecEnv surface on the block-per-128-games kernel instead of the persistent one (A/B runs).
        want   : names of ewn_step_args outputs to allocate; `out` may carry pre-allocated ones.
        Returns the dict of output tensors.
        """
        n = self._n(state)
        xp = self.xp
        res = dict(out or {})
        shapes = {"reward_f32": ((n,), "f32"), "reward_f64": ((n,), "f64"), "terminated": ((n,), "u8"), "truncated": ((n,), "u8"),
                  "code": ((n,), "i32"), "action_out": ((n,), "u8"), "obs_board": ((n, self.N, self.N), "i16"), "obs_dice": ((n,), "i64"),
                  "term_board": ((n, self.N, self.N), "i16"), "term_dice": ((n,), "i64"), "term_code": ((n,), "u8")}
        for name in want:
            if name not in res:
                res[name] = xp.empty(*shapes[name])
        a = StepArgs()
        a.state, a.n, a.board_size, a.cube_layer = xp.ptr(state), n, self.N, self.L
        keep = []
        if action is not None:
            import numpy as np
            dtn = str(action.dtype).replace("torch.", "")
            kind = {"int64": "i64", "int32": "i32", "uint8": "u8"}.get(dtn, "i64")
            act = xp.asarray(action, kind)          # on the engine's device and contiguous (a no-op for tensors that already are)
            keep.append(act)
            setattr(a, "action_" + kind, xp.ptr(act))
        for name in shapes:
            if name in res and res[name] is not None:
                setattr(a, name, xp.ptr(res[name]))
        if stats is not None:
            a.stats = xp.ptr(stats)
        if term_compact is not None:               # (count int32[1] zeroed by the caller, index int32[n])
            a.term_count, a.term_index = xp.ptr(term_compact[0]), xp.ptr(term_compact[1])
            if len(term_compact) > 2:              # counter to clear for the next step
                a.term_count_clear = xp.ptr(term_compact[2])
        if shaped:
            a.tolerance, a.prev_score = xp.ptr(tolerance), xp.ptr(prev_score)
        if inject is not None:
            d0, oa, d1 = (xp.asarray(inject[0], "i32"), xp.asarray(inject[1], "i32"), xp.asarray(inject[2], "i32"))
            keep += [d0, oa, d1]
            a.inject_dice_opp, a.inject_opp_action, a.inject_dice_next = xp.ptr(d0), xp.ptr(oa), xp.ptr(d1)
        a.reward, a.illegal_move_reward = float(reward), float(illegal_move_reward)
        a.seed, a.step, a.env_id0 = seed, step, env_id0
        a.agent_random, a.autoreset, a.shaped = int(agent_random), int(autoreset), int(shaped)
        a.flags = (1 if block_kernel else 0) | (2 if no_pdl else 0)
        if step_dev is not None:                   # int64[1] device counter added to `step` at run time (CUDA-graph replays)
            a.step_dev = xp.ptr(step_dev)
        self._last_keep = keep                     # StepPlan holds on to these: its launches reuse the pointers
        self._chk(self.lib.ewn_step(C.byref(a), xp.stream()), "ewn_step")
        return res

    def step_dice(self, n: int, seed: int, step: int, env_id0: int = 0, step_dev=None):
        """-> (dice_opp i32[n], dice_next i32[n]): the dice step `step` draws (envs/ewn.py:458 and :483), for external opponents"""
        d0, d1 = self.xp.empty((n,), "i32"), self.xp.empty((n,), "i32")
        self._chk(self.lib.ewn_step_dice(self.xp.ptr(d0), self.xp.ptr(d1), n, self.N, self.L, seed, step, env_id0, self.xp.ptr(step_dev),
                                         self.xp.stream()), "ewn_step_dice")
        return d0, d1

    def counter_add(self, counter, k: int):
        """*counter += k on the current stream (see ewn_step_args.step_dev)"""
        self._chk(self.lib.ewn_counter_add(self.xp.ptr(counter), int(k), self.xp.stream()), "ewn_counter_add")

    # ------------------------------------------------------------------ rollout collection with a policy in the loop
    def policy_net(self, module):
        """ewn_policy_net over the parameters of a train.MultiInputActorCritic, read in place (fp32, contiguous, on this device)"""
        from ._lib import PolicyNet
        lin = {"pi_w1": module.pi[0].weight, "pi_b1": module.pi[0].bias, "pi_w2": module.pi[2].weight, "pi_b2": module.pi[2].bias,
               "vf_w1": module.vf[0].weight, "vf_b1": module.vf[0].bias, "vf_w2": module.vf[2].weight, "vf_b2": module.vf[2].bias,
               "action_w": module.action_net.weight, "action_b": module.action_net.bias,
               "value_w": module.value_net.weight, "value_b": module.value_net.bias}
        net = PolicyNet()
        net.hidden = int(module.pi[0].weight.shape[0])
        for k, t in lin.items():
            if str(t.dtype) != "torch.float32" or not t.is_contiguous() or not t.is_cuda:
                raise ValueError(f"policy parameter {k}: fp32, contiguous, on the GPU expected")
            setattr(net, k, t.data_ptr())
        net._keep = list(lin.values())
        return net

    def policy_forward(self, net, board, dice, *, seed=0, step=0, step_dev=None, env_id0=0, deterministic=False, actions=None, values=None,
                       logp=None, logits=None, want=("actions", "values", "logp")):
        """ActorCriticPolicy.forward for n observations in one launch (ewn_policy_forward) -> dict of the outputs asked for"""
        xp = self.xp
        n = int(board.shape[0])
        out = {"actions": actions, "values": values, "logp": logp, "logits": logits}
        shapes = {"actions": ((n, 2), "i64"), "values": ((n,), "f32"), "logp": ((n,), "f32"), "logits": ((n, 5), "f32")}
        for k in want:
            if out[k] is None:
                out[k] = xp.empty(*shapes[k])
        self._chk(self.lib.ewn_policy_forward(C.byref(net), xp.ptr(board), xp.ptr(dice), n, self.N, self.L, seed, step, xp.ptr(step_dev), env_id0,
                                              int(bool(deterministic)), xp.ptr(out["actions"]), xp.ptr(out["values"]), xp.ptr(out["logp"]),
                                              xp.ptr(out["logits"]), xp.stream()), "ewn_policy_forward")
        return {k: v for k, v in out.items() if v is not None}

    def rollout_plan(self, net, state, T, bufs, *, seed, env_id0=0, step_dev=None, policy_seed=0, reward=1.0, shaped=False,
                     illegal_move_reward=-1.0, tolerance=None, prev_score=None, stats=None, gamma=0.99, gae_lambda=1.0, agent_random=False):
        """A bound ewn_rollout call (the whole T-step collection in one launch) -> callable(step) that enqueues it.  `bufs`: dict with
        board [T+1,n,N,N] i16, dice [T+1,n] i64, actions [T,n,2] i64, rewards / values / log_probs / adv / ret [T,n] f32, dones [T,n] u8,
        last_values [n] f32.  net=None: no policy, the actions are read from bufs["actions"] (values / log_probs / ... not needed)."""
        from ._lib import RolloutArgs
        xp = self.xp
        a = RolloutArgs()
        a.state, a.n, a.board_size, a.cube_layer, a.T, a.shaped = xp.ptr(state), self._n(state), self.N, self.L, int(T), int(bool(shaped))
        if net is not None:
            a.net = C.pointer(net)
        for field, key in (("board", "board"), ("dice", "dice"), ("actions", "actions"), ("rewards", "rewards"), ("dones", "dones"),
                           ("values", "values"), ("logp", "log_probs"), ("last_values", "last_values"), ("advantages", "adv"), ("returns", "ret")):
            setattr(a, field, xp.ptr(bufs.get(key)))
        a.tolerance, a.prev_score, a.stats = xp.ptr(tolerance), xp.ptr(prev_score), xp.ptr(stats)
        a.reward, a.illegal_move_reward = float(reward), float(illegal_move_reward)
        a.seed, a.step, a.env_id0, a.step_dev, a.policy_seed = int(seed), 0, int(env_id0), xp.ptr(step_dev), int(policy_seed)
        a.gamma, a.gae_lambda = float(gamma), float(gae_lambda)
        a.agent_random = int(bool(agent_random))           # net=None only: the in-kernel uniform-over-legal agent; bufs["actions"] receives its moves
        keep = (net, state, bufs, tolerance, prev_score, stats, step_dev)
        fn, ref = self.lib.ewn_rollout, C.byref(a)

        def launch(step=0, _keep=keep, _a=a):
            _a.step = int(step)
            self._chk(fn(ref, xp.stream()), "ewn_rollout")
        return launch

    def gae(self, rewards, values, dones, last_values, gamma=0.99, gae_lambda=1.0, advantages=None, returns=None):
        """RolloutBuffer.compute_returns_and_advantage over [T][n] tensors (ewn_gae) -> (advantages, returns)"""
        xp = self.xp
        T, n = int(rewards.shape[0]), int(rewards.shape[1])
        adv = xp.empty((T, n), "f32") if advantages is None else advantages
        ret = xp.empty((T, n), "f32") if returns is None else returns
        self._chk(self.lib.ewn_gae(xp.ptr(rewards), xp.ptr(values), xp.ptr(dones), xp.ptr(last_values), T, n, float(gamma), float(gae_lambda),
                                   xp.ptr(adv), xp.ptr(ret), xp.stream()), "ewn_gae")
        return adv, ret

    def step_plan(self, state, action=None, **kw):
        """Pre-bind a step (pointers, sizes, flags) so that a rollout loop pays one ctypes call per launch:
        ``plan = e.step_plan(state, action, out=...); plan.launch(step)``.  Same keywords as ``step``."""
        return StepPlan(self, state, action, **kw)

    # ------------------------------------------------------------------ search
    def evaluate(self, state, heuristic="hybrid"):
        n = self._n(state)
        v = self.xp.empty((n,), "f64")
        self._chk(self.lib.ewn_evaluate(self.xp.ptr(state), self.xp.ptr(v), n, self.N, self.L, HEURISTICS[heuristic], self.xp.stream()),
                  "ewn_evaluate")
        return v

    def sim_winrate(self, state, num_simulations: int = 100, seed: int = 0, unit0: int = 0):
        """MinimaxEnv.evaluate('sim_winrate') = simulate() (minimax_ewn.py:215-238) for every position -> f64[n]"""
        n = self._n(state)
        v = self.xp.empty((n,), "f64")
        self._chk(self.lib.ewn_sim_winrate(self.xp.ptr(state), n, self.N, self.L, int(num_simulations), seed, unit0, self.xp.ptr(v),
                                           self.xp.stream()), "ewn_sim_winrate")
        return v

    def expectiminimax(self, state, depth: int, heuristic="hybrid", pruned=True, replay_path=False, out=None, num_simulations: int = 100,
                       seed: int = 0):
        """-> (value f64[n], action i8[n,2]); action (-1,-1) = None for finished roots (minimax.py:22-23).
        replay_path: run depth 3 / 4 on the record + replay kernels (the depth >= 5 path) instead of the fused kernel (A/B).
        heuristic 'sim_winrate': leaves are evaluated by num_simulations random playouts (Philox key `seed`), depth <= 4."""
        n = self._n(state)
        v, a = out if out is not None else (self.xp.empty((n,), "f64"), self.xp.empty((n, 2), "i8"))
        if heuristic == "sim_winrate":
            self._chk(self.lib.ewn_expectiminimax_sim(self.xp.ptr(state), n, self.N, self.L, depth, int(bool(pruned)), int(num_simulations), seed,
                                                      self.xp.ptr(v), self.xp.ptr(a), self.xp.stream()), "ewn_expectiminimax_sim")
            return v, a
        self._chk(self.lib.ewn_expectiminimax(self.xp.ptr(state), n, self.N, self.L, depth, HEURISTICS[heuristic],
                                              int(bool(pruned)) | (2 if replay_path else 0),
                                              self.xp.ptr(v), self.xp.ptr(a), None, 0, self.xp.stream()), "ewn_expectiminimax")
        return v, a

    def playouts(self, state, seed: int, first: int, count: int, root_ids=None, wins=None, plies=None):
        n = self._n(state)
        w = self.xp.zeros((n,), "u32") if wins is None else wins
        rid = self._opt(root_ids, "u32")
        self._chk(self.lib.ewn_playouts(self.xp.ptr(state), self.xp.ptr(rid), n, self.N, self.L, seed, first, count, self.xp.ptr(w),
                                        self.xp.ptr(plies), self.xp.stream()), "ewn_playouts")
        return w

    def mc_root_wins(self, state, seed: int, first: int, count: int, root_ids=None, wins=None, plies=None):
        """-> (wins i32[n,6] by legal-list position, n_children i32[n])"""
        n = self._n(state)
        w = self.xp.zeros((n, 6), "u32") if wins is None else wins
        nch = self.xp.empty((n,), "i32")
        rid = self._opt(root_ids, "u32")
        self._chk(self.lib.ewn_mc_root_wins(self.xp.ptr(state), self.xp.ptr(rid), n, self.N, self.L, seed, first, count, self.xp.ptr(w),
                                            self.xp.ptr(nch), self.xp.ptr(plies), None, 0, self.xp.stream()), "ewn_mc_root_wins")
        return w, nch

    def mc_predict(self, state, seed: int, first: int, count: int, root_ids=None, wins=None, plies=None, want_action=True):
        """MctsAgent.predict in one call (ewn_mc_predict): `wins` is overwritten, no memset launches, expand -> playouts -> arg-max chained
        with programmatic dependent launch.  -> (action i8[n,2] or None, wins i32[n,6], n_children i32[n])"""
        n = self._n(state)
        w = self.xp.empty((n, 6), "u32") if wins is None else wins
        a = self.xp.empty((n, 2), "i8") if want_action else None
        nch = self.xp.empty((n,), "i32")
        rid = self._opt(root_ids, "u32")
        self._chk(self.lib.ewn_mc_predict(self.xp.ptr(state), self.xp.ptr(rid), n, self.N, self.L, seed, first, count, self.xp.ptr(w), self.xp.ptr(a),
                                          self.xp.ptr(nch), self.xp.ptr(plies), None, 0, self.xp.stream()), "ewn_mc_predict")
        return a, w, nch

    def mc_predict_static(self, state, seed: int, first: int, count: int, wins, action, n_children=None, root_ids=None, plies=None):
        """`mc_predict` through ewn_mc_predict_static: the workspace is the library's file-scope device array, the call allocates nothing
        (the caller brings every output: wins u32[n,6], action i8[n,2], optionally n_children i32[n]).  At most
        ewn_mc_static_max_roots() roots per call; calls that overlap on the GPU share -- and race on -- the one array."""
        n = self._n(state)
        rc = self.lib.ewn_mc_predict_static(self.xp.ptr(state), self.xp.ptr(root_ids), n, self.N, self.L, seed, first, count, self.xp.ptr(wins),
                                            self.xp.ptr(action), self.xp.ptr(n_children), self.xp.ptr(plies), self.xp.stream())
        if rc == _lib.EWN_ESPACE:
            raise ValueError(self.lib.ewn_last_cuda_error().decode())
        self._chk(rc, "ewn_mc_predict_static")
        return action, wins, n_children

    def uct_search(self, state, seed: int, rounds: int, batch: int = 32, c: float = 1.41, root_ids=None):
        """UCT at the root over the legal moves with batched random playouts (ewn_uct_search; ucth_temp.py:6-89's intent).
        -> (action i8[n,2], wins i32[n,6], visits i32[n,6])"""
        n = self._n(state)
        w, v, a = self.xp.zeros((n, 6), "u32"), self.xp.zeros((n, 6), "u32"), self.xp.empty((n, 2), "i8")
        rid = self._opt(root_ids, "u32")
        self._chk(self.lib.ewn_uct_search(self.xp.ptr(state), self.xp.ptr(rid), n, self.N, self.L, seed, int(rounds), int(batch), float(c),
                                          self.xp.ptr(w), self.xp.ptr(v), self.xp.ptr(a), None, 0, self.xp.stream()), "ewn_uct_search")
        return a, w, v

    def mc_best_action(self, state, wins):
        n = self._n(state)
        a = self.xp.empty((n, 2), "i8")
        self._chk(self.lib.ewn_mc_best_action(self.xp.ptr(state), self.xp.ptr(wins), self.xp.ptr(a), n, self.N, self.L, self.xp.stream()),
                  "ewn_mc_best_action")
        return a


class StepPlan:
    """A bound ``ewn_step`` call: everything but the step index is fixed.  Keeps its tensors alive."""

    def __init__(self, engine, state, action=None, **kw):
        self.engine = engine
        captured = {}
        real = engine.lib.ewn_step

        class _Capture:                      # record the argument block Engine.step builds instead of launching
            def __getattr__(self_inner, name):
                return getattr(engine.lib, name)

            def ewn_step(self_inner, args_ref, stream):
                captured["args"] = args_ref._obj
                captured["stream"] = stream
                return 0

        lib, engine.lib = engine.lib, _Capture()
        try:
            self.out = engine.step(state, action, **kw)
        finally:
            engine.lib = lib
        self.args, self.stream = captured["args"], captured["stream"]
        self._keep = (state, action, kw, list(getattr(engine, "_last_keep", ())))    # every tensor whose pointer the argument block holds
        self._fn, self._ref = real, C.byref(self.args)

    def launch(self, step: int):
        self.args.step = step
        rc = self._fn(self._ref, self.stream)
        if rc:
            _lib.check(rc, "ewn_step", self.engine.lib)
        return self.out
