"""A2C rollout collection and training on the batched env (the reference's train.py:35-63,126-155 drives SB3's A2C with
"MultiInputPolicy" over a SubprocVecEnv of MiniMaxHeuristicEnv).  stable_baselines3 is not installed in this image, so this module is
the torch-only counterpart of the pieces train.py uses, with SB3's defaults and names:

    env = EWNDeviceVecEnv(num_envs, board_size=7, cube_layer=3, shaped=True)        # train.py:126-134 (EWN-v0 = MiniMaxHeuristicEnv)
    env.seed(config["env_seed"]); env.reset()                                       # train.py:135-136
    model = return_model(config, env)                                               # train.py:35-63  -> A2C("MultiInputPolicy", env, n_steps, lr, Tanh)
    model.learn(total_timesteps=config["timesteps_per_epoch"], reset_num_timesteps=False)        # train.py:144-155
    win_rate, avg_score = evaluate(config, model)                                   # train.py:66-117: greedy play vs opponent_policy

Observations, actions, rewards and dones never leave the GPU: the policy reads the tensors the step kernel wrote and writes its
actions into the tensor the next step kernel reads (`env.action_buffer`).  What is reproduced of SB3's A2C (v2 defaults): features =
flattened board + one-hot dice (CombinedExtractor over Dict{Box, Discrete}), separate 64-64 Tanh towers for policy and value,
MultiDiscrete([2, 3]) = two independent categoricals, n_steps = 5, gamma 0.99, gae_lambda 1.0, vf_coef 0.5, ent_coef 0, RMSprop(alpha
0.99, eps 1e-5), max_grad_norm 0.5, advantages not normalised, episode boundaries from `dones` (the env auto-resets).
"""
from __future__ import annotations

from typing import Any, Dict

import numpy as np
import torch as th
from torch import nn


class MultiInputActorCritic(nn.Module):
    """SB3 "MultiInputPolicy" for obs = {"board": Box int16 (N, N), "dice_roll": Discrete(cube_num + 1, start=1)} and
    action = MultiDiscrete([2, 3]) (envs/ewn.py:61-69)."""

    def __init__(self, board_size: int, cube_num: int, activation_fn=nn.Tanh, net_arch=(64, 64)):
        super().__init__()
        self.board_size, self.cube_num = int(board_size), int(cube_num)
        feat = board_size * board_size + cube_num + 1               # Flatten(Box) + one-hot(Discrete(cube_num + 1)): envs/ewn.py:66-68

        def tower():
            layers, d = [], feat
            for h in net_arch:
                layers += [nn.Linear(d, h), activation_fn()]
                d = h
            return nn.Sequential(*layers), d

        self.pi, d_pi = tower()
        self.vf, d_vf = tower()
        self.action_net = nn.Linear(d_pi, 2 + 3)                    # logits of the two categoricals, concatenated (SB3 MultiCategorical)
        self.value_net = nn.Linear(d_vf, 1)
        for m, gain in ((self.pi, np.sqrt(2)), (self.vf, np.sqrt(2)), (self.action_net, 0.01), (self.value_net, 1.0)):
            for layer in (m.modules() if isinstance(m, nn.Sequential) else [m]):
                if isinstance(layer, nn.Linear):                    # SB3 ortho_init
                    nn.init.orthogonal_(layer.weight, gain=gain)
                    nn.init.zeros_(layer.bias)

    def features(self, obs) -> th.Tensor:
        board = obs["board"].reshape(obs["board"].shape[0], -1).float()          # SB3 does not rescale a non-image Box
        dice = nn.functional.one_hot((obs["dice_roll"].long() - 1).clamp_(0, self.cube_num), self.cube_num + 1).float()   # start = 1
        return th.cat([board, dice], 1)

    def _dists(self, f):
        logits = self.action_net(self.pi(f))
        # validate_args=False: the default validation reads the logits back on the host (a device synchronisation per step)
        return (th.distributions.Categorical(logits=logits[:, :2], validate_args=False),
                th.distributions.Categorical(logits=logits[:, 2:], validate_args=False))

    def forward(self, obs, deterministic: bool = False):
        """-> (actions int64[n, 2], values f32[n], log_prob f32[n])"""
        f = self.features(obs)
        d0, d1 = self._dists(f)
        a0 = d0.probs.argmax(1) if deterministic else d0.sample()
        a1 = d1.probs.argmax(1) if deterministic else d1.sample()
        return th.stack([a0, a1], 1), self.value_net(self.vf(f)).squeeze(1), d0.log_prob(a0) + d1.log_prob(a1)

    def evaluate_actions(self, obs, actions):
        """-> (values, log_prob, entropy) of stored actions (SB3 ActorCriticPolicy.evaluate_actions)"""
        f = self.features(obs)
        d0, d1 = self._dists(f)
        return (self.value_net(self.vf(f)).squeeze(1), d0.log_prob(actions[:, 0]) + d1.log_prob(actions[:, 1]), d0.entropy() + d1.entropy())

    def predict_values(self, obs):
        return self.value_net(self.vf(self.features(obs))).squeeze(1)

    @th.no_grad()
    def predict(self, obs, deterministic: bool = True):
        """SB3 BasePolicy.predict for a batch (or one un-batched observation, as train.py:87 calls it) -> (action, None)"""
        single = np.ndim(obs["dice_roll"]) == 0
        dev = next(self.parameters()).device
        o = {"board": th.as_tensor(np.asarray(obs["board"]) if not th.is_tensor(obs["board"]) else obs["board"], device=dev),
             "dice_roll": th.as_tensor(np.asarray(obs["dice_roll"]) if not th.is_tensor(obs["dice_roll"]) else obs["dice_roll"], device=dev)}
        if single:
            o = {"board": o["board"][None], "dice_roll": o["dice_roll"].reshape(1)}
        a, _, _ = self.forward(o, deterministic)
        return (a[0].cpu().numpy() if single else a), None


class _LazyLog(dict):
    """logger values may be device scalars (the fused collector leaves its sums on the GPU): read back when somebody looks"""

    def __getitem__(self, k):
        v = dict.__getitem__(self, k)
        return float(v) if th.is_tensor(v) else v

    def get(self, k, default=None):
        return self[k] if k in self else default

    def values(self):
        return [self[k] for k in self]

    def items(self):
        return [(k, self[k]) for k in self]


def _t(x, device):
    """env output (CUDA tensor, or numpy array from the host-simulator test backend) -> tensor on `device`, no copy when possible"""
    return x if th.is_tensor(x) else th.as_tensor(np.asarray(x), device=device)


class A2C:
    """stable_baselines3.A2C("MultiInputPolicy", env, ...) on a batched env: collect_rollouts (policy forward -> env.step -> buffer,
    train.py:144-155 via OnPolicyAlgorithm.learn) and the A2C update.  `env` follows the VecEnv contract (EWNDeviceVecEnv / EWNVecEnv)."""

    def __init__(self, policy: str, env, n_steps: int = 5, learning_rate: float = 7e-4, gamma: float = 0.99, gae_lambda: float = 1.0,
                 ent_coef: float = 0.0, vf_coef: float = 0.5, max_grad_norm: float = 0.5, policy_kwargs: Dict[str, Any] | None = None,
                 seed: int | None = None, device=None, fused: bool | None = None, persistent: bool | None = None, **_ignored):
        assert policy == "MultiInputPolicy", "the reference trains MultiInputPolicy only (train.py:41,53)"
        self.env, self.n_steps, self.gamma, self.gae_lambda = env, int(n_steps), float(gamma), float(gae_lambda)
        self.ent_coef, self.vf_coef, self.max_grad_norm = float(ent_coef), float(vf_coef), float(max_grad_norm)
        if seed is not None:
            th.manual_seed(int(seed))
        if device is None:
            st = getattr(env, "state", None)
            device = st.device if th.is_tensor(st) else (getattr(getattr(env, "e", None), "xp", None) and getattr(env.e.xp, "device", None)) or "cpu"
        self.device = th.device(device)
        pk = dict(policy_kwargs or {})
        self.policy = MultiInputActorCritic(env.board_size, env.cube_num, activation_fn=pk.get("activation_fn", nn.Tanh),
                                            net_arch=tuple(pk.get("net_arch", (64, 64)))).to(self.device)
        self.optimizer = th.optim.RMSprop(self.policy.parameters(), lr=learning_rate, alpha=0.99, eps=1e-5)
        self.num_timesteps = 0
        self._last_obs, self._last_episode_starts = None, None
        n, T, N = env.num_envs, self.n_steps, env.board_size
        dev = self.device
        self.buf = {"board": th.zeros((T, n, N, N), dtype=th.int16, device=dev), "dice": th.zeros((T, n), dtype=th.int64, device=dev),
                    "actions": th.zeros((T, n, 2), dtype=th.int64, device=dev), "rewards": th.zeros((T, n), device=dev),
                    "starts": th.zeros((T, n), device=dev), "values": th.zeros((T, n), device=dev), "log_probs": th.zeros((T, n), device=dev)}
        self.logger: Dict[str, float] = _LazyLog()
        self.policy_seed = int(seed) if seed is not None else 0
        # fused collection (policy forward and env step as two launches per step, the whole rollout one CUDA graph) when the env is the
        # device-resident one with its in-kernel opponent and the policy has the default 64-wide towers; the generic loop otherwise
        self.fused = fused if fused is not None else self._can_fuse()
        # ... and, for the six-cube layouts, the whole rollout as ONE launch (ewn_rollout: policy forward and env step alternate on chip;
        # bit-identical buffers); persistent=False keeps the two-launches-per-step graph
        self.persistent = self.fused and (persistent if persistent is not None else env.cube_layer <= 3)
        self._graph = None

    # ------------------------------------------------------------------ fused collection (K5: ewn_policy_forward + ewn_step + ewn_gae)
    def _can_fuse(self) -> bool:
        env = self.env
        return (self.device.type == "cuda" and hasattr(env, "bind_step") and getattr(env, "opponent", 1) is None
                and getattr(getattr(env, "e", None), "xp", None) is not None and getattr(env.e.xp, "torch", None) is not None
                and self.policy.pi[0].weight.shape[0] == 64 and self.policy.pi[2].weight.shape[0] == 64
                and isinstance(self.policy.pi[1], nn.Tanh))

    def _build_fused(self):
        """One CUDA graph for a whole rollout: per step ewn_policy_forward (reads observation slot t, writes actions / values / log-probs
        of slot t) and the env's step kernel (reads those actions, writes reward / done of slot t and the observation of slot t + 1
        straight into the buffer -- no copies); then V of the last observation, GAE, and the step-counter bump for the next replay."""
        env, e, T, n, N, dev = self.env, self.env.e, self.n_steps, self.env.num_envs, self.env.board_size, self.device
        if env.state is None:
            env.reset()
        last = env._obs(env._last)
        f = {"board": th.zeros((T + 1, n, N, N), dtype=th.int16, device=dev), "dice": th.zeros((T + 1, n), dtype=th.int64, device=dev),
             "actions": th.zeros((T, n, 2), dtype=th.int64, device=dev), "rewards": th.zeros((T, n), device=dev),
             "dones": th.zeros((T, n), dtype=th.uint8, device=dev), "trunc": th.zeros((n,), dtype=th.uint8, device=dev),
             "values": th.zeros((T, n), device=dev), "log_probs": th.zeros((T, n), device=dev), "last_values": th.zeros((n,), device=dev),
             "adv": th.zeros((T, n), device=dev), "ret": th.zeros((T, n), device=dev), "sums": th.zeros((2,), device=dev)}
        f["board"][T].copy_(last["board"]); f["dice"][T].copy_(last["dice_roll"])      # a replay starts by moving slot T to slot 0
        self.fb = f
        self.buf = {"board": f["board"][:T], "dice": f["dice"][:T], "actions": f["actions"], "rewards": f["rewards"], "values": f["values"],
                    "log_probs": f["log_probs"], "dones": f["dones"]}
        self.advantages, self.returns = f["adv"], f["ret"]
        self._ctr = th.zeros(1, dtype=th.int64, device=dev)
        self._ctr_host = None
        self._net = e.policy_net(self.policy)
        if self.persistent:
            self._roll = e.rollout_plan(self._net, env.state, T, f, seed=env.seed_value, env_id0=env.env_id0, step_dev=self._ctr,
                                        policy_seed=self.policy_seed, reward=env.reward, shaped=env.shaped,
                                        illegal_move_reward=env.illegal_move_reward, tolerance=getattr(env, "_tol", None),
                                        prev_score=getattr(env, "_prev", None), stats=env.stats, gamma=self.gamma, gae_lambda=self.gae_lambda)
            self._graph = "persistent"
            return
        # first launches load the kernels: not inside a capture
        e.policy_forward(self._net, f["board"][0], f["dice"][0], actions=f["actions"][0], values=f["values"][0], logp=f["log_probs"][0])
        e.gae(f["rewards"], f["values"], f["dones"], f["last_values"], self.gamma, self.gae_lambda, f["adv"], f["ret"])
        kw = env._step_kw(0)                             # the env's own step configuration, played once on COPIES of its games
        kw.pop("term_compact", None)
        kw.update(stats=None, want=("reward_f32", "terminated", "truncated", "obs_board", "obs_dice"),
                  out={"reward_f32": f["rewards"][0], "terminated": f["dones"][0], "truncated": f["trunc"], "obs_board": f["board"][1],
                       "obs_dice": f["dice"][1]})
        if env.shaped:
            kw.update(tolerance=env._tol.clone(), prev_score=env._prev.clone())
        e.step(env.state.clone(), f["actions"][0], **kw)
        th.cuda.synchronize(dev)
        side = th.cuda.Stream(device=dev)
        with th.cuda.stream(side):                      # plans bind the stream that is current when they are built
            plans = [env.bind_step(f["actions"][t], {"reward_f32": f["rewards"][t], "terminated": f["dones"][t], "truncated": f["trunc"],
                                                     "obs_board": f["board"][t + 1], "obs_dice": f["dice"][t + 1]}, step=t, step_dev=self._ctr)
                     for t in range(T)]

            def body():
                f["board"][0].copy_(f["board"][T]); f["dice"][0].copy_(f["dice"][T])
                for t in range(T):
                    e.policy_forward(self._net, f["board"][t], f["dice"][t], seed=self.policy_seed, step=t, step_dev=self._ctr, env_id0=env.env_id0,
                                     actions=f["actions"][t], values=f["values"][t], logp=f["log_probs"][t])
                    plans[t].launch(t)
                e.policy_forward(self._net, f["board"][T], f["dice"][T], values=f["last_values"], want=("values",))
                e.gae(f["rewards"], f["values"], f["dones"], f["last_values"], self.gamma, self.gae_lambda, f["adv"], f["ret"])
                f["sums"][0] = f["dones"].sum(); f["sums"][1] = f["rewards"].mean()
                e.counter_add(self._ctr, T)

            graph = th.cuda.CUDAGraph()
            with th.cuda.graph(graph, stream=side):
                body()
        th.cuda.synchronize(dev)
        self._plans, self._graph = plans, graph

    @th.no_grad()
    def _collect_fused(self):
        if self._graph is not None and self._bound_seed != self.env.seed_value:
            self._graph = None                          # env.seed(s) + reset(): the bound launches carry the old Philox key
        if self._graph is None:
            self._build_fused()
            self._bound_seed = self.env.seed_value
        env, f, T = self.env, self.fb, self.n_steps
        if self._ctr_host != env.step_index:            # the env was stepped or reset behind the graph's back: re-seat the counter and
            self._ctr.fill_(env.step_index)             # take the env's latest observation as the rollout's first
            if self._ctr_host is not None:
                last = env._obs(env._last)
                f["board"][T].copy_(last["board"]); f["dice"][T].copy_(last["dice_roll"])
        if self.persistent:
            self._roll(0)                               # one launch: all T steps of every env, V(last), GAE
            f["sums"][0] = f["dones"].sum(); f["sums"][1] = f["rewards"].mean()
            e_ = env.e
            e_.counter_add(self._ctr, T)
        else:
            self._graph.replay()
        env.advance(T, obs=(f["board"][T], f["dice"][T]))
        self._ctr_host = env.step_index
        self._last_obs = {"board": f["board"][T], "dice_roll": f["dice"][T]}
        self._last_episode_starts = f["dones"][T - 1].float()
        self.num_timesteps += env.num_envs * T
        self.logger["rollout/episodes"] = f["sums"][0]
        self.logger["rollout/mean_reward"] = f["sums"][1]

    # ------------------------------------------------------------------ OnPolicyAlgorithm.collect_rollouts
    @th.no_grad()
    def collect_rollouts(self):
        if self.fused:
            return self._collect_fused()
        env, b, dev = self.env, self.buf, self.device
        if self._last_obs is None:
            obs = env.reset()
            self._last_obs = {"board": _t(obs["board"], dev), "dice_roll": _t(obs["dice_roll"], dev)}
            self._last_episode_starts = th.ones(env.num_envs, device=dev)
        ep_done = 0
        for t in range(self.n_steps):
            actions, values, log_probs = self.policy(self._last_obs)
            ab = getattr(env, "action_buffer", None)
            if th.is_tensor(ab):
                ab.copy_(actions)                                   # the step kernel reads the policy's output in place
                new_obs, rewards, dones, infos = env.step(ab)
            else:
                new_obs, rewards, dones, infos = env.step(actions.cpu().numpy())
            # the buffer takes step t's INPUT observation after env.step() has run -- the env's outputs stay intact for one more step
            b["board"][t].copy_(self._last_obs["board"]); b["dice"][t].copy_(self._last_obs["dice_roll"])
            b["actions"][t].copy_(actions); b["values"][t].copy_(values); b["log_probs"][t].copy_(log_probs)
            b["rewards"][t].copy_(_t(rewards, dev).float()); b["starts"][t].copy_(self._last_episode_starts)
            d = _t(dones, dev)
            # every truncation of this env is also a termination (envs/ewn.py:445-478): TimeLimit.truncated is always False, nothing to bootstrap
            self._last_obs = {"board": _t(new_obs["board"], dev), "dice_roll": _t(new_obs["dice_roll"], dev)}
            self._last_episode_starts = d.float()
            ep_done = ep_done + d.sum()
            self.num_timesteps += env.num_envs
        last_values = self.policy.predict_values(self._last_obs)
        # RolloutBuffer.compute_returns_and_advantage (GAE; lambda = 1 gives the n-step return A2C uses)
        adv = th.zeros_like(b["rewards"])
        last_gae = th.zeros(env.num_envs, device=dev)
        for t in reversed(range(self.n_steps)):
            nonterminal = 1.0 - (self._last_episode_starts if t == self.n_steps - 1 else b["starts"][t + 1])
            next_values = last_values if t == self.n_steps - 1 else b["values"][t + 1]
            delta = b["rewards"][t] + self.gamma * next_values * nonterminal - b["values"][t]
            last_gae = delta + self.gamma * self.gae_lambda * nonterminal * last_gae
            adv[t] = last_gae
        self.advantages, self.returns = adv, adv + b["values"]
        self.logger["rollout/episodes"] = float(ep_done)
        self.logger["rollout/mean_reward"] = float(b["rewards"].mean())

    # ------------------------------------------------------------------ A2C.train
    def train(self):
        b = self.buf
        T, n = b["rewards"].shape
        obs = {"board": b["board"].reshape(T * n, *b["board"].shape[2:]), "dice_roll": b["dice"].reshape(T * n)}
        values, log_prob, entropy = self.policy.evaluate_actions(obs, b["actions"].reshape(T * n, 2))
        adv = self.advantages.reshape(-1)
        policy_loss = -(adv * log_prob).mean()
        value_loss = nn.functional.mse_loss(self.returns.reshape(-1), values)
        entropy_loss = -entropy.mean()
        loss = policy_loss + self.ent_coef * entropy_loss + self.vf_coef * value_loss
        self.optimizer.zero_grad()
        loss.backward()
        nn.utils.clip_grad_norm_(self.policy.parameters(), self.max_grad_norm)
        self.optimizer.step()
        self.logger.update({"train/policy_loss": float(policy_loss.detach()), "train/value_loss": float(value_loss.detach()),
                            "train/entropy_loss": float(entropy_loss.detach())})

    def learn(self, total_timesteps: int, reset_num_timesteps: bool = True, **_ignored):
        if reset_num_timesteps:
            self.num_timesteps = 0
        target = self.num_timesteps + int(total_timesteps) if not reset_num_timesteps else int(total_timesteps)
        while self.num_timesteps < target:
            self.collect_rollouts()
            self.train()
        return self

    def predict(self, obs, deterministic: bool = False, **_ignored):
        return self.policy.predict(obs, deterministic)

    def save(self, path):
        th.save({"policy": self.policy.state_dict(), "optimizer": self.optimizer.state_dict(), "num_timesteps": self.num_timesteps}, path)

    def load(self, path):
        ck = th.load(path, map_location=self.device)
        self.policy.load_state_dict(ck["policy"]); self.optimizer.load_state_dict(ck["optimizer"]); self.num_timesteps = ck["num_timesteps"]
        return self


def return_model(config: Dict[str, Any], env) -> A2C:
    """train.py:35-63 for the A2C branch (PPO is SB3's; not provided here)"""
    assert config.get("algorithm", "A2C") == "A2C", "only the A2C branch of train.py:51-62 has a torch-only counterpart"
    return A2C("MultiInputPolicy", env, n_steps=config.get("n_steps", 5), learning_rate=config.get("learning_rate", 7e-4),
               policy_kwargs=dict(activation_fn=nn.Tanh), seed=config.get("model_seed"))


@th.no_grad()
def evaluate(config: Dict[str, Any], model, engine=None):
    """train.py:66-117: eval_episode_num greedy episodes (seeds 0..) on EinsteinWuerfeltNichtEnv against `opponent_policy`
    (train.py:80-82 hard-codes minimax, max_depth 5), all episodes in lock step -> (win rate, average final reward)."""
    from .device_env import EWNDeviceVecEnv
    n = int(config["eval_episode_num"])
    env = EWNDeviceVecEnv(n, config["board_size"], config["cube_layer"], seed=config.get("eval_seed", 0), engine=engine,
                          opponent_policy=config.get("eval_opponent_policy", "minimax"), max_depth=config.get("eval_max_depth", 5))
    dev = model.device
    obs = env.reset()
    final = th.zeros(n, device=dev)
    seen = th.zeros(n, dtype=th.bool, device=dev)
    for _ in range(int(config.get("eval_max_steps", 400))):
        a, _ = model.predict({"board": _t(obs["board"], dev), "dice_roll": _t(obs["dice_roll"], dev)}, deterministic=True)
        obs, rew, done, _ = env.step(a if th.is_tensor(env.action_buffer) else a.cpu().numpy())
        rew, done = _t(rew, dev).float(), _t(done, dev).bool()
        first = done & ~seen                                        # the reward that ends an env's FIRST episode is its score
        final = th.where(first, rew, final)
        seen |= done
        if bool(seen.all()):
            break
    return float((final > 0).float().mean()), float(final.mean())


def train(config: Dict[str, Any], log=print):
    """train.py:120-160 without W&B: SubprocVecEnv([make_env] * num_envs) -> one EWNDeviceVecEnv; env.seed; env.reset; epochs of learn +
    evaluate; the best model by win rate is saved to config["save_dir"] when given."""
    from .device_env import EWNDeviceVecEnv
    env = EWNDeviceVecEnv(config["num_envs"], config["board_size"], config["cube_layer"], shaped=True,
                          illegal_move_reward=config.get("illegal_move_reward", -1.0), illegal_move_tolerance=config.get("illegal_move_tolerance", 10))
    env.seed(config.get("env_seed", 9487))
    env.reset()
    model = return_model(config, env)
    if config.get("checkpoint"):
        model.load(config["checkpoint"])
    best = 0.0
    for epoch in range(int(config["epoch_num"])):
        model.learn(total_timesteps=config["timesteps_per_epoch"], reset_num_timesteps=False)
        win_rate, avg = evaluate(config, model, engine=None)
        log(f"epoch {epoch}: win rate {win_rate:.3f}, avg score {avg:.3f}, timesteps {model.num_timesteps}")
        if win_rate > best:
            best = win_rate
            if config.get("save_dir"):
                import os
                os.makedirs(config["save_dir"], exist_ok=True)
                model.save(os.path.join(config["save_dir"], f"{epoch}.pt"))
    return model, best
