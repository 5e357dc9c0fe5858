"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference (jchen8tw/ewn-gym).

Imports the reference from ``/root/reference`` (present only in the build container, never on the
GPU box) after planting three stub modules in ``sys.modules``: ``gymnasium``, ``stable_baselines3``
and ``pygame`` are not installed in this image and cannot be installed (no network).  The stubs carry
no game logic -- only the names the reference's import chain touches:

  envs/ewn.py:1-2,7                      gymnasium, gymnasium.error.DependencyNotInstalled, stable_baselines3.A2C
  classical_policies/alpha_zero/ewn/EWNPlayers.py:2   pygame

Used by ``tests/golden/make_golden.py`` to produce the committed golden vectors and by the
container-only tests that run the oracle directly against the live reference.  Nothing in the
product package (``ewn_gym_b200``) may import this file.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("EWN_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "envs", "ewn.py"))


def _install_stubs() -> None:
    import numpy as np

    if "gymnasium" not in sys.modules:
        gym = types.ModuleType("gymnasium")

        class Env:  # gymnasium.Env: no-op base
            metadata: dict = {}

            def __init__(self, *a, **k):
                pass

        class _Space:
            def __init__(self, *a, **k):
                self.args, self.kwargs = a, k
                self._rng = np.random.default_rng()

            def seed(self, seed=None):
                self._rng = np.random.default_rng(seed)
                return [seed]

        class MultiDiscrete(_Space):
            def __init__(self, nvec, **k):
                super().__init__(nvec, **k)
                self.nvec = np.asarray(nvec)

            def sample(self):
                return (self._rng.random(self.nvec.shape) * self.nvec).astype(np.int64)

        class Box(_Space):
            def __init__(self, low, high, shape=None, dtype=np.float32, **k):
                super().__init__(low, high, shape, dtype)
                self.low, self.high, self.shape, self.dtype = low, high, shape, dtype

        class Discrete(_Space):
            def __init__(self, n, start=0, **k):
                super().__init__(n, start)
                self.n, self.start = n, start

        class Dict(_Space):
            def __init__(self, spaces=None, **k):
                super().__init__(spaces)
                self.spaces = dict(spaces or {})

            def __getitem__(self, key):
                return self.spaces[key]

        spaces = types.ModuleType("gymnasium.spaces")
        spaces.MultiDiscrete, spaces.Box, spaces.Discrete, spaces.Dict = MultiDiscrete, Box, Discrete, Dict
        error = types.ModuleType("gymnasium.error")

        class DependencyNotInstalled(Exception):
            pass

        error.DependencyNotInstalled = DependencyNotInstalled
        envs_mod = types.ModuleType("gymnasium.envs")
        registration = types.ModuleType("gymnasium.envs.registration")
        _registry: dict = {}

        def register(id, entry_point, **kw):
            _registry[id] = entry_point

        def make(id, **kwargs):
            mod_name, cls_name = _registry[id].split(":")
            mod = __import__(mod_name, fromlist=[cls_name])
            return getattr(mod, cls_name)(**kwargs)

        registration.register = register
        envs_mod.registration = registration
        gym.Env, gym.spaces, gym.error, gym.envs, gym.make = Env, spaces, error, envs_mod, make
        sys.modules["gymnasium"] = gym
        sys.modules["gymnasium.spaces"] = spaces
        sys.modules["gymnasium.error"] = error
        sys.modules["gymnasium.envs"] = envs_mod
        sys.modules["gymnasium.envs.registration"] = registration

    if "stable_baselines3" not in sys.modules:
        sb3 = types.ModuleType("stable_baselines3")

        class _Algo:
            @classmethod
            def load(cls, *a, **k):
                raise RuntimeError("stable_baselines3 is stubbed in the oracle harness")

        sb3.A2C = type("A2C", (_Algo,), {})
        sb3.PPO = type("PPO", (_Algo,), {})
        sys.modules["stable_baselines3"] = sb3

    if "pygame" not in sys.modules:
        sys.modules["pygame"] = types.ModuleType("pygame")


_loaded = None


def load_reference():
    """Return a namespace with the reference's classes, imported unmodified from REFERENCE_ROOT."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT} (expected on the GPU box)")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import envs as ref_envs  # noqa: E402  (reference package)
    import classical_policies as ref_pol  # noqa: E402
    import constants as ref_const  # noqa: E402

    ns = types.SimpleNamespace(
        EinsteinWuerfeltNichtEnv=ref_envs.EinsteinWuerfeltNichtEnv,
        MinimaxEnv=ref_envs.MinimaxEnv,
        MiniMaxHeuristicEnv=ref_envs.MiniMaxHeuristicEnv,
        RandomAgent=ref_pol.RandomAgent,
        ExpectiMinimaxAgent=ref_pol.ExpectiMinimaxAgent,
        MctsAgent=ref_pol.MctsAgent,
        Player=ref_const.Player,
        ClassicalPolicy=ref_const.ClassicalPolicy,
    )
    _loaded = ns
    return ns
