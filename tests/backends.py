"""Engine factories for the parity tests.

  "cuda"    : the product -- ewn_gym_b200/csrc/libewn_b200.so on torch CUDA tensors (tests marked `gpu`)
  "hostsim" : TEST ONLY -- tests/hostsim/hostsim.cpp = the same per-game code (csrc/ewn_core.cuh) compiled with g++ and
              driven through the same C signatures on numpy arrays, so the packed-state logic is checked on the CPU tier.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOSTSIM_DIR = os.path.join(ROOT, "tests", "hostsim")
HOSTSIM_SO = os.path.join(HOSTSIM_DIR, "libewn_hostsim.so")


class NumpyArrays:
    _dt = {"i8": np.int8, "u8": np.uint8, "i16": np.int16, "i32": np.int32, "u32": np.int32, "i64": np.int64, "u64": np.int64,
           "f32": np.float32, "f64": np.float64}

    def empty(self, shape, dt):
        return np.zeros(shape, dtype=self._dt[dt])

    zeros = empty

    def asarray(self, x, dt):
        a = np.asarray(x)
        if a.dtype == np.uint32:
            a = a.view(np.int32)
        return np.ascontiguousarray(a, dtype=self._dt[dt])

    def ptr(self, a):
        return None if a is None else a.ctypes.data

    def stream(self):
        return None

    def guard(self):
        import contextlib
        return contextlib.nullcontext()


def build_hostsim():
    src = os.path.join(HOSTSIM_DIR, "hostsim.cpp")
    core = os.path.join(ROOT, "ewn_gym_b200", "csrc", "ewn_core.cuh")
    hdr = os.path.join(ROOT, "include", "ewn_b200.h")
    if not os.path.exists(HOSTSIM_SO) or any(os.path.getmtime(HOSTSIM_SO) < os.path.getmtime(f) for f in (src, core, hdr)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-Wno-unknown-pragmas", "-o", HOSTSIM_SO, src],
                       check=True, cwd=HOSTSIM_DIR)
    return HOSTSIM_SO


_hostsim_lib = None


def make_engine(backend, N=5, L=3):
    from ewn_gym_b200 import _lib
    from ewn_gym_b200.engine import Engine
    if backend == "cuda":
        return Engine(N, L)
    global _hostsim_lib
    if _hostsim_lib is None:
        _hostsim_lib = _lib.bind(build_hostsim(), optional=_lib._VEC_ONLY)
        assert _hostsim_lib.ewn_version() < 0
    return Engine(N, L, arrays=NumpyArrays(), library=_hostsim_lib)


def to_np(x):
    if x is None or isinstance(x, np.ndarray):
        return x
    return x.detach().cpu().numpy()


BACKENDS = [pytest.param("hostsim", id="hostsim"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]
