"""ctypes binding of libewn_b200.so (include/ewn_b200.h).

The library is built in-tree (``ewn_gym_b200/csrc/Makefile`` or ``__graft_entry__.build()``).  There is no CPU
fallback: a missing library raises ``ImportError`` with the build command, and compute calls without a CUDA
device return ``EWN_ECUDA`` which ``check()`` turns into ``RuntimeError``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
# EWN_B200_LIB: load another build of the same sources instead (the -DEWN_CHECKS build of tools/checked_build.sh, experiment variants)
LIB_PATH = os.environ.get("EWN_B200_LIB") or os.path.join(CSRC, "libewn_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-Xlinker", "--exclude-libs,ALL"]

EWN_OK, EWN_EARG, EWN_ECUDA, EWN_ENOSUP, EWN_ESPACE = 0, -1, -2, -3, -4

p_void = C.c_void_p
i64, i32, u64, u32 = C.c_int64, C.c_int32, C.c_uint64, C.c_uint32


class StepArgs(C.Structure):
    """struct ewn_step_args"""
    _fields_ = [
        ("state", p_void), ("n", i64), ("board_size", i32), ("cube_layer", i32),
        ("action_i64", p_void), ("action_i32", p_void), ("action_u8", p_void),
        ("reward_f32", p_void), ("reward_f64", p_void), ("terminated", p_void), ("truncated", p_void), ("code", p_void),
        ("action_out", p_void), ("obs_board", p_void), ("obs_dice", p_void), ("term_board", p_void), ("term_dice", p_void),
        ("term_count", p_void), ("term_index", p_void), ("term_index_base", i64), ("term_count_clear", p_void), ("stats", p_void),
        ("tolerance", p_void), ("prev_score", p_void),
        ("inject_dice_opp", p_void), ("inject_opp_action", p_void), ("inject_dice_next", p_void),
        ("reward", C.c_double), ("illegal_move_reward", C.c_double),
        ("seed", u64), ("step", u64), ("env_id0", u64),
        ("agent_random", i32), ("autoreset", i32), ("shaped", i32), ("flags", i32), ("step_dev", p_void), ("term_code", p_void),
    ]


class PolicyNet(C.Structure):
    """struct ewn_policy_net"""
    _fields_ = [("hidden", i32), ("reserved", i32)] + [(k, p_void) for k in (
        "pi_w1", "pi_b1", "pi_w2", "pi_b2", "vf_w1", "vf_b1", "vf_w2", "vf_b2", "action_w", "action_b", "value_w", "value_b")]


class RolloutArgs(C.Structure):
    """struct ewn_rollout_args"""
    _fields_ = [("state", p_void), ("n", i64), ("board_size", i32), ("cube_layer", i32), ("T", i32), ("shaped", i32), ("net", C.POINTER(PolicyNet)),
                ("board", p_void), ("dice", p_void), ("actions", p_void), ("rewards", p_void), ("dones", p_void), ("values", p_void), ("logp", p_void),
                ("last_values", p_void), ("advantages", p_void), ("returns", p_void), ("tolerance", p_void), ("prev_score", p_void), ("stats", p_void),
                ("reward", C.c_double), ("illegal_move_reward", C.c_double), ("seed", u64), ("step", u64), ("env_id0", u64), ("step_dev", p_void),
                ("policy_seed", u64), ("gamma", C.c_double), ("gae_lambda", C.c_double), ("agent_random", i32), ("reserved", i32)]


class VecConfig(C.Structure):
    """struct ewn_vec_config"""
    _fields_ = [
        ("n", i64), ("board_size", i32), ("cube_layer", i32), ("reward", C.c_double), ("illegal_move_reward", C.c_double),
        ("seed", u64), ("env_id0", u64), ("shaped", i32), ("illegal_move_tolerance", i32), ("chunks", i32), ("device", i32), ("flags", i32), ("reserved", i32),
    ]


# name -> (restype, argtypes); every symbol include/ewn_b200.h declares
PROTOTYPES = {
    "ewn_version": (C.c_int, []),
    "ewn_strerror": (C.c_char_p, [C.c_int]),
    "ewn_last_cuda_error": (C.c_char_p, []),
    "ewn_device_count": (C.c_int, []),
    "ewn_device_pci_bus_id": (C.c_int, [C.c_int, C.c_char_p, C.c_int]),
    "ewn_state_words": (C.c_int, [C.c_int]),
    "ewn_state_words_cfg": (C.c_int, [C.c_int, C.c_int]),
    "ewn_launch_count": (i64, []),
    "ewn_reset": (C.c_int, [p_void, i64, C.c_int, C.c_int, u64, u64, u64, p_void]),
    "ewn_reset_dice": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_encode": (C.c_int, [p_void, p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_decode": (C.c_int, [p_void, p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_legal": (C.c_int, [p_void, C.c_int, p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_check_win": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_ply": (C.c_int, [p_void, C.c_int, p_void, p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_set_dice": (C.c_int, [p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_decode_onehot": (C.c_int, [p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_numpy_random_games": (C.c_int, [p_void, i64, C.c_int, C.c_int, p_void, p_void, p_void, p_void, p_void]),
    "ewn_numpy_stream": (C.c_int, [p_void, i64, C.c_int, p_void, p_void]),
    "ewn_numpy_draw": (C.c_int, [p_void, p_void, p_void, C.c_int, i64, C.c_int, C.c_int, C.c_int, C.c_int, p_void, p_void]),
    "ewn_flip": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_peer_alloc": (C.c_int, [C.c_size_t, C.POINTER(p_void), p_void]),
    "ewn_peer_open": (C.c_int, [p_void, C.POINTER(p_void)]),
    "ewn_peer_close": (C.c_int, [p_void]),
    "ewn_peer_free": (C.c_int, [p_void]),
    "ewn_mc_best_action_peers": (C.c_int, [p_void, p_void, p_void, C.c_int, C.c_int, C.c_uint32, C.c_uint32, p_void, p_void, i64, C.c_int,
                                           C.c_int, p_void]),
    "ewn_random_action": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u64, u64, p_void]),
    "ewn_step": (C.c_int, [C.POINTER(StepArgs), p_void]),
    "ewn_counter_add": (C.c_int, [p_void, u64, p_void]),
    "ewn_step_dice": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u64, u64, p_void, p_void]),
    "ewn_policy_forward": (C.c_int, [C.POINTER(PolicyNet), p_void, p_void, i64, C.c_int, C.c_int, u64, u64, p_void, u64, C.c_int, p_void, p_void, p_void,
                                     p_void, p_void]),
    "ewn_rollout": (C.c_int, [C.POINTER(RolloutArgs), p_void]),
    "ewn_gae": (C.c_int, [p_void, p_void, p_void, p_void, C.c_int, i64, C.c_double, C.c_double, p_void, p_void, p_void]),
    "ewn_evaluate": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, C.c_int, p_void]),
    "ewn_expectiminimax_workspace": (C.c_size_t, [i64, C.c_int]),
    "ewn_expectiminimax": (C.c_int, [p_void, i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, p_void, p_void, p_void, C.c_size_t, p_void]),
    "ewn_sim_winrate": (C.c_int, [p_void, i64, C.c_int, C.c_int, C.c_int, u64, u32, p_void, p_void]),
    "ewn_expectiminimax_sim": (C.c_int, [p_void, i64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u64, p_void, p_void, p_void]),
    "ewn_playouts": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u32, u32, p_void, p_void, p_void]),
    "ewn_mc_workspace": (C.c_size_t, [i64]),
    "ewn_mc_root_wins": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u32, u32, p_void, p_void, p_void, p_void, C.c_size_t, p_void]),
    "ewn_uct_workspace": (C.c_size_t, [i64]),
    "ewn_uct_search": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, C.c_int, u32, C.c_double, p_void, p_void, p_void, p_void, C.c_size_t, p_void]),
    "ewn_mc_predict": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u32, u32, p_void, p_void, p_void, p_void, p_void, C.c_size_t, p_void]),
    "ewn_mc_static_max_roots": (i64, []),
    "ewn_mc_predict_static": (C.c_int, [p_void, p_void, i64, C.c_int, C.c_int, u64, u32, u32, p_void, p_void, p_void, p_void, p_void]),
    "ewn_mc_best_action": (C.c_int, [p_void, p_void, p_void, i64, C.c_int, C.c_int, p_void]),
    "ewn_vec_create": (C.c_int, [C.POINTER(VecConfig), C.POINTER(p_void)]),
    "ewn_vec_destroy": (None, [p_void]),
    "ewn_vec_reset": (C.c_int, [p_void, p_void, p_void]),
    "ewn_vec_step": (C.c_int, [p_void] * 11),
    "ewn_vec_step2": (C.c_int, [p_void] * 12),
    "ewn_vec_set_seed": (C.c_int, [p_void, u64]),
    "ewn_host_alloc": (C.c_int, [C.POINTER(p_void), C.c_size_t]),
    "ewn_host_free": (C.c_int, [p_void]),
    "ewn_vec_state": (p_void, [p_void]),
    "ewn_vec_step_index": (u64, [p_void]),
}
# entry points a host simulator (tests/hostsim) does not provide
_VEC_ONLY = {"ewn_vec_create", "ewn_vec_destroy", "ewn_vec_reset", "ewn_vec_step", "ewn_vec_step2", "ewn_vec_set_seed", "ewn_host_alloc", "ewn_host_free",
             "ewn_vec_state", "ewn_vec_step_index", "ewn_device_pci_bus_id",
             "ewn_mc_static_max_roots", "ewn_mc_predict_static", "ewn_policy_forward", "ewn_gae", "ewn_rollout"}          # (device-memory entry points: the test-only host build has none)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libewn_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, "ewn_kernels.cu"), os.path.join(CSRC, "ewn_core.cuh"),
            os.path.join(os.path.dirname(_HERE), "include", "ewn_b200.h")]
    if os.environ.get("EWN_B200_LIB"):
        return LIB_PATH
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-o", LIB_PATH, srcs[0]]
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB_PATH


def bind(path: str, optional=()):
    """dlopen `path` and attach the prototypes above. Raises if a required symbol is missing."""
    lib = C.CDLL(path)
    for name, (res, args) in PROTOTYPES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if name in optional:
                continue
            raise ImportError(f"{path} does not export {name} (include/ewn_b200.h)")
        fn.restype, fn.argtypes = res, args
    return lib


_lib = None


def lib():
    """The product library. Never falls back to anything else."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {CSRC}` or `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). ewn_gym_b200 has no CPU fallback.")
        _lib = bind(LIB_PATH)
    return _lib


def check(rc: int, what: str = "ewn call", library=None):
    if rc == EWN_OK:
        return
    l = library or lib()
    msg = l.ewn_strerror(rc).decode()
    if rc == EWN_ECUDA:
        msg += ": " + l.ewn_last_cuda_error().decode()
    raise RuntimeError(f"{what} failed ({rc}): {msg}")
