"""CPU tier: the product library builds for sm_100a, loads, and exports exactly what include/ewn_b200.h declares."""
import ctypes as C
import os
import re
import subprocess

import pytest

from ewn_gym_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "ewn_b200.h")).read()
    return sorted(set(re.findall(r"^EWN_API[^;(]*?\b(ewn_\w+)\s*\(", txt, flags=re.M)))


def test_header_matches_binding_table():
    syms = header_symbols()
    assert len(syms) >= 30
    assert syms == sorted(_lib.PROTOTYPES), "include/ewn_b200.h and ewn_gym_b200/_lib.py PROTOTYPES must list the same entry points"


def test_library_builds_loads_and_exports_every_symbol():
    path = _lib.build()
    lib = _lib.bind(path)                         # raises ImportError on a missing symbol
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert lib.ewn_version() >= 100
    assert [lib.ewn_state_words(n) for n in (2, 3, 5, 6, 9, 10)] == [_lib.EWN_EARG, 1, 1, 2, 2, _lib.EWN_EARG]
    assert lib.ewn_expectiminimax_workspace(10, 4) == 10 * 222 * 8 and lib.ewn_expectiminimax_workspace(1, 6) == 7998 * 8
    assert lib.ewn_mc_workspace(0) == 0
    assert lib.ewn_strerror(_lib.EWN_ESPACE) == b"workspace too small"
    # only ewn_* is exported: the static CUDA runtime inside must not leak into the process
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
    exported = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert exported and all(s.startswith("ewn_") for s in exported), [s for s in exported if not s.startswith("ewn_")][:5]


def test_struct_layouts_match_header():
    """ctypes mirrors of ewn_step_args / ewn_vec_config have the C compiler's size."""
    src = '#include "ewn_b200.h"\n#include <stdio.h>\nint main(){printf("%zu %zu\\n", sizeof(ewn_step_args), sizeof(ewn_vec_config));return 0;}\n'
    exe = os.path.join(ROOT, "tests", "hostsim", "_sizes")
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src, text=True, check=True)
    a, b = map(int, subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split())
    os.remove(exe)
    assert C.sizeof(_lib.StepArgs) == a and C.sizeof(_lib.VecConfig) == b


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to run (and says so) instead of computing on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from ewn_gym_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(5, 3)
    lib = _lib.lib()
    assert lib.ewn_device_count() in (0, _lib.EWN_ECUDA)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ewn_gym_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.lower(), os.path.join(dirpath, f)


def test_numa_binding_helper(tmp_path):
    """bind_host_to_device reads the GPU's NUMA node from sysfs; a node that covers every allowed CPU changes nothing"""
    import os
    from ewn_gym_b200.vec_env import _parse_cpulist, bind_host_to_device
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11} and _parse_cpulist("\n") == set()

    class FakeLib:
        def __init__(self, rc=0):
            self.rc = rc

        def ewn_device_pci_bus_id(self, device, buf, n):
            buf.value = b"0000:1B:00.0"
            return self.rc

    dev = tmp_path / "bus" / "pci" / "devices" / "0000:1b:00.0"
    dev.mkdir(parents=True)
    node = tmp_path / "devices" / "system" / "node" / "node1"
    node.mkdir(parents=True)
    allowed = sorted(os.sched_getaffinity(0))
    (node / "cpulist").write_text(",".join(map(str, allowed)) + "\n")
    (dev / "numa_node").write_text("1\n")
    r = bind_host_to_device(FakeLib(), 0, sysfs=str(tmp_path))
    assert r == {"pci": "0000:1b:00.0", "node": 1, "cpus": len(allowed), "changed": False}
    assert sorted(os.sched_getaffinity(0)) == allowed
    (dev / "numa_node").write_text("-1\n")
    assert bind_host_to_device(FakeLib(), 0, sysfs=str(tmp_path)) is None
    assert bind_host_to_device(FakeLib(rc=-2), 0, sysfs=str(tmp_path)) is None
    assert bind_host_to_device(FakeLib(), 0, sysfs=str(tmp_path / "nowhere")) is None
