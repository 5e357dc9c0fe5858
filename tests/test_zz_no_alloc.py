"""ewn_mc_predict_static: the Monte-Carlo decision with its workspace in a file-scope device array of the library -- an entry point
for callers that cannot allocate (it has no counterpart in the reference; see include/ewn_b200.h).  Called through the C ABI with raw
device pointers and checked, bit for bit (integer win counts, int8 actions), against ewn_mc_predict on two calls in a row and more.

This file is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _roots(e, n, seed):
    st = e.reset(n, seed=seed)
    for s in range(1, 6):
        e.step(st, None, seed=seed, step=s, agent_random=True, autoreset=(s != 5), want=())
    return st


def _ptr(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("N", [5, 7])
def test_static_workspace_entry_point_equals_mc_predict(N):
    import torch
    from ewn_gym_b200 import _lib
    from ewn_gym_b200.engine import Engine
    lib = _lib.lib()
    dev = torch.device("cuda", 0)
    e = Engine(N, 3, device=dev)
    cap = int(lib.ewn_mc_static_max_roots())
    assert cap == 16384
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call, (n, seed, first, count) in enumerate([(4000, 3, 0, 24), (cap, 4, 5, 16), (1, 5, 0, 40), (4000, 3, 0, 24)]):
        st = _roots(e, n, seed=20 + call)
        want_a, want_w, want_nch = e.mc_predict(st, seed=seed, first=first, count=count)              # the existing entry point
        wins = torch.full((n, 6), 0x5A5A5A5A, dtype=torch.int32, device=dev)                           # dirty: the call overwrites
        action = torch.full((n, 2), 99, dtype=torch.int8, device=dev)
        nch = torch.full((n,), -7, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        rc = lib.ewn_mc_predict_static(_ptr(st), None, n, N, 3, seed, first, count, _ptr(wins), _ptr(action), _ptr(nch), None, stream)
        assert rc == _lib.EWN_OK, (rc, lib.ewn_last_cuda_error())
        torch.cuda.synchronize()
        if call:                                                                                       # (the first call loads the kernels)
            assert torch.cuda.mem_get_info(dev)[0] == free0, "the call changed the device's free memory"
        assert torch.equal(wins, want_w.view(torch.int32)), call
        assert torch.equal(action, want_a) and torch.equal(nch, want_nch), call
        assert int(want_w.view(torch.int32).sum()) > 0
    # the Engine method goes through the same symbol
    a2, w2, _ = e.mc_predict_static(st, seed, first, count, wins=torch.empty_like(wins), action=torch.empty_like(action))
    assert torch.equal(w2, wins) and torch.equal(a2, action)


def test_static_workspace_rejects_more_roots_than_it_holds():
    import torch
    from ewn_gym_b200 import _lib
    from ewn_gym_b200.engine import Engine
    lib = _lib.lib()
    dev = torch.device("cuda", 0)
    e = Engine(5, 3, device=dev)
    n = int(lib.ewn_mc_static_max_roots()) + 1
    st = e.reset(n, seed=1)
    wins = torch.full((n, 6), 7, dtype=torch.int32, device=dev)
    action = torch.full((n, 2), 99, dtype=torch.int8, device=dev)
    rc = lib.ewn_mc_predict_static(_ptr(st), None, n, 5, 3, 1, 0, 8, _ptr(wins), _ptr(action), None, None,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == _lib.EWN_ESPACE
    msg = lib.ewn_last_cuda_error().decode()
    assert "EWN_MC_STATIC_MAX_ROOTS" in msg and str(n) in msg
    torch.cuda.synchronize()
    assert int((wins != 7).sum()) == 0 and int((action != 99).sum()) == 0                              # nothing was launched
    with pytest.raises(ValueError, match="exceed the static workspace"):
        e.mc_predict_static(st, 1, 0, 8, wins=wins, action=action)
