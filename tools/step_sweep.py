"""GPU experiment: where does k_step's time go?  Sweeps output sets / working-set rotation and prints us per launch."""
import sys, os, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200 import _lib
if os.environ.get("EWN_LIB_VARIANT"):          # experiment builds (build/variants/*.so); the product always loads csrc/libewn_b200.so
    _lib.LIB_PATH = os.path.abspath(os.environ["EWN_LIB_VARIANT"])
from ewn_gym_b200.engine import Engine

dev = torch.device("cuda", 0)
e = Engine(5, 3)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
SEED = 9487
g = torch.Generator(device=dev); g.manual_seed(1)
shapes = {"reward_f32": ((n,), "f32"), "terminated": ((n,), "u8"), "truncated": ((n,), "u8"), "obs_board": ((n, 5, 5), "i16"),
          "obs_dice": ((n,), "i64"), "term_board": ((n, 5, 5), "i16"), "term_dice": ((n,), "i64")}


full = list(shapes)
vec = ["reward_f32", "terminated", "truncated", "obs_board", "obs_dice"]


def run(label, nsets, want, iters=200, agent_random=False, stats=True, flush=False, compact=False, block_kernel=False, no_pdl=False):
    sets = []
    st = torch.zeros(4, dtype=torch.int64, device=dev) if stats else None
    for s in range(nsets):
        S = {"state": e.reset(n, seed=SEED + s),
             "act": torch.stack([torch.randint(0, 2, (n,), generator=g, device=dev), torch.randint(0, 3, (n,), generator=g, device=dev)], 1).contiguous()}
        S["out"] = {k: e.xp.empty(*shapes[k]) for k in want}
        S["cnt"] = torch.zeros(2, dtype=torch.int32, device=dev); S["tidx"] = torch.empty(n, dtype=torch.int32, device=dev)
        S["plans"] = [e.step_plan(S["state"], None if agent_random else S["act"], seed=SEED + s, autoreset=True, stats=st, want=tuple(want), out=S["out"],
                                  agent_random=agent_random, block_kernel=block_kernel, no_pdl=no_pdl,
                                  term_compact=(S["cnt"][b:b + 1], S["tidx"], S["cnt"][1 - b:2 - b]) if compact else None) for b in (0, 1)]
        sets.append(S)
    fl = torch.empty(256 << 20, dtype=torch.int8, device=dev) if flush else None
    for t in range(20):
        sets[t % nsets]["plans"][(t // nsets) & 1].launch(t // nsets + 1)
    torch.cuda.synchronize()
    if flush:
        tot = 0.0
        for t in range(20, 20 + 50):
            fl.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); sets[t % nsets]["plans"][(t // nsets) & 1].launch(t // nsets + 1); b.record(); torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        us = tot / 50 * 1e3
    else:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for t in range(20, 20 + iters):
            sets[t % nsets]["plans"][(t // nsets) & 1].launch(t // nsets + 1)
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / iters * 1e3
    print(f"{label:60s} sets={nsets} {us:8.2f} us/launch  {n / us * 1e6:.3e} steps/s", flush=True)


if os.environ.get("EWN_SWEEP_SHORT"):
    run("vec + compact term obs, PERSISTENT kernel", 4, full, compact=True)
    run("vec + compact term obs, PERSISTENT, no PDL", 4, full, compact=True, no_pdl=True)
    run("vec + compact term obs, PERSISTENT, 1 set (true dependency)", 1, full, compact=True)
    run("vec + compact term obs, block kernel", 4, full, compact=True, block_kernel=True)
    run("vec surface, no term obs, PERSISTENT", 4, vec)
    run("vec surface, no term obs, no stats, PERSISTENT", 4, vec, stats=False)
    sys.exit(0)
run("vec surface + DENSE term obs", 4, full)
run("vec + compact term obs, block kernel", 4, full, compact=True, block_kernel=True)
run("vec surface + COMPACT term obs (bench)", 4, full, compact=True)
run("vec surface + COMPACT term obs, 1 set (L2 resident)", 1, full, compact=True)
run("vec surface, no term obs", 4, vec)
run("vec surface, no term obs, 1 set", 1, vec)
run("no board obs (reward/done/dice only)", 4, ["reward_f32", "terminated", "truncated", "obs_dice"])
run("no outputs at all (state only)", 4, [])
run("no outputs at all, 1 set", 1, [])
run("no outputs, in-kernel random agent", 4, [], agent_random=True)
run("vec surface, no stats", 4, vec, stats=False)
run("vec surface + compact term obs, L2 flush by write, per-launch events", 1, full, flush=True, compact=True)
# copy roofline reference: torch copy of the same byte volume (96 MB algorithmic: 48 MB read + 48 MB write)
x = torch.empty(48 << 20, dtype=torch.int8, device=dev); y = torch.empty_like(x)
for _ in range(5): y.copy_(x)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(100): y.copy_(x)
b.record(); torch.cuda.synchronize()
print(f"torch copy 48 MiB -> 48 MiB: {a.elapsed_time(b) / 100 * 1e3:.2f} us  ({96 * 1.048576 / (a.elapsed_time(b) / 100):.0f} GB/s)")
