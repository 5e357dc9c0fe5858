"""GPU experiment: is the k_step loop host-launch-bound?  us per launch as a function of n (tiny n = host floor)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200.engine import Engine

dev = torch.device("cuda", 0)
e = Engine(5, 3)
SEED = 9487
g = torch.Generator(device=dev); g.manual_seed(1)


def run(n, want, iters=400, nsets=4, graph=False):
    shapes = {"reward_f32": ((n,), "f32"), "terminated": ((n,), "u8"), "truncated": ((n,), "u8"), "obs_board": ((n, 5, 5), "i16"),
              "obs_dice": ((n,), "i64")}
    sets = []
    for s in range(nsets):
        S = {"state": e.reset(n, seed=SEED + s),
             "act": torch.stack([torch.randint(0, 2, (n,), generator=g, device=dev), torch.randint(0, 3, (n,), generator=g, device=dev)], 1).contiguous()}
        S["out"] = {k: e.xp.empty(*shapes[k]) for k in want}
        S["plan"] = e.step_plan(S["state"], S["act"], seed=SEED + s, autoreset=True, want=tuple(want), out=S["out"])
        sets.append(S)
    for t in range(20):
        sets[t % nsets]["plan"].launch(t // nsets + 1)
    torch.cuda.synchronize()
    if graph:
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for S in sets:      # re-bind on the capture stream
                S["plan"] = e.step_plan(S["state"], S["act"], seed=SEED, autoreset=True, want=tuple(want), out=S["out"])
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                for t in range(iters):
                    sets[t % nsets]["plan"].launch(t // nsets + 1)
            gr.replay(); torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); gr.replay(); b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / iters * 1e3
        host = 0.0
    else:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        a.record()
        for t in range(20, 20 + iters):
            sets[t % nsets]["plan"].launch(t // nsets + 1)
        b.record()
        host = (time.perf_counter() - h0) / iters * 1e6
        torch.cuda.synchronize()
        us = a.elapsed_time(b) / iters * 1e3
    print(f"n={n:9d} outs={len(want)} graph={int(graph)} {us:8.2f} us/launch (host enqueue {host:6.2f} us)  {n / us * 1e6:.3e} steps/s  "
          f"{n * (16 + 16 + (4 + 2 + 50 + 8 if want else 0)) / us / 1e3:7.1f} GB/s", flush=True)


vec = ["reward_f32", "terminated", "truncated", "obs_board", "obs_dice"]
for n in (256, 1 << 16, 1 << 18, 1 << 20, 1 << 22, 1 << 23):
    run(n, vec)
    run(n, [])
for n in (1 << 18, 1 << 20, 1 << 22):
    run(n, vec, graph=True, iters=100)
    run(n, [], graph=True, iters=100)
