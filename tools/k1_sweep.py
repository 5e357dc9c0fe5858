"""GPU experiment driver for K1 (k_vstep): times the bench.py workload (1 Mi 5x5 games, 4 env sets round-robin, compacted terminal
observations + stats) for one or more builds of the library, launched from a stream (as bench.py does) and replayed from a CUDA graph.
  python tools/k1_sweep.py [lib.so ...]        (default: the product library)
Each variant runs in its own process (a library is loaded once per process)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if os.environ.get("EWN_K1_CHILD"):
    import torch
    from ewn_gym_b200 import _lib
    if os.environ.get("EWN_LIB_VARIANT"):
        _lib.LIB_PATH = os.path.abspath(os.environ["EWN_LIB_VARIANT"])
    from ewn_gym_b200.engine import Engine
    dev = torch.device("cuda", 0)
    e = Engine(5, 3)
    n, SEED, NS = int(os.environ.get("EWN_K1_N", 1 << 20)), 9487, 4
    ar = os.environ.get("EWN_K1_AR") == "1"
    g = torch.Generator(device=dev); g.manual_seed(1234)
    shapes = {"reward_f32": ((n,), "f32"), "terminated": ((n,), "u8"), "truncated": ((n,), "u8"), "obs_board": ((n, 5, 5), "i16"),
              "obs_dice": ((n,), "i64"), "term_board": ((n, 5, 5), "i16"), "term_dice": ((n,), "i64")}
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    ctr = torch.zeros(1, dtype=torch.int64, device=dev)
    sets = []
    for s_i in range(NS):
        S = {"state": e.reset(n, seed=SEED + s_i), "out": {k: e.xp.empty(sh, d) for k, (sh, d) in shapes.items()},
             "act": torch.stack([torch.randint(0, 2, (n,), generator=g, device=dev), torch.randint(0, 3, (n,), generator=g, device=dev)], 1).contiguous(),
             "cnt": torch.zeros(2, dtype=torch.int32, device=dev), "tidx": torch.empty(n, dtype=torch.int32, device=dev)}
        sets.append(S)

    def plans(step_dev=None):
        return [[e.step_plan(S["state"], None if ar else S["act"], seed=SEED + s_i, step=0, autoreset=True, stats=stats, want=tuple(shapes), out=S["out"],
                             agent_random=ar, step_dev=step_dev, term_compact=(S["cnt"][b:b + 1], S["tidx"], S["cnt"][1 - b:2 - b]),
                             block_kernel=os.environ.get("EWN_K1_BLOCK") == "1") for b in (0, 1)]
                for s_i, S in enumerate(sets)]

    P = plans()
    def one(t):
        k = t // NS
        P[t % NS][k & 1].launch(k + 1)
    t = 0
    for _ in range(40): one(t); t += 1
    torch.cuda.synchronize()
    res = []
    for rep in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200): one(t); t += 1
        b.record(); torch.cuda.synchronize()
        res.append(a.elapsed_time(b) / 200 * 1e3)
    line = f"stream {min(res):7.2f} us (runs {' '.join(f'{x:.2f}' for x in res)})"
    if os.environ.get("EWN_K1_TIMELINE"):                  # -DEWN_EXP_TIMELINE build: per-warp clock stamps of the last launch
        import ctypes, numpy as np
        tl = np.zeros(148 * 64 * 16, dtype=np.uint64)
        rc = ctypes.CDLL(_lib.LIB_PATH).ewn_debug_timeline(tl.ctypes.data_as(ctypes.c_void_p))
        np.save(os.environ["EWN_K1_TIMELINE"], tl.reshape(-1, 16))
        line += f"   timeline rc={rc}"
    if os.environ.get("EWN_K1_GRAPH", "1") == "1":
        side = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(side):
            PG = plans(step_dev=ctr)
            def cycle():                                   # 2 parities x NS sets = 8 steps, then bump the device step counter
                for k in (0, 1):
                    for s_i in range(NS):
                        PG[s_i][k].launch(k + 1)
                e.counter_add(ctr, 2)
            cycle(); torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                cycle()
            for _ in range(5): graph.replay()
            torch.cuda.synchronize()
            gres = []
            for rep in range(3):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(25): graph.replay()
                b.record(); torch.cuda.synchronize()
                gres.append(a.elapsed_time(b) / 200 * 1e3)
        line += f"   graph {min(gres):7.2f} us (runs {' '.join(f'{x:.2f}' for x in gres)})"
    print(f"{os.environ.get('EWN_K1_LABEL', 'product'):44s} {line}   finished/step {int(stats[3].item()) / max(1, t + 0):.0f}", flush=True)
    sys.exit(0)

libs = sys.argv[1:] or [""]
for lib in libs:
    env = dict(os.environ, EWN_K1_CHILD="1", EWN_K1_LABEL=os.path.basename(lib) or "product")
    if lib:
        env["EWN_LIB_VARIANT"] = lib
    r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=300)
    print(r.stdout.strip() or ("FAILED " + lib + ": " + r.stderr.strip()[-400:]), flush=True)
