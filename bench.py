#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: env steps/s (and random playouts/s) for 5x5 EinStein wuerfelt nicht.

  python bench.py --gpus N --steps K --warmup W            # our arm (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  # the reference algorithm on the host cores (CPU port)

A "step" is one VecEnv.step over `--envs` concurrent games per GPU (BASELINE.json configs[1]: 1,048,576 games, random
actions): one launch of the fused kernel k_step (agent ply, Philox die, random opponent ply, die, win tests, auto-reset,
reward / done, int16 board + int64 dice observation, terminal observations).  `value` is timed with CUDA events on the
launching stream with inputs resident in HBM and L2 flushed between steps; `e2e` goes through EWNVecEnv.step with host
(page-locked) action and observation buffers.  DESIGN.md "Measurement" has the byte accounting.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_BOARD, N_LAYER = 5, 3
N_SETS = 4
SEED = 9487
# algorithmic bytes per env step, 5x5 gym surface: state 8 R + 8 W, action int64x2 16 R, reward f32 4 W,
# terminated + truncated 2 W, board int16[25] 50 W, dice int64 8 W.
BYTES_PER_STEP = 8 + 8 + 16 + 4 + 2 + 50 + 8
# ... plus, per FINISHED game, its compacted terminal observation (info["terminal_observation"] of the VecEnv contract):
# board int16[25] 50 W, dice int64 8 W, env index int32 4 W.  Counted from the episodes that really finished in the timed steps.
BYTES_PER_TERMINAL_ROW = 50 + 8 + 4
BYTES_PER_STEP_AGENT_RANDOM = BYTES_PER_STEP - 16            # the in-kernel agent reads no action array


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index, self.t0 = [], None, index, 0.0

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def mark(self):
        self.t0 = time.perf_counter()

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        for ts, r in self.rows:
            if ts < self.t0 + 0.15:                 # a sample reports the period before it
                continue
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_steps(n_envs, min_seconds, threads):
    """The reference algorithm (envs/ewn.py:436-486 with a random opponent) as restated in oracle/ewn_oracle.c, all host
    threads, uniform random actions, auto-reset: env steps per second."""
    from oracle import oracle as O
    board, dice = O.reset_philox(n_envs, SEED, 0)
    rng = np.random.default_rng(1)
    acts = [np.stack([rng.integers(0, 2, n_envs), rng.integers(0, 3, n_envs)], 1).astype(np.int64) for _ in range(4)]
    O.step_philox(board[:1024], dice[:1024], acts[0][:1024], SEED, 0)          # build + warm
    steps, t0, t = 0, time.perf_counter(), 0
    while True:
        out = O.step_philox(board, dice, acts[t % 4], SEED, t + 1, threads=threads)
        board, dice = out["board"], out["dice"]
        t += 1; steps += n_envs
        el = time.perf_counter() - t0
        if el >= min_seconds:
            return steps / el, t, el


def reference_python(seconds=10.0):
    """The reference's OWN Python (eval_random.py:17-51 loop, MctsAgent.simulate, ExpectiMinimaxAgent(4).predict), one process per
    core.  Timed live when a reference tree travels with the repo (baseline/_ref, or EWN_REFERENCE_ROOT); otherwise the committed
    measurement of the same script in the build container is quoted (profiles/r2_reference_python.json) and labelled as such --
    /root/reference does not exist on the GPU box and bench.py never reads it."""
    root = os.environ.get("EWN_REFERENCE_ROOT") or os.path.join(ROOT, "baseline", "_ref")
    if os.path.isfile(os.path.join(root, "envs", "ewn.py")):
        os.environ["EWN_REFERENCE_ROOT"] = root
        try:
            from oracle import time_reference
            out = time_reference.measure(seconds)
            out["where"] = "this machine, live"
            return out
        except Exception as err:                                   # noqa: BLE001
            return {"unavailable": f"reference at {root} failed to run: {err}"[:300]}
    try:
        out = json.load(open(os.path.join(ROOT, "profiles", "r2_reference_python.json")))
        out["where"] = "build container, committed measurement (profiles/r2_reference_python.json): the pure-Python reference does not travel to the GPU box"
        return out
    except Exception:
        return {"unavailable": "no reference tree on this machine and no committed measurement"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU algorithm for the same workload (same number of concurrent games) on the box's host
    cores: the C restatement oracle/ewn_oracle.c ("port", ours) timed on all cores, with the reference's own Python quoted beside it
    (`cpu_baseline.reference_python`).  Rank 0 only."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_s = args.envs
    from oracle import oracle as O
    board, dice = O.reset_philox(n_s, SEED, 0)
    rng = np.random.default_rng(1)
    acts = [np.stack([rng.integers(0, 2, n_s), rng.integers(0, 3, n_s)], 1).astype(np.int64) for _ in range(4)]
    t_all = []
    for k in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        out = O.step_philox(board, dice, acts[k % 4], SEED, k + 1, threads=cores)
        board, dice = out["board"], out["dice"]
        if k >= args.warmup:
            t_all.append(time.perf_counter() - t0)
    ms = 1e3 * float(np.mean(t_all))
    v = n_s / (ms / 1e3)
    line = {"impl": "reference", "metric": "env_steps_per_sec", "value": v, "unit": "env steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic", "config": workload_config(args, n_s),
            "cpu_baseline": {"value": v, "unit": "env steps/s", "cores": cores, "kind": "port", "label": "ours: oracle/ewn_oracle.c on all host threads",
                             "sample": f"{n_s} concurrent 5x5 games per step, uniform random actions, auto-reset, {args.steps} steps",
                             "reference_python": None if args.skip_reference_python else reference_python()},
            "e2e": {"value": v, "unit": "env steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, n_envs):
    return {"workload": "5x5 batched env step (BASELINE.json configs[1]): concurrent games per GPU, uniform random actions over "
                        "MultiDiscrete([2,3]), random opponent in-kernel, auto-reset, VecEnv observation surface",
            "board_size": N_BOARD, "cube_layer": N_LAYER, "envs_per_gpu": n_envs, "l2": f"no flush: {N_SETS} independent env sets of envs_per_gpu games stepped round-robin, working set {N_SETS} x ~154 MB > 126 MB L2",
            "bytes_per_env_step": BYTES_PER_STEP}


def run_ours(args, rank, world, local_rank):
    import torch
    from ewn_gym_b200 import _lib
    from ewn_gym_b200.engine import Engine

    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()
    clk = ClockSampler(local_rank).__enter__()      # nvidia-smi takes ~1 s to start: launch it before the set-up work
    e = Engine(N_BOARD, N_LAYER, device=dev)
    n = args.envs
    env_id0 = rank * n
    g = torch.Generator(device=dev); g.manual_seed(1234 + rank)
    # N_SETS independent env sets stepped round-robin: a step never finds its inputs in L2 (N_SETS x ~154 MB >> 126 MB)
    shapes = {"reward_f32": ((n,), "f32"), "terminated": ((n,), "u8"), "truncated": ((n,), "u8"), "obs_board": ((n, N_BOARD, N_BOARD), "i16"),
              "obs_dice": ((n,), "i64"), "term_board": ((n, N_BOARD, N_BOARD), "i16"), "term_dice": ((n,), "i64")}
    sets = []
    for s_i in range(N_SETS):
        sets.append({
            "state": e.reset(n, seed=SEED + s_i, step=0, env_id0=env_id0),
            "act": torch.stack([torch.randint(0, 2, (n,), generator=g, device=dev), torch.randint(0, 3, (n,), generator=g, device=dev)], 1).contiguous(),
            "out": {k: e.xp.empty(sh, d) for k, (sh, d) in shapes.items()}})
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    want = tuple(shapes)

    # pre-bound launches (one ctypes call per step in the timed loop).  Terminal observations are appended compacted;
    # each set alternates two append counters, the launch that uses one clears the other (no memset launches).
    for s_i, S in enumerate(sets):
        S["cnt"] = torch.zeros(2, dtype=torch.int32, device=dev)
        S["tidx"] = torch.empty(n, dtype=torch.int32, device=dev)
        S["plans"] = [e.step_plan(S["state"], S["act"], seed=SEED + s_i, step=0, env_id0=env_id0, autoreset=True, stats=stats, want=want,
                                  out=S["out"], term_compact=(S["cnt"][b:b + 1], S["tidx"], S["cnt"][1 - b:2 - b])) for b in (0, 1)]

    def one_step(t):
        k = t // N_SETS
        sets[t % N_SETS]["plans"][k & 1].launch(k + 1)

    # the same surface with the agent = uniform over its LEGAL list inside the kernel (first variant): games run
    # to real wins and losses (captures, nearest-cube fallback, corner / elimination endings) instead of dying on illegal openings
    stats_ar = torch.zeros(4, dtype=torch.int64, device=dev)
    for s_i, S in enumerate(sets):
        S["state_ar"] = e.reset(n, seed=SEED + 16 + s_i, step=0, env_id0=env_id0)
        S["plans_ar"] = [e.step_plan(S["state_ar"], None, seed=SEED + 16 + s_i, step=0, env_id0=env_id0, autoreset=True, stats=stats_ar, want=want,
                                     out=S["out"], agent_random=True, term_compact=(S["cnt"][b:b + 1], S["tidx"], S["cnt"][1 - b:2 - b])) for b in (0, 1)]

    def one_step_ar(t):
        k = t // N_SETS
        sets[t % N_SETS]["plans_ar"][k & 1].launch(k + 1)

    t = 0
    for _ in range(max(args.warmup, 3)):
        one_step(t); t += 1
    torch.cuda.synchronize()
    stats_before = stats.clone()
    if dist: dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = lib.ewn_launch_count()
    wall0 = time.perf_counter()
    clk.mark()
    ev0.record()
    for k in range(args.steps):
        one_step(t); t += 1
    ev1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - wall0
    launches = lib.ewn_launch_count() - launches0
    if dist: dist.barrier()
    ms_step = ev0.elapsed_time(ev1) / args.steps
    kernel_ms = ms_step
    finished_timed = int((stats - stats_before)[3].item())      # episodes that ended inside the timed steps of this rank
    # ---- agent = uniform over the legal list: mid-game mix first (40 untimed steps: games are at every stage), then the same protocol
    ar_run = None
    if not args.skip_agent_random:
        for S in sets:
            S["cnt"].zero_()                           # the two phases share the append counters
        ta = 0
        for _ in range(40 * N_SETS + max(args.warmup, 3)):
            one_step_ar(ta); ta += 1
        torch.cuda.synchronize()
        st0 = stats_ar.clone()
        if dist: dist.barrier()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        for k in range(args.steps):
            one_step_ar(ta); ta += 1
        b1.record(); torch.cuda.synchronize()
        ar_ms = b0.elapsed_time(b1) / args.steps
        st_ar = stats_ar - st0
        if dist:
            tt = torch.tensor([ar_ms], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); ar_ms_max = float(tt.item())
            dist.all_reduce(st_ar)
        else:
            ar_ms_max = ar_ms
        sa = st_ar.cpu().tolist()
        rows = sa[3] / (args.steps * world)
        ar_bytes = BYTES_PER_STEP_AGENT_RANDOM * n + BYTES_PER_TERMINAL_ROW * rows
        hbm_ar, _ = peaks()
        ar_run = {"value": world * n / (ar_ms_max / 1e3), "unit": "env steps/s", "ms_per_step": ar_ms_max,
              "config": {"agent": "uniform over legal list (in-kernel, random_policy.py:11-15 on Philox word 3)", "envs_per_gpu": n,
                         "bytes_per_env_step": BYTES_PER_STEP_AGENT_RANDOM, "kernel": "k_vstep<P5,5,AR>"},
              "episodes": {"agent_wins": sa[0], "agent_losses": sa[1], "illegal_endings": sa[2], "finished": sa[3], "steps_timed": args.steps},
              "roofline": {"bound": "hbm", "achieved": ar_bytes / (ar_ms / 1e3) / 1e9, "peak": hbm_ar, "unit": "GB/s",
                           "frac": ar_bytes / (ar_ms / 1e3) / 1e9 / hbm_ar, "algorithmic_bytes_per_launch": ar_bytes,
                           "terminal_rows_per_launch": rows}}
    # the timed region is milliseconds long; keep the identical load running so that nvidia-smi (100 ms period) sees it
    t_end = time.perf_counter() + (0.0 if args.skip_clocks else max(0.0, 1.5 - wall))
    while time.perf_counter() < t_end:
        for _ in range(50):
            one_step(t); t += 1
        torch.cuda.synchronize()
    clk.__exit__()
    if dist:
        tt = torch.tensor([ms_step], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_step = float(tt.item())
        dist.all_reduce(stats)                    # the only data-path collective: win / loss / episode counters
    value = world * n / (ms_step / 1e3)

    # ---- secondary metric: random playouts/s (BASELINE.json configs[2], flat Monte-Carlo, 1024 playouts per root child)
    playout = None
    if not args.skip_playouts:
        roots_n = 4096
        roots = e.reset(roots_n, seed=SEED + 1)
        for s in range(1, 5):                     # a few random plies so that roots are mid-game positions
            e.step(roots, None, seed=SEED + 1, step=s, agent_random=True, autoreset=True, want=())
        per = 1024 // world
        iters = 8                                 # Monte-Carlo decisions back to back (fresh playout streams each), as an arena makes them
        wins = torch.zeros(roots_n, 6, dtype=torch.int32, device=dev)
        plies = torch.zeros(1, dtype=torch.int64, device=dev)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def decide_nccl(it, count, pl=None):
            # one rank: ewn_mc_predict = expand (clears wins) -> playouts -> arg-max, three PDL-chained launches, no memsets
            a_, _, k = e.mc_predict(roots, seed=SEED + it, first=rank * per, count=count, wins=wins, plies=pl, want_action=not dist)
            if dist:
                dist.all_reduce(wins)             # root statistics merge
                a_ = e.mc_best_action(roots, wins)
            return a_, k

        def timed(fn):
            fn(0, 8)                              # warm-up
            torch.cuda.synchronize()
            if dist: dist.barrier()
            a0.record()
            for it in range(iters):
                out = fn(it, per)
            a1.record(); torch.cuda.synchronize()
            return a0.elapsed_time(a1) / iters, out

        pms, (best, nch) = timed(decide_nccl)
        decide_nccl(iters - 1, per, plies)        # plies of one decision (untimed)
        tot_ref = wins.clone()
        merge = {"kind": "none (one rank)"}
        if dist:
            # the same decisions with the merge done by ONE kernel over NVLink peer memory instead of NCCL all-reduce + arg-max
            # (ewn_gym_b200/peer.py); both are timed, the product path is the peer kernel, and the moves must be identical
            from ewn_gym_b200.peer import PeerExchange
            nccl_ms = pms
            try:
                ex = PeerExchange(e, roots_n)     # raises on every rank or on none
            except Exception as err:              # noqa: BLE001 - e.g. CUDA IPC not permitted on this box: the NCCL numbers stand
                ex, merge = None, {"kind": "NCCL all-reduce + k_mc_best", "ms": nccl_ms, "peer_memory": f"unavailable: {err}"[:300]}
            if ex is not None:
                def decide_peers(it, count):
                    e.mc_predict(roots, seed=SEED + it, first=rank * per, count=count, wins=ex.wins_buffer(zero=False), want_action=False)
                    return ex.best_action(roots)

                decide_peers(0, 8); decide_peers(0, 8)    # maps the peers' memory on first touch
                peer_ms, (best_p, tot_p) = timed(decide_peers)
                try:
                    ex.check(); waited = 0.0
                except RuntimeError:
                    waited = 1.0                   # a wait for the peers hit its time limit: results of that decision are void
                same = bool(torch.equal(best_p, best)) and bool(torch.equal(tot_p, tot_ref))
                tt = torch.tensor([nccl_ms, peer_ms, 0.0 if same else 1.0, waited], device=dev, dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                nccl_ms, peer_ms, bad, waited = (float(x) for x in tt.tolist())
                if waited:
                    merge = {"kind": "NCCL all-reduce + k_mc_best", "ms": nccl_ms, "peer_memory": "a wait for the peers timed out"}
                elif bad:
                    raise SystemExit("peer-memory merge differs from all-reduce + arg-max")
                else:
                    pms = peer_ms
                    merge = {"kind": "one kernel over NVLink peer memory (ewn_mc_best_action_peers)", "ms": peer_ms,
                             "nccl_allreduce_argmax_ms": nccl_ms, "identical_to_nccl": True}
                dist.barrier(); ex.close()
            dist.all_reduce(plies)
            # SURVEY 8d config 3 "1-vs-8-GPU counts identical": rank 0 replays a fixed sub-batch of the LAST timed decision alone -- all
            # 1024 playouts of every root action, same Philox keys (root id, action, playout index) -- and the merged N-rank counters
            # (tot_ref: the sum over ranks of 1024 / N playouts each) must equal them, or no line is printed
            sub = 512
            if rank == 0:
                w1, _ = e.mc_root_wins(roots[:sub], seed=SEED + iters - 1, first=0, count=per * world)
                bad = 0.0 if bool(torch.equal(w1, tot_ref[:sub])) else 1.0
            else:
                bad = 0.0
            tt = torch.tensor([bad], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            if float(tt.item()):
                raise SystemExit("Monte-Carlo win counts of the N-rank run differ from the single-rank run with the same Philox keys")
            merge["identical_to_single_rank"] = True
            merge["identity_check"] = f"rank 0 alone: {sub} roots x all actions x {per * world} playouts == merged counters of {world} ranks"
        n_pairs = int(nch.sum().item())
        playout = {"value": n_pairs * 1024 / (pms / 1e3), "unit": "playouts/s", "ms": pms, "roots": roots_n, "root_children": n_pairs,
                   "playouts_per_child": 1024, "decisions_timed": iters, "launches_per_decision": 3, "plies_per_playout": float(plies.item()) / max(1, n_pairs * 1024), "merge": merge}
        try:    # on-chip kernel: the roofline is the SM issue rate; instructions per ply come from the committed ncu capture
            prof = json.load(open(os.path.join(ROOT, "profiles", "k_playouts_issue.json")))
            wi = prof["warp_instructions_per_ply"] * float(plies.item()) / (pms / 1e3)
            playout["issue_roofline"] = {"bound": "issue", "achieved": wi, "peak": prof["issue_peak_warp_instructions_per_s"] * world,
                                         "unit": "warp-instructions/s", "frac": wi / (prof["issue_peak_warp_instructions_per_s"] * world),
                                         "warp_instructions_per_ply": prof["warp_instructions_per_ply"], "alu_pipe_pct_ncu": prof["alu_pipe_pct"],
                                         "note": "includes launch + root-statistics merge + arg-max inside the timed region"}
        except Exception:
            pass

    # ---- secondary: expectiminimax depth 4 (configs[3])
    emm = None
    if not args.skip_search:
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sizes = {}
        for roots_n in (65536, 1 << 20):            # the per-root replay is a dependent chain: it needs roots in flight to hide latency
            roots = e.reset(roots_n, seed=SEED + 2)
            for s in range(1, 5):
                e.step(roots, None, seed=SEED + 2, step=s, agent_random=True, autoreset=True, want=())
            e.expectiminimax(roots, 4)              # warm-up at full size: the library's scratch allocation happens here
            torch.cuda.synchronize(); a0.record()
            v, a = e.expectiminimax(roots, 4)
            a1.record(); torch.cuda.synchronize()
            ems = a0.elapsed_time(a1)
            sizes[roots_n] = {"roots_per_s": roots_n / (ems / 1e3), "ms": ems}
            del roots, v, a
        emm = {"value": sizes[1 << 20]["roots_per_s"], "unit": "roots/s (depth 4, hybrid, pruned semantics)", "ms": sizes[1 << 20]["ms"],
               "roots": 1 << 20, "per_gpu": True, "at_65536_roots": sizes[65536], "kernel": "k_emm_fused<P5,hybrid> (one launch, tree on chip)"}
        try:    # on-chip kernel: bound by the SM issue rate; instructions per root from the committed ncu capture of this kernel
            prof = json.load(open(os.path.join(ROOT, "profiles", "k_emm_issue.json")))
            wi = prof["warp_instructions_per_root"] * emm["value"]
            emm["roofline"] = {"bound": "issue", "achieved": wi, "peak": prof["issue_peak_warp_instructions_per_s"], "unit": "warp-instructions/s",
                               "frac": wi / prof["issue_peak_warp_instructions_per_s"], "warp_instructions_per_root": prof["warp_instructions_per_root"],
                               "alu_pipe_pct_ncu": prof["alu_pipe_pct"],
                               "hbm": {"achieved_gbs": 18 * emm["value"] / 1e9, "algorithmic_bytes_per_root": 18,
                                       "note": "8 B packed state in, 8 B value + 2 B action out; no node records leave the chip at depth 3 / 4"}}
        except Exception:
            pass

    # ---- secondary: the headline batch with T = 32 env steps per LAUNCH (ewn_rollout without a network: pooled uniform actions read from a
    # [T][n][2] tensor, the games stay in registers, every step's observation / reward / done written as the step launches write them)
    rollout_5x5 = None
    if not args.skip_rollout:
        try:
            T5 = 32
            g5 = torch.Generator(device=dev); g5.manual_seed(77 + rank)
            acts5 = torch.stack([torch.randint(0, 2, (T5, n), generator=g5, device=dev), torch.randint(0, 3, (T5, n), generator=g5, device=dev)], 2).contiguous()
            st5 = e.reset(n, seed=SEED + 9, env_id0=rank * n)
            buf5 = {"board": torch.empty((T5 + 1, n, 5, 5), dtype=torch.int16, device=dev), "dice": torch.empty((T5 + 1, n), dtype=torch.int64, device=dev),
                    "actions": acts5, "rewards": torch.empty((T5, n), device=dev), "dones": torch.empty((T5, n), dtype=torch.uint8, device=dev)}
            ctr5, stats5 = torch.zeros(1, dtype=torch.int64, device=dev), torch.zeros(4, dtype=torch.int64, device=dev)
            roll5 = e.rollout_plan(None, st5, T5, buf5, seed=SEED + 9, env_id0=rank * n, step_dev=ctr5, stats=stats5)

            def one5():
                roll5(1); e.counter_add(ctr5, T5)
            for _ in range(2): one5()
            torch.cuda.synchronize()
            if dist: dist.barrier()
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record()
            for _ in range(4): one5()
            r1.record(); torch.cuda.synchronize()
            ms5 = r0.elapsed_time(r1) / 4
            if dist:
                tt = torch.tensor([ms5], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); ms5 = float(tt[0])
            byt5 = 79 + 58 / T5
            pk5, _src5 = peaks()
            rollout_5x5 = {"workload": f"5x5, {n} games per GPU x {T5} env steps per launch, pooled uniform actions (ewn_rollout, net = NULL: k_rollout_env)",
                           "ms_per_launch": ms5, "us_per_step_of_the_batch": 1e3 * ms5 / T5, "value": n * world * T5 / (ms5 / 1e3), "unit": "env steps/s",
                           "roofline": {"bound": "hbm", "bytes_per_env_step": byt5, "achieved": n * T5 * byt5 / (ms5 / 1e3) / 1e9, "peak": pk5, "unit": "GB/s",
                                        "frac": n * T5 * byt5 / (ms5 / 1e3) / 1e9 / pk5,
                                        "note": "action 16 R + reward 4 + done 1 + board 50 + dice 8 W per env step, one more observation per launch; the "
                                                "packed state is read and written once per launch, not per step"}}
            del buf5, acts5, st5
        except Exception as err:                     # noqa: BLE001
            rollout_5x5 = {"unavailable": str(err)[:200]}

    # ---- secondary: configs[4], 7x7 MiniMaxHeuristicEnv rollout collection, 4096 envs x 128 steps over all ranks, counters all-reduced
    rollout = None
    if not args.skip_rollout:
        e7 = Engine(7, 3, device=dev)
        n7, T = 4096 // world, 128
        st7 = e7.reset(n7, seed=SEED + 3, env_id0=rank * n7)
        tol = torch.full((n7,), 10, dtype=torch.int32, device=dev)
        prev = e7.evaluate(st7, "hybrid")
        # policy stub (SB3 is not installed): uniform MultiDiscrete([2,3]) samples.  The pool is a function of the GLOBAL env id (same
        # generator on every rank, each rank takes its slice), so that the N-rank job plays the same games as one rank would
        g7 = torch.Generator(device=dev); g7.manual_seed(4242)
        pool_global = [torch.stack([torch.randint(0, 2, (n7 * world,), generator=g7, device=dev),
                                    torch.randint(0, 3, (n7 * world,), generator=g7, device=dev)], 1).contiguous() for _ in range(8)]
        pool = [p_[rank * n7:(rank + 1) * n7].contiguous() for p_ in pool_global]
        want7 = ("reward_f32", "terminated", "truncated", "obs_board", "obs_dice")
        out7 = {k: None for k in want7}
        st_roll = torch.zeros(4, dtype=torch.int64, device=dev)
        ctr = torch.zeros(1, dtype=torch.int64, device=dev)
        first = e7.step(st7, pool[0], seed=SEED + 3, step=1, env_id0=rank * n7, shaped=True, illegal_move_reward=-1.0, tolerance=tol,
                        prev_score=prev, stats=st_roll, want=want7, step_dev=ctr)
        side = torch.cuda.Stream(device=dev)
        torch.cuda.synchronize()
        with torch.cuda.stream(side):              # plans bind the stream that is current when they are built
            plans = [e7.step_plan(st7, pool[t % 8], seed=SEED + 3, step=t + 1, env_id0=rank * n7, shaped=True, illegal_move_reward=-1.0,
                                  tolerance=tol, prev_score=prev, stats=st_roll, want=want7, out=first, step_dev=ctr) for t in range(T)]

            def rollout_stream():
                for t in range(T):
                    plans[t].launch(t + 1)
                e7.counter_add(ctr, T)

            rollout_stream(); torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                rollout_stream()                    # 128 step kernels + the counter bump: the next replay draws fresh Philox counters
            graph.replay(); torch.cuda.synchronize()
        if dist:                                    # first NCCL call on this stream sets up the communicator: not part of a rollout
            with torch.cuda.stream(side):
                dist.all_reduce(st_roll.clone())
            torch.cuda.synchronize()
        res = {}
        for name, fn in (("stream", rollout_stream), ("graph", graph.replay)):
            reps = 10
            st_roll.zero_()
            torch.cuda.synchronize()
            if dist: dist.barrier()
            a0, a1, a2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            with torch.cuda.stream(side):
                a0.record()
                for _ in range(reps):
                    fn()
                a1.record()
                tot = st_roll.clone()
                if dist: dist.all_reduce(tot)       # wins / losses / illegal endings / episodes of the whole job
                a2.record()                         # a1..a2 = the collective PLUS the wait for the slowest rank's rollouts
                if dist:
                    a3, a4 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a3.record(); dist.all_reduce(tot.clone()); a4.record()     # ranks are aligned now: the collective alone
            torch.cuda.synchronize()
            ms, ar = a0.elapsed_time(a1) / reps, a0.elapsed_time(a2) - a0.elapsed_time(a1)
            if dist:
                tt = torch.tensor([ms, ar], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); ms, ar = tt.tolist()
            res[name] = {"ms_per_rollout": ms, "env_steps_per_s": n7 * world * T / (ms / 1e3), "allreduce_us": 1e3 * ar,
                         "allreduce_alone_us": 1e3 * a3.elapsed_time(a4) if dist else None,
                         "episodes": int(tot[3].item()) // reps}
        # the same collection as ONE launch (ewn_rollout without a network: the pooled actions are read from a [T][n][2] tensor, the games stay
        # in registers for all 128 steps; identical to the step launches: tests/test_policy_kernel.py::test_rollout_without_a_policy_equals_step_calls)
        try:
            st7p = e7.reset(n7, seed=SEED + 3, env_id0=rank * n7)
            tolp, prevp = torch.full((n7,), 10, dtype=torch.int32, device=dev), e7.evaluate(st7p, "hybrid")
            stp, ctrp = torch.zeros(4, dtype=torch.int64, device=dev), torch.zeros(1, dtype=torch.int64, device=dev)
            bufp = {"board": torch.zeros((T + 1, n7, 7, 7), dtype=torch.int16, device=dev), "dice": torch.zeros((T + 1, n7), dtype=torch.int64, device=dev),
                    "actions": torch.stack([pool[t % 8] for t in range(T)]).contiguous(), "rewards": torch.zeros((T, n7), device=dev),
                    "dones": torch.zeros((T, n7), dtype=torch.uint8, device=dev)}
            with torch.cuda.stream(side):
                roll = e7.rollout_plan(None, st7p, T, bufp, seed=SEED + 3, env_id0=rank * n7, step_dev=ctrp, shaped=True, illegal_move_reward=-1.0,
                                       tolerance=tolp, prev_score=prevp, stats=stp)

                def rollout_one_launch():
                    roll(1); e7.counter_add(ctrp, T)
                rollout_one_launch(); torch.cuda.synchronize()
                stp.zero_(); torch.cuda.synchronize()
                if dist: dist.barrier()
                q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                q0.record()
                for _ in range(20):
                    rollout_one_launch()
                q1.record()
            torch.cuda.synchronize()
            qms = q0.elapsed_time(q1) / 20
            if dist:
                tt = torch.tensor([qms], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); qms = float(tt[0])
                tot = stp.clone(); dist.all_reduce(tot)
            else:
                tot = stp
            res["one_launch"] = {"ms_per_rollout": qms, "env_steps_per_s": n7 * world * T / (qms / 1e3), "episodes": int(tot[3].item()) // 20,
                                 "kernel": "k_rollout without a network (ewn_rollout, net = NULL): 128 steps of every game in one launch"}
        except Exception as err:                     # noqa: BLE001
            res["one_launch"] = {"unavailable": str(err)[:200]}
        rollout = {"workload": "7x7 MiniMaxHeuristicEnv (tolerance 10, illegal_move_reward -1) rollout collection, 4096 envs x 128 steps, "
                               "policy stub = pooled uniform actions, stats all-reduce after the rollouts", "envs_per_gpu": n7, **res}

        # the same collection with a POLICY between the steps (train.py:144-155 -> A2C.collect_rollouts: MultiInputPolicy forward, sample,
        # env.step, buffer add), everything on the device: what an SB3-style trainer pays per rollout on this env
        try:
            from ewn_gym_b200.device_env import EWNDeviceVecEnv
            from ewn_gym_b200.train import A2C
            for key, fused, persistent in (("with_policy", True, True), ("with_policy_graph", True, False), ("with_policy_eager", False, False)):
                penv = EWNDeviceVecEnv(n7, 7, 3, seed=SEED + 7, shaped=True, env_id0=rank * n7, engine=e7)
                a2c = A2C("MultiInputPolicy", penv, n_steps=T, seed=rank, fused=fused, persistent=persistent)
                a2c.collect_rollouts(); a2c.train(); torch.cuda.synchronize()
                if dist: dist.barrier()
                reps = 20 if fused else 3
                p0, p1, p2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                p0.record()
                for _ in range(reps):
                    a2c.collect_rollouts()
                p1.record()
                a2c.train()
                p2.record(); torch.cuda.synchronize()
                pms, tms = p0.elapsed_time(p1) / reps, p1.elapsed_time(p2)
                if dist:
                    tt = torch.tensor([pms, tms], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); pms, tms = tt.tolist()
                rollout[key] = {"ms_per_rollout": pms, "env_steps_per_s": n7 * world * T / (pms / 1e3), "a2c_update_ms": tms,
                                "policy": "MultiInputPolicy (board + one-hot dice -> 64-64 Tanh towers)",
                                "collector": ("ONE launch per rollout (ewn_rollout): policy forward + categorical draw and env step alternate on chip for all "
                                              "128 steps, games in registers, parameters staged once, then V(last) and GAE; buffers bit-identical to the "
                                              "two-launches-per-step graph" if persistent else
                                              "one CUDA graph per rollout: per step ewn_policy_forward (forward + categorical draw, one launch) and the env "
                                              "step kernel writing straight into the rollout buffer, then V(last), ewn_gae" if fused else
                                              "eager torch modules between env.step calls (kept for policies of another shape)"),
                                "launches_per_rollout": 6 if persistent else 2 * T + 7 if fused else None,
                                "episodes_last_rollout": a2c.logger.get("rollout/episodes")}
                penv.close()
            if world > 1:
                # A rollout is 128 dependent steps of ~8 us each: giving a rank FEWER envs does not shorten it (configs[4]'s 4096 envs over
                # 8 GPUs = 512 per rank: the same 1.0 ms).  What scales is envs: the same collection with 4096 envs on EVERY rank.
                penv = EWNDeviceVecEnv(4096, 7, 3, seed=SEED + 7, shaped=True, env_id0=rank * 4096, engine=e7)
                a2c = A2C("MultiInputPolicy", penv, n_steps=T, seed=rank)
                a2c.collect_rollouts(); torch.cuda.synchronize()
                dist.barrier()
                p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                p0.record()
                for _ in range(20):
                    a2c.collect_rollouts()
                p1.record(); torch.cuda.synchronize()
                tt = torch.tensor([p0.elapsed_time(p1) / 20], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                rollout["with_policy_4096_per_gpu"] = {"ms_per_rollout": float(tt[0]), "envs_per_gpu": 4096,
                                                       "env_steps_per_s": 4096 * world * T / (float(tt[0]) / 1e3), "scaling": "weak"}
                penv.close()
        except Exception as err:                     # noqa: BLE001
            rollout["with_policy"] = {"unavailable": str(err)[:200]}

        def rollout_stats(lo, hi):                  # one rollout of the games with global ids [lo, hi) from a fresh reset -> stats[4]
            m = hi - lo
            st = e7.reset(m, seed=SEED + 5, env_id0=lo)
            tl_, pv_ = torch.full((m,), 10, dtype=torch.int32, device=dev), None
            pv_ = e7.evaluate(st, "hybrid")
            acc = torch.zeros(4, dtype=torch.int64, device=dev)
            chk = torch.zeros(1, dtype=torch.int64, device=dev)
            for t in range(T):
                o = e7.step(st, pool_global[t % 8][lo:hi].contiguous(), seed=SEED + 5, step=t + 1, env_id0=lo, shaped=True, illegal_move_reward=-1.0,
                            tolerance=tl_, prev_score=pv_, stats=acc, want=("reward_f32", "terminated"))
                chk += (o["reward_f32"].double() * 1024).round().long().sum() + o["terminated"].long().sum() * (t + 1)
            return torch.cat([acc, chk])

        if dist:
            # env ids are global, so the N-rank rollout statistics (all-reduced) must equal the single-rank ones -- rank 0
            # plays all 4,096 games alone and compares {wins, losses, illegal endings, episodes, reward/done checksum}
            mine_ = rollout_stats(rank * n7, (rank + 1) * n7)
            dist.all_reduce(mine_)
            bad = 0.0
            if rank == 0:
                alone = rollout_stats(0, n7 * world)
                bad = 0.0 if bool(torch.equal(alone, mine_)) else 1.0
            tt = torch.tensor([bad], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            if float(tt.item()):
                raise SystemExit("rollout statistics of the N-rank run differ from the single-rank run of the same games")
            rollout["identical_to_single_rank"] = True
            rollout["identity_check"] = f"rank 0 alone: {n7 * world} games x {T} steps == all-reduced stats + reward/done checksum of {world} ranks"

    # ---- e2e through the public host-buffer API (EWNVecEnv.step: pinned host actions in, host observations out), every rank at once
    from ewn_gym_b200.vec_env import EWNVecEnv
    e2e = None
    if not args.skip_e2e:
        env = EWNVecEnv(n, N_BOARD, N_LAYER, seed=SEED, device=local_rank, env_id0=env_id0, staged=args.e2e_staged, chunks=args.e2e_chunks,
                        numa_bind=world > 1 and os.environ.get("EWN_NUMA_BIND", "1") != "0")
        env.reset()
        rng = np.random.default_rng(7 + rank)
        pool = [np.stack([rng.integers(0, 2, n), rng.integers(0, 3, n)], 1).astype(np.int64) for _ in range(2)]
        abuf = env.action_buffer
        for k in range(3):
            abuf[...] = pool[k % 2]; env.step(abuf)
        ksteps = max(3, min(args.steps, 20))
        tot, term_rows = 0.0, 0
        if dist: dist.barrier()
        for k in range(ksteps):
            abuf[...] = pool[k % 2]                  # the policy's output lands in page-locked memory (untimed)
            t0 = time.perf_counter()
            obs, rew, done, infos = env.step(abuf)   # H2D actions + kernel + D2H obs/reward/done/terminal obs, synchronous
            tot += time.perf_counter() - t0
            term_rows += env.terminal_count
        if dist:                                     # whole job: all ranks step their own games over their own PCIe link; slowest rank counts
            tt = torch.tensor([tot], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); tot = float(tt.item())
        nn = N_BOARD * N_BOARD
        e2e = {"value": world * n / (tot / ksteps), "unit": "env steps/s", "h2d_bytes_per_step": 16 * n * world,
               "d2h_bytes_per_step": ((2 * nn + 8 + 4 + 1 + 1) * n + (2 * nn + 8 + 4) * (term_rows // ksteps) + 4) * world, "ms_per_step": 1e3 * tot / ksteps,
               "steps": ksteps, "n_gpus": world, "terminal_observations_per_step": term_rows // ksteps, "numa_bind": env.numa,
               "note": "host wall clock around EWNVecEnv.step on every rank (max over ranks): H2D copy of the int64 actions from page-locked memory, k_vstep in 8 chunks on 3 streams, "
                       + ("results staged in HBM and copied back (D2H)" if args.e2e_staged else "results stored by the kernel straight into the page-locked host buffers over PCIe")
                       + "; PCIe-bound"}
        env.close()
        # the ceiling of this line: page-locked D2H of the same bytes with a concurrent H2D of the actions, every rank at once (what the
        # host side of the box moves when all its PCIe links are busy together)
        d_bytes, h_bytes = int(e2e["d2h_bytes_per_step"] // world), int(e2e["h2d_bytes_per_step"] // world)
        db, hb = torch.empty(d_bytes, dtype=torch.uint8, device=dev), torch.empty(d_bytes, dtype=torch.uint8).pin_memory()
        db2, hb2 = torch.empty(h_bytes, dtype=torch.uint8, device=dev), torch.empty(h_bytes, dtype=torch.uint8).pin_memory()
        s2 = torch.cuda.Stream(device=dev)
        for _ in range(3):
            hb.copy_(db, non_blocking=True)
        torch.cuda.synchronize()
        if dist: dist.barrier()
        reps = 10
        t0 = time.perf_counter()
        for _ in range(reps):
            hb.copy_(db, non_blocking=True)
            with torch.cuda.stream(s2):
                db2.copy_(hb2, non_blocking=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        if dist:
            tt = torch.tensor([dt], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); dt = float(tt.item())
        ceil_gbs = world * (d_bytes + h_bytes) / dt / 1e9
        got_gbs = (e2e["d2h_bytes_per_step"] + e2e["h2d_bytes_per_step"]) / (e2e["ms_per_step"] / 1e3) / 1e9
        e2e["host_ceiling_gbs"] = ceil_gbs
        e2e["achieved_gbs"] = got_gbs
        e2e["frac_of_ceiling"] = got_gbs / ceil_gbs
        e2e["host_ceiling_note"] = (f"cudaMemcpyAsync of the step's {d_bytes} B D2H + {h_bytes} B H2D per rank between page-locked and device "
                                    f"memory, {world} rank(s) concurrently, max over ranks; the step's results are STORED by the kernel over PCIe "
                                    "(no copy engine), so frac_of_ceiling can exceed 1 where the copy engines, not the host memory, limit the memcpy rate")
        del db, hb, db2, hb2

    # ---- the same step through the device-resident VecEnv class (CUDA tensors in and out: what a policy on the GPU calls)
    device_env = None
    if not args.skip_e2e:
        from ewn_gym_b200.device_env import EWNDeviceVecEnv
        denv = EWNDeviceVecEnv(n, N_BOARD, N_LAYER, seed=SEED, env_id0=env_id0, engine=e)
        denv.reset()
        denv.action_buffer.copy_(sets[0]["act"])
        for _ in range(5):
            denv.step(denv.action_buffer)
        ksteps = max(10, min(args.steps, 100))
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        if dist: dist.barrier()
        d0.record()
        for _ in range(ksteps):
            denv.step(denv.action_buffer)        # the policy wrote its actions into the env's action tensor
        d1.record(); torch.cuda.synchronize()
        dms = d0.elapsed_time(d1) / ksteps
        if dist:
            tt = torch.tensor([dms], device=dev, dtype=torch.float64); dist.all_reduce(tt, op=dist.ReduceOp.MAX); dms = float(tt.item())
        device_env = {"value": world * n / (dms / 1e3), "unit": "env steps/s", "ms_per_step": dms, "steps": ksteps,
                      "note": "EWNDeviceVecEnv.step(action tensor) -> (obs, reward, done, infos) as CUDA tensors, one env set of 154 MB (> L2), "
                              "Python call overhead included, no host copies"}
        denv.close()

    if rank != 0:
        if dist: dist.destroy_process_group()
        return

    hbm, peak_src = peaks()
    rows_per_launch = finished_timed / args.steps
    alg_bytes = BYTES_PER_STEP * n + BYTES_PER_TERMINAL_ROW * rows_per_launch
    achieved = alg_bytes / (kernel_ms / 1e3) / 1e9
    obs_only = BYTES_PER_STEP * n / (kernel_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": "k_vstep<P5,5>", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                "traffic": None, "peak_source": peak_src, "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes,
                "bytes": {"per_env_step": BYTES_PER_STEP, "per_terminal_row": BYTES_PER_TERMINAL_ROW, "terminal_rows_per_launch": rows_per_launch,
                          "note": "96 B per env step (state R+W, action R, reward, flags, board, dice) + 62 B per finished game (its compacted "
                                  "terminal observation: board, dice, env index), rows counted from the episodes that ended in the timed steps"},
                "without_terminal_rows": {"achieved": obs_only, "frac": obs_only / hbm, "note": "round-1 accounting (96 B x games only)"}}
    prof = os.path.join(ROOT, "profiles", "k_step_traffic.json")
    if os.path.exists(prof):
        try:
            roofline["traffic"] = json.load(open(prof)).get("dram_bytes_per_launch")
        except Exception:
            pass
    cpu = None
    if not args.skip_cpu:
        cores = os.cpu_count() or 1
        v, nsteps, el = cpu_port_steps(n, 10.0, cores)
        cpu = {"value": v, "unit": "env steps/s", "cores": cores, "kind": "port", "label": "ours: oracle/ewn_oracle.c on all host threads",
               "sample": f"{n} concurrent 5x5 games x {nsteps} steps ({el:.1f} s) of oracle/ewn_oracle.c, uniform random actions, auto-reset",
               "reference_python": None if args.skip_reference_python else reference_python()}
    st = stats.cpu().tolist()
    line = {"metric": "env_steps_per_sec", "value": value, "unit": "env steps/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": workload_config(args, n), "clocks": clk.summary(), "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roofline, "cpu_baseline": cpu, "agent_uniform_legal": ar_run, "playouts": playout, "expectiminimax": emm, "rollout_7x7": rollout,
            "rollout_5x5": rollout_5x5, "device_env": device_env,
            "episodes": {"agent_wins": st[0], "agent_losses": st[1], "illegal_endings": st[2], "finished": st[3],
                         "agent": "uniform over MultiDiscrete([2,3]) (host-supplied int64 actions; most games end on an illegal opening move -- "
                                  "the in-kernel legal agent run is `agent_uniform_legal`)"},
            "wall_s_timed_region": wall}
    print(json.dumps(line), flush=True)
    if dist: dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=1 << 20, help="concurrent games per GPU")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-agent-random", action="store_true")
    ap.add_argument("--skip-reference-python", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-playouts", action="store_true")
    ap.add_argument("--skip-search", action="store_true")
    ap.add_argument("--skip-rollout", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=0, help="pipeline depth of EWNVecEnv.step (0 = library default)")
    ap.add_argument("--e2e-staged", action="store_true", help="e2e through device staging arrays + D2H copies instead of direct host stores")
    ap.add_argument("--skip-clocks", action="store_true", help="no extended load phase for the clock sampler (profiling runs)")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
