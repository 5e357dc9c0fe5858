"""K5 timing on one B200: ewn_policy_forward alone, the 7x7 shaped step kernel alone, the two alternating (one rollout step), and the
fused collector's graph; us per launch from CUDA events.   python tools/k5_profile.py [n_envs]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200 import _lib
if os.environ.get("EWN_LIB_VARIANT"):
    _lib.LIB_PATH = os.path.abspath(os.environ["EWN_LIB_VARIANT"])
from ewn_gym_b200.device_env import EWNDeviceVecEnv
from ewn_gym_b200.train import A2C
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = 128
env = EWNDeviceVecEnv(n, 7, 3, seed=1, shaped=True)
m = A2C("MultiInputPolicy", env, n_steps=T, seed=0, persistent=False)      # the graph collector: its plans are timed one by one below
for _ in range(3): m.collect_rollouts()
torch.cuda.synchronize()
e, f = env.e, m.fb


def timed(fn, reps=200):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


pol = lambda: e.policy_forward(m._net, f["board"][0], f["dice"][0], seed=1, step=1, actions=f["actions"][0], values=f["values"][0], logp=f["log_probs"][0])
plan = m._plans[0]
stp = lambda: plan.launch(0)
def both(): pol(); stp()
def graph_of(fn, k=100):
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        if fn is stp:
            pl = env.bind_step(f["actions"][0], {"reward_f32": f["rewards"][0], "terminated": f["dones"][0], "truncated": f["trunc"],
                                                 "obs_board": f["board"][1], "obs_dice": f["dice"][1]})
            call = lambda: pl.launch(0)
        else:
            call = fn
        call(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(k): call()
    return lambda: g.replay(), k
gp, k = graph_of(pol); gs, _ = graph_of(stp)
print(f"n={n}  in a graph of {k}: policy {timed(gp, 20) / k:7.2f} us   step {timed(gs, 20) / k:7.2f} us", flush=True)
print(f"n={n}  policy alone {timed(pol):7.2f} us   step alone {timed(stp):7.2f} us   alternating {timed(both):7.2f} us per pair   "
      f"graph {timed(m._graph.replay, 20) / T:7.2f} us per step", flush=True)
