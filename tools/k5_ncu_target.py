"""ncu target for K5 / K5p: a few fused rollouts of BASELINE configs[4] (4096 envs x 128 steps, 7x7 shaped), persistent and graph collectors"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewn_gym_b200.device_env import EWNDeviceVecEnv
from ewn_gym_b200.train import A2C
for persistent in (True, False):
    env = EWNDeviceVecEnv(4096, 7, 3, seed=1, shaped=True)
    m = A2C("MultiInputPolicy", env, n_steps=128, seed=0, persistent=persistent)
    for _ in range(3): m.collect_rollouts()
    torch.cuda.synchronize()
