import sys, os
sys.path.insert(0, '.')
import torch
from ewn_gym_b200.device_env import EWNDeviceVecEnv
from ewn_gym_b200.train import A2C
env = EWNDeviceVecEnv(4096, 7, 3, seed=1, shaped=True)
m = A2C("MultiInputPolicy", env, n_steps=4, seed=0)
m.collect_rollouts(); m.collect_rollouts()
torch.cuda.synchronize()
