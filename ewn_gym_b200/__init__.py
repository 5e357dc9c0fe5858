"""ewn_gym_b200 -- a B200-native (sm_100a) batched EinStein-wuerfelt-nicht engine behind the ewn-gym surface.

Public surface (mirrors jchen8tw/ewn-gym):
  envs      EinsteinWuerfeltNichtEnv, MinimaxEnv, MiniMaxHeuristicEnv, EWNVecEnv
  policies  PolicyBase, RandomAgent, ExpectiMinimaxAgent, MctsAgent
  engine    Engine: batched device-tensor calls over the C ABI (include/ewn_b200.h)
There is no CPU fallback: importing the compute modules needs ewn_gym_b200/csrc/libewn_b200.so, running them a GPU.
"""
__version__ = "0.1.0"
