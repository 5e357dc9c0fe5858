"""Helpers shared by the oracle (CPU) and engine (GPU) parity tests: load golden traces, index them per step."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def game_traces(name):
    """-> dict with per-step arrays plus `t` (step index inside its game) and `first` (row of the game's step 0)."""
    z = load(name)
    steps = z["steps"]
    game = steps[:, 0]
    # rows are grouped by game in play order
    change = np.r_[True, game[1:] != game[:-1]]
    first_row = np.flatnonzero(change)
    gidx = np.cumsum(change) - 1
    t = np.arange(len(game)) - first_row[gidx]
    return dict(N=int(z["N"]), L=int(z["L"]), steps=steps, t=t, gidx=gidx, first_row=first_row,
                rewards=z["rewards"].astype(np.float64), boards=z["boards"].astype(np.int16),
                game_rewards=z["game_rewards"], game_steps=z["game_steps"], init_dice=z["init_dice"].astype(np.int32))
