"""classical_policies surface (ewn_gym_b200/policies.py): predict(obs) -> (action, None) against the reference's recorded
answers (ExpectiMinimaxAgent) and against the oracle with the same Philox keys (MctsAgent)."""
import numpy as np
import pytest

from oracle import oracle as O
from tests.backends import BACKENDS, make_engine
from tests.golden_util import load

from ewn_gym_b200.policies import ExpectiMinimaxAgent, MctsAgent, PolicyBase, RandomAgent

pytestmark = pytest.mark.parametrize("backend", BACKENDS)


@pytest.mark.parametrize("name", ["minimax_5x5.npz", "minimax_7x7.npz"])
def test_expectiminimax_agent_predict(backend, name):
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    eng = make_engine(backend, N, L)
    boards, dice = z["boards"].astype(np.int16), z["dice"].astype(np.int32)
    for depth in (1, 3, 4):
        agent = ExpectiMinimaxAgent(max_depth=depth, cube_layer=L, board_size=N, engine=eng, some_cli_flag=True)   # extra kwargs are swallowed
        for i in range(0, 40, 3):                                     # the reference's one-observation-at-a-time interface
            action, state = agent.predict({"board": boards[i], "dice_roll": int(dice[i])})
            want = z[f"action_d{depth}"][i]
            assert state is None
            if want[0] < 0:
                assert action is None                                 # finished position: minimax.py:22-23 returns action None
            else:
                assert isinstance(action, list) and action == [int(want[0]), int(want[1])]
            assert agent.get_action({"board": boards[i], "dice_roll": int(dice[i])})[0] == action
        acts, vals = agent.predict_batch(boards[:500], dice[:500])     # the batched form
        assert np.array_equal(vals.view(np.uint64), z[f"value_d{depth}"][:500]) and np.array_equal(acts, z[f"action_d{depth}"][:500])


def test_mcts_agent_predict(backend):
    z = load("minimax_5x5.npz")
    boards, dice = z["boards"][:64].astype(np.int16), z["dice"][:64].astype(np.int32)
    eng = make_engine(backend, 5, 3)
    agent = MctsAgent(cube_layer=3, board_size=5, num_simulations=6, num_env_copies=4, engine=eng, seed=11)
    assert isinstance(agent, PolicyBase) and agent.playouts_per_action == 24
    acts, wins = agent.predict_batch(boards, dice)
    ow, onch, _ = O.mc_root_wins(boards, dice, np.arange(64, dtype=np.uint32), 11, 0, 24)
    fin = O.check_win(boards) == 1                                     # finished roots have no children (predict -> None)
    onch = np.where(fin, 0, onch); ow[fin] = 0
    assert np.array_equal(wins.astype(np.uint32), ow)
    for i in range(64):
        if onch[i] == 0:
            assert tuple(acts[i]) == (-1, -1)
            continue
        best = int(np.argmax(ow[i, :onch[i]]))                        # np.argmax = first maximum (mcts.py:68)
        legal = O.legal(boards[i:i + 1], dice[i:i + 1], 1)[1][0]
        assert tuple(acts[i]) == tuple(legal[best])
    a, st = agent.predict({"board": boards[3], "dice_roll": int(dice[3])})
    assert st is None and (a is None or (a.shape == (2,) and a[0] in (0, 1) and a[1] in (0, 1, 2)))


def test_random_agent_uses_global_numpy_stream(backend):
    from ewn_gym_b200.envs import EinsteinWuerfeltNichtEnv
    env = EinsteinWuerfeltNichtEnv(engine=make_engine(backend, 5, 3))
    env.reset(seed=0)
    agent = RandomAgent(env)
    np.random.seed(123)
    a1, _ = agent.predict(None)
    np.random.seed(123)
    k = np.random.randint(0, 3)                                        # three legal moves at the initial position
    assert list(a1) == env.get_legal_actions(env.current_player)[k] and isinstance(a1, np.ndarray)
