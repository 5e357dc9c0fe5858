"""The single-environment gym surface (ewn_gym_b200/envs.py) against traces recorded from the reference.

The env classes hold no rules logic: every decision is an engine call, so these tests run once against the test-only host
build of the kernel bodies (CPU tier) and once against the CUDA library (GPU tier)."""
import numpy as np
import pytest

from tests.backends import BACKENDS, make_engine
from tests.golden_util import game_traces, load

from ewn_gym_b200.constants import ClassicalPolicy, Player
from ewn_gym_b200.envs import EinsteinWuerfeltNichtEnv, EinsteinWurfeltNichtEnv, MiniMaxHeuristicEnv, MinimaxEnv
from ewn_gym_b200.policies import RandomAgent


@pytest.mark.parametrize("backend", BACKENDS)
def test_eval_random_plumbing_reproduces_reference_games(backend):
    """eval_random.py:17-51 with our env + RandomAgent: same seeds -> the reference's games, observation by observation
    (dice and random moves come from the global numpy stream in the reference's order)."""
    tr = game_traces("random_games_5x5.npz")
    env = EinsteinWuerfeltNichtEnv(board_size=5, cube_layer=3, engine=make_engine(backend, 5, 3))
    agent = RandomAgent(env)
    for seed in range(25):
        obs, info = env.reset(seed=seed)
        assert info == {} and obs["dice_roll"] == tr["init_dice"][seed]
        row, done, steps = tr["first_row"][seed], False, 0
        while not done:
            action, _ = agent.predict(obs)
            s = tr["steps"][row]
            assert obs["dice_roll"] == s[1] and list(action) == [s[2], s[3]]
            obs, reward, done, trunc, info = env.step(action)
            assert np.array_equal(obs["board"], tr["boards"][row]) and reward == tr["rewards"][row]
            assert done == bool(s[8]) and trunc == bool(s[9])
            if not done:
                assert obs["dice_roll"] == s[7]
            row += 1; steps += 1
        assert steps == tr["game_steps"][seed] and reward == tr["game_rewards"][seed]
        assert info["message"] == ("You won!" if reward > 0 else "You lost!")


class _Replay(MiniMaxHeuristicEnv):
    """dice and opponent answers come from the recorded trace instead of the RNG"""
    dice_feed, opp_feed = None, None

    def roll_dice(self):
        if self.dice_feed:
            self.dice_roll = int(self.dice_feed.pop(0))
        else:
            super().roll_dice()

    def opponent_action(self):
        return np.array(self.opp_feed.pop(0)) if self.opp_feed else super().opponent_action()


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("name,n_envs,n_steps", [("shaped_5x5.npz", 6, 128), ("shaped_7x7.npz", 3, 96)])
def test_minimax_heuristic_env_traces(backend, name, n_envs, n_steps):
    """MiniMaxHeuristicEnv.step (training_ewn.py:40-99): shaped fp64 rewards bit for bit, lifetime tolerance budget."""
    z = load(name)
    N, L = int(z["N"]), int(z["L"])
    eng = make_engine(backend, N, L)
    for i in range(n_envs):
        env = _Replay(board_size=N, cube_layer=L, seed=1000 + i, illegal_move_reward=-1.0, illegal_move_tolerance=int(z["tolerance0"]),
                      goal_reward=10., engine=eng)
        assert env.reward == 1.0 and isinstance(env.opponent_policy, RandomAgent)
        assert np.array_equal(env.board, z["init_board"][i]) and env.dice_roll == z["init_dice"][i]
        assert np.float64(env.prev_score).view(np.uint64) == z["init_prev"][i]
        for t in range(n_steps):
            r = z["rows"][i, t]
            assert env.dice_roll == r[0]
            env.dice_feed, env.opp_feed = [d for d in (r[3], r[6]) if d > 0], ([(r[4], r[5])] if r[4] >= 0 else [])
            obs, reward, term, trunc, info = env.step(np.array([r[1], r[2]]))
            assert np.array_equal(obs["board"], z["boards"][i, t])
            assert np.float64(reward).view(np.uint64) == z["rewards"][i, t], (i, t, reward)
            assert (term, trunc) == (bool(r[7]), bool(r[8])) and env.illegal_move_tolerance == r[10]
            if term:
                env.dice_feed = [r[9]]
                env.reset()


@pytest.mark.parametrize("backend", BACKENDS)
def test_helper_methods(backend):
    env = MinimaxEnv(board_size=5, cube_layer=3, engine=make_engine(backend, 5, 3))
    assert EinsteinWurfeltNichtEnv is EinsteinWuerfeltNichtEnv
    env.reset(seed=0)                                                    # board rows and dice 5 after reset(seed=0)
    assert env.board.tolist() == [[1, 2, 4, 0, 0], [3, 5, 0, 0, 0], [6, 0, 0, 0, -6], [0, 0, 0, -5, -3], [0, 0, -4, -2, -1]]
    assert env.dice_roll == 5 and env.current_player == Player.TOP_LEFT and not env.check_win()
    for d in range(1, 7):
        env.set_dice_roll(d)
        assert env.get_legal_actions(Player.TOP_LEFT) == [[0, 0], [0, 1], [0, 2]] == env.get_legal_actions(Player.BOTTOM_RIGHT)
    pos = env.cube_pos
    assert tuple(pos[0]) == (0, 0) and tuple(pos[5]) == (2, 0) and tuple(pos[-1]) == (4, 4) and tuple(pos[-6]) == (2, 4)
    env.set_dice_roll(5)
    assert env.find_cube_to_move(True) == 4 and env.find_cube_to_move(False, Player.BOTTOM_RIGHT) == -5
    assert env.get_cube_legal_directions(0) == [0, 1, 2] and env.get_cube_legal_directions(-1) == [0, 1, 2]
    # TL cube 1 captures its own cube 2 by moving right (envs/ewn.py:254-258), then 2 is masked and neighbours are looked up
    assert env.execute_move(0, 0) and env.board[0, 1] == 1 and env.board[0, 0] == 0
    assert env.cube_pos.mask[1][0]
    env.set_dice_roll(2)
    assert env.find_cube_to_move(True) == 2 and env.find_cube_to_move(False) == 0
    assert env.find_near_cube(1, True) == 2 and env.find_near_cube(1, False) == 0 and env.find_near_cube(0, False) is None
    assert env.find_near_cube(-6, True, Player.BOTTOM_RIGHT) is None and env.find_near_cube(-6, False, Player.BOTTOM_RIGHT) == -5
    assert env.get_legal_actions(Player.TOP_LEFT) == [[1, 0], [1, 1], [1, 2], [0, 0], [0, 1], [0, 2]]
    assert env.execute_move(-1, 0, dry_run=True) and env.board[4, 4] == -1          # dry run leaves the state alone
    # make / undo (envs/ewn.py:377-434)
    before = env.board.copy()
    env.make_simulated_action(Player.BOTTOM_RIGHT, [0, 2])
    assert not np.array_equal(env.board, before)
    env.undo_simulated_action()
    assert np.array_equal(env.board, before)
    # off-board move: state untouched, False (:262-263)
    env.reset(seed=1)
    b1 = env.board.copy()
    assert env.update_position(4, 4, 2, -1) == (3, 3) and not env.is_within_board(5, 0)
    # a direction outside {0,1,2} matches no branch of :229-247: the cube stays, the move counts as valid
    assert env.execute_move(-1, 5) and np.array_equal(env.board, b1)
    # evaluate: fresh MiniMaxHeuristicEnv value of SURVEY 8c, terminal +-10
    assert env.evaluate() == -0.49999999999999994 and env.evaluate("attk") == -12.0
    w = env.simulate()
    assert 0.0 <= w <= 1.0


@pytest.mark.parametrize("backend", BACKENDS)
def test_step_illegal_and_opponents(backend):
    eng = make_engine(backend, 5, 3)
    env = EinsteinWuerfeltNichtEnv(engine=eng, reward=2.0)
    obs, _ = env.reset(seed=3)
    b0 = obs["board"].copy()
    obs, r, term, trunc, info = env.step(np.array([0, 7]))                   # direction out of range: a valid no-op in the reference
    assert (r, term, trunc) == (0, False, False) and (obs["board"] > 0).sum() == 6 and np.array_equal(obs["board"] * (obs["board"] > 0), b0 * (b0 > 0))
    for _ in range(40):                                                        # walk right until a cube leaves the board (:445-448)
        obs, r, term, trunc, info = env.step(np.array([0, 0]))
        if term:
            break
    assert term and r in (2.0, -2.0)
    if trunc:
        assert r == -2.0 and info["message"] == "Invalid move for player! End the game."
    # minimax and Monte-Carlo opponents plug in through the same ctor keywords as the reference (envs/ewn.py:265-283)
    for pol, kw in ((ClassicalPolicy.minimax, dict(max_depth=2)), ("mcts", dict(num_simulations=4, num_env_copies=2))):
        env = EinsteinWuerfeltNichtEnv(engine=eng, opponent_policy=pol, **kw)
        obs, _ = env.reset(seed=5)
        agent, done, n = RandomAgent(env), False, 0
        while not done and n < 60:
            a, _ = agent.predict(obs)
            obs, r, done, trunc, info = env.step(a)
            assert not trunc
            n += 1
        assert done and r in (1.0, -1.0)
    with pytest.raises(AssertionError):
        EinsteinWuerfeltNichtEnv(board_size=4, cube_layer=3, engine=eng)
