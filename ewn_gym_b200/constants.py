"""Player / ClassicalPolicy enums with the reference's values and helpers (constants/player.py:4-17, policy.py:4-20)."""
from enum import Enum


class Player(Enum):
    TOP_LEFT = 1
    BOTTOM_RIGHT = 2
    CHANCE = 3

    @classmethod
    def get_opponent(cls, player: "Player") -> "Player":
        if player == Player.TOP_LEFT:
            return Player.BOTTOM_RIGHT
        if player == Player.BOTTOM_RIGHT:
            return Player.TOP_LEFT
        raise ValueError("Invalid player")


class ClassicalPolicy(Enum):
    random = "random"
    minimax = "minimax"
    uct = "uct"
    alpha_zero = "alpha_zero"
    mcts = "mcts"

    def __str__(self):
        return self.value

    @staticmethod
    def from_string(s):
        """constants/policy.py:15-20: unknown names are returned unchanged (they are paths of SB3 checkpoints)."""
        try:
            return ClassicalPolicy[s]
        except KeyError:
            return s
