"""RandomMCTS mirror (RandomMCTS.py:4-25): the opponent of Blackbird.TestRandom.

It never searches, so it needs no engine: one draw from the global ``np.random`` stream picks a
legal action, a second one is the reported win rate (same order of draws as the reference, so an
arena game seeded with ``random.seed`` / ``np.random.seed`` replays move for move).  The state it
is handed is a DeviceBoard; the move is applied through the batched rule kernels (Copy +
ApplyAction of the mirror game classes), not on the host.
"""
import numpy as np

from .MCTS import MCTS


class RandomMCTS(MCTS):
    def __init__(self, *args, **kwargs):
        # RandomMCTS.py:5-6 skips MCTS.__init__ altogether; the mirror only needs the root flags
        # DropRoot / MoveRoot look at.
        self._eng = None
        self._has_root = False
        self._root_view = None

    def FindMove(self, state, *args, **kwargs):
        legal = state.LegalActions()                                   # RandomMCTS.py:10-11
        action = np.random.choice([i for i in range(len(legal)) if legal[i] == 1])
        winRate = np.random.random()                                   # RandomMCTS.py:12
        childProbability = state.LegalActions()                        # RandomMCTS.py:13-16
        s = sum(childProbability)
        if s > 0:
            childProbability /= s
        self._last_action = int(action)
        return self._applyAction(state, action), winRate, childProbability

    def ResetRoot(self, *args, **kwargs):
        return

    def MoveRoot(self, *args, **kwargs):
        return
