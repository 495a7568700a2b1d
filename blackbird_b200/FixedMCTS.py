"""FixedMCTS mirror (FixedMCTS.py:1-34): every simulation descends exactly maxDepth plies from the
current Root, expanding unexpanded nodes on the way (engine search kind BB_FIXED)."""
from .MCTS import MCTS, Node  # noqa: F401


class FixedMCTS(MCTS):
    _KIND = 'fixed'

    def __init__(self, **kwargs):
        self.MaxDepth = kwargs.get('maxDepth')
        explorationRate = kwargs.get('explorationRate')
        timeLimit = kwargs.get('timeLimit')
        playLimit = kwargs.get('playLimit')
        kwargs.get('threads', 1)                     # read and dropped by the reference too (FixedMCTS.py:13)
        if self.MaxDepth <= 0:
            raise ValueError('MaxDepth for MCTS must be > 0.')
        super().__init__(explorationRate, timeLimit, playLimit)
