"""DynamicMCTS mirror (DynamicMCTS.py:1-34): a simulation descends to the first unexpanded node,
expands it and evaluates THAT node (engine search kind BB_DYNAMIC)."""
from .MCTS import MCTS, Node  # noqa: F401


class DynamicMCTS(MCTS):
    _KIND = 'dynamic'

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
