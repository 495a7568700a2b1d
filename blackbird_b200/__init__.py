"""blackbird_b200 -- B200-native self-play engine behind BlackBird's GameState / MCTS API.

Layout: csrc/ (CUDA kernels + C-ABI, built into libbbengine.so), _lib.py (ctypes binding),
engine.py (batched handle), records.py (wire format) and the reference-facing mirror modules
(GameState, TicTacToe, Connect4, DragonChess, MCTS, FixedMCTS, DynamicMCTS, Blackbird,
DataManager).  Importing the package does not need a GPU; computing anything does.
"""
__version__ = '0.1.0'
