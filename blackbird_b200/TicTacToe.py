"""TicTacToe mirror: the reference's TicTacToe.BoardState (TicTacToe.py:9-138) over the packed
GPU state (csrc/games.cuh, struct TicTacToe).  Same class attributes, methods, return types and
error messages; the rules run in the CUDA library."""
import numpy as np

from .GameState import DeviceBoard


class BoardState(DeviceBoard):
    Players = {0: ' ', 1: 'X', 2: 'O'}
    Size = 3
    InARow = 3
    Dirs = [(0, 1), (1, 1), (1, 0), (1, -1)]
    BoardShape = np.array([Size, Size], dtype=np.int8)
    LegalMoves = Size ** 2
    GameType = 'TicTacToe'

    @property
    def Board(self):
        """int8 (3, 3, 2) one-hot planes, decoded from the packed state (a copy: edit through
        ApplyAction or FromArrays)."""
        cells = self._unpack()[0].reshape(self.Size, self.Size)
        return np.stack([cells == 1, cells == 2], axis=-1).astype(np.int8)

    @classmethod
    def FromArrays(cls, board, player, previousPlayer=None, device=0):
        from . import engine as _engine
        cells = (board[:, :, 0] + 2 * board[:, :, 1]).astype(np.int8).reshape(1, -1)
        return cls(_packed=_engine.pack(cls.GameType, cells, [player], [previousPlayer or 0], None, device=device))

    def LegalActionShape(self):
        return self.BoardShape                                   # TicTacToe.py:38-39

    def EvalToString(self, eval):
        return str(eval.reshape(3, 3))                           # TicTacToe.py:78-80

    def _coordsToIndex(self, coords):
        return coords[0] * self.Size + coords[1]

    def _indexToCoords(self, index):
        return (index // self.Size, index % self.Size)

    def __str__(self):
        cells = self._unpack()[0].reshape(self.Size, self.Size)
        s = ''
        for i in range(self.Size):
            s += '[ ' + '|'.join(' {} '.format(self.Players[int(c)]) for c in cells[i]) + ']\n'
        return s
