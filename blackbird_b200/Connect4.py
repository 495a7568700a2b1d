"""Connect4 mirror: the reference's Connect4.BoardState (Connect4.py:9-135) over the packed GPU
state (csrc/games.cuh, struct Connect4: two 7-bit-stride column bitboards)."""
import numpy as np

from .GameState import DeviceBoard


class BoardState(DeviceBoard):
    Players = {0: ' ', 1: 'X', 2: 'O'}
    Width = 7
    Height = 6
    InARow = 4
    Dirs = [(0, 1), (1, 1), (1, 0), (1, -1)]
    BoardShape = np.array([Width, Height], dtype=np.int8)
    LegalMoves = Width
    GameType = 'Connect4'

    @property
    def Board(self):
        """int8 (6, 7, 2), row 0 = bottom (Connect4.py:20, 45-51); a decoded copy."""
        cells = self._unpack()[0].reshape(self.Height, self.Width)
        return np.stack([cells == 1, cells == 2], axis=-1).astype(np.int8)

    @classmethod
    def FromArrays(cls, board, player, previousPlayer=None, device=0):
        from . import engine as _engine
        cells = (board[:, :, 0] + 2 * board[:, :, 1]).astype(np.int8).reshape(1, -1)
        return cls(_packed=_engine.pack(cls.GameType, cells, [player], [previousPlayer or 0], None, device=device))

    def LegalActionShape(self):
        return np.array([self.Width], dtype=np.int8)             # Connect4.py:38-39

    def __str__(self):
        cells = self._unpack()[0].reshape(self.Height, self.Width)
        s = ''
        for i in reversed(range(self.Height)):                   # printed top row first (Connect4.py:119)
            s += '[ ' + '|'.join(' {} '.format(self.Players[int(c)]) for c in cells[i]) + ']\n'
        return s
