"""DragonChess mirror: the reference's DragonChess.BoardState (DragonChess.py:10-371) over the
packed GPU state (csrc/games.cuh, struct DragonChess: 8 rows of nibbles + castle/turn bits).
White = K + 3 P and moves twice per turn; promotions and castling are never legal (SURVEY.md
A.18-A.20); 4032 from->to actions."""
import numpy as np

from .GameState import DeviceBoard


class BoardState(DeviceBoard):
    piece_map = {1: 10, -1: 11, 2: 0, -2: 1, 3: 4, -3: 5, 4: 6, -4: 7, 5: 2, -5: 3, 6: 8, -6: 9}
    int_to_letter = {0: ' ', 1: 'K', 2: 'P', 3: 'N', 4: 'B', 5: 'R', 6: 'Q',
                     -1: 'k', -2: 'p', -3: 'n', -4: 'b', -5: 'r', -6: 'q'}
    letter_to_int = {v: k for k, v in int_to_letter.items() if k != 0}
    GameType = 'DragonChess'
    LegalMoves = 4032
    COPY_KEEPS_PREV = True                                       # DragonChess.py:66-76 copies PreviousPlayer
    fen = 'rnbqkbnr/pppppppp/8/8/8/8/3PPP2/4K3 w kq - 0 1'
    # action id <-> 'from to' strings in the reference's enumeration order (DragonChess.py:26-34)
    possible_moves = [str(a) + ' ' + str(b) for a in range(64) for b in range(64) if a != b]
    move_to_int = {m: i for i, m in enumerate(possible_moves)}
    int_to_move = {i: m for i, m in enumerate(possible_moves)}

    @property
    def board(self):
        """float64 (8, 8) signed piece codes, row 0 = White's home rank (DragonChess.py:39-44); a copy."""
        return self._unpack()[0].reshape(8, 8).astype(np.float64)

    Board = board

    def _castle(self, bit):
        return bool((self._unpack()[3] >> bit) & 1)

    _white_castle_kingside = property(lambda self: self._castle(0))
    _white_castle_queenside = property(lambda self: self._castle(1))
    _black_castle_kingside = property(lambda self: self._castle(2))
    _black_castle_queenside = property(lambda self: self._castle(3))

    @classmethod
    def FromArrays(cls, board, player, previousPlayer=None, castle=(False, False, True, True), device=0):
        from . import engine as _engine
        bits = sum(int(bool(b)) << i for i, b in enumerate(castle))
        cells = np.asarray(board).astype(np.int8).reshape(1, -1)
        return cls(_packed=_engine.pack(cls.GameType, cells, [player], [previousPlayer or 0], [bits], device=device))

    def LegalActionShape(self):
        return np.array([0 for _ in range(4212)], dtype=np.int8)  # DragonChess.py:108-109

    def Winner(self, prevAction=None):
        code = self._winner_code(None)                            # python ints here (DragonChess.py:161-167)
        return None if code < 0 else code

    def __str__(self):
        fboard = np.flip(self._unpack()[0].reshape(8, 8), axis=0)
        rep = '-----------------\n'
        for row in range(8):
            rep += '|' + ''.join('{}|'.format(self.int_to_letter[int(fboard[row, col])]) for col in range(8)) + '\n'
        return rep + '-----------------'
