"""Oracle game rules: numpy restatement of the reference's three GameState classes.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Each class keeps the
reference's duck-typed surface so the oracle search in
``oracle/search_np.py`` -- and the reference's own MCTS -- can drive it.

Behaviours that look odd are deliberate and cited: they define the results
the CUDA path must reproduce bit-exactly.
"""
import numpy as np

_DIRS = ((0, 1), (1, 1), (1, 0), (1, -1))   # Connect4.py:14, TicTacToe.py:13


class GameState(object):
    """Abstract interface, GameState.py:1-29."""
    Board = None
    Player = None
    PreviousPlayer = None


def _collapse(planes):
    """Connect4.py:110-114 / TicTacToe.py:113-117: float64 board of 0/1/2."""
    cells = np.zeros(planes.shape[:2])
    cells[planes[:, :, 0] == 1] = 1
    cells[planes[:, :, 1] == 1] = 2
    return cells


def _line_through(cells, i, j, need):
    """_checkVictory, Connect4.py:94-108 / TicTacToe.py:85-105.

    Counts the run of cells equal to cells[i, j] through (i, j) in each of the
    four directions (both ways); returns the cell value (np.float64, possibly
    0.0 for an empty cell!) on the first direction with a run >= need.
    """
    rows, cols = cells.shape
    p = cells[i, j]
    for di, dj in _DIRS:
        run = 0
        r = 0
        while r * di + i < rows and 0 <= r * dj + j < cols and cells[r * di + i, r * dj + j] == p:
            run += 1
            r += 1
        r = -1
        while r * di + i >= 0 and 0 <= r * dj + j < cols and cells[r * di + i, r * dj + j] == p:
            run += 1
            r -= 1
        if run >= need:
            return p
    return None


class _LineGame(GameState):
    Rows = 0
    Cols = 0
    InARow = 0
    Players = {0: ' ', 1: 'X', 2: 'O'}
    Dirs = list(_DIRS)

    def __init__(self):
        self.Board = np.zeros((self.Rows, self.Cols, 2), dtype=np.int8)
        self.Player = 1
        self.PreviousPlayer = None

    def Copy(self):
        # PreviousPlayer is NOT copied (Connect4.py:24-28, TicTacToe.py:23-27)
        twin = type(self)()
        twin.Player = self.Player
        twin.Board = self.Board.copy()
        return twin

    def AsInputArray(self):
        # Connect4.py:55-60 / TicTacToe.py:50-55
        out = np.zeros((1, self.Rows, self.Cols, 3), dtype=np.int8)
        out[0, :, :, :2] = self.Board
        out[0, :, :, 2] = 1 if self.Player == 1 else -1
        return out

    def EvalToString(self, eval):
        return str(eval)

    def _advance_turn(self):
        self.PreviousPlayer = self.Player
        self.Player = 1 if self.Player == 2 else 2

    def __eq__(self, other):
        # PreviousPlayer never compared (Connect4.py:129-132)
        return other.Player == self.Player and bool((other.Board == self.Board).all())

    def __hash__(self):
        return hash((self.Player, self.Board.tobytes()))


class Connect4State(_LineGame):
    """Connect4.py:9-135.  Board[row, col, player-1], row 0 = bottom."""
    Rows, Cols, InARow = 6, 7, 4
    Width, Height = 7, 6
    BoardShape = np.array([7, 6], dtype=np.int8)
    LegalMoves = 7
    GameType = 'Connect4'

    def LegalActions(self):
        # Connect4.py:30-36: top cell of the column empty; ignores game-over (A.2)
        top = self.Board[self.Rows - 1].sum(axis=1)
        return (top == 0).astype(np.float64)

    def LegalActionShape(self):
        return np.array([self.Cols], dtype=np.int8)

    def ApplyAction(self, action):
        # Connect4.py:41-53
        filled = self.Board[:, action, :].sum(axis=1) != 0
        if filled[self.Rows - 1]:
            raise ValueError('Tried to make an illegal move.')
        height = int(filled.nonzero()[0].max()) + 1 if filled.any() else 0
        self.Board[height, action, self.Player - 1] = 1
        self._advance_turn()

    def Winner(self, prevAction=None):
        # Connect4.py:62-83
        cells = _collapse(self.Board)
        if prevAction is not None:
            # top stone of that column; an EMPTY column leaves i = 0 (A.5)
            i = 0
            for i in reversed(range(self.Rows)):
                if cells[i, prevAction] != 0:
                    break
            hit = _line_through(cells, i, prevAction, self.InARow)
            if hit is not None:
                return hit
        else:
            for i in range(self.Rows):
                for j in range(self.Cols):
                    if cells[i, j] == 0:
                        continue
                    hit = _line_through(cells, i, j, self.InARow)
                    if hit is not None:
                        return hit
        if (cells[self.Rows - 1] != 0).all():      # _isOver, Connect4.py:88-92
            return 0
        return None


class TicTacToeState(_LineGame):
    """TicTacToe.py:9-138.  action = row*3 + col."""
    Rows, Cols, InARow = 3, 3, 3
    Size = 3
    BoardShape = np.array([3, 3], dtype=np.int8)
    LegalMoves = 9
    GameType = 'TicTacToe'

    def LegalActions(self):
        # TicTacToe.py:29-36
        return (self.Board.sum(axis=2) == 0).astype(np.float64).ravel()

    def LegalActionShape(self):
        return self.BoardShape

    def ApplyAction(self, action):
        # TicTacToe.py:41-48
        i, j = action // 3, action % 3
        if self.Board[i, j].sum() != 0:
            raise ValueError('Tried to make an illegal move.')
        self.Board[i, j, self.Player - 1] = 1
        self._advance_turn()

    def Winner(self, prevAction=None):
        # TicTacToe.py:57-76
        cells = _collapse(self.Board)
        if prevAction is not None:
            hit = _line_through(cells, prevAction // 3, prevAction % 3, self.InARow)
            if hit is not None:
                return hit
        else:
            for i in range(3):
                for j in range(3):
                    if cells[i, j] == 0:
                        continue
                    hit = _line_through(cells, i, j, self.InARow)
                    if hit is not None:
                        return hit
        if (cells > 0).sum() == 9:                 # _isOver, TicTacToe.py:82-83
            return 0
        return None


# ---------------------------------------------------------------------------
# DragonChess (DragonChess.py:10-371; rules summarised in SURVEY.md Appendix D)
# ---------------------------------------------------------------------------
K, P, N, B, R, Q = 1, 2, 3, 4, 5, 6
_PLANE_OF = {P: 0, -P: 1, R: 2, -R: 3, N: 4, -N: 5, B: 6, -B: 7, Q: 8, -Q: 9, K: 10, -K: 11}  # :11


def dc_action_id(sq1, sq2):
    """move_to_int, DragonChess.py:26-34: ordered pairs of distinct squares."""
    return sq1 * 63 + sq2 - (1 if sq2 > sq1 else 0)


def dc_action_squares(action):
    sq1, rem = divmod(action, 63)
    sq2 = rem + (1 if rem >= sq1 else 0)
    return sq1, sq2


class DragonChessState(GameState):
    GameType = 'DragonChess'
    LegalMoves = 4032

    def __init__(self):
        # FEN 'rnbqkbnr/pppppppp/8/8/8/8/3PPP2/4K3 w kq - 0 1' (:36), row 0 = White home rank (:44)
        b = np.zeros((8, 8))
        b[0, 4] = K
        b[1, 3:6] = P
        b[6, :] = -P
        b[7, :] = [-R, -N, -B, -Q, -K, -B, -N, -R]
        self.board = b
        self.Player = 1
        self.PreviousPlayer = None
        self.castle = [False, False, True, True]    # WK, WQ, BK, BQ  (:57-60)

    @property
    def Board(self):
        return self.board

    def Copy(self):
        # DragonChess.py:66-76: PreviousPlayer IS copied here (A.6)
        twin = DragonChessState.__new__(DragonChessState)
        twin.board = self.board.copy()
        twin.Player = self.Player
        twin.PreviousPlayer = self.PreviousPlayer
        twin.castle = list(self.castle)
        return twin

    # -- move legality ------------------------------------------------------
    def _pre_check(self, r1, c1, r2, c2):
        # _sanity_check, DragonChess.py:242-259 (targets are always on-board here)
        src, dst = self.board[r1, c1], self.board[r2, c2]
        if (r1, c1) == (r2, c2):
            return False
        if src > 0:
            if dst > 0:
                return False
        elif dst < 0:
            return False
        if src < 0 and self.Player == 1:
            return False
        if src > 0 and self.Player == 2:
            return False
        return src != 0

    def _clear_between(self, r1, c1, r2, c2):
        dr, dc = np.sign(r2 - r1), np.sign(c2 - c1)
        r, c = r1 + dr, c1 + dc
        while (r, c) != (r2, c2):
            if self.board[r, c] != 0:
                return False
            r, c = r + dr, c + dc
        return True

    def _pawn(self, r1, c1, r2, c2):
        # _is_legal_move_pawn, DragonChess.py:280-319.  Promotion shapes return the
        # integer 0 in the reference => never legal (A.19).
        b = self.board
        if b[r1, c1] > 0:
            fwd, home, last = 1, 1, 6
            enemy = lambda v: v < 0
        else:
            fwd, home, last = -1, 6, 1
            enemy = lambda v: v > 0
        goal = last + fwd
        if r1 == last and c1 == c2 and r2 == goal and b[goal, c2] == 0:
            return False
        if r1 == last and abs(c1 - c2) == 1 and r2 == goal and enemy(b[goal, c2]):
            return False
        if r1 == home:
            if c1 == c2:
                if r2 == home + 2 * fwd and b[home + 2 * fwd, c2] == 0 and b[home + fwd, c2] == 0:
                    return True
                return bool(r2 == home + fwd and b[home + fwd, c2] == 0)
        elif 1 < r1 < 6:
            if c1 == c2 and b[r2, c2] == 0 and r2 == r1 + fwd:
                return True
        return bool(abs(c1 - c2) == 1 and r2 == r1 + fwd and enemy(b[r2, c2]))

    def _geometry(self, r1, c1, r2, c2):
        kind = abs(self.board[r1, c1])
        dr, dc = abs(r1 - r2), abs(c1 - c2)
        if kind == K:
            return dr <= 1 and dc <= 1              # castling shapes return 0 (:356-363)
        if kind == P:
            return self._pawn(r1, c1, r2, c2)
        if kind == N:
            return (dr, dc) in ((1, 2), (2, 1))
        diag = dr == dc and self._clear_between(r1, c1, r2, c2)
        line = (dr == 0 or dc == 0) and self._clear_between(r1, c1, r2, c2)
        if kind == B:
            return diag
        if kind == R:
            return line
        return diag or line

    def _legal(self, r1, c1, r2, c2):
        return self._pre_check(r1, c1, r2, c2) and bool(self._geometry(r1, c1, r2, c2))

    def LegalActions(self):
        # DragonChess.py:78-106
        out = np.zeros(4032)
        for sq1 in range(64):
            if self.board[sq1 // 8, sq1 % 8] == 0:
                continue
            for sq2 in range(64):
                if sq2 != sq1 and self._legal(sq1 // 8, sq1 % 8, sq2 // 8, sq2 % 8):
                    out[dc_action_id(sq1, sq2)] = 1
        return out

    def LegalActionShape(self):
        return np.zeros(4212, dtype=np.int8)        # :108-109

    def ApplyAction(self, action):
        # DragonChess.py:127-137,158-159 -> Move :172-214
        if action >= 4032:
            raise ValueError('Tried to make an illegal move.')   # promotion/castle ids: never legal (A.19)
        sq1, sq2 = dc_action_squares(int(action))
        r1, c1, r2, c2 = sq1 // 8, sq1 % 8, sq2 // 8, sq2 % 8
        if not self._legal(r1, c1, r2, c2):
            raise ValueError('Tried to make an illegal move.')
        b = self.board
        b[r2, c2] = b[r1, c1]
        b[r1, c1] = 0
        if self.PreviousPlayer == 1 and self.Player == 1:         # double-move rule :192-197
            self.Player = 2
            self.PreviousPlayer = 1
        else:
            self.PreviousPlayer = self.Player
            self.Player = 1
        wk, wq, bk, bq = self.castle                               # :199-212
        if b[0, 4] != K:
            wk = wq = False
        elif b[0, 7] != R:
            wk = False
        if b[0, 0] != R:
            wq = False
        if b[7, 4] != -K:
            bk = bq = False
        elif b[7, 7] != -R:
            bk = False
        if b[7, 0] != -R:
            bq = False
        self.castle = [wk, wq, bk, bq]

    def Winner(self, prevAction=None):
        # DragonChess.py:161-167
        if not (self.board == -K).any():
            return 1
        if not (self.board == K).any():
            return 2
        return None

    def AsInputArray(self):
        # DragonChess.py:111-125
        out = np.zeros((1, 8, 8, 17), dtype=np.int8)
        for r in range(8):
            for c in range(8):
                v = int(self.board[r, c])
                if v:
                    out[0, r, c, _PLANE_OF[v]] = 1
        for k in range(4):
            out[0, :, :, 12 + k] = int(self.castle[k])
        out[0, :, :, 16] = 1 if (self.Player == 1 and self.PreviousPlayer == 1) else 0
        return out

    def EvalToString(self, eval):
        return str(eval)

    def __eq__(self, other):
        # DragonChess.py:226-237: PreviousPlayer ignored (A.21)
        return (other.Player == self.Player and list(other.castle) == list(self.castle)
                and bool((other.board == self.board).all()))

    def __hash__(self):
        return hash((self.Player, self.board.tobytes()))      # castle flags ignored, as :239-240


GAMES = {'TicTacToe': TicTacToeState, 'Connect4': Connect4State, 'DragonChess': DragonChessState}
