"""GameState -- the abstract game interface of the reference (GameState.py:1-29) and the shared
device-backed implementation behind the three game mirrors (TicTacToe.py, Connect4.py,
DragonChess.py in this package).

A mirror object owns ONE packed state on the GPU (uint8 tensor [1, state_bytes]) and answers every
method of the reference's BoardState through the batched rule kernels of the C-ABI library with
n = 1 (bb_game_legal / apply / winner / planes / equal / pack / unpack).  There is no CPU
fallback: without the CUDA library these classes raise.  The heavy paths (self-play, search) never
go through these per-object calls; they exist so that main.py / Blackbird.py keep working when the
import is switched.
"""
import numpy as np
import torch

from . import engine as _engine


class GameState(object):
    """GameState.py:1-29 (same method set; NumericRepresentation / SerializeState are declared
    but implemented by no game in the reference either)."""

    def __init__(self):
        self.Board = None
        self.Player = None
        self.PreviousPlayer = None

    def Copy(self):
        raise NotImplementedError

    def LegalActions(self):
        raise NotImplementedError

    def LegalActionShape(self):
        raise NotImplementedError

    def ApplyAction(self, action):
        raise NotImplementedError

    def Winner(self, prevAction=None):
        raise NotImplementedError

    def NumericRepresentation(self):
        raise NotImplementedError

    def EvalToString(self, eval):
        return str(eval)

    def SerializeState(self, state, policy, eval):
        raise NotImplementedError


class DeviceBoard(GameState):
    """Common machinery of the three mirrors.  Subclasses set GameType / LegalMoves / COPY_KEEPS_PREV
    and the array conversions."""
    GameType = None
    LegalMoves = None
    COPY_KEEPS_PREV = False            # TicTacToe / Connect4 Copy() drops PreviousPlayer
    DEVICE = 0

    def __init__(self, _packed=None):
        # no GameState.__init__: Board / Player / PreviousPlayer are properties over the packed state
        self._st = _engine.init_states(self.GameType, 1, self.DEVICE) if _packed is None else _packed

    # -- packed <-> reference arrays ------------------------------------------------------------
    def _unpack(self):
        cells, pl, pv, ca = _engine.unpack(self.GameType, self._st)
        return cells.cpu().numpy()[0], int(pl.item()), int(pv.item()), int(ca.item())

    def _repack(self, cells=None, player=None, prev=None, castle=None):
        c0, p0, v0, k0 = self._unpack()
        cells = c0 if cells is None else cells
        player = p0 if player is None else player
        prev = v0 if prev is None else prev
        castle = k0 if castle is None else castle
        self._st = _engine.pack(self.GameType, np.asarray(cells, dtype=np.int8).reshape(1, -1), [player], [prev], [castle],
                                device=self._st.device.index or 0)

    @classmethod
    def FromPacked(cls, packed_bytes, device=0):
        """Build a state object around packed bytes (numpy uint8[state_bytes]) as the engine exports them."""
        t = torch.as_tensor(np.ascontiguousarray(packed_bytes, dtype=np.uint8).reshape(1, -1)).to(torch.device('cuda', device))
        return cls(_packed=t)

    def Packed(self):
        return self._st.cpu().numpy()[0]

    @property
    def Player(self):
        return self._unpack()[1]

    @Player.setter
    def Player(self, value):
        self._repack(player=int(value))

    @property
    def PreviousPlayer(self):
        return self._unpack()[2] or None

    @PreviousPlayer.setter
    def PreviousPlayer(self, value):
        self._repack(prev=int(value or 0))

    # -- reference methods ----------------------------------------------------------------------
    def Copy(self):
        new = type(self)(_packed=self._st.clone())
        if not self.COPY_KEEPS_PREV:
            new._repack(prev=0)
        return new

    def LegalActions(self):
        return _engine.legal(self.GameType, self._st).cpu().numpy()[0]

    def ApplyAction(self, action):
        ok = _engine.apply(self.GameType, self._st, torch.tensor([int(action)], dtype=torch.int32))
        if int(ok.item()) == 0:
            raise ValueError('Tried to make an illegal move.')

    def _winner_code(self, prevAction):
        prev = None if prevAction is None else torch.tensor([int(prevAction)], dtype=torch.int32)
        return int(_engine.winner(self.GameType, self._st, prev).item())

    def Winner(self, prevAction=None):
        code = self._winner_code(prevAction)
        # TicTacToe / Connect4 read the winner off a float array (np.float64 1.0 / 2.0), a draw is the
        # python int 0, the empty-cell artefact is np.float64(0.0)
        if code < 0:
            return None
        if code == 0:
            return 0
        if code == 3:
            return np.float64(0.0)
        return np.float64(code)

    def AsInputArray(self):
        return _engine.planes(self.GameType, self._st).cpu().numpy()

    def __eq__(self, other):
        if not isinstance(other, DeviceBoard) or other.GameType != self.GameType:
            return False
        return bool(_engine.equal(self.GameType, self._st, other._st.to(self._st.device)).item())

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return "{0}{1}".format(self.Player, str(self)).__hash__()
