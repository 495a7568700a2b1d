"""Generate golden vectors by running the UNMODIFIED reference (/root/reference/src) in the
build container.  Not importable on the GPU box (the reference does not travel); the JSON
fixtures it writes next to itself are what the tests read.

Run:  PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python python tests/golden/make_golden.py

Shims: pure-python protobuf, numpy.float alias, a MagicMock
`tensorflow` module, and Network.Network's four TF-touching methods replaced by the
deterministic integer evaluator of Appendix B.2.  Nothing else of the reference is altered.
"""
import hashlib
import json
import os
import sys
import tempfile
from unittest.mock import MagicMock

os.environ.setdefault('PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION', 'python')
HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/src'
sys.path.insert(0, REF)

import numpy as np  # noqa: E402

np.float = float
sys.modules['tensorflow'] = MagicMock()
os.chdir(tempfile.mkdtemp(prefix='bbgold_'))       # the reference creates ./data/blackbird.db

import Network  # noqa: E402
import TicTacToe  # noqa: E402
import Connect4  # noqa: E402
import DragonChess  # noqa: E402
from FixedMCTS import FixedMCTS  # noqa: E402
from DynamicMCTS import DynamicMCTS  # noqa: E402

GAMES = {'TicTacToe': TicTacToe.BoardState, 'Connect4': Connect4.BoardState,
         'DragonChess': DragonChess.BoardState}


# --- Appendix B.2 integer evaluator, patched over the TF session wrapper ------------------
def _h(planes):
    x = planes.astype(np.int64).ravel()
    return int(sum((j + 1) * int(v) for j, v in enumerate(x)) % 1000003)


def _get_policy(self, planes):
    h = _h(planes)
    n = [1 + (h * (i + 1)) % 97 for i in range(self._nActions)]
    return np.array(n, dtype=np.float32) / np.float32(sum(n))


def _get_evaluation(self, planes):
    h = _h(planes)
    return np.float32((h % 1009) * 2 - 1008) / np.float32(1009)


def _net_init(self, name, networkConstructor=None, tensorflowConfig={}):
    self._nActions = networkConstructor.policyShape


Network.Network.__init__ = _net_init
Network.Network.__del__ = lambda self: None
Network.Network.getPolicy = _get_policy
Network.Network.getEvaluation = _get_evaluation

import Blackbird  # noqa: E402

NETCFG = {'policy': {'dirichlet': {'alpha': 0.2, 'epsilon': 0.0}}, 'hasTeacher': False}


def make_model(game, kind, playLimit, uncached=False):
    cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': playLimit}
    if kind == 'fixed':                                   # A.24: re-point the base class
        cls = type('Model', (FixedMCTS, Network.Network), dict(Blackbird.Model.__dict__))
        saved = Blackbird.MCTS
        Blackbird.MCTS = FixedMCTS
        try:
            m = cls(GAMES[game], 'gold', cfg, NETCFG, {})
        finally:
            Blackbird.MCTS = saved
    else:
        cls = type('Model', (DynamicMCTS, Network.Network), dict(Blackbird.Model.__dict__))
        m = cls(GAMES[game], 'gold', cfg, NETCFG, {})
    if uncached:                                          # A.21: DragonChess __eq__ aliases states
        m.GetPriors = lambda s: cls.GetPriors.__wrapped__(m, s)
        m.SampleValue = lambda s, p: cls.SampleValue.__wrapped__(m, s, p)
    return m


def wbits(v):
    return 0 if isinstance(v, int) else np.float32(v).tobytes().hex()


def wcode(w):
    """None -> -1, int 0 (draw via _isOver) -> 0, float 1.0/2.0 -> 1/2, float 0.0 -> 3 (A.5)."""
    if w is None:
        return -1
    if isinstance(w, (int,)) and not isinstance(w, np.generic):
        return int(w)
    return 3 if float(w) == 0.0 else int(w)


# --- 1. random legal playouts (Appendix B.1) -------------------------------------------------
def playouts():
    out = {}
    for game, seeds in (('TicTacToe', range(8)), ('Connect4', range(8)), ('DragonChess', range(3))):
        rows = []
        for seed in seeds:
            rs = np.random.RandomState(seed)
            s = GAMES[game]()
            trace = []
            while s.Winner() is None and len(trace) < 600:
                idx = np.flatnonzero(s.LegalActions() == 1)
                a = int(idx[rs.randint(idx.size)])
                trace.append((int(idx.size), a))
                s.ApplyAction(a)
            w = None if s.Winner() is None else int(s.Winner())
            planes = s.AsInputArray().tobytes().hex()
            digest = hashlib.sha256(repr((trace, w, planes)).encode()).hexdigest()[:16]
            rows.append({'seed': seed, 'plies': len(trace), 'winner': w,
                         'sum_legal': sum(t[0] for t in trace), 'digest': digest,
                         'trace': trace, 'planes': planes})
        out[game] = rows
    return out


# --- 2. synthetic positions: arbitrary bit patterns exercise scan order (A.4, A.5) ----------
def line_positions():
    out = {}
    for game, n in (('TicTacToe', 160), ('Connect4', 200)):
        rs = np.random.RandomState(7)
        cls = GAMES[game]
        rows = []
        for k in range(n):
            s = cls()
            H, W = s.Board.shape[:2]
            fill = rs.uniform(0.1, 1.0)
            cells = rs.choice(3, size=(H, W), p=[1 - fill, fill / 2, fill / 2])
            if game == 'Connect4' and k % 2 == 0:         # gravity-consistent half
                for j in range(W):
                    col = cells[:, j][cells[:, j] > 0]
                    cells[:, j] = 0
                    cells[:len(col), j] = col
            s.Board[:, :, 0] = cells == 1
            s.Board[:, :, 1] = cells == 2
            s.Player = int(rs.randint(1, 3))
            nA = cls.LegalMoves
            row = {'cells': cells.astype(int).ravel().tolist(), 'player': s.Player,
                   'legal': s.LegalActions().astype(int).tolist(),
                   'winner': wcode(s.Winner()),
                   'winner_prev': [wcode(s.Winner(a)) for a in range(nA)],
                   'planes': s.AsInputArray().tobytes().hex()}
            applied = []
            for a in range(nA):
                t = s.Copy()
                try:
                    t.ApplyAction(a)
                    applied.append(t.AsInputArray().tobytes().hex())
                except ValueError:
                    applied.append(None)
            row['applied'] = applied
            rows.append(row)
        out[game] = rows
    return out


def dragon_positions():
    """Positions along random playouts (every 9th ply) + mutated boards (random piece soup)."""
    rows = []
    cls = GAMES['DragonChess']

    def snap(s):
        legal = np.flatnonzero(s.LegalActions() == 1).astype(int).tolist()
        row = {'board': s.board.astype(int).ravel().tolist(), 'player': s.Player,
               'prev': s.PreviousPlayer,
               'castle': [int(s._white_castle_kingside), int(s._white_castle_queenside),
                          int(s._black_castle_kingside), int(s._black_castle_queenside)],
               'legal': legal, 'winner': wcode(s.Winner()),
               'planes_sha': hashlib.sha256(s.AsInputArray().tobytes()).hexdigest()[:16]}
        after = {}
        for a in legal[:6] + legal[-3:]:
            t = s.Copy()
            t.ApplyAction(a)
            after[str(a)] = {'board': t.board.astype(int).ravel().tolist(), 'player': t.Player,
                             'prev': t.PreviousPlayer,
                             'castle': [int(t._white_castle_kingside), int(t._white_castle_queenside),
                                        int(t._black_castle_kingside), int(t._black_castle_queenside)]}
        row['after'] = after
        return row

    for seed in range(2):
        rs = np.random.RandomState(100 + seed)
        s = cls()
        ply = 0
        while s.Winner() is None and ply < 400:
            if ply % 9 == 0:
                rows.append(snap(s))
            idx = np.flatnonzero(s.LegalActions() == 1)
            s.ApplyAction(int(idx[rs.randint(idx.size)]))
            ply += 1
        rows.append(snap(s))
    rs = np.random.RandomState(5)
    for k in range(24):                                    # piece soup: pawns on odd ranks, blocked sliders ...
        s = cls()
        s.board[:] = 0
        n = rs.randint(4, 28)
        sq = rs.choice(64, size=n, replace=False)
        s.board.ravel()[sq] = rs.choice([1, 2, 2, 2, 3, 4, 5, 6, -1, -2, -2, -2, -3, -4, -5, -6], size=n)
        s.Player = int(rs.randint(1, 3))
        s.PreviousPlayer = [None, 1, 2][rs.randint(3)]
        s._white_castle_kingside, s._white_castle_queenside = bool(rs.randint(2)), bool(rs.randint(2))
        s._black_castle_kingside, s._black_castle_queenside = bool(rs.randint(2)), bool(rs.randint(2))
        rows.append(snap(s))
    return rows


# --- 3. MCTS known answers through the real Blackbird.Model (Appendix B.2) ------------------
def mcts_kats():
    out = []
    cases = [('TicTacToe', 'dynamic', 200, 3), ('TicTacToe', 'fixed', 50, 3),
             ('Connect4', 'dynamic', 200, 3), ('Connect4', 'fixed', 50, 3),
             ('Connect4', 'dynamic', 800, 2), ('Connect4', 'fixed', 100, 4),
             ('DragonChess', 'dynamic', 60, 4), ('DragonChess', 'fixed', 12, 4)]
    for game, kind, P, moves in cases:
        m = make_model(game, kind, P, uncached=(game == 'DragonChess'))
        np.random.seed(0)
        s = GAMES[game]()
        rows = []
        for _ in range(moves):
            nxt, v, probs = m.FindMove(s, 1)
            plays = m.Root.ChildPlays()
            nz = np.flatnonzero(plays).astype(int).tolist()
            action = next(a for a in range(len(plays)) if m.Root.Children[a] is not None
                          and m.Root.Children[a].State == nxt)
            rows.append({'plays_idx': nz, 'plays': plays[nz].astype(int).tolist(), 'action': action,
                         'winrate': wbits(v), 'root_plays': int(m.Root.Plays),
                         'probs_sha': hashlib.sha256(probs.tobytes()).hexdigest()[:16]})
            s = nxt
            m.MoveRoot(s)
        out.append({'game': game, 'kind': kind, 'playLimit': P, 'seed': 0, 'moves': rows})
    return out


# --- 4. whole self-play games through GenerateTrainingSamples, records captured -------------
def record_games():
    out = []
    cases = [('Connect4', 'dynamic', 50, [0, 1, 2]), ('Connect4', 'fixed', 30, [0, 1]),
             ('TicTacToe', 'dynamic', 40, [0, 1]), ('TicTacToe', 'fixed', 10, [0, 1, 2])]
    for game, kind, P, seeds in cases:
        m = make_model(game, kind, P)
        captured = []
        m.Conn.PutGames = lambda name, version, gameType, games: captured.append((gameType, list(games)))
        for seed in seeds:
            np.random.seed(seed)
            Blackbird.GenerateTrainingSamples(m, 1, 1)
            gameType, recs = captured[-1]
            out.append({'game': game, 'kind': kind, 'playLimit': P, 'seed': seed, 'gameType': gameType,
                        'records': [r.hex() for r in recs]})
    return out


# --- 5. arena games against RandomMCTS through the unmodified Blackbird.TestRandom (8f-2) ----
def arena_games():
    import random
    out = []
    cases = [('TicTacToe', 'dynamic', 30, 0.1, 4, 0), ('TicTacToe', 'fixed', 12, 0.1, 3, 1),
             ('Connect4', 'dynamic', 40, 0.1, 3, 2), ('Connect4', 'fixed', 20, 1, 2, 3)]
    for game, kind, P, temp, n, seed in cases:
        m = make_model(game, kind, P)
        stats_rows, boards = [], []
        m.Conn.PutTrainingStatistic = lambda *a: stats_rows.append(list(a))
        move_root = m.MoveRoot

        def logged(state, _mr=move_root):
            boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                         + bytes([int(state.Player)])).hexdigest()[:12])
            return _mr(state)
        m.MoveRoot = logged
        random.seed(seed)
        np.random.seed(seed)
        stats = Blackbird.TestRandom(m, temp, n)
        out.append({'game': game, 'kind': kind, 'playLimit': P, 'temp': temp, 'numTests': n, 'seed': seed,
                    'stats': dict(stats), 'rows': stats_rows, 'boards': boards})
    return out


# --- 6. the base-class evaluator (all-ones priors, a random rollout per leaf visit): FindMove KATs and arena games
#        against the 'good' opponent of Blackbird.TestGood, with playLimit in place of its wall-clock limit (8f-2) ----
def vcode(v):
    """Root.Value / WinRate of the rollout evaluator: exact Python int / float -> repr."""
    return repr(float(v))


def rollout_kats():
    out = []
    cases = [('TicTacToe', 'fixed', 40, 4, 0), ('TicTacToe', 'dynamic', 60, 4, 1), ('Connect4', 'fixed', 30, 4, 2),
             ('Connect4', 'dynamic', 80, 4, 3), ('Connect4', 'fixed', 135, 3, 4), ('DragonChess', 'fixed', 6, 3, 5),
             ('DragonChess', 'dynamic', 10, 3, 6)]
    for game, kind, P, moves, seed in cases:
        cls = FixedMCTS if kind == 'fixed' else DynamicMCTS
        m = cls(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=P)
        np.random.seed(seed)
        s = GAMES[game]()
        rows = []
        for _ in range(moves):
            nxt, v, probs = m.FindMove(s, 1)
            plays = m.Root.ChildPlays()
            nz = np.flatnonzero(plays).astype(int).tolist()
            action = next(a for a in range(len(plays)) if m.Root.Children[a] is not None and m.Root.Children[a].State == nxt)
            rows.append({'plays_idx': nz, 'plays': plays[nz].astype(int).tolist(), 'action': action,
                         'child_values': [vcode(m.Root.Children[a].Value) for a in nz],
                         'winrate': vcode(v), 'root_plays': int(m.Root.Plays), 'root_value': vcode(m.Root.Value),
                         'rng_after': int(np.random.get_state()[2]),
                         'probs_sha': hashlib.sha256(probs.tobytes()).hexdigest()[:16]})
            s = nxt
            m.MoveRoot(s)
        out.append({'game': game, 'kind': kind, 'playLimit': P, 'seed': seed, 'moves': rows,
                    'next_double': float(np.random.random_sample())})
    return out


def arena_good_games():
    import random
    out = []
    cases = [('TicTacToe', 'dynamic', 30, 25, 0.1, 3, 0), ('Connect4', 'fixed', 20, 30, 0.1, 2, 1), ('Connect4', 'dynamic', 40, 60, 1, 2, 2)]
    for game, kind, P, Pgood, temp, n, seed in cases:
        m = make_model(game, kind, P)
        boards, results = [], []
        move_root = m.MoveRoot

        def logged(state, _mr=move_root):
            boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                         + bytes([int(state.Player)])).hexdigest()[:12])
            return _mr(state)
        m.MoveRoot = logged
        random.seed(seed)
        np.random.seed(seed)
        for _ in range(n):
            # Blackbird.TestGood (Blackbird.py:146-174) with the opponent's timeLimit=1 replaced by a play limit: the
            # wall-clock limit is the only non-deterministic input (A.25); everything else is the reference's own code
            good = FixedMCTS(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=Pgood)
            results.append(int(Blackbird.TestModels(m, good, temp, numTests=1)))
        out.append({'game': game, 'kind': kind, 'playLimit': P, 'goodPlayLimit': Pgood, 'temp': temp, 'numTests': n, 'seed': seed,
                    'results': results, 'boards': boards, 'next_double': float(np.random.random_sample())})
    return out


def main():
    def dump(name, obj):
        with open(os.path.join(HERE, name), 'w') as f:
            json.dump(obj, f, separators=(',', ':'))
        print(name, os.path.getsize(os.path.join(HERE, name)), 'bytes')

    dump('playouts.json', playouts())
    dump('line_positions.json', line_positions())
    dump('dragon_positions.json', dragon_positions())
    dump('mcts_kat.json', mcts_kats())
    dump('records.json', record_games())
    dump('arena.json', arena_games())
    dump('rollout_kat.json', rollout_kats())
    dump('arena_good.json', arena_good_games())


if __name__ == '__main__':
    main()
