"""CPU: pin the numpy oracle against golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  No GPU, no CUDA calls."""
import hashlib

import numpy as np
import pytest

from oracle import rules_np, search_np, selfplay_np
from oracle.rules_np import GAMES


def wcode(w):
    if w is None:
        return -1
    if isinstance(w, int):
        return w
    return 3 if float(w) == 0.0 else int(w)


@pytest.mark.parametrize('game', ['TicTacToe', 'Connect4', 'DragonChess'])
def test_playout_digests(golden, game):
    for row in golden('playouts.json')[game]:
        rs = np.random.RandomState(row['seed'])
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
        assert digest == row['digest'], (game, row['seed'])
        assert w == row['winner'] and len(trace) == row['plies']


def line_state(game, row):
    s = GAMES[game]()
    H, W = s.Board.shape[:2]
    cells = np.array(row['cells']).reshape(H, W)
    s.Board[:, :, 0] = cells == 1
    s.Board[:, :, 1] = cells == 2
    s.Player = row['player']
    return s


@pytest.mark.parametrize('game', ['TicTacToe', 'Connect4'])
def test_line_positions(golden, game):
    for row in golden('line_positions.json')[game]:
        s = line_state(game, row)
        assert s.LegalActions().astype(int).tolist() == row['legal']
        assert wcode(s.Winner()) == row['winner']
        assert [wcode(s.Winner(a)) for a in range(s.LegalMoves)] == row['winner_prev']
        assert s.AsInputArray().tobytes().hex() == row['planes']
        for a, want in enumerate(row['applied']):
            t = s.Copy()
            if want is None:
                with pytest.raises(ValueError):
                    t.ApplyAction(a)
            else:
                t.ApplyAction(a)
                assert t.AsInputArray().tobytes().hex() == want


def dragon_state(row):
    s = rules_np.DragonChessState()
    s.board = np.array(row['board'], dtype=float).reshape(8, 8)
    s.Player = row['player']
    s.PreviousPlayer = row['prev']
    s.castle = [bool(v) for v in row['castle']]
    return s


def test_dragon_positions(golden):
    for row in golden('dragon_positions.json'):
        s = dragon_state(row)
        assert np.flatnonzero(s.LegalActions() == 1).tolist() == row['legal']
        assert wcode(s.Winner()) == row['winner']
        assert hashlib.sha256(s.AsInputArray().tobytes()).hexdigest()[:16] == row['planes_sha']
        for a, want in row['after'].items():
            t = s.Copy()
            t.ApplyAction(int(a))
            assert t.board.astype(int).ravel().tolist() == want['board']
            assert (t.Player, t.PreviousPlayer) == (want['player'], want['prev'])
            assert [int(v) for v in t.castle] == want['castle']


def wbits(v):
    return 0 if isinstance(v, int) else np.float32(v).tobytes().hex()


def test_mcts_kats(golden):
    for case in golden('mcts_kat.json'):
        if case['game'] == 'DragonChess' and case['kind'] == 'dynamic':
            continue                      # slow in pure python; covered on the GPU side + fixed case
        game = GAMES[case['game']]
        ev = search_np.KatEvaluator(game.LegalMoves)
        np.random.seed(case['seed'])
        m = search_np.NetSearch(ev, case['kind'], maxDepth=10, explorationRate=0.85,
                                playLimit=case['playLimit'])
        s = game()
        for mv in case['moves']:
            nxt, v, probs = m.FindMove(s, 1)
            plays = m.Root.ChildPlays()
            nz = np.flatnonzero(plays).tolist()
            assert nz == mv['plays_idx'] and plays[nz].astype(int).tolist() == mv['plays']
            assert m.LastAction == mv['action']
            assert wbits(v) == mv['winrate'] and m.Root.Plays == mv['root_plays']
            assert hashlib.sha256(probs.tobytes()).hexdigest()[:16] == mv['probs_sha']
            s = nxt
            m.MoveRoot(s)


def test_selfplay_records(golden):
    for case in golden('records.json'):
        game = GAMES[case['game']]
        ev = search_np.KatEvaluator(game.LegalMoves)
        m = search_np.NetSearch(ev, case['kind'], maxDepth=10, explorationRate=0.85,
                                playLimit=case['playLimit'])
        np.random.seed(case['seed'])
        recs, winner, _ = selfplay_np.play_game(m, game, 1)
        assert [r.hex() for r in recs] == case['records'], (case['game'], case['kind'], case['seed'])


def test_random_arena(golden):
    """oracle arena_random == the unmodified Blackbird.TestRandom, position for position (arena.json)."""
    import random
    for case in golden('arena.json'):
        game = GAMES[case['game']]
        ev = search_np.KatEvaluator(game.LegalMoves)
        m = search_np.NetSearch(ev, case['kind'], maxDepth=10, explorationRate=0.85, playLimit=case['playLimit'])
        boards = []

        def seen(state):
            boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                         + bytes([int(state.Player)])).hexdigest()[:12])
        random.seed(case['seed'])
        np.random.seed(case['seed'])
        stats, results = selfplay_np.arena_random(m, game, case['temp'], case['numTests'], seen)
        assert boards == case['boards'], (case['game'], case['kind'])
        assert results == [r[0] for r in case['rows']]
        assert stats == case['stats']


def vrepr(v):
    return repr(float(v))


def test_rollout_kats(golden):
    """oracle RolloutSearch (all-ones priors, a fresh random rollout per leaf visit, MCTS.py:346-383) == the unmodified
    FixedMCTS / DynamicMCTS base classes: visit counts, exact child values, chosen moves and the state of numpy's
    global MT19937 stream after every move (rollout_kat.json)."""
    for case in golden('rollout_kat.json'):
        if case['game'] == 'DragonChess':
            continue                      # minutes in pure python; the GPU tests cover these cases
        game = GAMES[case['game']]
        np.random.seed(case['seed'])
        m = search_np.RolloutSearch(case['kind'], maxDepth=10, explorationRate=0.85, playLimit=case['playLimit'])
        s = game()
        for mv in case['moves']:
            nxt, v, probs = m.FindMove(s, 1)
            plays = m.Root.ChildPlays()
            nz = np.flatnonzero(plays).tolist()
            assert nz == mv['plays_idx'] and plays[nz].astype(int).tolist() == mv['plays']
            assert [vrepr(m.Root.Children[a].Value) for a in nz] == mv['child_values']
            assert m.LastAction == mv['action'] and vrepr(v) == mv['winrate']
            assert m.Root.Plays == mv['root_plays'] and vrepr(m.Root.Value) == mv['root_value']
            assert int(np.random.get_state()[2]) == mv['rng_after']
            assert hashlib.sha256(probs.tobytes()).hexdigest()[:16] == mv['probs_sha']
            s = nxt
            m.MoveRoot(s)
        assert float(np.random.random_sample()) == case['next_double']


def test_good_arena(golden):
    """oracle arena against the rollout opponent == the unmodified Blackbird.TestModels(model, FixedMCTS(maxDepth=10,
    explorationRate=0.85, playLimit=...)) -- TestGood with a play limit in place of its wall-clock limit (arena_good.json)."""
    import random
    for case in golden('arena_good.json'):
        game = GAMES[case['game']]
        ev = search_np.KatEvaluator(game.LegalMoves)
        m = search_np.NetSearch(ev, case['kind'], maxDepth=10, explorationRate=0.85, playLimit=case['playLimit'])
        boards = []

        def seen(state):
            boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                         + bytes([int(state.Player)])).hexdigest()[:12])
        random.seed(case['seed'])
        np.random.seed(case['seed'])
        results = []
        for _ in range(case['numTests']):
            good = search_np.RolloutSearch('fixed', maxDepth=10, explorationRate=0.85, playLimit=case['goodPlayLimit'])
            results.append(selfplay_np.arena_game(m, good, game, case['temp'], seen))
        assert boards == case['boards'], (case['game'], case['kind'])
        assert results == case['results']
        assert float(np.random.random_sample()) == case['next_double']
