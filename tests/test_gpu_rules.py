"""GPU: game-rule kernels (through the C-ABI) vs golden vectors and vs the numpy oracle."""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
from oracle import rules_np  # noqa: E402
from oracle.rules_np import GAMES  # noqa: E402


@pytest.fixture(scope='module')
def E():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine
    return engine


def to_dev(E, game, states_np):
    return torch.as_tensor(np.ascontiguousarray(states_np)).cuda()


def pack_oracle(E, game, states):
    """oracle state objects -> packed device states"""
    if game == 'DragonChess':
        cells = np.stack([s.board.astype(np.int8).ravel() for s in states])
        castle = [sum(int(b) << i for i, b in enumerate(s.castle)) for s in states]
    else:
        cells = np.stack([(s.Board[:, :, 0] + 2 * s.Board[:, :, 1]).astype(np.int8).ravel() for s in states])
        castle = None
    return E.pack(game, cells, [s.Player for s in states], [s.PreviousPlayer or 0 for s in states], castle)


def check_against_oracle(E, game, states):
    A = GAMES[game].LegalMoves
    dev = pack_oracle(E, game, states)
    legal = E.legal(game, dev).cpu().numpy()
    win = E.winner(game, dev).cpu().numpy()
    planes = E.planes(game, dev).cpu().numpy()
    for i, s in enumerate(states):
        assert np.array_equal(legal[i], s.LegalActions()), (game, i)
        w = s.Winner()
        code = -1 if w is None else (w if isinstance(w, int) else (3 if float(w) == 0.0 else int(w)))
        assert win[i] == code, (game, i)
        assert np.array_equal(planes[i][None], s.AsInputArray()), (game, i)


@pytest.mark.parametrize('game', ['TicTacToe', 'Connect4', 'DragonChess'])
def test_playout_goldens(E, golden, game):
    """Replay the reference's random playouts on the GPU: legal counts, planes, winners, digests."""
    for row in golden('playouts.json')[game]:
        st = E.init_states(game, 1)
        trace = []
        for n_legal, a in row['trace']:
            assert int(E.winner(game, st).item()) == -1
            assert int(E.legal(game, st).sum().item()) == n_legal
            trace.append((n_legal, a))
            ok = E.apply(game, st, torch.tensor([a]))
            assert int(ok.item()) == 1
        w = int(E.winner(game, st).item())
        w = None if w == -1 else w
        planes = E.planes(game, st).cpu().numpy().tobytes().hex()
        assert planes == row['planes']
        digest = hashlib.sha256(repr((trace, w, planes)).encode()).hexdigest()[:16]
        assert digest == row['digest'] and w == row['winner']


@pytest.mark.parametrize('game', ['TicTacToe', 'Connect4'])
def test_line_position_goldens(E, golden, game):
    rows = golden('line_positions.json')[game]
    A = GAMES[game].LegalMoves
    cells = np.array([r['cells'] for r in rows], dtype=np.int8)
    players = [r['player'] for r in rows]
    dev = E.pack(game, cells, players, [0] * len(rows))
    n = len(rows)
    assert E.legal(game, dev).cpu().numpy().astype(int).tolist() == [r['legal'] for r in rows]
    assert E.winner(game, dev).cpu().tolist() == [r['winner'] for r in rows]
    got_planes = E.planes(game, dev).cpu().numpy()
    for i, r in enumerate(rows):
        assert got_planes[i].tobytes().hex() == r['planes']
    for a in range(A):
        prev = torch.full((n,), a, dtype=torch.int32)
        assert E.winner(game, dev, prev).cpu().tolist() == [r['winner_prev'][a] for r in rows], a
        nxt = dev.clone()
        ok = E.apply(game, nxt, prev).cpu().numpy()
        pl = E.planes(game, nxt).cpu().numpy()
        for i, r in enumerate(rows):
            assert bool(ok[i]) == (r['applied'][a] is not None)
            if ok[i]:
                assert pl[i].tobytes().hex() == r['applied'][a]
            else:
                assert torch.equal(nxt[i], dev[i])          # untouched on illegal moves


def test_dragon_position_goldens(E, golden):
    rows = golden('dragon_positions.json')
    cells = np.array([r['board'] for r in rows], dtype=np.int8)
    castle = [sum(b << i for i, b in enumerate(r['castle'])) for r in rows]
    dev = E.pack('DragonChess', cells, [r['player'] for r in rows], [r['prev'] or 0 for r in rows], castle)
    legal = E.legal('DragonChess', dev).cpu().numpy()
    win = E.winner('DragonChess', dev).cpu().tolist()
    planes = E.planes('DragonChess', dev).cpu().numpy()
    for i, r in enumerate(rows):
        assert np.flatnonzero(legal[i]).tolist() == r['legal'], i
        assert win[i] == r['winner']
        assert hashlib.sha256(planes[i].tobytes()).hexdigest()[:16] == r['planes_sha']
        for a, want in r['after'].items():
            one = dev[i:i + 1].clone()
            assert int(E.apply('DragonChess', one, torch.tensor([int(a)])).item()) == 1
            c, pl, pv, ca = E.unpack('DragonChess', one)
            assert c[0].cpu().tolist() == want['board']
            assert (int(pl.item()), int(pv.item()) or None) == (want['player'], want['prev'])
            assert [(int(ca.item()) >> k) & 1 for k in range(4)] == want['castle']
        # every illegal action id is rejected and leaves the state untouched
        bad = np.flatnonzero(legal[i] == 0)[:: max(1, 4032 // 40)]
        many = dev[i:i + 1].repeat(len(bad), 1).clone()
        ok = E.apply('DragonChess', many, torch.as_tensor(bad.astype(np.int32)))
        assert int(ok.sum().item()) == 0 and torch.equal(many, dev[i:i + 1].repeat(len(bad), 1))


@pytest.mark.parametrize('game,n_games,max_ply', [('TicTacToe', 200, 9), ('Connect4', 200, 42), ('DragonChess', 6, 60)])
def test_random_playouts_vs_oracle(E, game, n_games, max_ply):
    """Seeded random legal playouts driven by the oracle; the GPU state tracks it ply by ply."""
    rs = np.random.RandomState(11)
    states = [GAMES[game]() for _ in range(n_games)]
    dev = E.init_states(game, n_games)
    for ply in range(max_ply):
        check_against_oracle(E, game, states)
        acts = []
        for s in states:
            idx = np.flatnonzero(s.LegalActions() == 1)
            a = int(idx[rs.randint(idx.size)]) if (idx.size and s.Winner() is None) else -1
            acts.append(a)
        if all(a < 0 for a in acts):
            break
        live = [i for i, a in enumerate(acts) if a >= 0]
        sub = dev[live].clone()
        ok = E.apply(game, sub, torch.tensor([acts[i] for i in live]))
        assert int(ok.sum().item()) == len(live)
        dev[live] = sub
        for i in live:
            states[i].ApplyAction(acts[i])
        assert bool(E.equal(game, dev, pack_oracle(E, game, states)).all().item())


def test_equal_ignores_previous_player(E):
    a = E.pack('Connect4', np.zeros((1, 42), np.int8), [1], [0])
    b = E.pack('Connect4', np.zeros((1, 42), np.int8), [1], [2])
    c = E.pack('Connect4', np.zeros((1, 42), np.int8), [2], [1])
    assert int(E.equal('Connect4', a, b).item()) == 1 and int(E.equal('Connect4', a, c).item()) == 0


def test_kat_evaluator_matches_oracle(E):
    from oracle.search_np import KatEvaluator
    for game in ('TicTacToe', 'Connect4', 'DragonChess'):
        rs = np.random.RandomState(3)
        states = []
        s = GAMES[game]()
        for _ in range(12):
            states.append(s.Copy())
            idx = np.flatnonzero(s.LegalActions() == 1)
            if s.Winner() is not None or idx.size == 0:
                break
            s.ApplyAction(int(idx[rs.randint(idx.size)]))
        dev = pack_oracle(E, game, states)
        pol, val = E.kat_forward(game, dev)
        ev = KatEvaluator(GAMES[game].LegalMoves)
        for i, st in enumerate(states):
            assert np.array_equal(pol[i].cpu().numpy(), ev.getPolicy(st.AsInputArray()))
            assert val[i].item() == float(ev.getEvaluation(st.AsInputArray()))
