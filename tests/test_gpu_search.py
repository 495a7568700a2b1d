"""GPU: tree search and self-play (through the C-ABI) vs the reference's golden vectors and vs
the numpy oracle -- visit counts, chosen moves, win-rate bits and record bytes, bit-exact."""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
from oracle import search_np, selfplay_np  # noqa: E402
from oracle.rules_np import GAMES  # noqa: E402


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine, records
    return engine, records


def step_move(eng, play_limit, u):
    """FindMove + MoveRoot on every slot; returns per-slot dicts like the golden rows."""
    eng.begin_move(play_limit)
    eng.run()
    eng.check_status()
    st = eng.root_stats()
    acts = eng.choose(1.0, u)
    eng.move_root(acts)
    rows = []
    for g in range(eng.n_games):
        k = int(st['n_children'][g])
        plays = np.zeros(eng.A)
        plays[st['actions'][g, :k]] = st['plays'][g, :k]
        wr = (np.float32(st['root_value'][g]) / np.float32(st['root_plays'][g])).tobytes().hex()
        total = sum(plays)
        probs = plays / total if total > 0 else np.zeros(eng.A)
        nz = np.flatnonzero(plays)
        rows.append({'plays_idx': nz.tolist(), 'plays': plays[nz].astype(int).tolist(), 'action': int(acts[g]),
                     'winrate': wr, 'root_plays': int(st['root_plays'][g]),
                     'probs_sha': hashlib.sha256(probs.tobytes()).hexdigest()[:16]})
    return rows


def test_mcts_goldens(mods, golden):
    """Appendix B.2: reference Blackbird.Model + integer evaluator, np.random.seed(0), temp 1."""
    engine, _ = mods
    for case in golden('mcts_kat.json'):
        eng = engine.Engine(case['game'], case['kind'], n_games=1, evaluator='kat', play_limit_hint=case['playLimit'])
        u = np.random.RandomState(case['seed']).random_sample(len(case['moves']))
        for i, want in enumerate(case['moves']):
            got = step_move(eng, case['playLimit'], u[i:i + 1])[0]
            assert got == want, (case['game'], case['kind'], i, got, want)
        eng.close()


def test_tie_break_kat(mods):
    """Uniform priors + constant value 0.5 (SURVEY.md B.2 footnote): child plays
    [29,29,29,28,28,28,28] for Connect4 dynamic, 200 sims -- served by the external evaluator."""
    engine, _ = mods
    eng = engine.Engine('Connect4', 'dynamic', n_games=1, evaluator='external', play_limit_hint=200)
    eng.begin_move(200)
    while True:
        n = eng.run()
        if n == 0:
            break
        k, states, _ = eng.eval_pending()
        pol = torch.full((k, 7), 1.0 / 7, dtype=torch.float32, device='cuda')
        val = torch.zeros((k,), dtype=torch.float32, device='cuda')      # (0+1)*0.5 = 0.5
        eng.eval_submit(pol, val)
    st = eng.root_stats()
    assert st['plays'][0, :7].tolist() == [29, 29, 29, 28, 28, 28, 28]


@pytest.mark.parametrize('game,kind,P,n_games', [
    ('Connect4', 'dynamic', 120, 24), ('Connect4', 'fixed', 40, 24),
    ('TicTacToe', 'dynamic', 60, 32), ('TicTacToe', 'fixed', 25, 32),
    ('DragonChess', 'dynamic', 16, 2), ('DragonChess', 'fixed', 4, 2),
])
def test_selfplay_vs_oracle(mods, game, kind, P, n_games):
    """Whole games on the device (bb_selfplay_run) vs the oracle's GenerateTrainingSamples
    restatement, per-game RandomState(1000+g) move sampling: record bytes, winners, moves."""
    engine, records = mods
    cap = 12 if game == 'DragonChess' else None
    eng = engine.Engine(game, kind, n_games=n_games, evaluator='kat', play_limit_hint=P, max_plies=cap)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(n_games)])
    eng.selfplay(P, 1.0, u)
    got = records.build_records(eng)
    raw = eng.records_raw()
    packed = records.split_records(*eng.records_serialized())       # the same bytes from the device-side packer
    assert packed == [r[0] for r in got]
    tot_sims = 0
    for g in range(n_games):
        s = search_np.NetSearch(search_np.KatEvaluator(eng.A), kind, maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        want, winner, trace = selfplay_np.play_game(s, GAMES[game], 1, max_plies=cap)
        recs, wcode, n_moves = got[g]
        assert n_moves == len(trace)
        assert raw['chosen'][g, :n_moves].tolist() == [t[1] for t in trace], g
        assert wcode == (-1 if winner is None else int(winner))
        assert recs == want, (game, kind, g)
        tot_sims += s.Stats['sims']
    c = eng.counters()
    assert c['sims'] == tot_sims and c['games_done'] == n_games
    eng.close()


@pytest.mark.parametrize('game,kind,P,slots,rows', [('Connect4', 'fixed', 30, 6, 20), ('TicTacToe', 'dynamic', 40, 8, 29),
                                                   ('Connect4', 'dynamic', 50, 24, 24)])
def test_more_games_than_slots(mods, game, kind, P, slots, rows):
    """bb_selfplay_games: `rows` whole games on `slots` slots (a slot that ends a game starts the next unplayed
    one).  Every game's records must equal the oracle's for that game, whichever slot played it."""
    engine, records = mods
    eng = engine.Engine(game, kind, n_games=slots, evaluator='kat', play_limit_hint=P)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(rows)])
    eng.selfplay(P, 1.0, u)
    assert eng.rec_rows == rows
    got = records.build_records(eng)
    assert len(got) == rows
    for g in range(rows):
        s = search_np.NetSearch(search_np.KatEvaluator(eng.A), kind, maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        want, winner, trace = selfplay_np.play_game(s, GAMES[game], 1)
        assert got[g][0] == want, (game, kind, g)
        assert got[g][1] == int(winner)
    buf, off, sz = eng.records_serialized()
    assert records.split_records(buf, off, sz) == [r[0] for r in got]
    assert list(records.iter_records(buf, off, sz)) == [x for r in got for x in r[0]] and int(off[-1]) == len(buf)
    assert eng.counters()['games_done'] == rows
    # the engine is reusable afterwards, with fewer rows again
    eng.selfplay(P, 1.0, u[:slots])
    assert eng.rec_rows == slots and records.build_records(eng)[0][0] == got[0][0]
    eng.close()


def test_selfplay_record_goldens(mods, golden):
    """Records of whole games produced by the UNMODIFIED reference (GenerateTrainingSamples)."""
    engine, records = mods
    for case in golden('records.json'):
        eng = engine.Engine(case['game'], case['kind'], n_games=1, evaluator='kat', play_limit_hint=case['playLimit'])
        u = np.random.RandomState(case['seed']).random_sample(eng.max_plies)[None]
        eng.selfplay(case['playLimit'], 1.0, u)
        recs = records.build_records(eng)[0][0]
        assert [r.hex() for r in recs] == case['records'], (case['game'], case['kind'], case['seed'])
        packed = records.split_records(*eng.records_serialized())[0]      # k_pack_records: SerializeState on the device
        assert [r.hex() for r in packed] == case['records'], (case['game'], case['kind'], case['seed'])
        eng.close()


def test_full_size_properties(mods):
    """BASELINE config 2 shape at reduced sims (Connect4 fixed, 4096 games): size-independent
    properties -- every game ends, visit counts per record sum to the simulations spent on that
    move, records replay legally to the recorded terminal state, slot 0 matches the oracle."""
    engine, records = mods
    G, P = 4096, 16
    eng = engine.Engine('Connect4', 'fixed', n_games=G, evaluator='kat', play_limit_hint=P)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.selfplay(P, 1.0, u)
    raw = eng.records_raw()
    assert (raw['winner'] >= 0).all() and (raw['n_moves'] >= 7).all() and (raw['n_moves'] <= 42).all()
    c = eng.counters()
    assert c['sims'] == int(raw['n_moves'].sum()) * P and c['games_done'] == G
    for g in range(0, G, 97):
        st = engine.init_states('Connect4', 1)
        for r in range(int(raw['n_moves'][g])):
            lo, hi = raw['rec_off'][g, r], raw['rec_off'][g, r + 1]
            visits = int(raw['pool_plays'][g, lo:hi].sum())
            assert visits >= P
            a = int(raw['chosen'][g, r])
            assert raw['pool_plays'][g, lo:hi][raw['pool_actions'][g, lo:hi] == a][0] > 0
            assert np.array_equal(st.cpu().numpy()[0], raw['states'][g, r])
            assert int(engine.apply('Connect4', st, torch.tensor([a])).item()) == 1
        assert int(engine.winner('Connect4', st).item()) == int(raw['winner'][g])
    s = search_np.NetSearch(search_np.KatEvaluator(7), 'fixed', maxDepth=10, explorationRate=0.85, playLimit=P,
                            rng=np.random.RandomState(1000))
    want, _, _ = selfplay_np.play_game(s, GAMES['Connect4'], 1)
    assert records.build_records(eng, raw)[0][0] == want
    eng.close()


def test_steady_state_stepping(mods):
    """bb_selfplay_step (barrier per move round) and bb_selfplay_stream (no barrier): every
    simulation is accounted for, slots refill in place, and a later bb_selfplay_run still
    reproduces the oracle (no state leaks from the half-arena copies)."""
    engine, records = mods
    G, P = 64, 20
    eng = engine.Engine('Connect4', 'dynamic', n_games=G, evaluator='kat', play_limit_hint=P)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.reset()
    eng.step(3, P, 1.0, u)
    eng.check_status()
    c = eng.counters(reset=True)
    assert c['moves'] == 3 * G and c['sims'] == 3 * G * P
    eng.stream_moves(40 * G, P)
    eng.check_status()
    c = eng.counters(reset=True)
    assert c['moves'] >= 40 * G and c['games_done'] > 0
    assert c['moves'] * P <= c['sims'] <= (c['moves'] + G) * P
    eng.selfplay(P, 1.0, u)
    got = records.build_records(eng)
    for g in range(0, G, 7):
        s = search_np.NetSearch(search_np.KatEvaluator(7), 'dynamic', maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        want, _, _ = selfplay_np.play_game(s, GAMES['Connect4'], 1)
        assert got[g][0] == want, g
    eng.close()


def test_arena_overflow_is_loud(mods):
    """A half-arena that is too small must raise, never truncate the tree silently."""
    engine, _ = mods
    eng = engine.Engine('Connect4', 'fixed', n_games=2, evaluator='kat', node_cap=64, play_limit_hint=50)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(2)])
    with pytest.raises(MemoryError):
        eng.selfplay(50, 1.0, u)
    eng.close()


def test_lanes_do_not_change_results(mods):
    """Lanes (sub-populations of slots ticking on their own streams) are a scheduling device:
    records, winners and counters are the same for 1 and 3 lanes."""
    engine, records = mods
    G, P = 48, 24
    u = np.stack([np.random.RandomState(1000 + g).random_sample(42) for g in range(G)])
    outs = []
    for lanes in (1, 3):
        eng = engine.Engine('Connect4', 'fixed', n_games=G, evaluator='kat', play_limit_hint=P, lanes=lanes)
        assert eng.lanes == lanes
        eng.selfplay(P, 1.0, u)
        c = eng.counters()
        outs.append((records.build_records(eng), c['sims'], c['evals'], c['nodes'], c['moves']))
        eng.close()
    assert outs[0] == outs[1]


def test_stalemated_dragonchess_root_trips_the_reference_assert(mods):
    """A root with no legal move and no winner (the reference never detects stalemate: DragonChess.py:161-167 only looks for
    the kings): _selectAction's `assert root.LegalActions[choice] == 1` (MCTS.py:341) fails in the reference; here the slot
    reports BB_EASSERT -> AssertionError.  White: king in the corner behind three pawns that cannot move (pawns on the last
    two ranks have no move shapes); Black: a lone king far away."""
    engine, _ = mods
    from oracle import rules_np
    s = GAMES['DragonChess']()
    s.board[:] = 0
    s.board[7, 7] = rules_np.K
    s.board[6, 6] = s.board[6, 7] = s.board[7, 6] = rules_np.P
    s.board[0, 0] = -rules_np.K
    s.castle = [False, False, False, False]
    assert s.Winner() is None and int(s.LegalActions().sum()) == 0
    cells = s.board.astype(np.int8).ravel()[None, :]
    dev = engine.pack('DragonChess', cells, [1], [0], [0])
    assert int(engine.legal('DragonChess', dev).sum()) == 0
    for kind in ('fixed', 'dynamic'):
        eng = engine.Engine('DragonChess', kind, n_games=1, evaluator='kat', max_plies=8, node_cap=4096)
        eng.reset(dev)
        with pytest.raises(AssertionError):
            eng.step(1, 8, 1.0, np.full((1, eng.max_plies), 0.5))
            eng.check_status()
        eng.close()
