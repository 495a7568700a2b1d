"""GPU: virtual-loss mode (bb_config.contexts = W; DESIGN.md "x1") -- W simulations of a game in flight, visits counted on the
way down, values on the way up, contexts resumed in strict cyclic order.

W = 1 runs the virtual-loss code path with one context and must reproduce the exact mode bit for bit (a visit counted on the
way down is never read before its value arrives).  W > 1 is a different search by design (north_star (2), SURVEY.md 7.3-2:
reported separately): what is checked is that it is deterministic, that every move still gets exactly playLimit simulations,
that the records are well-formed games, and that trees stay consistent (visit counts add up)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
from oracle.rules_np import GAMES  # noqa: E402


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine, netinit, records
    return engine, netinit, records


def _play(mods, game, kind, P, G, contexts, max_plies=None, node_cap=None, seed=5):
    engine, netinit, records = mods
    w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=seed)
    eng = engine.Engine(game, kind, n_games=G, evaluator='net', play_limit_hint=P, max_plies=max_plies, node_cap=node_cap, contexts=contexts)
    eng.load_network(w, 4, 16, 16, precision='bf16')
    assert eng.streaming
    u = np.stack([np.random.RandomState(7000 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.selfplay(P, 1.0, u)
    c = eng.counters()
    raw = eng.records_raw()
    recs = records.build_records(eng)
    eng.close()
    return recs, c, raw


@pytest.mark.parametrize('game,kind,P,G', [('Connect4', 'fixed', 48, 300), ('Connect4', 'dynamic', 64, 200), ('TicTacToe', 'dynamic', 40, 200),
                                           ('DragonChess', 'dynamic', 10, 24)])
def test_one_context_equals_exact_mode(mods, game, kind, P, G):
    kw = dict(max_plies=40, node_cap=16384) if game == 'DragonChess' else {}
    exact, c0, _ = _play(mods, game, kind, P, G, 0, **kw)
    one, c1, _ = _play(mods, game, kind, P, G, 1, **kw)
    assert one == exact
    assert (c1['sims'], c1['nodes'], c1['moves'], c1['evals'] + c1['cache_hits']) == (c0['sims'], c0['nodes'], c0['moves'], c0['evals'] + c0['cache_hits'])


@pytest.mark.parametrize('game,kind,P,G,W', [('Connect4', 'fixed', 64, 256, 4), ('Connect4', 'dynamic', 100, 256, 8), ('TicTacToe', 'fixed', 30, 128, 3),
                                             ('DragonChess', 'dynamic', 16, 24, 4)])
def test_many_contexts_are_deterministic_and_well_formed(mods, game, kind, P, G, W):
    engine, _, _ = mods
    kw = dict(max_plies=40, node_cap=16384) if game == 'DragonChess' else {}
    a, ca, raw = _play(mods, game, kind, P, G, W, **kw)
    b, cb, _ = _play(mods, game, kind, P, G, W, **kw)
    assert a == b and ca['sims'] == cb['sims'] and ca['nodes'] == cb['nodes']        # no dependence on when answers arrive
    n_moves, rec_off, plays, acts, chosen = raw['n_moves'], raw['rec_off'], raw['pool_plays'], raw['pool_actions'], raw['chosen']
    assert ca['sims'] == int(n_moves.sum()) * P                                      # every move got exactly playLimit simulations
    for g in range(G):
        s = GAMES[game]()
        for ply in range(int(n_moves[g])):
            lo, hi = int(rec_off[g, ply]), int(rec_off[g, ply + 1])
            legal = np.flatnonzero(s.LegalActions() == 1)
            assert np.array_equal(acts[g, lo:hi], legal)                            # the root's children are the legal moves, in order
            visits = plays[g, lo:hi].astype(np.int64)
            # a fresh root's children hold playLimit - 1 visits (the first simulation expands the root); a carried root's more
            assert visits.sum() >= P - 1 and visits[list(legal).index(int(chosen[g, ply]))] > 0
            s.ApplyAction(int(chosen[g, ply]))
        if game != 'DragonChess':
            assert s.Winner() is not None


def test_exactly_play_limit_simulations_when_contexts_exceed_what_is_left(mods):
    """playLimit smaller than, and not a multiple of, the number of contexts."""
    _, c, raw = _play(mods, 'Connect4', 'dynamic', 5, 64, 8)
    assert c['sims'] == int(raw['n_moves'].sum()) * 5
    _, c, raw = _play(mods, 'TicTacToe', 'fixed', 7, 64, 3)
    assert c['sims'] == int(raw['n_moves'].sum()) * 7


def test_contexts_need_the_streaming_kernels(mods):
    engine, netinit, _ = mods
    with pytest.raises(ValueError):
        engine.Engine('Connect4', 'fixed', n_games=4, evaluator='kat', contexts=2)
    eng = engine.Engine('Connect4', 'fixed', n_games=4, evaluator='net', contexts=2)
    eng.load_network(netinit.random_weights(3, 7, seed=1), 4, 16, 16, precision='fp32')      # fp32 network: tick mode only
    with pytest.raises(ValueError):
        eng.selfplay(10, 1.0, np.random.RandomState(0).random_sample((4, eng.max_plies)))
    eng.close()
