"""GPU: built-in policy-value network forward vs the float64 restatement of NetworkFactory.py
(oracle/network_np.py; parity unpinned by the reference, TensorFlow is unavailable), and
search-with-network vs the oracle search fed the SAME (replayed) network outputs, bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
from oracle import network_np, search_np, selfplay_np  # noqa: E402
from oracle.rules_np import GAMES  # noqa: E402

REL_TOL = 1e-5      # north_star: "1e-5 rel in fp32"


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine, records
    return engine, records


def playout_states(game, n, seed):
    rs = np.random.RandomState(seed)
    out = []
    while len(out) < n:
        s = GAMES[game]()
        while s.Winner() is None and len(out) < n:
            out.append(s.Copy())
            idx = np.flatnonzero(s.LegalActions() == 1)
            if idx.size == 0:
                break
            s.ApplyAction(int(idx[rs.randint(idx.size)]))
            if game == 'DragonChess' and rs.rand() < 0.1:
                break
    return out


def pack(engine, game, states):
    if game == 'DragonChess':
        cells = np.stack([s.board.astype(np.int8).ravel() for s in states])
        castle = [sum(int(b) << i for i, b in enumerate(s.castle)) for s in states]
    else:
        cells = np.stack([(s.Board[:, :, 0] + 2 * s.Board[:, :, 1]).astype(np.int8).ravel() for s in states])
        castle = None
    return engine.pack(game, cells, [s.Player for s in states], [s.PreviousPlayer or 0 for s in states], castle)


@pytest.mark.parametrize('game,n', [('TicTacToe', 333), ('Connect4', 1001), ('DragonChess', 150)])
@pytest.mark.parametrize('jitter', [0.0, 0.2])
def test_forward_vs_fp64(mods, game, n, jitter):
    engine, _ = mods
    C = engine.plane_shape(game)[2]
    A = engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=4, filters=16, dense=16, seed=5, jitter=jitter)
    eng = engine.Engine(game, 'dynamic', n_games=1, evaluator='net')
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16)
    states = playout_states(game, n, seed=1)
    dev = pack(engine, game, states)
    pol, val = eng.net_forward(dev)
    planes = np.concatenate([s.AsInputArray() for s in states])
    want_p, want_v = network_np.forward(w, planes, dtype=np.float64)
    got_p, got_v = pol.cpu().numpy().astype(np.float64), val.cpu().numpy().astype(np.float64)
    assert np.abs(got_p.sum(axis=1) - 1).max() < 1e-6
    # 1e-5 relative; the 1e-8 absolute floor only matters for softmax tails below 1e-3 (random-init
    # heads give logits of magnitude ~20, where one fp32 ulp of the logit is already 2e-6 relative)
    assert (np.abs(got_p - want_p) <= REL_TOL * np.abs(want_p) + 1e-8).all(), \
        (np.abs(got_p - want_p) / np.maximum(np.abs(want_p), 1e-30)).max()
    # value = tanh(x): compare with tolerance relative to the [0,1] win-rate scale used downstream
    assert np.abs(got_v - want_v).max() < REL_TOL
    # run-to-run determinism (fixed reduction order)
    pol2, val2 = eng.net_forward(dev)
    assert torch.equal(pol, pol2) and torch.equal(val, val2)
    eng.close()


BF16_TOL = 2e-2     # north_star: "2e-2 in bf16" -- absolute, on probabilities (scale [0,1]) and on the value (scale [-1,1])
BF16_EMU_TOL = 2e-3  # against the CPU restatement of the kernel's own rounding points (accumulation order is the only difference)
TC_NP = {'TicTacToe': 32, 'Connect4': 8, 'DragonChess': 3}     # positions per batch of one pipeline (net_tc.cuh TcBatch)


def _check_bf16(tag, w, planes, got_p, got_v):
    """bf16 tcgen05 outputs vs (1) the fp64 graph at the stated tolerance and (2) the CPU restatement of the kernel's
    rounding points, which pins taps, BN folding, skip connections and head wiring an order of magnitude tighter."""
    got_p, got_v = got_p.astype(np.float64), got_v.astype(np.float64)
    want_p, want_v = network_np.forward(w, planes, dtype=np.float64)
    emu_p, emu_v = network_np.forward_bf16_emulated(w, planes)
    e64p, e64v = np.abs(got_p - want_p).max(), np.abs(got_v - want_v).max()
    big = want_p > 1e-3
    rel = (np.abs(got_p - want_p) / np.maximum(want_p, 1e-30))[big].max() if big.any() else 0.0
    eep, eev = np.abs(got_p - emu_p).max(), np.abs(got_v - emu_v).max()
    print('[bf16 parity] %s n=%d: vs fp64 oracle policy abs %.3e (rel %.3e where p>1e-3) value abs %.3e; vs bf16 restatement policy abs %.3e value abs %.3e'
          % (tag, planes.shape[0], e64p, rel, e64v, eep, eev))
    assert np.abs(got_p.sum(axis=1) - 1).max() < 1e-5
    assert e64p <= BF16_TOL and e64v <= BF16_TOL, (tag, e64p, e64v)
    assert eep <= BF16_EMU_TOL and eev <= BF16_EMU_TOL, (tag, eep, eev)
    return e64p, rel, e64v


@pytest.mark.parametrize('game,n_states', [('TicTacToe', 1000), ('Connect4', 1200), ('DragonChess', 1000)])
@pytest.mark.parametrize('jitter,seed', [(0.0, 0), (0.2, 4)])
def test_bf16_launch_mode_vs_fp64(mods, game, n_states, jitter, seed):
    """The benchmarked kernel, k_net_forward_tc, in launch mode (bb_net_forward, precision bf16) against the fp64
    oracle (NetworkFactory.py:22-183) on >= 1000 playout positions, at batch sizes that exercise a single position,
    partial last tiles (NP-1, NP+1 positions per pipeline batch) and batches that fill every pipeline of the grid."""
    engine, _ = mods
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=4, filters=16, dense=16, seed=seed, jitter=jitter)
    eng = engine.Engine(game, 'dynamic', n_games=1, evaluator='net')
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16, precision='bf16')
    states = playout_states(game, n_states, seed=1)
    dev_all = pack(engine, game, states)
    planes_all = np.concatenate([s.AsInputArray() for s in states])
    NP, pipes = TC_NP[game], 296
    worst = (0.0, 0.0, 0.0)
    for n in (1, NP - 1, NP, NP + 1, n_states, (NP - 1) * pipes + 3, NP * pipes, NP * pipes + NP + 1):
        idx = np.arange(n) % n_states                        # large batches repeat positions at other tile offsets
        if n > n_states and game == 'DragonChess':
            idx = idx % 200                                  # 4032-wide rows: keep the oracle's work bounded
        dev = dev_all[torch.as_tensor(idx, device=dev_all.device)]
        pol, val = eng.net_forward(dev)
        uniq = np.unique(idx)
        gp, gv = pol.cpu().numpy(), val.cpu().numpy()
        # the same position gives the same bits wherever it sits in a batch
        first = {int(i): k for k, i in reversed(list(enumerate(idx)))}
        firsts = np.array([first[int(i)] for i in idx])
        assert np.array_equal(gp, gp[firsts]) and np.array_equal(gv, gv[firsts])
        sel = np.array([first[int(i)] for i in uniq])
        r = _check_bf16('%s launch jitter=%g batch=%d' % (game, jitter, n), w, planes_all[uniq], gp[sel], gv[sel])
        worst = tuple(max(a, b) for a, b in zip(worst, r))
        pol2, val2 = eng.net_forward(dev)
        assert torch.equal(pol, pol2) and torch.equal(val, val2)
    print('[bf16 parity] %s jitter=%g WORST vs fp64: policy abs %.3e rel %.3e value abs %.3e' % ((game, jitter) + worst))
    eng.close()


WIDE_CASES = [('TicTacToe', 64, 2, 300), ('Connect4', 64, 2, 400), ('DragonChess', 64, 1, 60),
              ('TicTacToe', 256, 1, 200), ('Connect4', 256, 1, 150), ('DragonChess', 256, 1, 24),
              ('Connect4', 32, 2, 300), ('Connect4', 128, 1, 200), ('DragonChess', 32, 1, 40), ('TicTacToe', 128, 1, 150)]


@pytest.mark.parametrize('game,filters,blocks,n', WIDE_CASES + [('Connect4', 40, 1, 100)])
def test_other_tower_widths_fp32_vs_fp64(mods, game, filters, blocks, n):
    """NetworkFactory.py:46-103 takes `filters` from the YAML: the fp32 path takes any width (the layer-by-layer SIMT
    twin of net_wide.cuh), same 1e-5 bar as the template width."""
    engine, _ = mods
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=blocks, filters=filters, dense=16, seed=7, jitter=0.2)
    eng = engine.Engine(game, 'dynamic', n_games=1, evaluator='net')
    eng.load_network(network_np.flatten_weights(w), blocks, filters, 16)
    states = playout_states(game, n, seed=2)
    dev = pack(engine, game, states)
    pol, val = eng.net_forward(dev)
    planes = np.concatenate([s.AsInputArray() for s in states])
    want_p, want_v = network_np.forward(w, planes, dtype=np.float64)
    got_p, got_v = pol.cpu().numpy().astype(np.float64), val.cpu().numpy().astype(np.float64)
    assert np.abs(got_p.sum(axis=1) - 1).max() < 1e-6
    assert (np.abs(got_p - want_p) <= REL_TOL * np.abs(want_p) + 1e-8).all(), \
        (np.abs(got_p - want_p) / np.maximum(np.abs(want_p), 1e-30)).max()
    assert np.abs(got_v - want_v).max() < REL_TOL
    pol2, val2 = eng.net_forward(dev)
    assert torch.equal(pol, pol2) and torch.equal(val, val2)
    eng.close()


@pytest.mark.parametrize('game,filters,blocks,n', WIDE_CASES)
def test_other_tower_widths_bf16_vs_fp64(mods, game, filters, blocks, n):
    """k_wide_conv<64 / 256> (tcgen05, M128 N64 / N256 K16, weights streamed through a shared-memory ring, activations in
    HBM between layers) against the fp64 graph at 2e-2 and against the CPU restatement of its rounding points; batch sizes
    cover one position, partial tiles / groups and more groups than CTAs would take in one pass."""
    engine, _ = mods
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=blocks, filters=filters, dense=16, seed=3, jitter=0.2)
    eng = engine.Engine(game, 'dynamic', n_games=1, evaluator='net')
    eng.load_network(network_np.flatten_weights(w), blocks, filters, 16, precision='bf16')
    states = playout_states(game, n, seed=2)
    dev_all = pack(engine, game, states)
    planes_all = np.concatenate([s.AsInputArray() for s in states])
    ppt = {'TicTacToe': 8, 'Connect4': 2, 'DragonChess': 1}[game]
    tpg = 4 if filters <= 64 else 2
    for m in (1, ppt * tpg - 1, ppt * tpg + 1, n, ppt * tpg * 148 * 2 + 5, 4096 + 3 * ppt + 1):
        idx = np.arange(m) % n
        dev = dev_all[torch.as_tensor(idx, device=dev_all.device)]
        pol, val = eng.net_forward(dev)
        gp, gv = pol.cpu().numpy(), val.cpu().numpy()
        first = {int(i): k for k, i in reversed(list(enumerate(idx)))}
        firsts = np.array([first[int(i)] for i in idx])
        assert np.array_equal(gp, gp[firsts]) and np.array_equal(gv, gv[firsts])      # same position, same bits, wherever it sits
        uniq = np.unique(idx)
        sel = np.array([first[int(i)] for i in uniq])
        got_p, got_v = gp[sel].astype(np.float64), gv[sel].astype(np.float64)
        want_p, want_v = network_np.forward(w, planes_all[uniq], dtype=np.float64)
        emu_p, emu_v = network_np.forward_bf16_emulated(w, planes_all[uniq], last_bf16=True)
        e64p, e64v = np.abs(got_p - want_p).max(), np.abs(got_v - want_v).max()
        eep, eev = np.abs(got_p - emu_p).max(), np.abs(got_v - emu_v).max()
        print('[bf16 parity, %d filters] %s batch=%d: vs fp64 policy abs %.3e value abs %.3e; vs bf16 restatement policy abs %.3e value abs %.3e'
              % (filters, game, m, e64p, e64v, eep, eev))
        assert np.abs(got_p.sum(axis=1) - 1).max() < 1e-5
        assert e64p <= BF16_TOL and e64v <= BF16_TOL, (m, e64p, e64v)
        assert eep <= BF16_EMU_TOL and eev <= BF16_EMU_TOL, (m, eep, eev)
        pol2, val2 = eng.net_forward(dev)
        assert torch.equal(pol, pol2) and torch.equal(val, val2)
    eng.close()


def test_search_with_a_wide_network(mods):
    """A 64-filter network behind the search (tick mode: the persistent kernel pair exists for the template width only):
    the search runs to its play limit, the root has its seven children, and sixteen identical roots grow identical trees
    (the answers of the layer-by-layer kernels do not depend on where a position sits in a batch)."""
    engine, _ = mods
    game = 'Connect4'
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=1, filters=64, dense=16, seed=11, jitter=0.2)
    eng = engine.Engine(game, 'fixed', n_games=16, evaluator='net', max_depth=6)
    eng.load_network(network_np.flatten_weights(w), 1, 64, 16, precision='bf16')
    assert not eng.streaming
    eng.reset()
    eng.begin_move(60)
    eng.run()
    st = eng.root_stats(A)
    plays = st['plays'][:, :int(st['n_children'][0])]
    assert (st['n_children'] == 7).all() and int(plays[0].sum()) >= 58
    assert (plays == plays[0]).all()                      # sixteen identical roots, deterministic network: identical trees
    eng.close()


@pytest.mark.parametrize('game,kind,P,G', [('Connect4', 'fixed', 64, 2048), ('TicTacToe', 'dynamic', 40, 1024), ('DragonChess', 'dynamic', 8, 96)])
def test_bf16_serve_mode_vs_fp64(mods, game, kind, P, G):
    """The same kernel in SERVE mode (persistent, batches claimed from the request ring by k_search_stream's tickets):
    after a streaming self-play run every slot holds the last position it asked about and the answer it got
    (bb_stream_last_answers).  Those answers must equal launch mode bit for bit and the fp64 oracle within tolerance."""
    engine, _ = mods
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, seed=0)
    eng = engine.Engine(game, kind, n_games=G, evaluator='net', play_limit_hint=P, max_plies=24 if game == 'DragonChess' else None,
                        node_cap=16384 if game == 'DragonChess' else None)
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16, precision='bf16')
    assert eng.streaming
    u = np.stack([np.random.RandomState(500 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.selfplay(P, 1.0, u)
    st, pol, val, seq = eng.stream_last_answers()
    asked = seq > 0                      # a slot whose every evaluation came out of the evaluation cache never asked the network
    assert asked.sum() >= 0.5 * G
    st, pol, val = st[asked], pol[asked], val[asked]
    dev = torch.as_tensor(st).cuda()
    pol_l, val_l = eng.net_forward(dev)
    assert np.array_equal(val_l.cpu().numpy(), val)
    planes = engine.planes(game, dev).cpu().numpy()
    if A > 32:
        # wide heads answer the engine at the LEGAL actions of the asked position only (soft-max over those: the normaliser
        # cancels in GetPriors' division, Blackbird.py:386-387); bring both sides to that form: p * legal / sum(p * legal)
        legal = engine.legal(game, dev).cpu().numpy() > 0
        dense = pol_l.cpu().numpy().astype(np.float64) * legal
        dense /= dense.sum(axis=1, keepdims=True)
        got = np.where(legal, pol.astype(np.float64), 0.0)
        assert np.abs(got.sum(axis=1) - 1).max() < 1e-5 and np.abs(got - dense).max() < 1e-5
        _, first = np.unique(st, axis=0, return_index=True)
        first = first[:400]
        want_p, want_v = network_np.forward(w, planes[first], dtype=np.float64)
        want_p = want_p * legal[first]
        want_p /= want_p.sum(axis=1, keepdims=True)
        ep, ev = np.abs(got[first] - want_p).max(), np.abs(val[first] - want_v).max()
        print('[bf16 parity] %s serve mode, legal-only head (%d distinct positions): priors abs %.3e value abs %.3e vs fp64 oracle' % (game, len(first), ep, ev))
        assert ep <= BF16_TOL and ev <= BF16_TOL
        eng.close()
        return
    assert np.array_equal(pol_l.cpu().numpy(), pol)
    _, first = np.unique(st, axis=0, return_index=True)
    first = first[:400] if game == 'DragonChess' else first
    assert len(first) >= 20
    _check_bf16('%s serve mode (%d distinct positions of %d slots)' % (game, len(first), G), w, planes[first], pol[first], val[first])
    c = eng.counters()
    print('[eval cache] %s: %d lookups, %d hits, %d network evaluations' % (game, c['cache_lookups'], c['cache_hits'], c['evals']))
    eng.close()


@pytest.mark.parametrize('game,kind,P,G', [('Connect4', 'fixed', 100, 512), ('TicTacToe', 'dynamic', 60, 256)])
def test_eval_cache_is_transparent(mods, game, kind, P, G, monkeypatch):
    """The evaluation cache of streaming mode (the engine-wide form of the reference's lru_cache on GetPriors / SampleValue,
    Blackbird.py:350,372) must not change a single record with a pure evaluator: cache on == cache off, byte for byte, and
    it must actually be hit (transpositions inside a game, shared openings between games)."""
    engine, records = mods
    from blackbird_b200 import netinit
    w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=1)
    out = {}
    for lg in ('20', '0', '6'):                               # 1 M entries / off / 64 entries (constant eviction, lock contention)
        monkeypatch.setenv('BB_EVAL_CACHE', lg)
        eng = engine.Engine(game, kind, n_games=G, evaluator='net', play_limit_hint=P)
        eng.load_network(w, 4, 16, 16, precision='bf16')
        assert eng.streaming
        u = np.stack([np.random.RandomState(9000 + g).random_sample(eng.max_plies) for g in range(G)])
        eng.selfplay(P, 1.0, u)
        c = eng.counters()
        out[lg] = (records.build_records(eng), c['sims'], c['moves'])
        print('[eval cache] %s 2^%s entries: %d lookups, %d hits (%.1f %%), %d network evaluations for %d simulations'
              % (game, lg, c['cache_lookups'], c['cache_hits'], 100.0 * c['cache_hits'] / max(1, c['cache_lookups']), c['evals'], c['sims']))
        if lg == '20':
            assert c['cache_hits'] > 0.3 * c['cache_lookups'] and c['cache_hits'] + c['evals'] == c['cache_lookups']
        if lg == '0':
            assert c['cache_lookups'] == 0
        eng.close()
    assert out['20'] == out['0'] and out['6'] == out['0']


class Replay(object):
    """Evaluator for the oracle search that replays the GPU network's own outputs (looked up by
    input planes): isolates tree-search parity from network rounding."""

    def __init__(self, engine, eng, game):
        self.engine, self.eng, self.game, self.cache = engine, eng, game, {}

    def _eval(self, planes):
        key = planes.tobytes()
        if key not in self.cache:
            H, W, C = planes.shape[1:]
            cells = (planes[0, :, :, 0] + 2 * planes[0, :, :, 1]).astype(np.int8).reshape(1, -1)
            player = 1 if planes[0, 0, 0, 2] == 1 else 2
            dev = self.engine.pack(self.game, cells, [player], [0])
            pol, val = self.eng.net_forward(dev)
            self.cache[key] = (pol[0].cpu().numpy(), np.float32(val[0].item()))
        return self.cache[key]

    def getPolicy(self, planes):
        return self._eval(planes)[0]

    def getEvaluation(self, planes):
        return self._eval(planes)[1]


@pytest.mark.parametrize('kind,P,precision,filters', [('dynamic', 64, 'fp32', 16), ('fixed', 24, 'fp32', 16),
                                                      ('dynamic', 64, 'bf16', 16), ('fixed', 24, 'bf16', 16),
                                                      ('fixed', 24, 'bf16', 64), ('dynamic', 40, 'bf16', 256), ('fixed', 16, 'fp32', 40)])
def test_selfplay_with_network_replay(mods, kind, P, precision, filters):
    """Self-play driven by the built-in network vs the oracle search fed with the SAME kernel's outputs
    (replayed per position): visit counts, moves and record bytes must agree exactly.  fp32 runs in tick
    mode, bf16 in streaming mode (persistent search + network kernels, request ring) at the template width and in
    tick mode behind the layer-by-layer kernels at 64 / 256 filters."""
    engine, records = mods
    game, G = 'Connect4', 6
    blocks = 4 if filters == 16 else 1
    w = network_np.init_weights(3, 7, blocks=blocks, filters=filters, dense=16, seed=2, jitter=0.1)
    eng = engine.Engine(game, kind, n_games=G, evaluator='net', play_limit_hint=P)
    eng.load_network(network_np.flatten_weights(w), blocks, filters, 16, precision=precision)
    assert eng.streaming == (precision == 'bf16' and filters == 16)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.selfplay(P, 1.0, u)
    got = records.build_records(eng)
    replay = Replay(engine, eng, game)
    for g in range(G):
        s = search_np.NetSearch(replay, kind, maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        want, winner, _ = selfplay_np.play_game(s, GAMES[game], 1)
        assert got[g][0] == want, (kind, g)
    eng.close()


@pytest.mark.parametrize('game,kind,P,G', [('Connect4', 'fixed', 48, 300), ('TicTacToe', 'dynamic', 40, 200), ('DragonChess', 'dynamic', 10, 24)])
def test_streaming_equals_tick_mode(mods, game, kind, P, G, monkeypatch):
    """The persistent kernel pair and the tick loop produce byte-identical records and counters."""
    engine, records = mods
    from blackbird_b200 import netinit
    cin = engine.plane_shape(game)[2]
    w = netinit.random_weights(cin, engine.n_actions(game), seed=5)
    out = {}
    for mode in ('1', '0'):
        monkeypatch.setenv('BB_STREAM', mode)
        eng = engine.Engine(game, kind, n_games=G, evaluator='net', play_limit_hint=P, max_plies=40 if game == 'DragonChess' else None,
                            node_cap=16384 if game == 'DragonChess' else None)
        eng.load_network(w, 4, 16, 16, precision='bf16')
        assert eng.streaming == (mode == '1')
        u = np.stack([np.random.RandomState(7000 + g).random_sample(eng.max_plies) for g in range(G)])
        eng.selfplay(P, 1.0, u)
        c = eng.counters()
        # streaming mode answers part of the evaluations from its state-keyed cache: computed + cached = what tick mode computes
        out[mode] = (records.build_records(eng), c['sims'], c['evals'] + c['cache_hits'], c['nodes'], c['moves'])
        eng.close()
    assert out['1'][1:] == out['0'][1:]
    assert out['1'][0] == out['0'][0]


@pytest.mark.parametrize('game,kind,P,slots,rows', [('TicTacToe', 'fixed', 10, 64, 4096), ('TicTacToe', 'dynamic', 24, 33, 1500),
                                                   ('Connect4', 'dynamic', 32, 64, 1024)])
def test_few_slots_many_tickets_streaming(mods, game, kind, P, slots, rows, monkeypatch):
    """Ring hand-back regression: with a few dozen slots the open tickets are not contiguous (a pipeline serves a
    late ticket while the other slots have issued a ring's worth of newer ones).  Thousands of games on few slots
    must end (no watchdog), every engine run must give the same records, and they must equal tick mode's."""
    engine, records = mods
    from blackbird_b200 import netinit
    w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=3)
    u = np.stack([np.random.RandomState(3000 + g).random_sample(engine._DEFAULT_PLIES[game]) for g in range(rows)])
    got = []
    for mode, repeats in (('1', 3), ('0', 1)):
        monkeypatch.setenv('BB_STREAM', mode)
        eng = engine.Engine(game, kind, n_games=slots, evaluator='net', play_limit_hint=P)
        eng.load_network(w, 4, 16, 16, precision='bf16')
        assert eng.streaming == (mode == '1')
        for _ in range(repeats):
            eng.selfplay(P, 1.0, u, n_rows=rows)
            r = eng.records_raw()
            played = np.arange(r['chosen'].shape[1])[None, :] < r['n_moves'][:, None]
            got.append((r['n_moves'].tobytes(), r['winner'].tobytes(), np.where(played, r['chosen'], 0).tobytes()))
        eng.close()
    assert all(g == got[0] for g in got[1:])


def test_more_games_than_slots_streaming(mods):
    """bb_selfplay_games through the persistent kernel pair: 40 games on 12 slots, every game equal to the oracle
    replay of the same kernel's outputs."""
    engine, records = mods
    game, slots, rows, P = 'Connect4', 12, 40, 24
    w = network_np.init_weights(3, 7, seed=2, jitter=0.1)
    eng = engine.Engine(game, 'fixed', n_games=slots, evaluator='net', play_limit_hint=P)
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16, precision='bf16')
    assert eng.streaming
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(rows)])
    eng.selfplay(P, 1.0, u)
    got = records.build_records(eng)
    assert len(got) == rows
    replay = Replay(engine, eng, game)
    for g in range(rows):
        s = search_np.NetSearch(replay, 'fixed', maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        want, winner, _ = selfplay_np.play_game(s, GAMES[game], 1)
        assert got[g][0] == want, g
    eng.close()


def test_full_size_streaming(mods):
    """BASELINE configs[1] at full size through the benchmarked path: 4096 Connect4 games, FixedMCTS, 800
    simulations per move, bf16 network, streaming mode.  Size-independent properties over all games (every game
    ends, the simulations spent equal 800 per recorded move, records replay legally to the recorded winner) and,
    for two slots, the first three moves' visit counts and chosen actions against the oracle search replaying the
    same kernel's outputs."""
    engine, records = mods
    from blackbird_b200 import netinit
    game, G, P = 'Connect4', 4096, 800
    eng = engine.Engine(game, 'fixed', n_games=G, evaluator='net', play_limit_hint=P)
    eng.load_network(netinit.random_weights(3, 7, seed=0), 4, 16, 16, precision='bf16')
    assert eng.streaming
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(G)])
    eng.selfplay(P, 1.0, u)
    raw = eng.records_raw()
    c = eng.counters()
    assert (raw['winner'] >= 0).all() and (raw['n_moves'] >= 7).all() and (raw['n_moves'] <= 42).all()
    assert c['games_done'] == G and c['sims'] == int(raw['n_moves'].sum()) * P
    for g in range(0, G, 131):
        st = engine.init_states(game, 1)
        for r in range(int(raw['n_moves'][g])):
            lo, hi = raw['rec_off'][g, r], raw['rec_off'][g, r + 1]
            assert int(raw['pool_plays'][g, lo:hi].sum()) >= P           # a carried subtree adds to the 800
            a = int(raw['chosen'][g, r])
            assert raw['pool_plays'][g, lo:hi][raw['pool_actions'][g, lo:hi] == a][0] > 0
            assert np.array_equal(st.cpu().numpy()[0], raw['states'][g, r])
            assert int(engine.apply(game, st, torch.tensor([a])).item()) == 1
        assert int(engine.winner(game, st).item()) == int(raw['winner'][g])
    replay = Replay(engine, eng, game)
    for g in (0, G - 1):
        s = search_np.NetSearch(replay, 'fixed', maxDepth=10, explorationRate=0.85, playLimit=P,
                                rng=np.random.RandomState(1000 + g))
        _, _, trace = selfplay_np.play_game(s, GAMES[game], 1, max_plies=3)
        for r, (plays, action, _, _) in enumerate(trace):
            lo, hi = raw['rec_off'][g, r], raw['rec_off'][g, r + 1]
            got = np.zeros(7)
            got[raw['pool_actions'][g, lo:hi]] = raw['pool_plays'][g, lo:hi]
            assert np.array_equal(got, plays), (g, r, got, plays)
            assert int(raw['chosen'][g, r]) == action
    eng.close()


@pytest.mark.parametrize('alpha', [0.2, 0.5, 0.03, 0.9])
def test_beta_sampler_distribution(mods, alpha):
    """The kernels' noise sampler is Beta(alpha, 1 - alpha) (NetworkFactory.py:176-180: Dirichlet([alpha, 1 - alpha])
    sample, first component): Kolmogorov-Smirnov against the Beta cdf, moments, and independence of the stream from the seed."""
    engine, _ = mods
    from scipy import stats
    n = 200000
    x = engine.beta_noise(alpha, 1234, n).cpu().numpy().astype(np.float64)
    assert (x >= 0).all() and (x <= 1).all()
    # Kolmogorov-Smirnov distance on a grid that stays inside fp32's reach: Beta(0.03, 0.97) puts 7 % of its mass below
    # 1e-38 and Beta(0.9, 0.1) 10 % within 1e-10 of 1, where a float can only say 0 or 1
    grid = np.concatenate([np.logspace(-6, np.log10(0.5), 200), 1 - np.logspace(-6, np.log10(0.5), 200)])
    emp = np.searchsorted(np.sort(x), grid, side='right') / n
    d = np.abs(emp - stats.beta(alpha, 1 - alpha).cdf(grid)).max()
    assert d < 0.005, (alpha, d)                              # 1.36 / sqrt(n) = 0.003 at the 5 % level
    assert abs(x.mean() - alpha) < 0.004 and abs(x.var() - alpha * (1 - alpha) / 2) < 0.004
    y = engine.beta_noise(alpha, 99, n).cpu().numpy()
    assert abs(np.corrcoef(x, y)[0, 1]) < 0.01 and not np.array_equal(x, y)
    assert np.array_equal(x, engine.beta_noise(alpha, 1234, n).cpu().numpy().astype(np.float64))


@pytest.mark.parametrize('precision', ['fp32', 'bf16'])
def test_noise_mix_against_sampler(mods, precision):
    """policy = ((1 - eps) softmax + eps noise) / sum (NetworkFactory.py:181-182), in both network kernels: with the
    softmax known, ratios of outputs give back the noise draws up to the common normaliser; their distribution must be
    the sampler's."""
    engine, _ = mods
    from scipy import stats
    w = network_np.init_weights(3, 7, seed=0)
    eng = engine.Engine('Connect4', 'dynamic', n_games=1, evaluator='net')
    eps, alpha = 0.3, 0.2
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16, epsilon=eps, alpha=alpha, seed=7, precision=precision)
    n = 30000
    dev = engine.init_states('Connect4', n)
    pol, _ = eng.net_forward(dev)
    p = pol.cpu().numpy().astype(np.float64)
    assert np.abs(p.sum(axis=1) - 1).max() < 1e-5
    soft = network_np.forward(w, GAMES['Connect4']().AsInputArray(), dtype=np.float64)[0][0]
    # the normalisation removes one degree of freedom, so the draws cannot be solved for; compare the distribution of
    # every output component with the same mix computed on the host from the sampler's own draws (another seed)
    nz = engine.beta_noise(alpha, 555, n * 7).cpu().numpy().astype(np.float64).reshape(n, 7)
    m = (1 - eps) * soft + eps * nz
    want = m / m.sum(axis=1, keepdims=True)
    for a in range(7):
        d = stats.ks_2samp(p[:, a], want[:, a]).statistic
        assert d < (0.02 if precision == 'fp32' else 0.03), (a, d)     # 1.36 * sqrt(2 / n) = 0.011 at the 5 % level
    eng.close()


def test_noise_statistics(mods):
    """epsilon > 0: policy = (1-eps)*softmax + eps*Beta(alpha, 1-alpha) draws, renormalised (A.17).
    TF's RNG cannot be matched; check moments: E[Beta(a,1-a)] = a, Var = a(1-a)/2."""
    engine, _ = mods
    w = network_np.init_weights(3, 7, seed=2)
    eng = engine.Engine('Connect4', 'dynamic', n_games=1, evaluator='net')
    eps, alpha = 0.3, 0.2
    eng.load_network(network_np.flatten_weights(w), 4, 16, 16, epsilon=eps, alpha=alpha, seed=7)
    n = 20000
    dev = engine.init_states('Connect4', n)
    pol, _ = eng.net_forward(dev)
    soft = network_np.forward(w, GAMES['Connect4']().AsInputArray(), dtype=np.float64)[0][0]
    p = pol.cpu().numpy().astype(np.float64)
    assert np.abs(p.sum(axis=1) - 1).max() < 1e-5
    # un-normalise: p_i = m_i / sum(m), m_i = (1-eps) soft_i + eps n_i  => ratios against E[m]
    mean_m = (1 - eps) * soft + eps * alpha
    est = p.mean(axis=0)
    assert np.abs(est - mean_m / mean_m.sum()).max() < 0.01
    pol2, _ = eng.net_forward(dev)
    assert not torch.equal(pol, pol2)          # fresh draw on every evaluation
    eng.close()


def test_dragonchess_selfplay_behind_a_wide_network(mods):
    """DragonChess self-play with a 64-filter network (tick mode; k_wide_heads answers at the legal actions only, a warp per
    position): two runs give the same records, every recorded root has exactly the legal moves as children (in order), and the
    chosen move was visited."""
    engine, records = mods
    game, G, P = 'DragonChess', 8, 10
    C, A = engine.plane_shape(game)[2], engine.n_actions(game)
    w = network_np.init_weights(C, A, blocks=1, filters=64, dense=16, seed=13, jitter=0.2)
    runs = []
    for _ in range(2):
        eng = engine.Engine(game, 'dynamic', n_games=G, evaluator='net', play_limit_hint=P, max_plies=24, node_cap=16384)
        eng.load_network(network_np.flatten_weights(w), 1, 64, 16, precision='bf16')
        assert not eng.streaming
        u = np.stack([np.random.RandomState(2000 + g).random_sample(eng.max_plies) for g in range(G)])
        eng.selfplay(P, 1.0, u)
        runs.append((records.build_records(eng), eng.records_raw()))
        eng.close()
    assert runs[0][0] == runs[1][0]
    raw = runs[0][1]
    for g in range(G):
        s = GAMES[game]()
        for ply in range(int(raw['n_moves'][g])):
            lo, hi = int(raw['rec_off'][g, ply]), int(raw['rec_off'][g, ply + 1])
            legal = np.flatnonzero(s.LegalActions() == 1)
            assert np.array_equal(raw['pool_actions'][g, lo:hi], legal)
            a = int(raw['chosen'][g, ply])
            assert raw['pool_plays'][g, lo:hi][list(legal).index(a)] > 0
            s.ApplyAction(a)
