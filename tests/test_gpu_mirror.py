"""GPU: the reference-facing Python surface (blackbird_b200.MCTS / FixedMCTS / DynamicMCTS /
Blackbird / DataManager) used the way main.py and Blackbird.py use the reference, checked against
golden vectors produced by the unmodified reference (tests/golden/make_golden.py)."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def bb():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    import blackbird_b200.Blackbird as Blackbird
    from blackbird_b200 import Connect4, DragonChess, TicTacToe
    from blackbird_b200.DynamicMCTS import DynamicMCTS
    from blackbird_b200.FixedMCTS import FixedMCTS
    games = {'TicTacToe': TicTacToe.BoardState, 'Connect4': Connect4.BoardState, 'DragonChess': DragonChess.BoardState}
    return dict(Blackbird=Blackbird, games=games, DynamicMCTS=DynamicMCTS, FixedMCTS=FixedMCTS)


# Appendix B.2 integer evaluator, written against the reference's Network interface (planes in,
# f32 policy / value out) exactly as tests/golden/make_golden.py patches it over the TF wrapper.
def _h(planes):
    x = planes.astype(np.int64).ravel()
    return int(sum((j + 1) * int(v) for j, v in enumerate(x)) % 1000003)


class KatNet(object):
    def getPolicy(self, planes):
        h = _h(planes)
        n = [1 + (h * (i + 1)) % 97 for i in range(self._nActions)]
        return np.array(n, dtype=np.float32) / np.float32(sum(n))

    def getEvaluation(self, planes):
        h = _h(planes)
        return np.float32((h % 1009) * 2 - 1008) / np.float32(1009)


def wbits(v):
    return 0 if isinstance(v, int) else np.float32(v).tobytes().hex()


@pytest.mark.parametrize('index', [0, 1, 2, 3, 5, 6, 7])
def test_findmove_moveroot_vs_reference(bb, golden, index):
    """np.random.seed(0); FindMove(state, 1); MoveRoot(next) ... on a search object whose evaluator
    is plain Python -- visit counts, chosen action, Root.WinRate() bits, probabilities."""
    case = golden('mcts_kat.json')[index]
    base = bb['DynamicMCTS'] if case['kind'] == 'dynamic' else bb['FixedMCTS']
    cls = type('KatSearch', (base, KatNet), {})
    m = cls(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=case['playLimit'])
    state_cls = bb['games'][case['game']]
    m._nActions = state_cls.LegalMoves
    np.random.seed(case['seed'])
    s = state_cls()
    for i, want in enumerate(case['moves']):
        nxt, v, probs = m.FindMove(s, 1)
        plays = m.Root.ChildPlays()
        nz = np.flatnonzero(plays).astype(int).tolist()
        action = next(a for a in range(len(plays)) if m.Root.Children[a] is not None and m._last_action == a)
        got = {'plays_idx': nz, 'plays': plays[nz].astype(int).tolist(), 'action': action, 'winrate': wbits(v),
               'root_plays': int(m.Root.Plays), 'probs_sha': hashlib.sha256(probs.tobytes()).hexdigest()[:16]}
        assert got == want, (case['game'], case['kind'], i, got, want)
        applied = s.Copy()
        applied.ApplyAction(action)
        assert applied == nxt
        s = nxt
        m.MoveRoot(s)
        assert m.Root is None or m.Root.State == s


def test_api_errors(bb):
    """Argument checks of MCTS.FindMove / FixedMCTS.__init__ (MCTS.py:170-182, FixedMCTS.py:15-16)."""
    with pytest.raises(ValueError, match='MaxDepth for MCTS must be > 0.'):
        bb['FixedMCTS'](maxDepth=0, explorationRate=0.85)
    m = type('S', (bb['DynamicMCTS'], KatNet), {})(explorationRate=0.85)
    m._nActions = 7
    with pytest.raises(TypeError, match='State not of type GameState'):
        m.FindMove(object())
    s = bb['games']['Connect4']()
    with pytest.raises(ValueError, match='Not enough information to decide a stop time.'):
        m.FindMove(s)
    nxt, v, probs = m.FindMove(s, 1, playLimit=20)
    assert probs.shape == (7,) and abs(probs.sum() - 1) < 1e-12 and m.Root.Plays == 20
    other = s.Copy()
    other.ApplyAction(3)
    with pytest.raises(AssertionError):                     # Root primed for another state (MCTS.py:193)
        m.FindMove(other, 1, playLimit=20)
    m.DropRoot()
    assert m.Root is None
    m.MoveRoot(other)                                       # no-op on a dropped root (MCTS.py:274-275)
    m.FindMove(other, 1, playLimit=20)
    # wall-clock mode (A.25): runs, result not deterministic
    m.DropRoot()
    nxt, v, probs = m.FindMove(s, 1, moveTime=0.05)
    assert m.Root.Plays >= 8


def test_generate_training_samples_records(bb, golden, tmp_path, monkeypatch):
    """Blackbird.GenerateTrainingSamples(model, 1, 1) -> Conn.PutGames -> sqlite -> GetGames: the bytes the
    unmodified reference wrote for the same seeds (records.json), through the mirror's public API."""
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    for case in golden('records.json'):
        base = B.Model if case['kind'] == 'dynamic' else B.FixedModel
        cls = type('KatModel', (base,), {'_SELFPLAY_EVALUATOR': 'kat'})
        cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': case['playLimit'],
               'temperature': {'exploration': 1, 'exploitation': 0.1}}
        net = {'blocks': 1, 'filters': 16, 'eval': {'dense': 16}, 'policy': {'dirichlet': {'alpha': 0.2, 'epsilon': 0.0}}}
        name = 'gold_%s_%s_%d' % (case['game'], case['kind'], case['seed'])
        m = cls(bb['games'][case['game']], name, cfg, net, {})
        np.random.seed(case['seed'])
        n = B.GenerateTrainingSamples(m, 1, 1)
        got = m.Conn.GetGames(m.Name, m.Version)
        assert n == len(got)
        assert [bytes(r).hex() for r in got] == case['records'], (case['game'], case['kind'], case['seed'])
        ex = B.ExampleState.FromSerialized(got[0])
        assert ex.Board.shape == tuple(m.Game().AsInputArray().shape) and ex.MctsPolicy.shape == (m.Game.LegalMoves,)
        assert B.ExampleState(ex.MctsEval[0], ex.MctsPolicy, ex.Board).SerializeState() == bytes(got[0])
        m.Conn.Close()
    assert os.path.isfile(tmp_path / 'data' / 'blackbird.db')


def test_generate_training_samples_more_games_than_slots(bb, tmp_path, monkeypatch):
    """nGames larger than the slot count the memory budget allows: one engine call, slots restart with the next
    unplayed game; the stored records are the same as with one slot per game."""
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    cls = type('KatModel', (B.FixedModel,), {'_SELFPLAY_EVALUATOR': 'kat'})
    cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': 20,
           'temperature': {'exploration': 1, 'exploitation': 0.1}}
    net = {'blocks': 1, 'filters': 16, 'eval': {'dense': 16}, 'policy': {'dirichlet': {'alpha': 0.2, 'epsilon': 0.0}}}
    u = np.random.RandomState(7).random_sample((11, 42))
    out = []
    for tag, slots in (('wide', None), ('narrow', 3)):
        if slots:
            monkeypatch.setattr(B, '_batch_slots', lambda *a, **k: slots)
        m = cls(bb['games']['Connect4'], 'rows_' + tag, cfg, net, {})
        n = B.GenerateTrainingSamples(m, 11, 1, uniforms=u)
        got = [bytes(r) for r in m.Conn.GetGames(m.Name, m.Version)]
        assert n == len(got)
        out.append(got)
        m.Conn.Close()
    assert out[0] == out[1]


def test_model_network_selfplay_and_arena(bb, tmp_path, monkeypatch):
    """main.py's sequence with the template config on TicTacToe (BASELINE configs[0]): Model(...),
    GenerateTrainingSamples with the built-in network, then a TestPrevious-style arena game."""
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': 10,
           'temperature': {'exploration': 1, 'exploitation': 0.1}}
    net = {'hasTeacher': False, 'blocks': 4, 'filters': 16, 'eval': {'dense': 16},
           'policy': {'dirichlet': {'epsilon': 0.3, 'alpha': 0.2}}}
    model = B.FixedModel(bb['games']['TicTacToe'], 'Blackbird', cfg, net, {'GPUOptions': {'per_process_gpu_memory_fraction': 0.25}})
    assert model.Version == 1
    n = B.GenerateTrainingSamples(model, 64, cfg['temperature']['exploration'])
    recs = model.Conn.GetGames(model.Name, model.Version)
    assert n == len(recs) and 64 * 6 <= n <= 64 * 10
    ex = B.ExampleState.FromSerialized(recs[0])
    assert ex.Board.shape == (1, 3, 3, 3) and ex.MctsPolicy.shape == (9,)
    assert abs(ex.MctsPolicy.sum() - 1) < 1e-12
    # network hooks answer in the reference's shapes
    s = model.Game()
    p, v = model.getPolicy(s.AsInputArray()), model.getEvaluation(s.AsInputArray())
    assert p.shape == (9,) and p.dtype == np.float32 and -1 <= float(v) <= 1
    pri = model.GetPriors(s)
    assert abs(pri.sum() - 1) < 1e-12
    assert 0 <= float(model.SampleValue(s, 1)) <= 1
    # arena: two engines follow one game
    res = B.TestModels(model, model.LastVersion(), 0.1, 1)
    assert res in (1, 0, -1)
    model.Conn.Close()


def test_train_with_examples_mirror(bb, tmp_path, monkeypatch):
    """main.py:26-31: GenerateTrainingSamples then TrainWithExamples(model, batchSize, learningRate) with the template's
    training block (adam, lr 1e-4, batch 50): the mirror trains on exactly the batches the reference's loop would form
    (same permutation from numpy's global stream), the weights match the oracle trainer fed the same batches, Version
    is bumped and registered, and the search afterwards uses the trained weights."""
    from oracle import network_np, train_np
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': 10,
           'temperature': {'exploration': 1, 'exploitation': 0.1}}
    net = {'hasTeacher': False, 'blocks': 2, 'filters': 16, 'eval': {'dense': 16},
           'policy': {'dirichlet': {'epsilon': 0.0, 'alpha': 0.2}}, 'loss': {'L2_norm': 0.001},
           'training': {'optimizer': 'adam', 'learning_rate': 1e-3, 'momentum': 0.9, 'batch_size': 50}}
    model = B.FixedModel(bb['games']['TicTacToe'], 'Trainee', cfg, net, {})
    B.GenerateTrainingSamples(model, 40, 1)
    w0 = model.getWeights().copy()
    recs = model.Conn.GetGames(model.Name, model.Version)
    np.random.seed(5)
    B.TrainWithExamples(model, batchSize=50, learningRate=1e-3)
    assert model.Version == 2 and model.Conn.GetLastVersion('TicTacToe', 'Trainee') == 2
    assert model.batchCount == len(recs) // 50 >= 4
    w1 = model.getWeights()
    assert not np.array_equal(w0, w1)
    # the oracle trainer on the batches the reference's loop forms
    np.random.seed(5)
    order = np.random.choice(len(recs), len(recs) - len(recs) % 50, replace=False)
    exs = [B.ExampleState.FromSerialized(r) for r in recs]
    tmpl = network_np.init_weights(3, 9, blocks=2, filters=16, dense=16, seed=0)
    cur, o = {}, 0
    for k, v in tmpl.items():
        cur[k] = np.asarray(w0[o:o + v.size], np.float64).reshape(v.shape)
        o += v.size
    opt = train_np.Optimizer('adam')
    first_loss = None
    for i in range(len(order) // 50):
        b = [exs[j] for j in order[i * 50:(i + 1) * 50]]
        cur, losses = train_np.train_step(cur, opt, np.vstack([e.Board for e in b]), np.hstack([e.MctsEval for e in b]),
                                          np.vstack([e.MctsPolicy for e in b]).astype(np.float32), 1e-3)
        first_loss = first_loss or losses
    want = network_np.flatten_weights(cur)
    moved = np.abs(want - w0) > 0
    # Adam: |update| <= ~lr per step whatever the gradient, so compare against the distance travelled
    assert np.abs(w1 - want).max() < 0.05 * np.abs(want - w0).max() + 1e-6
    assert np.array_equal(w1[~moved], w0[~moved])                   # frozen BN statistics
    # the model keeps working with the new weights (search reloads them)
    s = model.Game()
    nxt, v, probs = model.FindMove(s, 1, playLimit=10)
    assert abs(probs.sum() - 1) < 1e-9
    with pytest.raises(NotImplementedError):
        model.train(s.AsInputArray(), np.zeros(1), np.zeros((1, 9)), 0.01, teacher=model)
    model.Conn.Close()


def test_main_sequence_with_a_64_filter_network(bb, tmp_path, monkeypatch):
    """`filters` comes from the YAML (NetworkFactory.py:46-103): main.py's sequence with a 64-filter network on Connect4 --
    self-play behind the layer-by-layer tensor-core kernels (tick mode), records in the database, one TrainWithExamples
    pass with the width-generic trainer kernels, and a search with the trained weights."""
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    cfg = {'maxDepth': 6, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': 16,
           'temperature': {'exploration': 1, 'exploitation': 0.1}}
    net = {'hasTeacher': False, 'blocks': 1, 'filters': 64, 'eval': {'dense': 16},
           'policy': {'dirichlet': {'epsilon': 0.3, 'alpha': 0.2}}, 'loss': {'L2_norm': 0.001},
           'training': {'optimizer': 'adam', 'learning_rate': 1e-3, 'momentum': 0.9, 'batch_size': 50}}
    model = B.FixedModel(bb['games']['Connect4'], 'Wide', cfg, net, {})
    n = B.GenerateTrainingSamples(model, 96, 1)
    recs = model.Conn.GetGames(model.Name, model.Version)
    assert n == len(recs) and n >= 96 * 8
    ex = B.ExampleState.FromSerialized(recs[0])
    assert ex.Board.shape == (1, 6, 7, 3) and abs(ex.MctsPolicy.sum() - 1) < 1e-12
    w0 = model.getWeights().copy()
    B.TrainWithExamples(model, batchSize=50, learningRate=1e-3)
    assert model.Version == 2 and not np.array_equal(w0, model.getWeights())
    assert all(np.isfinite(x) for x in model.lastLosses)
    s = model.Game()
    nxt, v, probs = model.FindMove(s, 1, playLimit=16)
    assert abs(probs.sum() - 1) < 1e-9
    model.Conn.Close()


@pytest.mark.parametrize('index', [0, 1, 2, 3])
def test_random_arena_vs_reference(bb, golden, index):
    """Blackbird.TestRandom (Blackbird.py:84-112) against the RandomMCTS mirror: with random.seed / np.random.seed
    as in tests/golden/make_golden.py:arena_games the games replay position for position (every state both
    players' MoveRoot saw), and the statistics rows and the win/draw/loss tally match the unmodified reference."""
    import random
    case = golden('arena.json')[index]
    B = bb['Blackbird']
    base = bb['DynamicMCTS'] if case['kind'] == 'dynamic' else bb['FixedMCTS']
    cls = type('KatSearch', (base, KatNet), {})
    m = cls(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=case['playLimit'])
    m.Game = bb['games'][case['game']]
    m._nActions = m.Game.LegalMoves
    m.Name, m.Version = 'gold', 1
    rows, boards = [], []

    class Conn(object):
        def PutTrainingStatistic(self, *a):
            rows.append(list(a))
    m.Conn = Conn()
    move_root = m.MoveRoot

    def logged(state):
        boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                     + bytes([int(state.Player)])).hexdigest()[:12])
        return move_root(state)
    m.MoveRoot = logged
    random.seed(case['seed'])
    np.random.seed(case['seed'])
    stats = B.TestRandom(m, case['temp'], case['numTests'])
    assert boards == case['boards'], (case['game'], case['kind'])
    assert rows == case['rows']
    assert dict(stats) == case['stats']


@pytest.mark.parametrize('index', [0, 1, 2, 3, 4, 5, 6])
def test_rollout_search_vs_reference(bb, golden, index):
    """Plain FixedMCTS / DynamicMCTS (the base class' own evaluator: all-ones priors, a fresh random rollout on every
    leaf visit, MCTS.py:346-383) on the device against the unmodified reference (rollout_kat.json): visit counts, exact
    child values, Root.Value / WinRate, chosen moves -- and numpy's global MT19937 stream ends up in the same state,
    i.e. the device consumed exactly the draws the reference's rollouts make."""
    case = golden('rollout_kat.json')[index]
    cls = bb['FixedMCTS'] if case['kind'] == 'fixed' else bb['DynamicMCTS']
    m = cls(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=case['playLimit'])
    np.random.seed(case['seed'])
    s = bb['games'][case['game']]()
    for mv in case['moves']:
        nxt, v, probs = m.FindMove(s, 1)
        plays = m.Root.ChildPlays()
        nz = np.flatnonzero(plays).tolist()
        assert nz == mv['plays_idx'] and plays[nz].astype(int).tolist() == mv['plays'], (case['game'], case['kind'])
        assert [repr(float(m.Root.Children[a].Value)) for a in nz] == mv['child_values']
        assert m._last_action == mv['action'] and repr(float(v)) == mv['winrate']
        assert m.Root.Plays == mv['root_plays'] and repr(float(m.Root.Value)) == mv['root_value']
        assert int(np.random.get_state()[2]) == mv['rng_after']
        assert hashlib.sha256(probs.tobytes()).hexdigest()[:16] == mv['probs_sha']
        s = nxt
        m.MoveRoot(s)
    assert float(np.random.random_sample()) == case['next_double']
    assert m._eng.evaluator == 'rollout'


@pytest.mark.parametrize('index', [0, 1, 2])
def test_good_arena_vs_reference(bb, golden, index):
    """Blackbird.TestGood (Blackbird.py:146-174) with the opponent's wall-clock second replaced by a play limit (the
    reference's only non-deterministic input): every game replays the unmodified reference's TestModels(model,
    FixedMCTS(maxDepth=10, explorationRate=0.85, playLimit=...)) position for position (arena_good.json)."""
    import random
    case = golden('arena_good.json')[index]
    B = bb['Blackbird']
    base = bb['DynamicMCTS'] if case['kind'] == 'dynamic' else bb['FixedMCTS']
    cls = type('KatSearch', (base, KatNet), {})
    m = cls(maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=case['playLimit'])
    m.Game = bb['games'][case['game']]
    m._nActions = m.Game.LegalMoves
    m.Name, m.Version = 'gold', 1
    rows, boards = [], []

    class Conn(object):
        def PutTrainingStatistic(self, *a):
            rows.append(list(a))
    m.Conn = Conn()
    move_root = m.MoveRoot

    def logged(state):
        boards.append(hashlib.sha256(np.asarray(state.Board, dtype=np.int8).tobytes()
                                     + bytes([int(state.Player)])).hexdigest()[:12])
        return move_root(state)
    m.MoveRoot = logged
    random.seed(case['seed'])
    np.random.seed(case['seed'])
    results = []
    for _ in range(case['numTests']):
        good = bb['FixedMCTS'](maxDepth=10, explorationRate=0.85, timeLimit=None, playLimit=case['goodPlayLimit'])
        stats = B.TestGood(m, case['temp'], 1, good=good)
        results.append({'wins': 1, 'draws': 0, 'losses': -1}[next(k for k, v in stats.items() if v)])
    assert boards == case['boards'], (case['game'], case['kind'])
    assert results == case['results'] and [r[0] for r in rows] == case['results'] and all(r[3] == 'MCTS' for r in rows)
    assert float(np.random.random_sample()) == case['next_double']


def test_good_opponent_wall_clock(bb):
    """The reference's own TestGood opponent: FixedMCTS(maxDepth=10, explorationRate=0.85, timeLimit=1)."""
    good = bb['FixedMCTS'](maxDepth=10, explorationRate=0.85, timeLimit=0.2)
    s = bb['games']['Connect4']()
    nxt, v, probs = good.FindMove(s, 0.1)
    assert good.Root.Plays >= 8 and abs(probs.sum() - 1) < 1e-12 and 0 <= float(v) <= 1
    assert float(good.SampleValue(s, 1)) in (0.0, 0.5, 1.0)


def test_checkpoint_bundle_interop(bb, tmp_path, monkeypatch):
    """Network.saveModel writes a TensorFlow Saver bundle (best.index / best.data-00000-of-00001, the reference graph's
    variable names) next to its own file; a model directory that holds ONLY such a bundle -- what the reference saves,
    minus the TensorFlow graph file -- loads with identical weights and evaluates identically."""
    from blackbird_b200 import tfbundle
    monkeypatch.chdir(tmp_path)
    B = bb['Blackbird']
    cfg = {'maxDepth': 10, 'explorationRate': 0.85, 'timeLimit': None, 'playLimit': 10}
    net = {'blocks': 4, 'filters': 16, 'eval': {'dense': 16}, 'policy': {'dirichlet': {'epsilon': 0.0, 'alpha': 0.2}}}
    m = B.Model(bb['games']['Connect4'], 'Ckpt', cfg, net, {})
    d = os.path.join('blackbird_models', 'Ckpt_1')
    assert sorted(os.listdir(d)) == ['best.data-00000-of-00001', 'best.index', 'best.npz']
    tensors = tfbundle.read_bundle(os.path.join(d, 'best'))
    assert tensors['resTower/block_3/conv_2/kernel'].shape == (3, 3, 16, 16) and tensors['policy/policy/bias'].shape == (7,)
    s = m.Game()
    p0, v0 = m.getPolicy(s.AsInputArray()), m.getEvaluation(s.AsInputArray())
    w0 = m.getWeights().copy()
    m.Conn.Close()
    os.remove(os.path.join(d, 'best.npz'))                             # only the TensorFlow bundle is left
    m2 = B.Model(bb['games']['Connect4'], 'Ckpt', cfg, net, {})
    assert np.array_equal(m2.getWeights(), w0)
    assert np.array_equal(m2.getPolicy(s.AsInputArray()), p0) and m2.getEvaluation(s.AsInputArray()) == v0
    m2.Conn.Close()
