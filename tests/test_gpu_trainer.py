"""GPU: the trainer step (csrc/train.cu, Network.train / NetworkFactory.py:185-245) against the oracle
(oracle/train_np.py: torch autograd in fp64 of the loss as written, pinned on CPU by finite differences)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')
from oracle import network_np, train_np  # noqa: E402
from tests.test_oracle_network import SHAPES  # noqa: E402
from tests.test_oracle_train import make_batch  # noqa: E402


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine, trainer
    return engine, trainer


def unflatten(template, flat):
    out, o = {}, 0
    for k, v in template.items():
        out[k] = np.asarray(flat[o:o + v.size], np.float64).reshape(v.shape)
        o += v.size
    assert o == flat.size
    return out


def rel_err(got, want):
    return np.abs(got - want).max() / max(np.abs(want).max(), 1e-30)


@pytest.mark.parametrize('game,B,shape', [('Connect4', 50, (4, 16, 16)), ('TicTacToe', 33, (2, 8, 5)), ('DragonChess', 6, (1, 16, 7)),
                                          ('Connect4', 7, (3, 32, 16))])
@pytest.mark.parametrize('eps', [0.0, 0.3])
def test_gradients_and_losses_vs_autograd(mods, game, B, shape, eps):
    """Loss terms and every gradient tensor of one step against torch autograd (fp64) of the reference's loss."""
    engine, trainer = mods
    cin, A = SHAPES[game]
    blocks, filters, dense = shape
    w = network_np.init_weights(cin, A, blocks=blocks, filters=filters, dense=dense, seed=7, jitter=0.15)
    planes, ev, pol = make_batch(game, B, seed=11)
    noise = np.random.RandomState(3).beta(0.2, 0.8, size=A).astype(np.float32) if eps else None
    flat = network_np.flatten_weights(w)
    tr = trainer.Trainer(game, flat, blocks, filters, dense, optimizer='sgd', epsilon=eps)
    tr.step(planes, ev, pol, lr=0.0, noise=noise)
    assert np.array_equal(tr.weights(), flat)                        # lr 0: weights untouched
    want_l, want_g = train_np.loss_and_grads_torch(w, planes, ev, pol.astype(np.float32), eps, noise)
    got_l = tr.losses()
    assert np.allclose(got_l, want_l, rtol=2e-5, atol=1e-6), (got_l, want_l)
    got_g = unflatten(w, tr.grads())
    worst = 0.0
    for k in train_np.trainable_names(w):
        e = rel_err(got_g[k], want_g[k])
        worst = max(worst, e)
        assert e < 2e-4, (k, e, np.abs(want_g[k]).max())
    for k in w:
        if k.endswith('.mean') or k.endswith('.var'):
            assert not got_g[k].any()
    print('[trainer parity] %s B=%d shape=%s eps=%g: loss %s, worst gradient tensor error %.2e (relative to the tensor max)' % (game, B, shape, eps, got_l, worst))
    tr.close()


@pytest.mark.parametrize('optimizer', ['adam', 'momentum', 'sgd'])
def test_steps_vs_oracle_optimizers(mods, optimizer):
    """Three consecutive steps from fixed weights on fixed batches: updated weights against the oracle's
    (tf.compat.v1.train Adam / Momentum / GradientDescent semantics).  Momentum and gradient descent are linear in the
    gradient: 1e-5 of the tensor's scale.  Adam's first steps move every coordinate by ~lr g / (|g| + 3e-7), which is
    insensitive to the gradient's size where |g| >> 3e-7 and arbitrary in sign where |g| is fp32 noise, so there the
    normalised update (theta' - theta) / lr is compared on the coordinates whose gradient is above 1e-4 of the tensor max."""
    engine, trainer = mods
    game, B = 'Connect4', 50
    w = network_np.init_weights(3, 7, seed=1, jitter=0.1)
    lr = {'adam': 1e-3, 'momentum': 1e-2, 'sgd': 5e-2}[optimizer]
    tr = trainer.Trainer(game, network_np.flatten_weights(w), 4, 16, 16, optimizer=optimizer, momentum=0.9)
    opt = train_np.Optimizer(optimizer, 0.9)
    cur = {k: np.asarray(v, np.float64) for k, v in w.items()}
    prev_got = unflatten(w, tr.weights())
    for step in range(3):
        planes, ev, pol = make_batch(game, B, seed=20 + step)
        pol = pol.astype(np.float32)
        tr.step(planes, ev, pol, lr)
        losses, grads = train_np.loss_and_grads_torch(cur, planes, ev, pol)
        nxt = opt.step(cur, grads, lr)
        assert np.allclose(tr.losses(), losses, rtol=5e-5, atol=1e-6)
        got = unflatten(w, tr.weights())
        for k in w:
            if k not in grads:
                assert np.array_equal(got[k], np.asarray(w[k], np.float64)), k          # frozen BN statistics
                continue
            if optimizer == 'adam':
                live = np.abs(grads[k]) > 1e-4 * max(np.abs(grads[k]).max(), 1e-30)
                du_got, du_want = (got[k] - prev_got[k]) / lr, (nxt[k] - cur[k]) / lr
                if live.any():
                    assert np.abs(du_got - du_want)[live].max() < 1e-2, (step, k, np.abs(du_got - du_want)[live].max())
                assert np.abs(du_got).max() < 1.2 * (1 if step == 0 else 4)
            else:
                scale = max(np.abs(nxt[k]).max(), 1e-3)
                assert np.abs(got[k] - nxt[k]).max() / scale < 1e-5, (optimizer, step, k)
        # keep the two trajectories together so that step 2 and 3 test one step each, not the accumulated drift
        if optimizer == 'adam':
            for k in grads:
                nxt[k] = got[k].copy()
            # the oracle's slots stay its own; they differ from the device's by fp32 rounding only
        cur, prev_got = nxt, got
    assert tr.steps == 3
    tr.close()


def test_step_from_record_stream(mods):
    """The trainer consumes the compact record stream on the device (no sqlite, no host parsing): a step on examples
    picked from the stream equals a step on the same examples parsed on the host (ExampleState.FromSerialized)."""
    engine, trainer = mods
    from blackbird_b200 import records
    from blackbird_b200.Blackbird import ExampleState
    eng = engine.Engine('Connect4', 'dynamic', n_games=16, evaluator='kat', play_limit_hint=30)
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(16)])
    eng.selfplay(30, 1.0, u)
    dev, off, sz = eng.records_serialized(to_host=False)
    host = dev.cpu().numpy()
    rec_off = np.concatenate([np.arange(int(off[g]), int(off[g + 1]), int(sz[g]), dtype=np.uint64) for g in range(16)])
    pick = np.random.RandomState(0).permutation(rec_off.size)[:50]
    w = network_np.flatten_weights(network_np.init_weights(3, 7, seed=4))
    a = trainer.Trainer('Connect4', w, 4, 16, 16, optimizer='adam')
    a.step_stream(dev, rec_off[pick], 1e-3)
    exs = [ExampleState.FromSerialized(bytes(host[int(o):int(o) + 200])) for o in rec_off[pick]]
    b = trainer.Trainer('Connect4', w, 4, 16, 16, optimizer='adam')
    b.step(np.vstack([e.Board for e in exs]), np.hstack([e.MctsEval for e in exs]), np.vstack([e.MctsPolicy for e in exs]), 1e-3)
    assert a.losses() == b.losses() and np.array_equal(a.weights(), b.weights())
    assert {float(e.MctsEval[0]) for e in exs} <= {-1.0, 0.0, 1.0}
    a.close(); b.close(); eng.close()


def test_training_reduces_the_loss(mods):
    engine, trainer = mods
    planes, ev, pol = make_batch('Connect4', 64, seed=5)
    w = network_np.flatten_weights(network_np.init_weights(3, 7, seed=2))
    tr = trainer.Trainer('Connect4', w, 4, 16, 16, optimizer='adam')
    first = None
    for _ in range(60):
        tr.step(planes, ev, pol, 1e-3)
        first = first or tr.losses()
    last = tr.losses()
    assert last[0] < first[0] and last[1] < first[1]
    tr.close()
