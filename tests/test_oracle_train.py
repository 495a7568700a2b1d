"""CPU: the trainer oracle (oracle/train_np.py: the reference's loss and optimizers, NetworkFactory.py:185-245) pinned
two ways -- the numpy loss equals the torch-op loss, and torch autograd's gradients equal central finite differences of
the numpy loss -- and the optimizer restatements against torch.optim where the semantics coincide."""
import numpy as np
import pytest

from oracle import network_np, train_np
from tests.test_oracle_network import SHAPES, playout_planes


def make_batch(game, B, seed):
    cin, A = SHAPES[game]
    rs = np.random.RandomState(seed)
    planes = playout_planes(game, B, seed=seed)
    pol = rs.dirichlet(np.ones(A) * 0.3, size=B)
    pol[rs.rand(B) < 0.2] = 0.0                              # terminal records carry an all-zero policy (Blackbird.py:256)
    ev = rs.choice([-1.0, 0.0, 1.0], size=B)
    return planes, ev, pol


@pytest.mark.parametrize('game,B', [('TicTacToe', 12), ('Connect4', 10), ('DragonChess', 3)])
@pytest.mark.parametrize('eps', [0.0, 0.3])
def test_loss_numpy_equals_torch_and_autograd_equals_finite_differences(game, B, eps):
    cin, A = SHAPES[game]
    w = network_np.init_weights(cin, A, blocks=2, filters=8, dense=6, seed=3, jitter=0.15, dtype=np.float64)
    planes, ev, pol = make_batch(game, B, seed=5)
    noise = np.random.RandomState(9).beta(0.2, 0.8, size=A) if eps else None
    ln = train_np.loss_np(w, planes, ev, pol, eps, noise)
    lt, grads = train_np.loss_and_grads_torch(w, planes, ev, pol, eps, noise)
    assert np.allclose(ln, lt, rtol=1e-12, atol=1e-12)
    rs = np.random.RandomState(1)
    for name in ['stem.w', 'stem.b', 'stem.gamma', 'block0.c1.beta', 'block1.c2.w', 'block1.c2.gamma', 'value.conv.w', 'value.d1.w',
                 'value.d1.b', 'value.d2.w', 'value.d2.b', 'policy.conv.beta', 'policy.d.w', 'policy.d.b']:
        flat = w[name].reshape(-1)
        for idx in rs.choice(flat.size, size=min(3, flat.size), replace=False):
            h = 1e-6
            old = flat[idx]
            flat[idx] = old + h
            up = train_np.loss_np(w, planes, ev, pol, eps, noise)[0]
            flat[idx] = old - h
            dn = train_np.loss_np(w, planes, ev, pol, eps, noise)[0]
            flat[idx] = old
            fd = (up - dn) / (2 * h)
            g = grads[name].reshape(-1)[idx]
            assert abs(fd - g) <= 1e-6 * max(1.0, abs(g)) + 1e-7, (name, idx, fd, g)
    # frozen statistics get no gradient, the others all do
    assert set(grads) == set(train_np.trainable_names(w)) and 'stem.mean' not in grads
    assert len(train_np.l2_names(w)) == 6 * 2 + 12


def test_policy_loss_is_the_all_pairs_mean_as_written():
    """-reduce_mean(tensordot(log(policy), transpose(policyLabel), axes=1)) is a mean over a B x B matrix."""
    w = network_np.init_weights(3, 7, blocks=1, filters=4, dense=4, seed=1, dtype=np.float64)
    planes, ev, pol = make_batch('Connect4', 6, seed=2)
    soft, _ = network_np.forward(w, planes)
    want = -np.mean([[np.dot(np.log(soft[i] / 6.0), pol[j]) for j in range(6)] for i in range(6)])
    assert abs(train_np.loss_np(w, planes, ev, pol)[2] - want) < 1e-12


@pytest.mark.parametrize('kind', ['adam', 'momentum', 'sgd'])
def test_optimizers_against_torch_optim(kind):
    import torch
    rs = np.random.RandomState(0)
    th0 = rs.randn(40)
    gs = [rs.randn(40) * 10.0 ** rs.randint(-9, 1) for _ in range(5)]
    opt = train_np.Optimizer(kind, momentum=0.9)
    w = {'x': th0.copy()}
    t = torch.tensor(th0.copy(), dtype=torch.float64, requires_grad=True)
    lr = 1e-2
    if kind == 'adam':
        to = torch.optim.Adam([t], lr=lr, betas=(0.9, 0.999), eps=1e-8)
    elif kind == 'momentum':
        to = torch.optim.SGD([t], lr=lr, momentum=0.9)
    else:
        to = torch.optim.SGD([t], lr=lr)
    for i, g in enumerate(gs):
        w = opt.step(w, {'x': g}, lr)
        t.grad = torch.tensor(g, dtype=torch.float64)
        to.step()
        if kind == 'adam':
            # TF1 Adam puts epsilon next to sqrt(v) BEFORE the bias correction ("epsilon hat"); torch after it.  They agree
            # where |g| >> epsilon; compare only those coordinates and check the closed form of the first step everywhere.
            if i == 0:
                assert np.allclose(w['x'], th0 - lr * g / (np.abs(g) + 1e-8 / np.sqrt(1 - 0.999)), rtol=1e-12, atol=0)
            big = np.abs(g) > 1e-3
            assert np.allclose(w['x'][big], t.detach().numpy()[big], rtol=1e-4, atol=1e-6)
            t.data = torch.tensor(w['x'])            # keep the two trajectories together
        else:
            assert np.allclose(w['x'], t.detach().numpy(), rtol=1e-12, atol=1e-15)
