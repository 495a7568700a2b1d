"""CPU: the network oracle (oracle/network_np.forward, the numpy restatement of NetworkFactory.py:22-183) against a
second, independent implementation of the same layer semantics -- torch.nn.functional conv2d(padding='same') /
batch_norm(training=False, eps=1e-3) / linear in fp64 -- and the CPU restatement of the bf16 tensor-core kernel's
rounding points against the fp64 graph.  TensorFlow itself is not installable here, so this is the strongest pin
available for the forward graph (DESIGN.md "oracle")."""
import numpy as np
import pytest

from oracle import network_np
from oracle.rules_np import GAMES

SHAPES = {'TicTacToe': (3, 9), 'Connect4': (3, 7), 'DragonChess': (17, 4032)}


def playout_planes(game, n, seed):
    rs = np.random.RandomState(seed)
    out = []
    while len(out) < n:
        s = GAMES[game]()
        while s.Winner() is None and len(out) < n:
            out.append(s.AsInputArray())
            idx = np.flatnonzero(s.LegalActions() == 1)
            if idx.size == 0:
                break
            s.ApplyAction(int(idx[rs.randint(idx.size)]))
            if game == 'DragonChess' and rs.rand() < 0.1:
                break
    return np.concatenate(out)


@pytest.mark.parametrize('game,n', [('TicTacToe', 200), ('Connect4', 200), ('DragonChess', 24)])
@pytest.mark.parametrize('jitter', [0.0, 0.2])
@pytest.mark.parametrize('shape', [(4, 16, 16), (2, 8, 5)])
def test_numpy_forward_equals_torch_functional(game, n, jitter, shape):
    """NetworkFactory.py:46-57 (stem), :59-103 (residual tower), :115-145 (value head), :155-172 (policy head)."""
    cin, A = SHAPES[game]
    blocks, filters, dense = shape
    w = network_np.init_weights(cin, A, blocks=blocks, filters=filters, dense=dense, seed=11, jitter=jitter)
    planes = playout_planes(game, n, seed=3)
    p, v = network_np.forward(w, planes, dtype=np.float64)
    pt, vt = network_np.forward_torch(w, planes)
    assert np.abs(p - pt).max() < 1e-12 and np.abs(v - vt).max() < 1e-12
    assert np.abs(p.sum(axis=1) - 1).max() < 1e-12


def test_torch_pin_catches_a_layout_slip():
    """The pin is not vacuous: transposing the spatial axes of one kernel (a HWIO slip) is seen."""
    w = network_np.init_weights(3, 7, seed=11)
    planes = playout_planes('Connect4', 50, seed=3)
    bad = dict(w)
    bad['block1.c1.w'] = np.ascontiguousarray(w['block1.c1.w'].transpose(1, 0, 2, 3))
    p, _ = network_np.forward(w, planes)
    pt, _ = network_np.forward_torch(bad, planes)
    assert np.abs(p - pt).max() > 1e-6


@pytest.mark.parametrize('game,n', [('TicTacToe', 200), ('Connect4', 200), ('DragonChess', 16)])
@pytest.mark.parametrize('jitter', [0.0, 0.2])
def test_bf16_restatement_stays_within_its_tolerance(game, n, jitter):
    """bf16 weights and activations (fp32 accumulation) against the fp64 graph: 2e-2 absolute on probabilities and on
    the value -- the tolerance the GPU test (tests/test_gpu_network.py) holds the tcgen05 kernel to."""
    cin, A = SHAPES[game]
    for seed in (0, 2):
        w = network_np.init_weights(cin, A, seed=seed, jitter=jitter)
        planes = playout_planes(game, n, seed=1)
        p, v = network_np.forward(w, planes)
        pb, vb = network_np.forward_bf16_emulated(w, planes)
        assert np.abs(pb - p).max() < 2e-2 and np.abs(vb - v).max() < 2e-2


def test_bf16_rounding_helper():
    x = np.array([1.0, 1.00390625, 1.001953125, 1.005859375, -3.14159, 0.0, 1e-30], np.float32)
    got = network_np._bf16_rn(x)
    import torch
    want = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    assert np.array_equal(got, want)
