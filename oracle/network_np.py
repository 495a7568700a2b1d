"""Oracle policy-value network forward: numpy restatement of NetworkFactory.py:22-183 as
evaluated by Network.getPolicy / getEvaluation (Network.py:48-64).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

PARITY UNPINNED: the reference builds this graph with TensorFlow, which is not installed in
the build image and cannot be installed (no network); the reference holds no test or golden
vector for any network output.  This file follows the graph line by line with the TF layer
defaults it relies on: conv2d NHWC/HWIO with bias, 'same' zero padding for
the 3x3 tower convs, batch_normalization(training=False, eps=1e-3) with frozen moving
statistics, dense on the last axis, heads reduced over H,W AFTER the per-cell dense.
"""
import numpy as np

BN_EPS = 1e-3            # tf.compat.v1.layers.batch_normalization default epsilon


def init_weights(cin, actions, blocks=4, filters=16, dense=16, seed=0, dtype=np.float32, jitter=0.0):
    """Random-init weights of the template shape (parameters_template.yaml:11-16).

    glorot-uniform kernels / zero biases / BN gamma 1, beta 0, mean 0, var 1 are what
    global_variables_initializer gives (Network.py:24).  ``jitter`` > 0 perturbs biases and BN
    parameters so parity tests exercise every term (a trained network has non-trivial ones).
    Returns an ordered dict name -> array; the order is the engine's flat weight layout
    (include/bbengine.h, bb_net_load).
    """
    rs = np.random.RandomState(seed)

    def glorot(*shape):
        rf = int(np.prod(shape[:-2])) if len(shape) > 2 else 1
        lim = np.sqrt(6.0 / (shape[-2] * rf + shape[-1] * rf))
        return rs.uniform(-lim, lim, size=shape).astype(dtype)

    def vec(n, base):
        return (base + jitter * rs.standard_normal(n)).astype(dtype)

    w = {}

    def conv_bn(name, k, ci, co):
        w[name + '.w'] = glorot(k, k, ci, co)
        w[name + '.b'] = vec(co, 0.0)
        w[name + '.gamma'] = vec(co, 1.0)
        w[name + '.beta'] = vec(co, 0.0)
        w[name + '.mean'] = vec(co, 0.0)
        w[name + '.var'] = np.abs(vec(co, 1.0)) + dtype(0.05) if jitter else vec(co, 1.0)

    conv_bn('stem', 3, cin, filters)
    for b in range(blocks):
        conv_bn('block%d.c1' % b, 3, filters, filters)
        conv_bn('block%d.c2' % b, 3, filters, filters)
    conv_bn('value.conv', 1, filters, 1)
    w['value.d1.w'] = glorot(1, dense)
    w['value.d1.b'] = vec(dense, 0.0)
    w['value.d2.w'] = glorot(dense, 1)
    w['value.d2.b'] = vec(1, 0.0)
    conv_bn('policy.conv', 1, filters, 2)
    w['policy.d.w'] = glorot(2, actions)
    w['policy.d.b'] = vec(actions, 0.0)
    return w


def _conv_same(x, k, b):
    """NHWC x HWIO 'same' convolution, stride 1 (NetworkFactory.py:46-49,70-74,85-89)."""
    kh, kw = k.shape[:2]
    n, h, wd, _ = x.shape
    ph, pw = kh // 2, kw // 2
    xp = np.zeros((n, h + 2 * ph, wd + 2 * pw, x.shape[3]), dtype=x.dtype)
    xp[:, ph:ph + h, pw:pw + wd] = x
    out = np.zeros((n, h, wd, k.shape[3]), dtype=x.dtype)
    for i in range(kh):
        for j in range(kw):
            out += xp[:, i:i + h, j:j + wd] @ k[i, j]
    return out + b


def _bn(x, w, name):
    return w[name + '.gamma'] * (x - w[name + '.mean']) / np.sqrt(w[name + '.var'] + x.dtype.type(BN_EPS)) \
        + w[name + '.beta']


def forward(weights, planes, eps=0.0, noise=None, dtype=np.float64):
    """planes int8 [N,H,W,C] -> (policy [N,A], value [N]).

    ``noise`` [N,A] stands for dist.sample([1,A])[0][:,0] (NetworkFactory.py:176-180, A.17);
    the final division is by the sum over the WHOLE batch tensor in the reference (:182) --
    the reference only ever runs batch 1, so rows are normalised independently here.
    """
    w = {k: v.astype(dtype) for k, v in weights.items()}
    x = planes.astype(dtype)
    relu = lambda t: np.maximum(t, 0)
    x = relu(_bn(_conv_same(x, w['stem.w'], w['stem.b']), w, 'stem'))            # :46-57
    b = 0
    while 'block%d.c1.w' % b in w:                                                # :59-103
        skip = x
        y = relu(_bn(_conv_same(x, w['block%d.c1.w' % b], w['block%d.c1.b' % b]), w, 'block%d.c1' % b))
        y = _bn(_conv_same(y, w['block%d.c2.w' % b], w['block%d.c2.b' % b]), w, 'block%d.c2' % b)
        x = relu(y + skip)
        b += 1
    # value head :115-145
    v = relu(_bn(_conv_same(x, w['value.conv.w'], w['value.conv.b']), w, 'value.conv'))
    v = v @ w['value.d1.w'] + w['value.d1.b']          # dense per cell, bias per cell
    v = relu(v.sum(axis=(1, 2)))
    v = v @ w['value.d2.w'] + w['value.d2.b']
    value = np.tanh(v.sum(axis=1))
    # policy head :155-182
    p = relu(_bn(_conv_same(x, w['policy.conv.w'], w['policy.conv.b']), w, 'policy.conv'))
    p = (p @ w['policy.d.w'] + w['policy.d.b']).sum(axis=(1, 2))
    p = p - p.max(axis=1, keepdims=True)
    e = np.exp(p)
    soft = e / e.sum(axis=1, keepdims=True)
    if noise is None:
        noise = np.zeros_like(soft)
    pol = (1 - dtype(eps)) * soft + dtype(eps) * noise.astype(dtype)
    pol = pol / pol.sum(axis=1, keepdims=True)
    return pol, value


class NetEvaluator(object):
    """Network.getPolicy/getEvaluation (Network.py:48-64) served by the restatement at batch 1,
    fp32, eps = 0 (no noise): the evaluator of the CPU baseline and of replay-parity tests."""

    def __init__(self, weights, eps=0.0, dtype=np.float32):
        self.W = weights
        self.Eps = eps
        self.Dtype = dtype

    def getPolicy(self, planes):
        return forward(self.W, planes, self.Eps, None, self.Dtype)[0][0].astype(np.float32)

    def getEvaluation(self, planes):
        return np.float32(forward(self.W, planes, self.Eps, None, self.Dtype)[1][0])


def flatten_weights(weights):
    """Flat fp32 vector in dict order: the layout bb_net_load() takes."""
    return np.concatenate([np.asarray(v, dtype=np.float32).ravel() for v in weights.values()])
