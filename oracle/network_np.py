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

    def __init__(self, weights, eps=0.0, dtype=np.float32, alpha=0.2, rng=None):
        self.W = weights
        self.Eps = eps
        self.Dtype = dtype
        self.Alpha = alpha
        self.Rng = rng                      # eps > 0 and an rng: a fresh Beta(alpha, 1-alpha) draw per action per call (A.17)

    def getPolicy(self, planes):
        noise = None
        if self.Eps > 0 and self.Rng is not None:
            A = self.W['policy.d.b'].shape[0]
            noise = self.Rng.beta(self.Alpha, 1.0 - self.Alpha, size=(planes.shape[0], A))
        return forward(self.W, planes, self.Eps, noise, self.Dtype)[0][0].astype(np.float32)

    def getEvaluation(self, planes):
        return np.float32(forward(self.W, planes, self.Eps, None, self.Dtype)[1][0])


def flatten_weights(weights):
    """Flat fp32 vector in dict order: the layout bb_net_load() takes."""
    return np.concatenate([np.asarray(v, dtype=np.float32).ravel() for v in weights.values()])


# ---- second implementation: the same graph through torch.nn.functional (CPU, fp64) -------------------
def forward_torch(weights, planes, dtype=None):
    """The graph of NetworkFactory.py:22-183 restated with torch's own layer kernels (conv2d with
    padding='same', batch_norm(training=False, eps=1e-3), linear) instead of the hand-written numpy
    loops above: an independent implementation of the layer semantics (NHWC/HWIO -> NCHW/OIHW
    conversion, zero padding, BN inference formula, dense on the last axis) that pins forward()
    against a layout or padding slip.  Test infrastructure only (tests/test_oracle_network.py)."""
    import torch
    import torch.nn.functional as F
    dt = dtype or torch.float64
    T = lambda a: torch.as_tensor(np.asarray(a, dtype=np.float64)).to(dt)

    def conv_bn(x, name):
        k = T(weights[name + '.w']).permute(3, 2, 0, 1).contiguous()          # HWIO -> OIHW
        y = F.conv2d(x, k, T(weights[name + '.b']), stride=1, padding='same')
        return F.batch_norm(y, T(weights[name + '.mean']), T(weights[name + '.var']), T(weights[name + '.gamma']),
                            T(weights[name + '.beta']), training=False, eps=BN_EPS)

    x = T(planes).permute(0, 3, 1, 2).contiguous()                             # NHWC -> NCHW
    x = F.relu(conv_bn(x, 'stem'))
    b = 0
    while 'block%d.c1.w' % b in weights:
        y = F.relu(conv_bn(x, 'block%d.c1' % b))
        y = conv_bn(y, 'block%d.c2' % b)
        x = F.relu(y + x)
        b += 1
    v = F.relu(conv_bn(x, 'value.conv')).permute(0, 2, 3, 1)                    # [N,H,W,1]
    v = F.linear(v, T(weights['value.d1.w']).t(), T(weights['value.d1.b']))     # dense on the last axis, bias per cell
    v = F.relu(v.sum(dim=(1, 2)))
    v = F.linear(v, T(weights['value.d2.w']).t(), T(weights['value.d2.b']))
    value = torch.tanh(v.sum(dim=1))
    p = F.relu(conv_bn(x, 'policy.conv')).permute(0, 2, 3, 1)                   # [N,H,W,2]
    p = F.linear(p, T(weights['policy.d.w']).t(), T(weights['policy.d.b'])).sum(dim=(1, 2))
    return torch.softmax(p, dim=1).numpy(), value.numpy()


# ---- the bf16 tensor-core kernel's arithmetic, restated on the CPU ------------------------------------
def _bf16_rn(a):
    """fp32 -> bf16 (round to nearest even) -> fp32, elementwise."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(np.shape(a))


def forward_bf16_emulated(weights, planes, last_bf16=False):
    """What blackbird_b200/csrc/net_tc.cuh computes, with its rounding points: BN folded into the
    convolutions in double and stored as bf16 tower weights / fp32 biases (net.cu net_build), bf16
    activations between tower layers (fp32 accumulation, + fp32 bias, + bf16 skip, ReLU, round), the last
    tower layer kept in fp32 for the heads, heads in fp32.  Only the accumulation ORDER differs from the
    kernel, so the kernel must agree with this far more tightly (~1e-3 of a probability: a rare bf16
    rounding flip of one activation) than with the fp64 graph (2e-2).  last_bf16: the layer-by-layer kernels of
    net_wide.cuh (other tower widths) store the last tower layer as bf16 too and feed the heads from that.
    Test infrastructure only."""
    f32 = np.float32
    w = {k: np.asarray(v, dtype=np.float64) for k, v in weights.items()}

    def fold(name):
        s = w[name + '.gamma'] / np.sqrt(w[name + '.var'] + 1e-3)
        return (w[name + '.w'] * s).astype(f32), ((w[name + '.b'] - w[name + '.mean']) * s + w[name + '.beta']).astype(f32)

    def conv(x, k, b):                      # x bf16-valued fp32 [N,H,W,Ci]; k bf16-valued; fp32 accumulate (products exact)
        return (_conv_same(x.astype(np.float64), k.astype(np.float64), 0.0)).astype(f32) + b

    relu = lambda t: np.maximum(t, f32(0))
    names = ['stem']
    b = 0
    while 'block%d.c1.w' % b in w:
        names += ['block%d.c1' % b, 'block%d.c2' % b]
        b += 1
    x = planes.astype(f32)                  # -1/0/1: exact in bf16
    act = None                              # fp32 value of the last layer
    for li, name in enumerate(names):
        k, bias = fold(name)
        y = conv(x, _bf16_rn(k), bias)
        if li > 0 and li % 2 == 0:          # conv_2 + BN + block input (the bf16 copy in shared memory)
            y = y + skip
        act = relu(y)
        x = _bf16_rn(act)
        if li % 2 == 0:
            skip = x
    x = _bf16_rn(act) if last_bf16 else act  # heads read the fp32 registers of the last epilogue (16 filters) / the stored bf16 activations
    kv, bv = fold('value.conv')
    kp, bp = fold('policy.conv')
    hv = relu(x @ kv[0, 0] + bv).sum(axis=(1, 2))[:, 0]                        # s = sum_hw relu(bn(conv))
    hp = relu(x @ kp[0, 0] + bp).sum(axis=(1, 2))                              # [N,2]
    hw = f32(planes.shape[1] * planes.shape[2])
    d1w, d1b = w['value.d1.w'].astype(f32)[0], w['value.d1.b'].astype(f32)
    d2w, d2b = w['value.d2.w'].astype(f32)[:, 0], w['value.d2.b'].astype(f32)[0]
    h = relu(hv[:, None] * d1w + hw * d1b)
    value = np.tanh((h * d2w).sum(axis=1) + d2b)
    pw, pb = w['policy.d.w'].astype(f32), w['policy.d.b'].astype(f32)
    z = hp[:, 0:1] * pw[0] + hp[:, 1:2] * pw[1] + hw * pb
    z = z - z.max(axis=1, keepdims=True)
    e = np.exp(z.astype(np.float64))
    return (e / e.sum(axis=1, keepdims=True)), value.astype(np.float64)
