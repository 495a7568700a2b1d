"""Random-init weights of the reference's network shape, in the engine's flat layout.

Stands for tf.compat.v1.global_variables_initializer() on the graph of NetworkFactory.py:22-183
(Network.py:24): glorot-uniform kernels, zero biases, BN gamma 1 / beta 0 / moving mean 0 /
moving variance 1.  Layout (DESIGN.md "weight layout"): for every conv+BN pair
w[k,k,ci,co] b gamma beta mean var; order stem, block{i}.c1, block{i}.c2, value.conv,
value.d1.w[1,D] .b[D], value.d2.w[D,1] .b[1], policy.conv, policy.d.w[2,A] .b[A].
"""
import numpy as np


def random_weights(cin, actions, blocks=4, filters=16, dense=16, seed=0):
    rs = np.random.RandomState(seed)
    parts = []

    def glorot(fan_in, fan_out, shape):
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        parts.append(rs.uniform(-lim, lim, size=shape).astype(np.float32).ravel())

    def conv_bn(k, ci, co):
        glorot(k * k * ci, k * k * co, (k, k, ci, co))
        for v in (0.0, 1.0, 0.0, 0.0, 1.0):          # bias, gamma, beta, mean, var
            parts.append(np.full(co, v, dtype=np.float32))

    conv_bn(3, cin, filters)
    for _ in range(2 * blocks):
        conv_bn(3, filters, filters)
    conv_bn(1, filters, 1)
    glorot(1, dense, (1, dense))
    parts.append(np.zeros(dense, np.float32))
    glorot(dense, 1, (dense, 1))
    parts.append(np.zeros(1, np.float32))
    conv_bn(1, filters, 2)
    glorot(2, actions, (2, actions))
    parts.append(np.zeros(actions, np.float32))
    return np.concatenate(parts)


def flops_per_eval(h, w, cin, actions, blocks=4, filters=16, dense=16):
    """2*MACs of one forward, BN folded, heads as sum-then-dense."""
    hw = h * w
    macs = hw * 9 * cin * filters + 2 * blocks * hw * 9 * filters * filters
    macs += hw * filters * 3                      # 1x1 convs: 1 value + 2 policy filters
    macs += dense + dense + 2 * actions           # dense(1->D), dense(D->1), dense(2->A)
    return 2 * macs
