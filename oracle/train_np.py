"""Oracle trainer step: restatement of the reference's loss and optimizers (NetworkFactory.py:185-245) over the
forward graph of network_np, as driven by Network.train (Network.py:66-84) and TrainWithExamples (Blackbird.py:271-312).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED by the reference (TensorFlow is not installable
here and the reference holds no test for any training output); the restatement is pinned instead by two independent
checks (tests/test_oracle_train.py): the loss in numpy equals the loss written with torch ops, and torch autograd's
gradients of it equal central finite differences of the numpy loss.

The loss AS WRITTEN (its quirks are restated, not fixed):
  * the policy fed to the loss is the NOISY one: ((1 - eps) softmax + eps noise[A]) / sum over the WHOLE [B, A] tensor
    (NetworkFactory.py:176-182; one noise vector shared by the batch) -- at batch B the rows sum to ~1/B;
  * lossPolicy = -mean over the B x B matrix tensordot(log(policy) [B, A], policyLabel^T [A, B]) (:190-194): every
    row's log-policy is paired with EVERY row's label, = -(1/B^2) sum_a (sum_i log p[i, a]) (sum_j label[j, a]);
  * lossEvaluation = mean((evaluation - label)^2) (:187-188);
  * lossParam = mean over the trainable variables without 'bias' in their name -- conv kernels, BN gamma AND beta, dense
    kernels -- of l2_loss(v) = sum(v^2) / 2 (:196-202);
  * BN runs with training=False everywhere (moving mean 0 / variance 1 are never updated), so it is an affine map
    whose gamma / beta are the only BN parameters that receive gradients;
  * the teacher term (:206-219) calls tf.log (removed in TF 2) and reduces a rank-1 tensor over axis 1: it cannot run
    in the reference's own TF 2 environment and is left out (hasTeacher is false in parameters_template.yaml:11).
Optimizers (:229-242, tf.compat.v1.train semantics): AdamOptimizer(lr) beta1 0.9, beta2 0.999, epsilon 1e-8 with
lr_t = lr sqrt(1 - beta2^t) / (1 - beta1^t) and theta -= lr_t m / (sqrt(v) + epsilon); MomentumOptimizer(lr, momentum):
accum = momentum accum + g, theta -= lr accum; GradientDescentOptimizer: theta -= lr g.
"""
import numpy as np

from . import network_np

BN_EPS = network_np.BN_EPS


def trainable_names(weights):
    """Variables the optimizer updates, in the engine's flat order; BN moving statistics are not trainable."""
    return [k for k in weights if not (k.endswith('.mean') or k.endswith('.var'))]


def l2_names(weights):
    """Trainable variables without 'bias' in their TF name: conv / dense kernels, BN gamma and beta."""
    return [k for k in trainable_names(weights) if not k.endswith('.b')]


def loss_np(weights, planes, evals, policies, eps=0.0, noise=None):
    """-> (loss, lossEvaluation, lossPolicy, lossParam) in float64, NetworkFactory.py:185-204."""
    w = {k: np.asarray(v, np.float64) for k, v in weights.items()}
    B = planes.shape[0]
    A = w['policy.d.b'].shape[0]
    soft, value = network_np.forward(w, planes, 0.0, None, np.float64)      # softmax and tanh value
    nz = np.zeros(A) if noise is None else np.asarray(noise, np.float64)
    u = (1.0 - eps) * soft + eps * nz[None, :]
    pol = u / u.sum()                                                         # tf.reduce_sum(policy): the whole tensor
    loss_eval = np.mean((value - np.asarray(evals, np.float64)) ** 2)
    loss_pol = -np.mean(np.log(pol) @ np.asarray(policies, np.float64).T)     # [B, B] matrix
    names = l2_names(w)
    loss_par = np.mean([0.5 * np.sum(w[k] ** 2) for k in names])
    return loss_eval + loss_pol + loss_par, loss_eval, loss_pol, loss_par


def loss_and_grads_torch(weights, planes, evals, policies, eps=0.0, noise=None):
    """The same loss written with torch ops in float64 and differentiated by autograd.
    -> ((loss, lossEvaluation, lossPolicy, lossParam), grads dict over trainable_names)."""
    import torch
    import torch.nn.functional as F
    dt = torch.float64
    P = {k: torch.tensor(np.asarray(v, np.float64), dtype=dt, requires_grad=not (k.endswith('.mean') or k.endswith('.var')))
         for k, v in weights.items()}

    def conv_bn(x, name):
        y = F.conv2d(x, P[name + '.w'].permute(3, 2, 0, 1), P[name + '.b'], padding='same')
        return F.batch_norm(y, P[name + '.mean'], P[name + '.var'], P[name + '.gamma'], P[name + '.beta'], training=False, eps=BN_EPS)

    x = torch.tensor(np.asarray(planes, np.float64), dtype=dt).permute(0, 3, 1, 2)
    x = F.relu(conv_bn(x, 'stem'))
    b = 0
    while 'block%d.c1.w' % b in P:
        y = F.relu(conv_bn(x, 'block%d.c1' % b))
        x = F.relu(conv_bn(y, 'block%d.c2' % b) + x)
        b += 1
    v = F.relu(conv_bn(x, 'value.conv')).permute(0, 2, 3, 1)
    v = F.relu((v @ P['value.d1.w'] + P['value.d1.b']).sum(dim=(1, 2)))
    value = torch.tanh((v @ P['value.d2.w'] + P['value.d2.b']).sum(dim=1))
    p = F.relu(conv_bn(x, 'policy.conv')).permute(0, 2, 3, 1)
    z = (p @ P['policy.d.w'] + P['policy.d.b']).sum(dim=(1, 2))
    soft = torch.softmax(z, dim=1)
    A = soft.shape[1]
    nz = torch.zeros(A, dtype=dt) if noise is None else torch.tensor(np.asarray(noise, np.float64), dtype=dt)
    u = (1.0 - eps) * soft + eps * nz[None, :]
    pol = u / u.sum()
    loss_eval = torch.mean((value - torch.tensor(np.asarray(evals, np.float64), dtype=dt)) ** 2)
    loss_pol = -torch.mean(torch.tensordot(torch.log(pol), torch.tensor(np.asarray(policies, np.float64), dtype=dt).t(), dims=1))
    names = l2_names(weights)
    loss_par = torch.stack([0.5 * (P[k] ** 2).sum() for k in names]).mean()
    loss = loss_eval + loss_pol + loss_par
    loss.backward()
    grads = {k: P[k].grad.numpy().copy() for k in trainable_names(weights)}
    return tuple(float(t.detach()) for t in (loss, loss_eval, loss_pol, loss_par)), grads


class Optimizer(object):
    """tf.compat.v1.train.{Adam,Momentum,GradientDescent}Optimizer state and update (NetworkFactory.py:229-242)."""

    def __init__(self, kind='adam', momentum=0.9):
        self.kind, self.momentum, self.t, self.slots = kind, momentum, 0, {}

    def step(self, weights, grads, lr):
        self.t += 1
        out = dict(weights)
        for k, g in grads.items():
            th = np.asarray(weights[k], np.float64)
            if self.kind == 'adam':
                m, v = self.slots.get(k, (np.zeros_like(th), np.zeros_like(th)))
                m = 0.9 * m + 0.1 * g
                v = 0.999 * v + 0.001 * g * g
                self.slots[k] = (m, v)
                lr_t = lr * np.sqrt(1.0 - 0.999 ** self.t) / (1.0 - 0.9 ** self.t)
                out[k] = th - lr_t * m / (np.sqrt(v) + 1e-8)
            elif self.kind == 'momentum':
                acc = self.momentum * self.slots.get(k, np.zeros_like(th)) + g
                self.slots[k] = acc
                out[k] = th - lr * acc
            else:
                out[k] = th - lr * g
        return out


def train_step(weights, opt, planes, evals, policies, lr, eps=0.0, noise=None):
    """One Network.train call (Network.py:66-84): sess.run(trainingOp).  -> (new weights, loss terms)."""
    losses, grads = loss_and_grads_torch(weights, planes, evals, policies, eps, noise)
    return opt.step(weights, grads, lr), losses
