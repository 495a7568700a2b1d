"""Thin Python handle over the C-ABI trainer (include/bbengine.h, csrc/train.cu): Network.train (Network.py:66-84) as
one GPU step -- forward with saved activations, the reference's loss as written, gradients, Adam / momentum / SGD
(NetworkFactory.py:185-245).  torch only carries the batch to the device."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check

OPTIMIZERS = {'adam': 0, 'momentum': 1, 'sgd': 2}


class Trainer(object):
    def __init__(self, game, flat_weights, blocks, filters, dense, optimizer='adam', momentum=0.9, epsilon=0.0, alpha=0.2,
                 seed=0, device=0):
        if not torch.cuda.is_available():
            raise _lib.EngineError('no CUDA device: blackbird_b200 has no CPU fallback')
        w = np.ascontiguousarray(flat_weights, dtype=np.float32)
        d = _lib.bb_train_desc()
        d.game, d.blocks, d.filters, d.dense = _lib.GAME_IDS[game], int(blocks), int(filters), int(dense)
        d.optimizer = OPTIMIZERS.get(optimizer, 2)               # NetworkFactory.py:229-242: anything else is plain gradient descent
        d.momentum, d.epsilon, d.alpha, d.noise_seed, d.device = float(momentum), float(epsilon), float(alpha), int(seed), int(device)
        self._h = C.c_void_p()
        check(lib.bb_trainer_create(C.byref(d), w.ctypes.data_as(C.c_void_p), w.size, C.byref(self._h)))
        self.n_params, self.device, self.game = w.size, device, game
        self.A = lib.bb_game_actions(_lib.GAME_IDS[game])

    def close(self):
        if getattr(self, '_h', None):
            lib.bb_trainer_destroy(self._h)
            self._h = None

    __del__ = close

    def step(self, planes, evals, policies, lr, noise=None):
        """planes int8 [B,H,W,C], evals [B], policies [B,A] (numpy or torch); one optimizer step."""
        dev = torch.device('cuda', self.device)
        p = torch.as_tensor(np.ascontiguousarray(planes, dtype=np.int8) if not torch.is_tensor(planes) else planes).to(dev, torch.int8).contiguous()
        e = torch.as_tensor(np.asarray(evals, dtype=np.float32) if not torch.is_tensor(evals) else evals).to(dev, torch.float32).contiguous()
        q = torch.as_tensor(np.asarray(policies, dtype=np.float32) if not torch.is_tensor(policies) else policies).to(dev, torch.float32).contiguous()
        B = p.shape[0]
        assert e.shape == (B,) and q.shape == (B, self.A), (p.shape, e.shape, q.shape)
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float32)
        torch.cuda.synchronize(dev)
        check(lib.bb_trainer_step(self._h, C.c_void_p(p.data_ptr()), C.c_void_p(e.data_ptr()), C.c_void_p(q.data_ptr()), B, float(lr),
                                  None if nz is None else nz.ctypes.data_as(C.c_void_p)))

    def step_stream(self, stream_dev, rec_offsets, lr, noise=None):
        """One step on the examples that start at rec_offsets (uint64 byte offsets) of a device-resident compact record
        stream (Engine.records_serialized(to_host=False) or parallel.allgather_serialized)."""
        off = np.ascontiguousarray(rec_offsets, dtype=np.uint64)
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float32)
        torch.cuda.synchronize(stream_dev.device)
        check(lib.bb_trainer_step_stream(self._h, C.c_void_p(stream_dev.data_ptr()), off.ctypes.data_as(C.c_void_p), int(off.size), float(lr),
                                         None if nz is None else nz.ctypes.data_as(C.c_void_p)))

    def losses(self):
        """(loss, lossEvaluation, lossPolicy, lossParam) of the last step's forward, before its update."""
        out = (C.c_float * 4)()
        check(lib.bb_trainer_losses(self._h, out))
        return tuple(float(x) for x in out)

    def weights(self):
        w = np.empty(self.n_params, np.float32)
        check(lib.bb_trainer_weights(self._h, w.ctypes.data_as(C.c_void_p), w.size))
        return w

    def set_weights(self, flat):
        w = np.ascontiguousarray(flat, dtype=np.float32)
        check(lib.bb_trainer_set_weights(self._h, w.ctypes.data_as(C.c_void_p), w.size))

    def grads(self):
        g = np.empty(self.n_params, np.float32)
        check(lib.bb_trainer_grads(self._h, g.ctypes.data_as(C.c_void_p), g.size))
        return g

    @property
    def steps(self):
        return int(lib.bb_trainer_steps(self._h))
