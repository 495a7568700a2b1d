"""Network mirror: the inference half of the reference's TF session wrapper (Network.py:9-64).

getPolicy / getEvaluation take the reference's input array (int8 [n, H, W, C] from
BoardState.AsInputArray) and return policy f32 [A] / value f32 of the first row, computed by the
engine's network kernels (bb_net_forward) -- there is no TensorFlow and no CPU path here.  Weights
are a flat fp32 vector in the layout of netinit.py / DESIGN.md "weight layout"; a fresh Network is
randomly initialised like tf.global_variables_initializer (Network.py:24).  train() is one GPU
step of csrc/train.cu; saveModel / loadModel keep the reference's directory convention but store
the flat vector (TF Saver bundles are SURVEY.md 8f-4).
"""
import os

import numpy as np
import torch

from . import engine as _engine
from . import netinit


def _entropy(bits):
    """Seed material from the OS: independent of Python's and numpy's global streams, which callers seed to replay games."""
    return int.from_bytes(os.urandom(8), 'little') >> (64 - bits)


def states_from_planes(game, planes, device=0):
    """Packed states whose AsInputArray() equals `planes` (int8 [n, H, W, C]).  The planes do not
    carry everything (e.g. whose move it is in DragonChess), only what the network sees."""
    planes = np.asarray(planes)
    n = planes.shape[0]
    if game == 'DragonChess':
        from .DragonChess import BoardState
        cells = np.zeros((n, 64), np.int8)
        for code, ch in BoardState.piece_map.items():
            cells[planes[:, :, :, ch].reshape(n, 64) == 1] = code
        castle = sum((planes[:, 0, 0, 12 + i] == 1).astype(np.uint8) << i for i in range(4))
        second = planes[:, 0, 0, 16] == 1                            # plane 16: White's second move (A.18)
        player = np.where(second, 1, 2).astype(np.uint8)
        prev = np.ones(n, np.uint8)
        return _engine.pack(game, cells, player, prev, castle, device=device)
    h, w = planes.shape[1:3]
    cells = (planes[:, :, :, 0] + 2 * planes[:, :, :, 1]).reshape(n, h * w).astype(np.int8)
    player = np.where(planes[:, 0, 0, 2] == 1, 1, 2).astype(np.uint8)
    return _engine.pack(game, cells, player, np.zeros(n, np.uint8), None, device=device)


class Network(object):
    PRECISION = 'fp32'                 # 'fp32': 1e-5 parity kernel; 'bf16': tcgen05 kernel (bulk self-play default)

    def __init__(self, name, networkConstructor=None, tensorflowConfig={}, game=None, seed=None, device=0):
        self._net_name = name
        self._net_game = game
        self._net_device = device
        self._net_eng = None
        self._net_loaded_weights = None
        self._net_trainer = None
        self._net_dirty = False
        self.lastLosses = None
        self.batchCount = 0
        if self.loadModel(name, networkConstructor):                  # Network.py:21: a saved network wins
            return
        if networkConstructor is None:
            raise FileNotFoundError('no saved network %r and no networkConfig to build one' % name)
        self._net_shape = networkConstructor
        cin = _engine.plane_shape(game)[2]
        self.setWeights(netinit.random_weights(cin, networkConstructor.policyShape, blocks=networkConstructor.blocks,
                                               filters=networkConstructor.filters, dense=networkConstructor.dense,
                                               seed=_entropy(31) if seed is None else seed))   # unseeded, like tf.global_variables_initializer
        self.saveModel(name)                                          # Network.py:25

    # -- weights ----------------------------------------------------------------------------------
    def setWeights(self, flat):
        self._net_loaded_weights = np.ascontiguousarray(flat, dtype=np.float32)
        if self._net_eng is not None:
            self._load_network_into(self._net_eng, precision=self.PRECISION)
        eng = getattr(self, '_eng', None)                             # a live search engine (MCTS mirror)
        if eng is not None and eng.evaluator == 'net':
            self._load_network_into(eng)

    def getWeights(self):
        self._sync_weights()
        return self._net_loaded_weights

    def _load_network_into(self, eng, precision=None, epsilon=None, seed=None):
        self._sync_weights()
        seed = _entropy(48) if seed is None else seed     # TF's noise is unseeded too; never drawn from `random` / `np.random` (the arena goldens replay those streams)
        s = self._net_shape
        precision = precision or self.PRECISION
        if precision == 'bf16' and s.filters not in (16, 32, 64, 128, 256):
            precision = 'fp32'                            # tensor-core kernels exist for 16 / 32 / 64 / 128 / 256 filters; the fp32 CUDA kernels take any width
        eng.load_network(self._net_loaded_weights, s.blocks, s.filters, s.dense,
                         epsilon=s.epsilon if epsilon is None else epsilon, alpha=s.alpha, seed=seed,
                         precision=precision)

    def _forward(self, planes):
        self._sync_weights()
        if self._net_eng is None:
            self._net_eng = _engine.Engine(self._net_game, 'dynamic', n_games=1, evaluator='net', node_cap=64,
                                           device=self._net_device, lanes=1)
            self._load_network_into(self._net_eng)
        st = states_from_planes(self._net_game, planes, device=self._net_device)
        pol, val = self._net_eng.net_forward(st)
        torch.cuda.synchronize()
        return pol.cpu().numpy(), val.cpu().numpy()

    # -- reference API (Network.py:48-64) ---------------------------------------------------------
    def getEvaluation(self, state):
        return self._forward(state)[1][0]

    def getPolicy(self, state):
        return self._forward(state)[0][0]

    def train(self, state, eval, policy, learningRate=0.01, teacher=None):
        """Network.py:66-84: one sess.run(trainingOp) on the batch (state int8 [B,H,W,C], eval [B], policy [B,A]) -- forward,
        the loss of NetworkFactory.py:185-204 as written, gradients and the configured optimizer, all on the GPU
        (csrc/train.cu).  The TensorBoard summary every 10th batch (:81-84) becomes self.lastLosses."""
        if teacher is not None:
            # NetworkFactory.py:206-219 calls tf.log (gone in TF 2) and reduces a rank-1 tensor over axis 1: the teacher term
            # cannot run in the reference's own environment (hasTeacher is false in parameters_template.yaml)
            raise NotImplementedError('the teacher cross-entropy term (NetworkFactory.py:206-219) does not run in the reference either')
        if self._net_trainer is None:
            from .trainer import Trainer
            s = self._net_shape
            self._net_trainer = Trainer(self._net_game, self._net_loaded_weights, s.blocks, s.filters, s.dense,
                                        optimizer=getattr(s, 'optimizer', 'sgd'), momentum=getattr(s, 'momentum', 0.9),
                                        epsilon=s.epsilon, alpha=s.alpha, seed=_entropy(48), device=self._net_device)
        self._net_trainer.step(state, eval, policy, learningRate)
        self._net_dirty = True
        self.batchCount += 1
        if self.batchCount % 10 == 0:
            self.lastLosses = self._net_trainer.losses()

    def train_stream(self, stream_dev, rec_offsets, learningRate=0.01):
        """One train() step whose batch is assembled on the GPU from a compact record stream (byte offsets of the examples)."""
        if self._net_trainer is None:
            from .trainer import Trainer
            s = self._net_shape
            self._net_trainer = Trainer(self._net_game, self._net_loaded_weights, s.blocks, s.filters, s.dense,
                                        optimizer=getattr(s, 'optimizer', 'sgd'), momentum=getattr(s, 'momentum', 0.9),
                                        epsilon=s.epsilon, alpha=s.alpha, seed=_entropy(48), device=self._net_device)
        self._net_trainer.step_stream(stream_dev, rec_offsets, learningRate)
        self._net_dirty = True
        self.batchCount += 1
        if self.batchCount % 10 == 0:
            self.lastLosses = self._net_trainer.losses()

    def _sync_weights(self):
        """Pull the trained weights back from the trainer before anything else uses them."""
        if self._net_dirty:
            self._net_dirty = False
            self.setWeights(self._net_trainer.weights())

    def saveModel(self, name=None):
        name = name or getattr(self, 'Name', None) or self._net_name   # Network.py:90-91 uses Model.Name
        self._sync_weights()
        saveDir = os.path.join('blackbird_models', name)
        os.makedirs(saveDir, exist_ok=True)
        s = self._net_shape
        # the reference's own format next to it: a TensorFlow checkpoint bundle (best.index + best.data-00000-of-00001) with the
        # reference graph's variable names, which tf.compat.v1.train.Saver.restore reads (Network.py:86-112; tfbundle.py)
        from . import tfbundle
        cin = _engine.plane_shape(self._net_game)[2]
        tfbundle.write_bundle(os.path.join(saveDir, 'best'),
                              tfbundle.flat_to_tensors(self._net_loaded_weights, cin, s.policyShape, s.blocks, s.filters, s.dense))
        np.savez(os.path.join(saveDir, 'best.npz'), weights=self._net_loaded_weights,
                 shape=np.array([s.blocks, s.filters, s.dense, s.policyShape]), noise=np.array([s.epsilon, s.alpha]),
                 training=np.array([str(getattr(s, 'optimizer', 'sgd')), str(getattr(s, 'momentum', 0.9))]))

    def loadModel(self, name, networkConstructor=None):
        path = os.path.join('blackbird_models', name, 'best.npz')
        if not os.path.isfile(path):
            # a checkpoint saved by the reference (TensorFlow Saver bundle, Network.py:86-99): weights and shape come from the
            # bundle, noise / optimizer settings from the caller's configuration (they live in the graph, not in the bundle)
            prefix = os.path.join('blackbird_models', name, 'best')
            if not os.path.isfile(prefix + '.index'):
                return False
            from . import tfbundle
            from .NetworkFactory import NetworkFactory
            flat, (cin, a, b, f, d) = tfbundle.tensors_to_flat(tfbundle.read_bundle(prefix))
            if cin != _engine.plane_shape(self._net_game)[2]:
                raise ValueError('checkpoint %s was trained on %d input planes, %s has %d' % (prefix, cin, self._net_game, _engine.plane_shape(self._net_game)[2]))
            cfg = dict(getattr(networkConstructor, 'networkConfig', {}) or {})
            cfg.update({'blocks': b, 'filters': f, 'eval': {'dense': d}})
            cfg.setdefault('policy', {'dirichlet': {'epsilon': 0.0, 'alpha': 0.2}})
            self._net_shape = NetworkFactory(cfg, a)
            self.setWeights(flat)
            return True
        from .NetworkFactory import NetworkFactory
        z = np.load(path)
        b, f, d, a = (int(x) for x in z['shape'])
        training = {'optimizer': str(z['training'][0]), 'momentum': float(z['training'][1])} if 'training' in z else {}
        self._net_shape = NetworkFactory({'blocks': b, 'filters': f, 'eval': {'dense': d}, 'training': training,
                                          'policy': {'dirichlet': {'epsilon': float(z['noise'][0]), 'alpha': float(z['noise'][1])}}}, a)
        self.setWeights(z['weights'])
        return True

    def __del__(self):
        eng = getattr(self, '_net_eng', None)
        if eng is not None:
            eng.close()
        tr = getattr(self, '_net_trainer', None)
        if tr is not None:
            tr.close()
