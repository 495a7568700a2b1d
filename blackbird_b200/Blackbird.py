"""Blackbird mirror: Model, GenerateTrainingSamples and ExampleState (Blackbird.py:21-81, 219-268,
315-389) over the CUDA engine.

GenerateTrainingSamples is the north-star path.  The reference plays nGames one after another on
one Python thread (Blackbird.py:238); here they are game slots of ONE engine and run concurrently
on the GPU -- FindMove (PUCT descent, expansion, network evaluation, backup), the move choice,
MoveRoot, Winner and the example bookkeeping all happen on the device (bb_selfplay_run) -- and the
finished (state, visit-distribution, outcome) records are serialized in the reference's protobuf
wire format and written through model.Conn.PutGames exactly as before.
"""
import random
import struct
from collections import defaultdict

import numpy as np

from . import engine as _engine
from . import records as _records
from .DataManager import Connection
from .DynamicMCTS import DynamicMCTS as MCTS
from .FixedMCTS import FixedMCTS
from .Network import Network
from .NetworkFactory import NetworkFactory
from .RandomMCTS import RandomMCTS

np.seterr(divide='ignore', invalid='ignore')          # Blackbird.py:17


class ExampleState(object):
    """Blackbird.py:21-81.  The wire format is proto/state.proto:3-9, written and parsed by hand."""

    def __init__(self, evaluation, policy, board, player=None):
        self.MctsPolicy = policy
        self.MctsEval = evaluation
        self.Board = board
        self.Player = player

    @classmethod
    def FromSerialized(cls, serialState):
        fields = {1: 0.0, 2: b'', 3: b'', 4: b'', 5: b''}
        i, buf = 0, bytes(serialState)
        while i < len(buf):
            tag = buf[i]
            i += 1
            number, wire = tag >> 3, tag & 7
            if wire == 5:
                fields[number] = struct.unpack_from('<f', buf, i)[0]
                i += 4
            elif wire == 2:
                n, shift = 0, 0
                while True:
                    b = buf[i]
                    i += 1
                    n |= (b & 0x7F) << shift
                    shift += 7
                    if not b & 0x80:
                        break
                fields[number] = buf[i:i + n]
                i += n
            else:
                raise ValueError('unexpected wire type %d in a State record' % wire)
        boardDims = np.frombuffer(fields[4], dtype=np.int8)
        policy = np.frombuffer(fields[2], dtype=np.float64)
        # policyDims is int8 in the reference and overflows for DragonChess (A.22): trust the payload length
        board = np.frombuffer(fields[3], dtype=np.int8).reshape(boardDims)
        return cls((fields[1],), policy, board)           # the reference keeps mctsEval as a 1-tuple (Blackbird.py:60)

    def SerializeState(self):
        return _records.encode_record(self.MctsEval, np.asarray(self.MctsPolicy, dtype=np.float64),
                                      np.asarray(self.Board, dtype=np.int8))


def _uniform_table(n_games, max_plies):
    """One double per possible move and game from the global numpy stream -- the source
    np.random.choice draws from in MCTS._selectAction (MCTS.py:336-338)."""
    return np.random.random_sample((n_games, max_plies))


# slots per engine batch: bounded by tree-pool memory (2 half-arenas x node_cap x 32 B per slot)
def _batch_slots(game, search, play_limit, max_depth, n_games, budget_bytes=48 << 30):
    per = 2 * 32 * _engine.default_node_cap(game, search, play_limit, max_depth)
    return int(max(1, min(n_games, budget_bytes // per, 65536)))


def GenerateTrainingSamples(model, nGames, temp, uniforms=None, precision='bf16', max_slots=None, _device_stream=None):
    """Blackbird.py:219-268.  Plays nGames self-play games with model's search settings and network
    and stores every example through model.Conn.PutGames.

    temp 1 (the self-play setting, main.py:26) and temp 0 choose moves on the device.  `uniforms`
    [nGames, max_plies] overrides the move-sampling doubles (tests use per-game RandomState streams).
    Returns the number of records written (the reference returns None)."""
    if nGames <= 0:
        raise ValueError('Use a positive integer for number of games.')
    if temp not in (0, 1):
        raise ValueError('device self-play supports temp 1 (sample by visit share) and 0 (PUCT argmax); got %r' % (temp,))
    game = model.Game.GameType
    search = model._KIND
    play_limit = model.PlayLimit
    if play_limit is None:
        raise ValueError('Not enough information to decide a stop time.')     # time-boxed search is per-object only (A.25)
    max_depth = getattr(model, 'MaxDepth', None) or 10
    slots = _batch_slots(game, search, play_limit, max_depth, nGames)
    if max_slots:
        slots = max(1, min(slots, int(max_slots)))         # concurrent games on the GPU (default: what the memory budget allows)
    evaluator = getattr(model, '_SELFPLAY_EVALUATOR', 'net')   # 'kat'  only in parity tests
    # networks wider than the template's 16 filters run tick by tick (no persistent kernel pair): eight lanes keep the tensor
    # cores busy with one lane's network launch while the others search (measured: 1.8 -> 2.6 M sims/s at 256 filters)
    wide = evaluator == 'net' and int(getattr(getattr(model, '_net_shape', None), 'filters', 16) or 16) != 16
    eng = _engine.Engine(game, search, n_games=slots, max_depth=max_depth, exploration_rate=model.ExplorationRate,
                         evaluator=evaluator, play_limit_hint=play_limit, device=model.DEVICE, lanes=8 if wide and slots >= 64 else 0)
    written = 0
    try:
        if evaluator == 'net':
            model._load_network_into(eng, precision=precision)      # noise seed from the OS (Network._entropy)
        # one engine call for all nGames: the slots (as many as fit the memory budget) start the next unplayed game
        # as soon as they end one (bb_selfplay_games), so the GPU stays full until the last games run out
        u = _uniform_table(nGames, eng.max_plies) if uniforms is None else np.asarray(uniforms)[:nGames]
        eng.selfplay(play_limit, float(temp), u)
        # ExampleState.SerializeState + the outcome labels run on the device (bb_records_serialize): one compact byte
        # stream comes back and goes to the sink in ONE transaction
        if _device_stream is not None:                      # parallel.generate_training_samples_sharded: keep the device copy for the all-gather
            dev, off, sz = eng.records_serialized(to_host=False)
            _device_stream['buf'] = dev
            buf = dev.cpu().numpy()
        else:
            buf, off, sz = eng.records_serialized()
        written += int(((off[1:] - off[:-1]) // sz).sum())
        model.Conn.PutGames(model.Name, model.Version, game, _records.iter_records(buf, off, sz))
    finally:
        eng.close()
    return written


def TrainWithExamples(model, batchSize, learningRate, epochs=1, teacher=None, model_override=None, version_override=None):
    """Blackbird.py:271-312: every stored example of (model, version) in a random order, truncated to a multiple of
    batchSize, one Network.train call per batch; then Version += 1 and the new version is registered (the weights are
    not saved, like the reference).  `epochs` is accepted and ignored, like the reference.  The permutation comes from
    numpy's global stream exactly as np.random.choice(examples, n, replace=False) draws it (permutation(len)[:n])."""
    states = model.Conn.GetGames(model_override if model_override is not None else model.Name,
                                 version_override if version_override is not None else model.Version)
    examples = [ExampleState.FromSerialized(state) for state in states]
    n = len(examples) - (len(examples) % batchSize)
    order = np.random.choice(len(examples), n, replace=False) if len(examples) else []
    for i in range(n // batchSize):
        batch = [examples[j] for j in order[i * batchSize:(i + 1) * batchSize]]
        model.train(np.vstack([b.Board for b in batch]), np.hstack([b.MctsEval for b in batch]),
                    np.vstack([b.MctsPolicy for b in batch]), learningRate, teacher)
    model.Version += 1
    model.Conn.PutModel(model.Game.GameType, model.Name, model.Version)


def TrainOnStream(model, stream_dev, game_off, rec_bytes, batchSize, learningRate, seed=None):
    """The same loop fed from a device-resident compact record stream (Engine.records_serialized(to_host=False), or
    the all-gathered streams of parallel.generate_training_samples_sharded) instead of sqlite: the batches are assembled
    on the GPU from byte offsets (bb_trainer_step_stream), no example is parsed on the host ("consuming
    the all-gathered buffer directly").  Returns the number of batches."""
    rec_off = np.concatenate([np.arange(int(game_off[g]), int(game_off[g + 1]), int(rec_bytes[g]), dtype=np.uint64)
                              for g in range(len(rec_bytes))]) if len(rec_bytes) else np.zeros(0, np.uint64)
    n = rec_off.size - (rec_off.size % batchSize)
    rs = np.random if seed is None else np.random.RandomState(seed)
    order = rs.choice(rec_off.size, n, replace=False) if n else []
    if n:
        model.train_stream(stream_dev, rec_off[order[:batchSize]], learningRate)     # creates the trainer
        for i in range(1, n // batchSize):
            model.train_stream(stream_dev, rec_off[order[i * batchSize:(i + 1) * batchSize]], learningRate)
    model.Version += 1
    model.Conn.PutModel(model.Game.GameType, model.Name, model.Version)
    return n // batchSize


def TestModels(model1, model2, temp, numTests):
    """Blackbird.py:177-216: both engines follow the game with FindMove / MoveRoot; +1 / 0 / -1 for model1."""
    for _ in range(numTests):
        model1ToMove = random.choice([True, False])
        model1Player = 1 if model1ToMove else 2
        winner = None
        model1.DropRoot()
        model2.DropRoot()
        state = model1.Game()
        while winner is None:
            mover = model1 if model1ToMove else model2
            (nextState, *_) = mover.FindMove(state, temp)
            state = nextState
            model1.MoveRoot(state)
            model2.MoveRoot(state)
            model1ToMove = not model1ToMove
            winner = state.Winner()
        if winner == model1Player:
            return 1
        elif winner == 0:
            return 0
        else:
            return -1


def TestPrevious(model, temp, numTests):
    """Blackbird.py:115-143."""
    oldModel = model.LastVersion()
    resultMap = {1: 'wins', 0: 'draws', -1: 'losses'}
    stats = defaultdict(int)
    for _ in range(numTests):
        result = TestModels(model, oldModel, temp, numTests=1)
        stats[resultMap.get(result, 'indeterminant')] += 1
        model.Conn.PutTrainingStatistic(result, model.Name, model.Version, oldModel.Name, oldModel.Version)
    return stats


def TestRandom(model, temp, numTests):
    """Blackbird.py:84-112: numTests arena games against RandomMCTS, one statistics row per game."""
    resultMap = {1: 'wins', 0: 'draws', -1: 'losses'}
    stats = defaultdict(int)
    for _ in range(numTests):
        result = TestModels(model, RandomMCTS(), temp, numTests=1)
        stats[resultMap.get(result, 'indeterminant')] += 1
        model.Conn.PutTrainingStatistic(result, model.Name, model.Version, 'RANDOM')
    return stats


def TestGood(model, temp, numTests, good=None):
    """Blackbird.py:146-174: numTests arena games against a plain rollout MCTS, FixedMCTS(maxDepth=10,
    explorationRate=0.85, timeLimit=1) -- all-ones priors and a random rollout per leaf visit (MCTS.py:346-383), searched
    on the device.  `good` overrides the opponent (the parity tests use a play limit instead of the wall-clock second,
    which is the reference's only non-deterministic input)."""
    good = good if good is not None else FixedMCTS(maxDepth=10, explorationRate=0.85, timeLimit=1)
    resultMap = {1: 'wins', 0: 'draws', -1: 'losses'}
    stats = defaultdict(int)
    for _ in range(numTests):
        result = TestModels(model, good, temp, numTests=1)
        stats[resultMap.get(result, 'indeterminant')] += 1
        model.Conn.PutTrainingStatistic(result, model.Name, model.Version, 'MCTS')
    return stats


class Model(MCTS, Network):
    """Blackbird.py:315-389: tree search driven by the policy-value network.  The reference's Model
    is a DynamicMCTS (Blackbird.py:1); FixedModel below is the same class over FixedMCTS (A.24)."""

    def __init__(self, game, name, mctsConfig, networkConfig={}, tensorflowConfig={}):
        self.Conn = Connection()
        self.Game = game
        self.Name = name
        self.Version = self.Conn.GetLastVersion(self.Game.GameType, self.Name)
        self._saveName = self.Name + '_' + str(self.Version)
        self.MCTSConfig = mctsConfig
        self.NetworkConfig = networkConfig
        self.TensorflowConfig = tensorflowConfig
        search_base = next(k for k in type(self).__mro__ if k in (MCTS, FixedMCTS))   # DynamicMCTS or FixedMCTS
        search_base.__init__(self, **mctsConfig)
        if networkConfig != {}:
            Network.__init__(self, self._saveName, NetworkFactory(networkConfig, game.LegalMoves), tensorflowConfig,
                             game=game.GameType, device=self.DEVICE)
        else:
            Network.__init__(self, self._saveName, tensorflowConfig=tensorflowConfig, game=game.GameType, device=self.DEVICE)

    def LastVersion(self):
        return type(self)(self.Game, self.Name, self.MCTSConfig, self.NetworkConfig, self.TensorflowConfig)

    def SampleValue(self, state, player):
        """Blackbird.py:350-370 (host restatement for callers; the search computes it on the device)."""
        value = self.getEvaluation(state.AsInputArray())
        value = (value + 1) * 0.5
        if state.Player != player:
            value = 1 - value
        assert value >= 0, 'Value: {}'.format(value)
        return value

    def GetPriors(self, state):
        """Blackbird.py:372-389."""
        policy = self.getPolicy(state.AsInputArray()) * state.LegalActions()
        policy /= np.sum(policy)
        return policy

    def __del__(self):
        MCTS.__del__(self)
        Network.__del__(self)


class FixedModel(FixedMCTS, Network):
    """Model with FixedMCTS as its search base ("FixedMCTS self-play", BASELINE configs 1, 2, 4)."""
    __init__ = Model.__init__
    LastVersion = Model.LastVersion
    SampleValue = Model.SampleValue
    GetPriors = Model.GetPriors

    def __del__(self):
        FixedMCTS.__del__(self)
        Network.__del__(self)
