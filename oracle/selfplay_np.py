"""Oracle self-play loop and record format: restatement of Blackbird.GenerateTrainingSamples
(Blackbird.py:219-268) and ExampleState.SerializeState (Blackbird.py:68-81, proto/state.proto:3-9).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
import struct

import numpy as np


def _varint(n):
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _bytes_field(number, payload):
    if not payload:
        return b''                                   # proto3: default values are not emitted
    return bytes([(number << 3) | 2]) + _varint(len(payload)) + payload


def dims_int8(shape):
    """np.array(shape, dtype=np.int8).tobytes() with legacy wrap-around.

    numpy 2 raises OverflowError for DragonChess' policyDims=[4032]; legacy
    numpy wrapped to 0xC0.  The engine emits the wrapped byte and so does the oracle.
    """
    return bytes((int(d) & 0xFF) for d in shape)


def serialize_record(evaluation, policy, planes):
    """State{float mctsEval=1; bytes mctsPolicy=2; bytes boardEncoding=3; bytes boardDims=4;
    bytes policyDims=5} in field order (state.proto:3-9)."""
    out = b''
    ev = float(evaluation)
    if struct.pack('<f', ev) != b'\x00\x00\x00\x00':
        out += b'\x0d' + struct.pack('<f', ev)
    out += _bytes_field(2, np.ascontiguousarray(policy, dtype=np.float64).tobytes())
    out += _bytes_field(3, np.ascontiguousarray(planes, dtype=np.int8).tobytes())
    out += _bytes_field(4, dims_int8(planes.shape))
    out += _bytes_field(5, dims_int8(policy.shape))
    return out


def play_game(search, game_cls, temp, max_plies=None):
    """One iteration of the loop at Blackbird.py:238-267.

    Returns (records: list[bytes], winner, trace) where trace holds, per move,
    (child plays f64[A], chosen action, Root.WinRate(), Root.Plays) for KAT tests.
    winner is None only when ``max_plies`` truncated the game (engine-side cap, A.20).
    """
    history = []
    trace = []
    state = game_cls()
    winner = None
    search.DropRoot()
    probs = None
    while winner is None:
        nxt, v, probs = search.FindMove(state, temp)
        root = search.Root
        trace.append((root.ChildPlays().copy(), search.LastAction, v, root.Plays))
        history.append([1 - v, probs, state.AsInputArray(), state.Player])
        state = nxt
        search.MoveRoot(state)
        winner = state.Winner(None)                  # lastAction is always None (Blackbird.py:242,253)
        if max_plies is not None and len(history) >= max_plies and winner is None:
            break
    history.append([None, np.zeros([len(probs)]), state.AsInputArray(), state.Player])
    for ex in history:                               # Blackbird.py:260-264 (A.27)
        if winner is None:
            ex[0] = 0                                # truncated: engine convention, not in the reference
        elif winner == 0:
            ex[0] = 0
        else:
            ex[0] = 1 if ex[3] == winner else -1
    return [serialize_record(e[0], e[1], e[2]) for e in history], winner, trace


class RandomSearch(object):
    """RandomMCTS.py:4-25: the opponent of Blackbird.TestRandom.  Two draws from the global
    numpy stream per move (the action, then the reported win rate); no tree."""

    def __init__(self, rng=None):
        self.Rng = rng if rng is not None else np.random
        self.Root = None

    def DropRoot(self):
        self.Root = None                              # inherited MCTS.DropRoot (MCTS.py:141-144)

    def FindMove(self, state, *args, **kwargs):
        legal = state.LegalActions()
        action = self.Rng.choice([i for i in range(len(legal)) if legal[i] == 1])    # RandomMCTS.py:10-11
        winRate = self.Rng.random()                                                 # RandomMCTS.py:12
        childProbability = state.LegalActions()
        s = sum(childProbability)
        if s > 0:
            childProbability /= s
        nxt = state.Copy()
        nxt.ApplyAction(action)
        return nxt, winRate, childProbability

    def MoveRoot(self, *args, **kwargs):
        return

    def ResetRoot(self, *args, **kwargs):
        return


def arena_game(model1, model2, game_cls, temp, on_state=None):
    """One iteration of Blackbird.TestModels (Blackbird.py:195-216): +1 / 0 / -1 for model1.
    ``on_state`` sees every position both players are moved to (test hook, not in the reference)."""
    import random
    model1ToMove = random.choice([True, False])       # Blackbird.py:196, the stdlib stream
    model1Player = 1 if model1ToMove else 2
    winner = None
    model1.DropRoot()
    model2.DropRoot()
    state = game_cls()
    while winner is None:
        mover = model1 if model1ToMove else model2
        nextState = mover.FindMove(state, temp)[0]
        state = nextState
        if on_state is not None:
            on_state(state)
        model1.MoveRoot(state)
        model2.MoveRoot(state)
        model1ToMove = not model1ToMove
        winner = state.Winner()
    if winner == model1Player:
        return 1
    elif winner == 0:
        return 0
    return -1


def arena_random(model, game_cls, temp, numTests, on_state=None):
    """Blackbird.TestRandom (Blackbird.py:84-112): (tally, per-game results)."""
    names = {1: 'wins', 0: 'draws', -1: 'losses'}
    stats, results = {}, []
    for _ in range(numTests):
        r = arena_game(model, RandomSearch(), game_cls, temp, on_state)
        stats[names[r]] = stats.get(names[r], 0) + 1
        results.append(r)
    return stats, results
