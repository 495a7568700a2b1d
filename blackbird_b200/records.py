"""(state, visit-distribution, outcome) records in the reference's wire format.

Replaces ExampleState.SerializeState (Blackbird.py:68-81; proto/state.proto:3-9) and the
labelling loop of GenerateTrainingSamples (Blackbird.py:256-266).  The engine keeps, per move,
the packed root state and the sparse child visit counts; planes are decoded on the GPU
(bb_game_planes), probabilities are plays / sum(plays) in IEEE f64 exactly as
Node.ChildProbability (MCTS.py:66-68), and the protobuf bytes are emitted by hand (five
fields, fixed order) so no protobuf runtime is needed.
"""
import struct

import numpy as np
import torch

from . import engine as _engine


def _varint(n):
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | 0x80 if n else b)
        if not n:
            return bytes(out)


def _field(number, payload):
    return (bytes([(number << 3) | 2]) + _varint(len(payload)) + payload) if payload else b''


def encode_record(evaluation, policy_f64, planes_i8):
    """State{float mctsEval=1; bytes mctsPolicy=2; bytes boardEncoding=3; bytes boardDims=4; bytes policyDims=5}.

    boardDims / policyDims are int8 in the reference; DragonChess' policyDims [4032] does not fit
    (numpy 2 raises, legacy numpy wrapped): the wrapped byte 0xC0 is emitted."""
    ev = struct.pack('<f', float(evaluation))
    out = (b'\x0d' + ev) if ev != b'\x00\x00\x00\x00' else b''
    out += _field(2, policy_f64.tobytes())
    out += _field(3, planes_i8.tobytes())
    out += _field(4, bytes((d & 0xFF) for d in planes_i8.shape))
    out += _field(5, bytes((d & 0xFF) for d in policy_f64.shape))
    return out


def build_records(eng, raw=None):
    """-> list over games of (records: list[bytes], winner code, n_moves).  `raw` defaults to this engine's
    buffers; the all-gathered buffers of several GPUs (parallel.allgather_raw) have more rows than eng.n_games."""
    raw = raw or eng.records_raw()
    game, A = eng.game, eng.A
    h, w, c = _engine.plane_shape(game)
    n_moves = raw['n_moves']
    # decode every recorded state's planes + player in one batched kernel call
    idx = [(g, r) for g in range(len(n_moves)) for r in range(int(n_moves[g]) + 1)]
    flat = np.stack([raw['states'][g, r] for g, r in idx]) if idx else np.zeros((0, eng.state_bytes), np.uint8)
    st = torch.as_tensor(flat).to(torch.device('cuda', eng.device))
    planes = _engine.planes(game, st).cpu().numpy()
    players = _engine.unpack(game, st)[1].cpu().numpy()
    out = []
    k = 0
    for g in range(len(n_moves)):
        winner = int(raw['winner'][g])
        recs = []
        for r in range(int(n_moves[g]) + 1):
            lo, hi = int(raw['rec_off'][g, r]), int(raw['rec_off'][g, r + 1])
            policy = np.zeros(A, dtype=np.float64)
            if hi > lo:
                plays = np.zeros(A, dtype=np.float64)
                plays[raw['pool_actions'][g, lo:hi]] = raw['pool_plays'][g, lo:hi]
                total = sum(plays)                          # builtin sum, as MCTS.py:66
                if total > 0:
                    policy = plays / total
            if winner <= 0:                                 # draw, or truncated by the ply cap
                label = 0
            else:
                label = 1 if int(players[k]) == winner else -1      # Blackbird.py:260-264
            recs.append(encode_record(label, policy, planes[k][None]))
            k += 1
        out.append((recs, winner, int(n_moves[g])))
    return out


def split_records(buf, game_off, rec_bytes):
    """Compact record stream (Engine.records_serialized) -> list over games of list[bytes] (what PutGames takes)."""
    mv = memoryview(np.ascontiguousarray(buf)).cast('B')
    out = []
    for g in range(len(rec_bytes)):
        lo, hi, sz = int(game_off[g]), int(game_off[g + 1]), int(rec_bytes[g])
        out.append([bytes(mv[o:o + sz]) for o in range(lo, hi, sz)])
    return out


def iter_records(buf, game_off, rec_bytes):
    """The same records as one flat generator of bytes objects (sqlite parameter stream), no per-game lists."""
    mv = memoryview(np.ascontiguousarray(buf)).cast('B')
    for g in range(len(rec_bytes)):
        sz = int(rec_bytes[g])
        for o in range(int(game_off[g]), int(game_off[g + 1]), sz):
            yield bytes(mv[o:o + sz])
