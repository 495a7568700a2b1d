"""CPU: host logic of the multi-GPU path (blackbird_b200/parallel.py) -- shard arithmetic, per-game
uniform streams, and the single all_gather of fixed-stride record buffers, run with world_size 2
over gloo."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_cover_all_games():
    from blackbird_b200.parallel import shard_range
    for n in (1, 7, 8, 4096, 16384, 131072, 13):
        for world in (1, 2, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes[0] == max(sizes)


def test_uniform_streams_depend_on_global_game_id_only():
    from blackbird_b200.parallel import game_uniforms, shard_range
    whole = game_uniforms(range(10), 42)
    for world in (2, 4):
        parts = [game_uniforms(range(*shard_range(10, r, world)), 42) for r in range(world)]
        assert np.array_equal(np.concatenate(parts), whole)
    assert np.array_equal(whole[3], np.random.RandomState(1003).random_sample(42))


def _fake_raw(rank, n=3, mp=5, sb=16, cap=35):
    rs = np.random.RandomState(77 + rank)
    return {'n_moves': rs.randint(1, mp, n).astype(np.int32), 'winner': rs.randint(-1, 3, n).astype(np.int32),
            'states': rs.randint(0, 256, (n, mp + 1, sb)).astype(np.uint8),
            'rec_off': rs.randint(0, cap, (n, mp + 2)).astype(np.uint32),
            'pool_actions': rs.randint(0, 7, (n, cap)).astype(np.uint16),
            'pool_plays': rs.randint(0, 2 ** 31, (n, cap)).astype(np.uint32) * 2 + 1,      # exercises the top bit
            'chosen': rs.randint(0, 4032, (n, mp)).astype(np.uint16)}


def _worker(rank, world, port, q):
    try:
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        from blackbird_b200.parallel import allgather_raw
        os.environ['MASTER_ADDR'] = '127.0.0.1'
        os.environ['MASTER_PORT'] = str(port)
        dist.init_process_group('gloo', rank=rank, world_size=world)
        got = allgather_raw(_fake_raw(rank))
        want = [_fake_raw(r) for r in range(world)]
        for k in got:
            exp = np.concatenate([w[k] for w in want])
            assert got[k].dtype == exp.dtype and got[k].shape == exp.shape, (k, got[k].dtype, got[k].shape)
            assert np.array_equal(got[k], exp), k
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, 'ok'))
    except Exception as e:       # noqa: BLE001
        q.put((rank, 'FAIL %r' % (e,)))


def test_allgather_of_record_buffers_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == [(0, 'ok'), (1, 'ok')], res


def test_allgather_single_process_is_identity():
    from blackbird_b200.parallel import allgather_raw
    raw = _fake_raw(0)
    got = allgather_raw(raw)
    for k in raw:
        assert np.array_equal(got[k], raw[k]) and got[k].dtype == raw[k].dtype


def test_selfplay_sharded_rows_slots_and_uniforms_with_a_stub_engine():
    """Host logic of selfplay_sharded without a GPU: game rows per rank, slot cap, per-game uniform rows, owner
    slices (single process; the engine is a stub that records the call)."""
    from blackbird_b200.parallel import game_uniforms, selfplay_sharded

    class StubEngine(object):
        max_plies = 42

        def __init__(self, slots):
            self.n_games = slots
            self.calls = []

        def selfplay(self, play_limit, temp, uniforms, n_rows=None):
            self.calls.append((play_limit, temp, np.array(uniforms), n_rows))

        def records_raw(self):
            return {'n_moves': np.zeros(self.calls[-1][3], np.int32)}

    for max_slots, want_slots in ((None, 10), (4, 4), (64, 10)):
        eng, raw, owners = selfplay_sharded(StubEngine, 10, 24, temp=1.0, seed0=1000, max_slots=max_slots)
        assert eng.n_games == want_slots and owners == [(0, 10)]
        (play_limit, temp, u, n_rows), = eng.calls
        assert (play_limit, temp, n_rows) == (24, 1.0, 10) and u.shape == (10, 42)
        assert np.array_equal(u, game_uniforms(range(10), 42, 1000))
        assert len(raw['n_moves']) == 10


def _worker_streams(rank, world, port, q):
    try:
        import torch
        import torch.distributed as dist
        sys.path.insert(0, ROOT)
        from blackbird_b200.parallel import allgather_serialized
        os.environ['MASTER_ADDR'] = '127.0.0.1'
        os.environ['MASTER_PORT'] = str(port)
        dist.init_process_group('gloo', rank=rank, world_size=world)
        mine = lambda r: np.random.RandomState(5 + r).randint(0, 256, 1000 + 613 * r).astype(np.uint8)   # ragged: every rank another length
        gathered, lens = allgather_serialized(torch.as_tensor(mine(rank)))
        assert lens.tolist() == [1000 + 613 * r for r in range(world)]
        assert gathered.shape[0] == world and gathered.shape[1] % 16 == 0 and gathered.shape[1] >= lens.max()
        for r in range(world):
            assert np.array_equal(gathered[r, :lens[r]].numpy(), mine(r)), r
            assert not gathered[r, lens[r]:].any()
        empty, lens0 = allgather_serialized(torch.empty((0,), dtype=torch.uint8) if rank else torch.as_tensor(mine(0)))
        assert lens0.tolist() == [1000] + [0] * (world - 1)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, 'ok'))
    except Exception as e:       # noqa: BLE001
        q.put((rank, 'FAIL %r' % (e,)))


def test_allgather_of_compact_record_streams_gloo_world2():
    """parallel.allgather_serialized: byte counts first, then the ragged streams padded to the longest."""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_streams, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(30)
    assert res == [(0, 'ok'), (1, 'ok')], res
