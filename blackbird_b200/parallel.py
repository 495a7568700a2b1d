"""Multi-GPU self-play: independent game shards, one collective to collect the examples.

The reference plays its games one after another in one process (Blackbird.py:238) and has no
distributed code; SURVEY.md 8(e) defines the sharding used here.  One process per GPU
(torch.distributed, NCCL over NVLink on the B200 box, gloo in the CPU tests):

* global game id g in [0, n_games) lives on rank g // per_rank (contiguous blocks); its move-sampling
  uniforms come from RandomState(seed0 + g), so every game's record bytes are identical for any
  number of GPUs (tests/test_parallel_host.py, tests/test_gpu_search.py);
* there is NO data-path collective during search -- the tree pools never talk to each other;
* after a batch of games one all_gather of the fixed-stride raw record buffers (bb_records_device
  views: packed states, sparse child visit counts, offsets, winners) delivers every example to
  every rank (the trainer rank reads them; DataManager.PutGames is called by rank 0 only).  The
  buffers are fixed-stride per slot, so no count exchange is needed; they stay on the GPU for NCCL.
"""
import numpy as np
import torch
import torch.distributed as dist

RAW_KEYS = ('n_moves', 'winner', 'states', 'rec_off', 'pool_actions', 'pool_plays', 'chosen')


def shard_range(n_games, rank, world):
    """Contiguous block of global game ids owned by `rank` (the last ranks may hold one game less)."""
    base, extra = divmod(n_games, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def game_uniforms(game_ids, max_plies, seed0=1000):
    """One RandomState stream per GLOBAL game id: the doubles np.random.choice would draw for the
    game's moves (MCTS.py:336-338)."""
    return np.stack([np.random.RandomState(seed0 + int(g)).random_sample(max_plies) for g in game_ids]) \
        if len(game_ids) else np.zeros((0, max_plies))


def _records_device_views(eng):
    """torch views (no copy) of the engine's record buffers on its GPU."""
    import ctypes as C
    from ._lib import lib, check
    ptrs = [C.c_void_p() for _ in range(7)]
    check(lib.bb_records_device(eng._h, *[C.byref(p) for p in ptrs]))
    n, mp, cap, sb = eng.rec_rows, eng.max_plies, lib.bb_records_pool_cap(eng._h), eng.state_bytes
    shapes = {'n_moves': ((n,), torch.int32), 'winner': ((n,), torch.int32), 'states': ((n, mp + 1, sb), torch.uint8),
              'rec_off': ((n, mp + 2), torch.int32), 'pool_actions': ((n, cap), torch.int16),
              'pool_plays': ((n, cap), torch.int32), 'chosen': ((n, mp), torch.int16)}
    out = {}
    for key, p in zip(RAW_KEYS, ptrs):
        shape, dtype = shapes[key]
        out[key] = _from_ptr(p.value, shape, dtype, eng.device)
    return out


def _from_ptr(ptr, shape, dtype, device):
    """Wrap a raw device pointer as a tensor through the CUDA array interface (no ownership)."""
    nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()

    class _Holder(object):
        __cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (int(ptr), False), 'version': 2}

    flat = torch.as_tensor(_Holder(), device=torch.device('cuda', device))
    return flat.view(dtype).view(*shape)


_NP_VIEW = {'rec_off': np.uint32, 'pool_actions': np.uint16, 'pool_plays': np.uint32, 'chosen': np.uint16}


def allgather_raw(raw, group=None, to_host=True):
    """raw: dict of per-rank buffers (torch tensors on this rank's device, or numpy arrays for the
    gloo/CPU tests), identical shapes on every rank.  Returns numpy arrays with a leading
    concatenation over ranks in rank order (= global game order for contiguous shards).

    One flat byte buffer per rank and ONE all_gather: the records are small (C4: ~0.9 KB per game),
    so the collective is latency-bound and a single launch is what matters.
    to_host=False returns the gathered [world, bytes] uint8 tensor without the device-to-host copy."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    tensors = {k: (torch.as_tensor(raw[k]) if not torch.is_tensor(raw[k]) else raw[k]) for k in RAW_KEYS}
    parts = [tensors[k].contiguous().view(-1).view(torch.uint8) for k in RAW_KEYS]
    flat = torch.cat(parts)
    if world > 1:
        gathered = torch.empty((world * flat.numel(),), dtype=torch.uint8, device=flat.device)
        dist.all_gather_into_tensor(gathered, flat, group=group)
        gathered = gathered.view(world, flat.numel())
    else:
        gathered = flat[None]
    if not to_host:                      # ranks that do not train keep the gathered buffer on the GPU
        return gathered
    host = gathered.cpu().numpy()
    out, off = {}, 0
    for k, p in zip(RAW_KEYS, parts):
        t = tensors[k]
        chunk = host[:, off:off + p.numel()]
        off += p.numel()
        arr = np.ascontiguousarray(chunk).view(torch.empty((), dtype=t.dtype).numpy().dtype)
        arr = arr.reshape((world * t.shape[0],) + tuple(t.shape[1:]))
        out[k] = arr.view(_NP_VIEW[k]) if k in _NP_VIEW else arr
    return out


def selfplay_sharded(make_engine, n_games, play_limit, temp=1.0, seed0=1000, group=None, gather=True, max_slots=None):
    """Play global games [0, n_games) across the ranks of `group` and all-gather the raw records.

    make_engine(n_slots) -> Engine on this rank's GPU (every rank must build the same configuration).
    Every rank plays the same number of game rows (the largest shard) so the gathered buffers are
    fixed-stride; surplus rows play a game that is dropped.  With max_slots a rank plays its rows on
    at most that many concurrent slots (a slot that ends a game takes the rank's next unplayed row,
    bb_selfplay_games): BASELINE configs[4b], 131072 games on 4096 slots per GPU.  Returns (engine,
    raw_all, owner_slices): raw_all as allgather_raw, owner_slices[r] = (first row of rank r in
    raw_all, games it owns)."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rows = shard_range(n_games, 0, world)[1]                # rank 0 holds the largest shard
    slots = rows if max_slots is None else max(1, min(rows, int(max_slots)))
    lo, hi = shard_range(n_games, rank, world)
    eng = make_engine(slots)
    u = np.ones((rows, eng.max_plies))
    u[:hi - lo] = game_uniforms(range(lo, hi), eng.max_plies, seed0)
    eng.selfplay(play_limit, temp, u, n_rows=rows)
    if gather and world > 1:
        raw_all = allgather_raw(_records_device_views(eng), group)
    else:
        raw_all = eng.records_raw()
    owners = []
    for r in range(world if (gather and world > 1) else 1):
        a, b = shard_range(n_games, r, world) if world > 1 else (0, n_games)
        owners.append((r * rows, b - a))
    return eng, raw_all, owners


def allgather_serialized(buf, group=None):
    """All-gather of the per-rank compact record streams (Engine.records_serialized(to_host=False): final protobuf
    bytes packed on the device).  The streams differ in length, so the ranks first exchange their byte counts (one
    8-byte all_gather), pad to the longest and gather once: the collective moves live record bytes plus at most the
    spread between the ranks ("exchange counts first or pad to max").  Returns (gathered uint8
    [world, max_len] on the device, lengths int64 [world] on the host); rank r's stream is gathered[r, :lengths[r]]."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return buf[None], np.array([buf.numel()], np.int64)
    n = torch.tensor([buf.numel()], dtype=torch.int64, device=buf.device)
    lens = torch.empty((world,), dtype=torch.int64, device=buf.device)
    dist.all_gather_into_tensor(lens, n, group=group)
    lens_h = lens.cpu().numpy()
    width = int(lens_h.max())
    width = (width + 15) // 16 * 16
    padded = torch.zeros((width,), dtype=torch.uint8, device=buf.device)
    padded[:buf.numel()] = buf
    gathered = torch.empty((world * width,), dtype=torch.uint8, device=buf.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return gathered.view(world, width), lens_h


def generate_training_samples_sharded(model, nGames, temp, group=None, seed0=None, max_slots=None, precision='bf16', gather=True):
    """Blackbird.GenerateTrainingSamples for one process per GPU: the nGames global games are split in contiguous
    blocks over the ranks (shard_range), every rank plays and serializes its block on its GPU and stores it through
    ITS OWN model.Conn (one sqlite sink per rank, written in parallel), and -- with gather -- the compact record
    streams are all-gathered so that every rank (the trainer) holds every example on its device.  seed0 keys the
    per-game move-sampling streams by global game id (records independent of the GPU count); None draws them from
    numpy's global stream like the reference.  Returns (records written by this rank, gathered, lengths)."""
    from . import Blackbird as _bb
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    lo, hi = shard_range(nGames, rank, world)
    gathered = lengths = None
    holder = {}
    u = None
    if seed0 is not None:
        from . import engine as _engine
        u = game_uniforms(range(lo, hi), _engine._DEFAULT_PLIES[model.Game.GameType], seed0)
    written = _bb.GenerateTrainingSamples(model, hi - lo, temp, uniforms=u, precision=precision, max_slots=max_slots,
                                          _device_stream=holder if gather else None) if hi > lo else 0
    if gather:
        buf = holder.get('buf')
        if buf is None:
            buf = torch.empty((0,), dtype=torch.uint8, device=torch.device('cuda', model.DEVICE))
        gathered, lengths = allgather_serialized(buf, group)
    return written, gathered, lengths
