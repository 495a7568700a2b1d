"""GPU: bb_records_serialize_static -- the allocation-free variant of bb_records_serialize (scratch in a file-scope
`static __device__` array; no counterpart in the reference) -- against the
existing entry point, through the C-ABI, on two calls in a row.  Byte work: the tolerance is equality.

Collected last on purpose: a failure here must not hide any other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip('torch')


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import _lib, engine
    return _lib, engine


def _static(lib, eng, cap_slack=64):
    """-> (bytes uint8 numpy, game_off uint64 numpy, rec_bytes uint32 numpy) through bb_records_serialize_static;
    every buffer is allocated by the caller BEFORE the call."""
    rows = eng.rec_rows
    n_rec, n_bytes = C.c_uint64(), C.c_uint64()
    off_h, sz_h = np.empty(rows + 1, np.uint64), np.empty(rows, np.uint32)
    lib.check(lib.lib.bb_records_serialize(eng._h, None, 0, C.byref(n_rec), C.byref(n_bytes), off_h.ctypes.data_as(C.c_void_p),
                                           sz_h.ctypes.data_as(C.c_void_p)))                 # sizes only (existing entry point)
    dev = torch.device('cuda', eng.device)
    out = torch.full((int(n_bytes.value) + cap_slack,), 0xAB, dtype=torch.uint8, device=dev)
    off = torch.full((rows + 1,), -1, dtype=torch.int64, device=dev)
    sz = torch.full((rows,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    r2, b2 = C.c_uint64(), C.c_uint64()
    lib.check(lib.lib.bb_records_serialize_static(eng._h, C.c_void_p(out.data_ptr()), out.numel(), C.c_void_p(off.data_ptr()),
                                                  C.c_void_p(sz.data_ptr()), C.byref(r2), C.byref(b2)))
    assert (r2.value, b2.value) == (n_rec.value, n_bytes.value)
    tail = out[int(b2.value):].cpu().numpy()
    assert (tail == 0xAB).all()                                        # nothing written past the stream
    return (out[:int(b2.value)].cpu().numpy(), off.cpu().numpy().view(np.uint64), sz.cpu().numpy().view(np.uint32))


def test_static_scratch_serialization_equals_the_allocating_one(mods):
    lib, engine = mods
    eng = engine.Engine('Connect4', 'fixed', n_games=12, evaluator='kat', play_limit_hint=24)
    for call, (rows, seed) in enumerate(((40, 1000), (17, 5000))):    # two calls in a row, the second with fewer, different games
        u = np.stack([np.random.RandomState(seed + g).random_sample(eng.max_plies) for g in range(rows)])
        eng.selfplay(24, 1.0, u, n_rows=rows)
        want, want_off, want_sz = eng.records_serialized()             # bb_records_serialize (cudaMalloc'ed scratch)
        got, got_off, got_sz = _static(lib, eng)
        assert np.array_equal(got_off, want_off) and np.array_equal(got_sz, want_sz), call
        assert got.tobytes() == np.asarray(want).tobytes(), call
        again, _, _ = _static(lib, eng)                                # the array keeps the last call's contents: same answer
        assert again.tobytes() == got.tobytes()
    eng.close()


def test_static_scratch_limits_are_reported(mods):
    lib, engine = mods
    assert lib.lib.bb_records_static_max_rows() == 65536
    eng = engine.Engine('TicTacToe', 'dynamic', n_games=64, evaluator='kat', play_limit_hint=6)
    u = np.random.RandomState(3).random_sample((64, eng.max_plies))
    eng.selfplay(6, 1.0, u)
    want, _, _ = eng.records_serialized()
    dev = torch.device('cuda', eng.device)
    small = torch.zeros((len(want) - 1,), dtype=torch.uint8, device=dev)
    need = C.c_uint64()
    rc = lib.lib.bb_records_serialize_static(eng._h, C.c_void_p(small.data_ptr()), small.numel(), None, None, None, C.byref(need))
    assert rc == lib.BB_EINVAL and need.value == len(want) and b'needed' in lib.lib.bb_last_error()
    assert not small.any().item()                                      # nothing was written into the short buffer
    rc = lib.lib.bb_records_serialize_static(eng._h, None, 0, None, None, None, None)
    assert rc == lib.BB_EINVAL
    eng.close()
    # more game rows than the array holds: rejected with a clear error, the allocating entry point still serves them
    rows = 65536 + 64
    eng = engine.Engine('TicTacToe', 'dynamic', n_games=4096, evaluator='kat', play_limit_hint=6)
    eng.selfplay(6, 1.0, np.random.RandomState(4).random_sample((rows, eng.max_plies)), n_rows=rows)
    big = torch.zeros((rows * 10 * 128,), dtype=torch.uint8, device=dev)
    rc = lib.lib.bb_records_serialize_static(eng._h, C.c_void_p(big.data_ptr()), big.numel(), None, None, None, None)
    assert rc == lib.BB_EINVAL and b'at most 65536' in lib.lib.bb_last_error()
    buf, off, sz = eng.records_serialized()
    assert int(off[-1]) == len(buf) and len(sz) == rows
    eng.close()
