"""GPU: multi-GPU self-play path (blackbird_b200/parallel.py): device views of the record buffers,
sharding invariance of the records, and (when the box has >= 2 GPUs) the NCCL all-gather under torchrun."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip('torch')
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def mods():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from blackbird_b200 import engine, parallel, records
    return engine, parallel, records


def test_device_views_match_host_export(mods):
    engine, parallel, records = mods
    eng = engine.Engine('Connect4', 'fixed', n_games=5, evaluator='kat', play_limit_hint=20)
    eng.selfplay(20, 1.0, parallel.game_uniforms(range(5), eng.max_plies))
    host = eng.records_raw()
    dev = parallel.allgather_raw(parallel._records_device_views(eng))
    for k in parallel.RAW_KEYS:
        assert dev[k].dtype == host[k].dtype and np.array_equal(dev[k], host[k]), k
    assert records.build_records(eng, dev) == records.build_records(eng)
    eng.close()


@pytest.mark.parametrize('game,kind,P', [('Connect4', 'dynamic', 40), ('TicTacToe', 'fixed', 12)])
def test_records_do_not_depend_on_sharding(mods, game, kind, P):
    """Global game g gets the same bytes whether 7 games run in one engine or as shards of 2 / 4 ranks."""
    engine, parallel, records = mods
    n = 7

    def play(ids):
        eng = engine.Engine(game, kind, n_games=len(ids), evaluator='kat', play_limit_hint=P)
        eng.selfplay(P, 1.0, parallel.game_uniforms(ids, eng.max_plies))
        out = records.build_records(eng)
        eng.close()
        return out

    whole = play(range(n))
    for world in (2, 4):
        parts = []
        for r in range(world):
            lo, hi = parallel.shard_range(n, r, world)
            parts += play(range(lo, hi))
        assert parts == whole, world


def test_single_process_script(mods):
    subprocess.check_call([sys.executable, os.path.join(ROOT, 'tests', 'dist_selfplay_check.py')], timeout=600)


def test_nccl_allgather_two_ranks(mods):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs (gpurun --gpus 2)')
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    subprocess.check_call([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
                           '--master-addr', '127.0.0.1', '--master-port', str(port),
                           os.path.join(ROOT, 'tests', 'dist_selfplay_check.py')], timeout=900)
