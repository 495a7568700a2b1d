"""Run under torchrun (one rank per GPU; NCCL) or as a single process: plays N_GAMES global Connect4
games sharded over the ranks, all-gathers the raw records, and checks on every rank that the gathered
records of every global game are byte-identical to the same games played by ONE engine on this rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

N_GAMES, P = 10, 24


def main():
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from blackbird_b200 import parallel, records
    from blackbird_b200.engine import Engine

    def make(slots):
        return Engine('Connect4', 'dynamic', n_games=slots, evaluator='kat', play_limit_hint=P, device=local)

    eng, raw_all, owners = parallel.selfplay_sharded(make, N_GAMES, P)
    got = records.build_records(eng, raw_all)
    flat = []
    for first, count in owners:
        flat += got[first:first + count]
    assert len(flat) == N_GAMES, (len(flat), owners)
    eng.close()
    # the same games on at most 2 slots per rank (more game rows than slots: bb_selfplay_games)
    eng, raw_all, owners = parallel.selfplay_sharded(make, N_GAMES, P, max_slots=2)
    assert eng.n_games == 2
    got = records.build_records(eng, raw_all)
    narrow = []
    for first, count in owners:
        narrow += got[first:first + count]
    assert narrow == flat, 'records depend on the slot count'
    eng.close()
    one = make(N_GAMES)
    one.selfplay(P, 1.0, parallel.game_uniforms(range(N_GAMES), one.max_plies))
    want = records.build_records(one)
    for g in range(N_GAMES):
        assert flat[g] == want[g], 'global game %d differs between %d-way sharding and one engine' % (g, world)
    one.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if int(os.environ.get('RANK', '0')) == 0:
        print('dist_selfplay_check ok: world=%d games=%d records=%d' % (world, N_GAMES, sum(len(r[0]) for r in flat)))


if __name__ == '__main__':
    main()
