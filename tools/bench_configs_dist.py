"""The two multi-GPU BASELINE.json configurations that are not the bench line, under torchrun (one rank per GPU,
NCCL) or as a single process; rank 0 prints one JSON line per configuration.

  cfg2   Connect4 DynamicMCTS 800 sims/move, 16384 whole games sharded across the ranks (8 GPUs: 2048 slots each)
  cfg4b  Connect4 DynamicMCTS 800 sims/move, 131072 whole games sharded across the ranks, at most 4096 slots per GPU

Both go through blackbird_b200.parallel.selfplay_sharded: global game id -> rank by contiguous blocks, per-game
uniform streams RandomState(1000 + global id) built on the host, whole games to completion (bb_selfplay_games), ONE
NCCL all-gather of the fixed-stride record buffers, rank 0 reads all records to the host.  The timed region (wall
clock between synchronize + barrier pairs, max over ranks) covers the H2D of the uniforms, the games, the all-gather
and the D2H; building the uniform table and the engine is outside.  Nothing here touches oracle/.

Usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           tools/bench_configs_dist.py [cfg2 cfg4b]
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from blackbird_b200 import engine, netinit, parallel
from blackbird_b200.engine import Engine

P = 800


def main():
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def make(slots):
        eng = Engine('Connect4', 'dynamic', n_games=slots, evaluator='net', play_limit_hint=P, device=local)
        w = netinit.random_weights(engine.plane_shape('Connect4')[2], engine.n_actions('Connect4'), seed=0,
                                   blocks=4, filters=16, dense=16)
        eng.load_network(w, 4, 16, 16, precision='bf16')
        return eng

    def run(name, n_games, max_slots):
        rows = parallel.shard_range(n_games, 0, world)[1]
        slots = min(rows, max_slots)
        lo, hi = parallel.shard_range(n_games, rank, world)
        eng = make(slots)
        u = np.ones((rows, eng.max_plies))
        u[:hi - lo] = parallel.game_uniforms(range(lo, hi), eng.max_plies)
        warm = min(rows, slots)
        eng.selfplay(P, 1.0, u[:warm], n_rows=warm)                    # untimed warm-up: one game per slot
        eng.counters(reset=True)
        barrier()
        t0 = time.perf_counter()
        eng.selfplay(P, 1.0, u, n_rows=rows)
        if world > 1:
            if rank == 0:
                raw = parallel.allgather_raw(parallel._records_device_views(eng))
            else:
                parallel.allgather_raw(parallel._records_device_views(eng), to_host=False)
                raw = None
        else:
            raw = eng.records_raw()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        c = eng.counters()
        tt = torch.tensor([dt], dtype=torch.float64, device='cuda')
        ss = torch.tensor([c['sims'], c['games_done'], c['evals']], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(ss, op=dist.ReduceOp.SUM)
        if rank == 0:
            nm = raw['n_moves']
            owned = np.concatenate([np.arange(r * rows, r * rows + (b - a)) for r in range(world)
                                    for a, b in [parallel.shard_range(n_games, r, world)]])
            win = raw['winner'][owned]
            print(json.dumps({
                'config': name, 'n_gpus': world, 'games': n_games, 'games_per_gpu': rows, 'slots_per_gpu': slots,
                'search': 'DynamicMCTS', 'sims_per_move': P, 'streaming': bool(getattr(eng, 'streaming', True)),
                'sims_per_sec': float(ss[0]) / float(tt), 'games_per_sec': float(ss[1]) / float(tt),
                'evals_per_sim': float(ss[2]) / max(1.0, float(ss[0])), 'seconds': float(tt),
                'records': int((nm[owned] + 1).sum()), 'mean_moves_per_game': float(nm[owned].mean()),
                'winners': {str(k): int((win == k).sum()) for k in np.unique(win)},
                'h2d_bytes_per_gpu': int(u.nbytes), 'd2h_bytes_rank0': int(sum(v.nbytes for v in raw.values())),
                'timed': 'H2D uniforms + whole games + one NCCL all-gather of the record buffers + D2H on rank 0; '
                         'wall clock, max over ranks'}), flush=True)
        eng.close()
        barrier()

    want = sys.argv[1:] or ['cfg2', 'cfg4b']
    if 'cfg2' in want:
        run('cfg2 Connect4 DynamicMCTS, 16384 games sharded', 16384, 1 << 30)
    if 'cfg4b' in want:
        run('cfg4b Connect4 DynamicMCTS, 131072 games sharded', 131072, 4096)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
    sys.stdout.flush()
    os._exit(0)
