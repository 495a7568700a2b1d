"""One steady-state move round of the bench workload (BASELINE configs[1]) between cudaProfilerStart/Stop, for
`ncu --replay-mode range`: range replay keeps the kernels of the range concurrent, which the persistent search +
network pair needs (kernel replay serialises them, see profiles/r01d_summary.md).

  ncu --replay-mode range --profile-from-start off --clock-control none \
      --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,... python tools/ncu_range_step.py

Prints the step's own counters (simulations, evaluations) so the range's bytes can be put per simulation.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from blackbird_b200 import engine, netinit
from blackbird_b200.engine import Engine

SLOTS = int(os.environ.get('BB_RANGE_SLOTS', '4096'))
P = 800


def main():
    eng = Engine('Connect4', 'fixed', n_games=SLOTS, evaluator='net', play_limit_hint=P)
    w = netinit.random_weights(engine.plane_shape('Connect4')[2], engine.n_actions('Connect4'), seed=0, blocks=4, filters=16, dense=16)
    eng.load_network(w, 4, 16, 16, precision='bf16')
    u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(SLOTS)])
    eng.reset()
    eng.stream_moves(1, 32, 1.0, u)
    eng.stream_moves(20 * SLOTS, 32)                       # desynchronise the slots
    eng.stream_moves(SLOTS, P)                             # warm-up round at full size
    eng.check_status()
    eng.counters(reset=True)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    eng.stream_moves(SLOTS, P)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    eng.check_status()
    c = eng.counters()
    print(json.dumps({'slots': SLOTS, 'streaming': eng.streaming, 'sims': c['sims'], 'evals': c['evals'], 'moves': c['moves'],
                      'kernel_launches': c['kernel_launches']}), flush=True)
    eng.close()


if __name__ == '__main__':
    main()
    sys.stdout.flush()
    os._exit(0)
