"""DRAM traffic of the two hot kernels, measured: the bench workload in TICK mode (the persistent pair cannot run under
ncu) between cudaProfilerStart/Stop, with the engine's own counters for exactly that region, so that

    ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
        --clock-control none --csv --log-file gpurun_out/r02_traffic.csv python tools/ncu_tick_traffic.py

gives bytes per k_advance launch and per k_net_forward_tc launch, and this script's last line gives the simulations and
evaluations those launches processed.  tools/traffic_from_ncu.py divides one by the other -> profiles/r02_traffic.json.
"""
import json
import os
import sys

os.environ['BB_STREAM'] = '0'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import __graft_entry__ as ge  # noqa: E402

ge.build()
from blackbird_b200 import netinit  # noqa: E402
from blackbird_b200.engine import Engine  # noqa: E402

G, P = 4096, 800
moves = int(sys.argv[1]) if len(sys.argv) > 1 else 256
eng = Engine('Connect4', 'fixed', n_games=G, max_depth=10, exploration_rate=0.85, evaluator='net', play_limit_hint=P)
eng.load_network(netinit.random_weights(3, 7, seed=0), 4, 16, 16, precision='bf16')
assert not eng.streaming
u = np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(G)])
eng.stream_moves(1, 32, 1.0, u)
eng.stream_moves(45 * G, 32)
eng.stream_moves(2 * G, P)
eng.check_status()
eng.counters(reset=True)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
eng.stream_moves(moves, P)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
c = eng.counters(reset=True)
print(json.dumps({'region': 'bb_selfplay_stream(%d moves) in tick mode, 4096 slots, steady state' % moves, **c}))
eng.close()
