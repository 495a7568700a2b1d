"""DragonChess bf16 network kernel in launch mode on a batch of 4096 positions (wide policy head: 4032 actions),
for `ncu --set full -k regex:k_net_forward_tc`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from blackbird_b200 import engine, netinit
from blackbird_b200.engine import Engine
from tools.bench_configs import playout_states

eng = Engine('DragonChess', 'fixed', n_games=4, evaluator='net', lanes=1)
w = netinit.random_weights(engine.plane_shape('DragonChess')[2], engine.n_actions('DragonChess'), seed=0, blocks=4, filters=16, dense=16)
eng.load_network(w, 4, 16, 16, precision='bf16')
st = playout_states('DragonChess', 4096, 6)
for _ in range(4):
    eng.net_forward(st)
torch.cuda.synchronize()
eng.close()
os._exit(0)
