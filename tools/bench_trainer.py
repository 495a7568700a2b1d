"""Trainer step and record packer timings on one B200 (profiles/r02_*): not bench lines, the same timing rules.

  trainer: Network.train steps (csrc/train.cu; forward + loss + gradients + Adam, one CUDA-graph replay per step) on
           synthetic Connect4 batches of the template shape, batch 50 (parameters_template.yaml) .. 4096; CUDA events
           around K steps after warm-up, batches rotated through a pool larger than L2 at the big sizes.
  packer:  bb_records_serialize (k_pack_records: SerializeState + labels on the device) over the records of 16384
           Connect4 games; algorithmic bytes = raw records read + protobuf bytes written.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from blackbird_b200 import netinit
from blackbird_b200.engine import Engine
from blackbird_b200.trainer import Trainer

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))
except Exception:
    pass


def out(**kw):
    print(json.dumps(kw), flush=True)


def trainer_bench():
    w = netinit.random_weights(3, 7, seed=0)
    fwd = netinit.flops_per_eval(6, 7, 3, 7)
    for B, steps in ((50, 200), (256, 200), (1024, 100), (4096, 40)):
        tr = Trainer('Connect4', w, 4, 16, 16, optimizer='adam')
        rs = np.random.RandomState(0)
        pool = 4
        batches = []
        for _ in range(pool):
            planes = torch.as_tensor(rs.randint(-1, 2, size=(B, 6, 7, 3)).astype(np.int8)).cuda()
            ev = torch.as_tensor(rs.choice([-1.0, 0.0, 1.0], size=B).astype(np.float32)).cuda()
            pol = torch.as_tensor(rs.dirichlet(np.ones(7), size=B).astype(np.float32)).cuda()
            batches.append((planes, ev, pol))
        for i in range(2 * pool):                                     # warm-up: one graph capture per batch buffer
            tr.step(*batches[i % pool], 1e-4)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(steps):
            tr.step(*batches[i % pool], 1e-4)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps
        # forward + backward-data + backward-weights of every convolution: 3 x the forward MACs (the stem has no data gradient)
        flops = 3.0 * fwd * B
        # activations are written once and read twice (backward data + backward weights), gradients once each way: fp32
        act_bytes = B * 42 * (3 + 9 * 16) * 4
        bytes_alg = 5.0 * act_bytes
        out(config='trainer step', game='Connect4', batch=B, optimizer='adam', us_per_step=dt * 1e6, examples_per_sec=B / dt,
            tflops_fp32=flops / dt / 1e12, algorithmic_GBps=bytes_alg / dt / 1e9, hbm_peak_GBps=PEAKS.get('hbm_gbs'),
            hbm_frac=bytes_alg / dt / 1e9 / PEAKS['hbm_gbs'] if PEAKS.get('hbm_gbs') else None,
            note='about 50 small SIMT launches replayed as one CUDA graph; launch- and latency-bound at the reference batch of 50',
            losses=tr.losses())
        tr.close()


def packer_bench():
    G, rows, P = 4096, 16384, 64
    eng = Engine('Connect4', 'fixed', n_games=G, evaluator='kat', play_limit_hint=P)
    u = np.random.RandomState(0).random_sample((rows, eng.max_plies))
    eng.selfplay(P, 1.0, u, n_rows=rows)
    raw = eng.records_raw()
    n_rec = int((raw['n_moves'] + 1).sum())
    live_in = n_rec * 16 + int(raw['rec_off'][np.arange(rows), raw['n_moves']].sum()) * 6 + rows * 8
    eng.records_serialized(to_host=False)                             # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        buf, off, sz = eng.records_serialized(to_host=False)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    t1 = time.perf_counter()
    host, off, sz = eng.records_serialized(to_host=True)
    dth = time.perf_counter() - t1
    out(config='record packer (bb_records_serialize)', games=rows, records=n_rec, bytes_out=int(buf.numel()), bytes_in_live=int(live_in),
        ms_device=dt * 1e3, GBps_device=(buf.numel() + live_in) / dt / 1e9, hbm_peak_GBps=PEAKS.get('hbm_gbs'),
        hbm_frac=(buf.numel() + live_in) / dt / 1e9 / PEAKS['hbm_gbs'] if PEAKS.get('hbm_gbs') else None,
        ms_with_d2h_to_pinned=dth * 1e3, records_per_sec_device=n_rec / dt,
        note='device time includes the size query (D2H of n_moves / winners, host prefix sum, H2D of the offsets)')
    eng.close()


if __name__ == '__main__':
    want = sys.argv[1:] or ['trainer', 'packer']
    if 'trainer' in want:
        trainer_bench()
    if 'packer' in want:
        packer_bench()
    sys.stdout.flush()
    os._exit(0)
