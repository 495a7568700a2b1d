"""The BASELINE.json configurations that are NOT the bench line, measured on one B200 (one JSON line each).

bench.py times configs[1] only (the configuration the metric is quoted on); this tool covers the rest so that
profiles/ holds a number for every row of SURVEY.md 8(d)'s table:

  cfg0   TicTacToe FixedMCTS, template parameters (playLimit 10), whole games through bb_selfplay_games
  cfg2   Connect4 DynamicMCTS 800 sims/move, 2048 slots (= one GPU's share of 16384 games on 8 GPUs), steady state
  cfg3   DragonChess FixedMCTS, playLimit 100 and 10 (template), 1024 slots, steady state
  cfg4a  network sweep, batch 256 .. 65536 x {TicTacToe, Connect4, DragonChess} shapes, fp32 SIMT and bf16 tcgen05
  cfg4b  Connect4 DynamicMCTS 800 sims/move, 131072 whole games on the 4096 slots of one GPU (host uniforms in,
         records out: the e2e definition of bench.py)

  mcts_only  Connect4 FixedMCTS / DynamicMCTS with the integer evaluator (no network): tree cost alone (not in the default set)

Usage: python tools/bench_configs.py [cfg0 cfg2 cfg3 cfg4a cfg4b mcts_only] > lines.jsonl
Timing: CUDA events / synchronize-bracketed wall clock after an untimed warm-up; synthetic data (random-init
weights seed 0, per-game uniforms RandomState(1000 + g)).  Nothing here touches oracle/.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from blackbird_b200 import engine, netinit
from blackbird_b200.engine import Engine

FLOPS = {'TicTacToe': 340516, 'Connect4': 1588700, 'DragonChess': 2694976}   # SURVEY.md 8(d), F=16


def out(**kw):
    print(json.dumps(kw), flush=True)


def make_engine(game, kind, slots, play_limit, evaluator='net', **kw):
    eng = Engine(game, kind, n_games=slots, evaluator=evaluator, play_limit_hint=play_limit, **kw)
    if evaluator != 'net':
        return eng
    w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=0, blocks=4, filters=16, dense=16)
    eng.load_network(w, 4, 16, 16, precision='bf16')
    return eng


def uniforms(eng, rows):
    return np.stack([np.random.RandomState(1000 + g).random_sample(eng.max_plies) for g in range(rows)])


def steady(name, game, kind, slots, play_limit, rounds, prelude=20, **kw):
    """Steady-state move rounds as in bench.py's `value`: finished games restart in place, no barrier between moves."""
    eng = make_engine(game, kind, slots, play_limit, **kw)
    eng.reset()
    eng.stream_moves(1, min(32, play_limit), 1.0, uniforms(eng, slots))
    eng.stream_moves(prelude * slots, min(32, play_limit))          # desynchronise the slots
    eng.stream_moves(slots, play_limit)                             # warm-up round at full size
    eng.check_status()
    eng.counters(reset=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.stream_moves(rounds * slots, play_limit)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    eng.check_status()
    c = eng.counters()
    out(config=name, game=game, search=kind, slots=slots, sims_per_move=play_limit, streaming=eng.streaming,
        sims_per_sec=c['sims'] / dt, moves_per_sec=c['moves'] / dt, games_per_sec=c['games_done'] / dt,
        evals_per_sim=c['evals'] / max(1, c['sims']), mean_depth=c['select_steps'] / max(1, c['sims']),
        children_per_expansion=c['nodes'] / max(1, c['expansions']), seconds=dt, rounds=rounds, **kw)
    eng.close()


def whole_games(name, game, kind, slots, rows, play_limit, **kw):
    """rows whole games on `slots` slots with HOST buffers: uniform table in, raw records out (bench.py's e2e)."""
    eng = make_engine(game, kind, slots, play_limit, **kw)
    u = uniforms(eng, rows)
    eng.reset(); eng.selfplay(play_limit, 1.0, u[:slots])            # warm-up: one generation of games
    eng.reset(); eng.counters(reset=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.selfplay(play_limit, 1.0, u, n_rows=rows)
    rec = eng.records_raw()
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    c = eng.counters()
    n_rec = int((rec['n_moves'] + 1).sum())
    out(config=name, game=game, search=kind, slots=slots, games=rows, sims_per_move=play_limit, streaming=eng.streaming,
        sims_per_sec=c['sims'] / dt, games_per_sec=rows / dt, records=n_rec, mean_moves_per_game=float(rec['n_moves'].mean()),
        h2d_bytes=int(u.nbytes), d2h_bytes=int(sum(v.nbytes for v in rec.values())), seconds=dt,
        winners={str(k): int((rec['winner'] == k).sum()) for k in np.unique(rec['winner'])}, **kw)
    eng.close()


def playout_states(game, n, plies, seed=0):
    """Positions a few random legal plies into a game (Appendix B.1's generator, on the GPU rule kernels)."""
    base = min(n, 4096)
    st = engine.init_states(game, base)
    gen = torch.Generator(device=st.device); gen.manual_seed(seed)
    for _ in range(plies):
        lg = engine.legal(game, st).float()
        live = (engine.winner(game, st) < 0) & (lg.sum(1) > 0)      # WIN_NONE is negative
        lg[~live] = 0; lg[~live, 0] = 1
        a = torch.multinomial(lg, 1, generator=gen).squeeze(1).int()
        nxt = st.clone()
        engine.apply(game, nxt, a)
        st = torch.where(live[:, None], nxt, st)
    reps = (n + base - 1) // base
    return st.repeat(reps, 1)[:n].contiguous()


def net_sweep():
    for game in ('TicTacToe', 'Connect4', 'DragonChess'):
        for prec in ('bf16', 'fp32'):
            eng = Engine(game, 'fixed', n_games=4, evaluator='net', lanes=1)
            w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=0, blocks=4, filters=16, dense=16)
            eng.load_network(w, 4, 16, 16, precision=prec)
            for n in (256, 1024, 4096, 16384, 65536):
                if game == 'DragonChess' and prec == 'fp32' and n > 16384:
                    continue                                          # bounded run time of the SIMT twin
                st = playout_states(game, n, {'TicTacToe': 3, 'Connect4': 8, 'DragonChess': 6}[game])
                s = torch.cuda.Stream()
                with torch.cuda.stream(s):
                    for _ in range(3):
                        eng.net_forward(st)
                    torch.cuda.synchronize()
                    reps = 10 if n <= 16384 else 4
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(s)
                    for _ in range(reps):
                        eng.net_forward(st)
                    e1.record(s)
                    torch.cuda.synchronize()
                us = e0.elapsed_time(e1) / reps * 1e3
                out(config='cfg4a network sweep', game=game, precision=prec, batch=n, us_per_batch=us,
                    positions_per_sec=n / us * 1e6, tflops=n * FLOPS[game] / us / 1e6, flops_per_position=FLOPS[game],
                    includes='policy / value output allocation and the launch, per call')
            eng.close()


def tower_flops(game, filters, blocks=4, dense=16):
    """2 x MACs of one evaluation (BN folded), the formula behind SURVEY.md 8(d)'s table."""
    h, w, c = engine.plane_shape(game)
    a = engine.n_actions(game)
    hw = h * w
    macs = hw * 9 * c * filters + 2 * blocks * hw * 9 * filters * filters + hw * 3 * filters + 2 * dense + 2 * a
    return 2 * macs


def net_sweep_widths():
    """SURVEY.md 8(d) row 5a, the optional widths: filters 64 and 256 (k_wide_conv, layer by layer), 4 blocks."""
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json'))) \
        if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')) else {}
    for game in ('Connect4', 'TicTacToe', 'DragonChess'):
        for filters in (32, 64, 128, 256):
            eng = Engine(game, 'fixed', n_games=4, evaluator='net', lanes=1)
            w = netinit.random_weights(engine.plane_shape(game)[2], engine.n_actions(game), seed=0, blocks=4, filters=filters, dense=16)
            eng.load_network(w, 4, filters, 16, precision='bf16')
            fl = tower_flops(game, filters)
            for n in (256, 1024, 4096, 16384, 65536):
                st = playout_states(game, n, {'TicTacToe': 3, 'Connect4': 8, 'DragonChess': 6}[game])
                s = torch.cuda.Stream()
                with torch.cuda.stream(s):
                    for _ in range(3):
                        eng.net_forward(st)
                    torch.cuda.synchronize()
                    reps = 10 if n <= 16384 else 4
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(s)
                    for _ in range(reps):
                        eng.net_forward(st)
                    e1.record(s)
                    torch.cuda.synchronize()
                us = e0.elapsed_time(e1) / reps * 1e3
                out(config='cfg4a network sweep, other widths', game=game, filters=filters, blocks=4, precision='bf16', batch=n, us_per_batch=us,
                    positions_per_sec=n / us * 1e6, tflops=n * fl / us / 1e6, flops_per_position=fl, peaks=peaks or None,
                    includes='policy / value output allocation, stage + 9 layer + heads launches per 4096-position chunk')
            eng.close()


def main():
    want = sys.argv[1:] or ['cfg0', 'cfg2', 'cfg3', 'cfg4a', 'cfg4b']
    if 'widths' in want:
        net_sweep_widths()
    if 'cfg0' in want:
        whole_games('cfg0 TicTacToe FixedMCTS template (playLimit 10)', 'TicTacToe', 'fixed', 4096, 16384, 10)
    if 'cfg2' in want:
        steady('cfg2 Connect4 DynamicMCTS, one GPU share of 16384 games / 8 GPUs', 'Connect4', 'dynamic', 2048, 800, 3)
        steady('cfg2 Connect4 DynamicMCTS, 4096 slots', 'Connect4', 'dynamic', 4096, 800, 3)
    if 'mcts_only' in want:
        # the tree cost alone -- the integer evaluator of Appendix B.2 (a few integer operations per
        # position, fused into the tick-mode evaluation kernel) instead of the network; one launch pair per tick
        steady('mcts_only Connect4 FixedMCTS, integer evaluator (no network), tick mode', 'Connect4', 'fixed', 4096, 800, 2,
               evaluator='kat')
        steady('mcts_only Connect4 DynamicMCTS, integer evaluator (no network), tick mode', 'Connect4', 'dynamic', 4096, 800, 2,
               evaluator='kat')
    if 'cfg3' in want:
        steady('cfg3 DragonChess FixedMCTS playLimit 100', 'DragonChess', 'fixed', 1024, 100, 2, prelude=4, max_plies=64)
        steady('cfg3 DragonChess FixedMCTS playLimit 10 (template)', 'DragonChess', 'fixed', 1024, 10, 4, prelude=4, max_plies=64)
    if 'cfg4a' in want:
        net_sweep()
    if 'cfg4b' in want:
        whole_games('cfg4b Connect4 DynamicMCTS 131072 games on one GPU', 'Connect4', 'dynamic', 4096, 131072, 800)


if __name__ == '__main__':
    main()
    sys.stdout.flush()
    os._exit(0)
