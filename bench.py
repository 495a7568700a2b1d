#!/usr/bin/env python
"""bench.py -- Connect4 MCTS simulations/s (BASELINE.json metric) on N B200s of one node.

Workload (N=1): BASELINE.json configs[1] -- Connect4, FixedMCTS maxDepth 10, 800 simulations per
move, 4096 concurrent games per GPU, template network (blocks 4, filters 16, dense 16), random-init
weights, synthetic data (games start from the empty board; move sampling from per-game uniforms).

A "step" is one move round: the game slots together play as many moves as there are slots (each
move = one FindMove of 800 simulations, every simulation a PUCT descent + lazy network evaluation
+ backup, then MoveRoot); there is no barrier between moves and finished games restart in place,
so all slots stay busy (steady state).  `value` = simulations of all ranks / device time
(max over ranks), inputs resident in HBM.  `e2e` = the same metric through the C-ABI self-play
call with HOST buffers: uniforms H2D, G whole games to completion, raw records D2H, all inside
the timed region.  `--impl reference` times the CPU port of the reference (oracle/) on the host
cores instead.  See DESIGN.md "measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GAME, SEARCH, MAX_DEPTH, C_PUCT = 'Connect4', 'fixed', 10, 0.85
NET = dict(blocks=4, filters=16, dense=16)
METRIC, UNIT = 'connect4_mcts_simulations_per_sec', 'sims/s'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--games', type=int, default=4096, help='concurrent games per GPU')
    ap.add_argument('--sims', type=int, default=800, help='simulations per move (playLimit)')
    ap.add_argument('--epsilon', type=float, default=0.0, help='policy noise weight (template: 0.3)')
    ap.add_argument('--cpu-seconds', type=float, default=12.0, help='CPU baseline sample length')
    ap.add_argument('--precision', default='bf16', choices=['bf16', 'fp32'], help='network kernel: bf16 tcgen05 or fp32 SIMT')
    ap.add_argument('--lanes', type=int, default=0, help='engine lanes (0 = engine default)')
    ap.add_argument('--prelude', type=int, default=45, help='untimed move rounds at 32 sims/move that desynchronise the game slots')
    ap.add_argument('--e2e-rounds', type=int, default=4, help='e2e plays this many times --games whole games per GPU on the --games slots')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    return ap.parse_args()


def config_dict(args, n_gpus):
    return {'workload': 'Connect4 FixedMCTS(maxDepth 10, c 0.85) self-play, %d sims/move, %d concurrent games/GPU, '
                        'network blocks4/filters16/dense16 (BASELINE configs[1])' % (args.sims, args.games),
            'games_per_gpu': args.games, 'sims_per_move': args.sims, 'search': 'FixedMCTS', 'max_depth': MAX_DEPTH,
            'step': 'one move round: the %d slots together play %d moves (FindMove+MoveRoot each), no barrier between moves, finished games restart in place' % (args.games, args.games),
            'prelude': '%d untimed move rounds at 32 sims/move to desynchronise the slots, then the warm-up steps' % args.prelude,
            'epsilon': args.epsilon, 'parallelism': 'games sharded over %d GPU(s), no data-path collective' % n_gpus,
            'l2_policy': 'working set (tree pool, GBs) larger than L2; no flush needed'}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference (Blackbird.GenerateTrainingSamples -> FindMove ...)
# ---------------------------------------------------------------------------------------------
def _cpu_worker(task):
    seed, seconds, sims, epsilon = task
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass
    import numpy as np
    from oracle import network_np, search_np
    from oracle.rules_np import GAMES
    w = network_np.init_weights(3, 7, seed=0, **NET)
    ev = network_np.NetEvaluator(w, eps=epsilon)
    s = search_np.NetSearch(ev, SEARCH, maxDepth=MAX_DEPTH, explorationRate=C_PUCT, playLimit=sims,
                            rng=np.random.RandomState(seed))
    t0 = time.time()
    end = t0 + seconds
    games = 0
    while time.time() < end:
        state = GAMES[GAME]()
        s.DropRoot()
        winner = None
        while winner is None and time.time() < end:
            state, _, _ = s.FindMove(state, 1, moveTime=max(end - time.time(), 1e-3), playLimit=sims)
            s.MoveRoot(state)
            winner = state.Winner(None)
        games += winner is not None
    return s.Stats['sims'], games, time.time() - t0


def cpu_sample(seconds, sims, epsilon, cores=None):
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    ctx = mp.get_context('fork')
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, [(1000 + i, seconds, sims, epsilon) for i in range(cores)])
    total = sum(r[0] for r in res)
    wall = max(r[2] for r in res)
    return total / wall, sum(r[1] for r in res), cores, wall


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    seconds = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_sample(min(seconds, 2.0), args.sims, args.epsilon)
    vals, tot_t = [], 0.0
    for _ in range(args.steps):
        v, _, cores, wall = cpu_sample(seconds, args.sims, args.epsilon)
        vals.append(v)
        tot_t += wall
    value = sum(vals) / len(vals)
    sample = '%d worker processes x %.0f s of Connect4 FixedMCTS self-play per step (oracle port, numpy fp32 network at batch 1)' % (cores, seconds)
    print_line(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * tot_t / max(1, args.steps), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config_dict(args, args.gpus),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0}))


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock / throttle-reason samples during the timed region, through NVML in-process (spawning
    nvidia-smi every 200 ms perturbs kernel launches of a latency-bound pipeline)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def run(self):
        nv = self.nv
        if nv is None or os.environ.get('BB_NO_CLOCKS'):
            return
        bits = {'hw_slowdown': nv.nvmlClocksThrottleReasonHwSlowdown,
                'hw_thermal_slowdown': nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                'sw_thermal_slowdown': nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                'sw_power_cap': nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag:
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((sm, [k for k, b in bits.items() if mask & b]))
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({k for r in self.rows for k in r[1]})
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': getattr(self, 'max_sm', None), 'reasons': reasons,
                'samples': len(self.rows), 'how': 'NVML in-process, 50 ms period'}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if local == 0:
        ge.build()
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
        dist.barrier()
    else:
        torch.cuda.set_device(local)
    from blackbird_b200 import netinit
    from blackbird_b200.engine import Engine, plane_shape

    G, P = args.games, args.sims
    h, w, cin = plane_shape(GAME)
    eng = Engine(GAME, SEARCH, n_games=G, max_depth=MAX_DEPTH, exploration_rate=C_PUCT, evaluator='net',
                 play_limit_hint=P, device=local, lanes=args.lanes)
    weights = netinit.random_weights(cin, 7, seed=0, **NET)
    eng.load_network(weights, NET['blocks'], NET['filters'], NET['dense'], epsilon=args.epsilon, alpha=0.2, seed=rank, precision=args.precision)
    # per-game move-sampling uniforms, game id = rank*G + slot (records identical for any GPU count)
    uniforms = np.stack([np.random.RandomState(1000 + rank * G + g).random_sample(eng.max_plies) for g in range(G)])
    stream = eng.stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    eng.reset()
    # Untimed prelude: all slots start on an empty board at once, so their game phases (and tree
    # sizes) are in lock-step for a whole game length.  Playing ~1.5 game lengths of cheap moves
    # (32 simulations each) spreads the slots over all game phases; the warm-up steps that follow
    # then rebuild full 800-simulation trees before anything is timed.
    eng.stream_moves(1, 32, 1.0, uniforms)           # uploads the uniform table
    if args.prelude > 0:
        eng.stream_moves(args.prelude * G, 32)
    for _ in range(max(1, args.warmup)):
        eng.stream_moves(G, P)
    eng.check_status()
    c0 = eng.counters(reset=True)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        eng.stream_moves(G, P)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.summary()
    eng.check_status()
    c = eng.counters(reset=True)
    t = torch.tensor([ms], dtype=torch.float64, device='cuda')
    cnt = torch.tensor([c['sims'], c['evals'], c['nodes'], c['expansions'], c['select_steps'], c['games_done'],
                        c['kernel_launches'], c['moves'], c['ticks']], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms = float(t.item())
    sims, evals, nodes, exps, sel, games_done, launches, moves, ticks = [float(x) for x in cnt.tolist()]
    value = sims / (ms * 1e-3)

    # kernel shares: one profiled step (CUDA events around every advance / network launch)
    streaming = eng.streaming
    st0 = eng.stream_stats()
    eng.profile(True, 16384)
    eng.stream_moves(G, P)
    prof = eng.profile_read()
    eng.profile(False)
    st1 = eng.stream_stats()
    cp = eng.counters(reset=True)
    flops = netinit.flops_per_eval(h, w, cin, 7, **NET)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    tensor_peak = float(peaks.get('bf16_tflops_sustained', 1400.0))
    hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'measured (MEASURED_PEAKS.json, sustained bf16)' if peaks else 'fallback'
    net_tflops = cp['evals'] * flops / (prof['eval_ms'] * 1e-3) / 1e12 if prof['eval_ms'] > 0 else 0.0
    # tree traffic model of SURVEY.md 8d with measured d, k, n_exp (Connect4: S folded out -- child states are not stored)
    d_bar = cp['select_steps'] / max(1, cp['sims'])
    nexp = cp['expansions'] / max(1, cp['sims'])
    k_bar = cp['nodes'] / max(1, cp['expansions'])
    # select: one 32-byte node per child per level; expand: k 32-byte nodes + the parent's header; evaluator
    # mailbox: state out/in + policy/value out/in; backup: 8-byte read-modify-write per path node
    b_sim = d_bar * 32 * k_bar + nexp * (k_bar * 32 + 16) + (cp['evals'] / max(1, cp['sims'])) * (2 * 16 + 2 * 4 * 8) + 16 * (d_bar + 1)
    tree_gbs = cp['sims'] * b_sim / (prof['advance_ms'] * 1e-3) / 1e9 if prof['advance_ms'] > 0 else 0.0
    launches_prof = max(1, cp['kernel_launches'] // 3) if streaming else max(1, prof['ticks'])
    if streaming:
        kname = 'k_net_forward_tc<Connect4> serve mode (bf16 tcgen05 tower, persistent: %d evals/launch, batches of %.2f)' % (
            cp['evals'] / launches_prof, cp['evals'] / max(1, prof['ticks']))
    elif args.precision == 'bf16':
        kname = 'k_net_forward_tc<Connect4> (bf16 tcgen05 tower, %d evals/launch avg)' % (cp['evals'] / max(1, prof['ticks']))
    else:
        kname = 'k_net_forward<Connect4> (fp32 SIMT tower, %d evals/launch avg)' % (cp['evals'] / max(1, prof['ticks']))
    roofline = {'bound': 'tensor', 'kernel': kname,
                'achieved': net_tflops, 'peak': tensor_peak, 'unit': 'TFLOP/s', 'frac': net_tflops / tensor_peak,
                'traffic': None, 'peak_source': peak_src, 'flops_per_eval': flops,
                'launch_ms_avg': prof['eval_ms'] / launches_prof,
                'step_share': prof['eval_ms'] / max(1e-9, prof['eval_ms'] + prof['advance_ms'])}
    roofline_tree = {'bound': 'hbm', 'kernel': ('k_search_stream<Connect4> (persistent select/expand/backup)' if streaming
                                                else 'k_advance<Connect4> (select/expand/backup)'), 'achieved': tree_gbs, 'peak': hbm_peak,
                     'unit': 'GB/s', 'frac': tree_gbs / hbm_peak, 'bytes_per_sim_model': b_sim, 'mean_depth': d_bar,
                     'expansions_per_sim': nexp, 'children_per_expansion': k_bar, 'evals_per_sim': cp['evals'] / max(1, cp['sims']),
                     'launch_ms_avg': prof['advance_ms'] / launches_prof}

    # the kernel with the larger share of the step is the one `roofline` reports; the other rides along
    roofline_tree['step_share'] = 1.0 - roofline['step_share']
    if streaming:
        # the two persistent kernels share every SM for the whole step: each one's launch spans the step, so its
        # share is 1 by construction; what tells them apart is how busy the network pipelines were (cycle
        # counters kept by the kernel itself: wait for tickets / convolution tower / heads + publish)
        dw, dt, dh = [st1[k] - st0[k] for k in ('cyc_wait', 'cyc_tower', 'cyc_heads')]
        nb = max(1, st1['batches'] - st0['batches'])
        roofline['step_share'] = roofline_tree['step_share'] = 1.0
        roofline['concurrent'] = roofline_tree['concurrent'] = 'both persistent kernels are resident on every SM for the whole step'
        roofline['pipeline_busy_frac'] = (dt + dh) / max(1, dw + dt + dh)
        roofline['cycles_per_batch'] = {'wait': dw / nb, 'tower': dt / nb, 'heads_publish': dh / nb}
        roofline['batches'] = nb
        roofline_tree['warp_busy_frac'] = (st1['cyc_search_busy'] - st0['cyc_search_busy']) / max(1, st1['cyc_search_total'] - st0['cyc_search_total'])
        roofline_tree['cycles_per_advance_call'] = (st1['cyc_search_busy'] - st0['cyc_search_busy']) / max(1, st1['search_calls'] - st0['search_calls'])
        dominant, other = roofline, roofline_tree           # the search warps wait on the network (busy_frac ~ 0.9)
    else:
        dominant, other = (roofline, roofline_tree) if roofline['step_share'] >= 0.5 else (roofline_tree, roofline)

    # e2e: whole games through the C-ABI call with host buffers (H2D uniforms, D2H records inside)
    e2e = None
    if not args.no_e2e:
        barrier()
        rows = G * max(1, args.e2e_rounds)
        # game id = rank * rows + row: one RandomState stream per global game, as in the timed loop
        u_e2e = np.stack([np.random.RandomState(1000 + rank * rows + g).random_sample(eng.max_plies) for g in range(rows)])
        pinned = torch.from_numpy(u_e2e).pin_memory()
        barrier()
        t0 = time.perf_counter()
        eng.selfplay(P, 1.0, pinned.numpy())
        if world > 1:                                        # every rank ends up with every game's records on its GPU;
            from blackbird_b200.parallel import allgather_raw, _records_device_views
            if rank == 0:                                    # the trainer rank reads all of them, the others their own shard
                raw = allgather_raw(_records_device_views(eng))
            else:
                allgather_raw(_records_device_views(eng), to_host=False)
                raw = eng.records_raw()
        else:
            raw = eng.records_raw()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        ce = eng.counters(reset=True)
        tt = torch.tensor([dt], dtype=torch.float64, device='cuda')
        ss = torch.tensor([ce['sims'], ce['games_done']], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(ss, op=dist.ReduceOp.SUM)
        d2h = sum(v.nbytes for v in raw.values())            # rank 0's bytes (the largest)
        e2e = {'value': float(ss[0].item()) / float(tt.item()), 'unit': UNIT, 'h2d_bytes_per_step': int(u_e2e.nbytes),
               'd2h_bytes_per_step': int(d2h), 'games_per_sec': float(ss[1].item()) / float(tt.item()),
               'seconds': float(tt.item()), 'what': 'bb_selfplay_games + %s: %d whole games/GPU to completion on %d slots' % (
                   'one NCCL all-gather of the record buffers + D2H' if world > 1 else 'bb_records_export', rows, G),
               'mean_moves_per_game': float(raw['n_moves'].mean()) if rank == 0 else None}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, games, cores, wall = cpu_sample(args.cpu_seconds, P, args.epsilon)
        cpu = {'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port',
               'sample': '%d worker processes x %.0f s of the same workload (oracle port of the reference, numpy fp32 network at batch 1)' % (cores, wall)}

    if rank == 0:
        print_line(json.dumps({
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms / max(1, args.steps), 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': args.precision, 'data': 'synthetic', 'config': config_dict(args, world),
            'games_per_sec': games_done / (ms * 1e-3), 'evals_per_sec': evals / (ms * 1e-3),
            'moves_per_sec': moves / (ms * 1e-3), 'ticks_per_step': ticks / max(1, args.steps) / world,
            'roofline': dominant, 'roofline_other': other, 'lanes': eng.lanes, 'streaming': bool(streaming), 'cpu_baseline': cpu, 'e2e': e2e,
            'gpu_launches': int(launches), 'clocks': clocks}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


_STDOUT = None


def print_line(text):
    """The bench line goes to the process's REAL stdout; everything else that writes to fd 1 (NCCL's version
    banner, build logs of child processes) was redirected to stderr in main(), so stdout is ONE JSON line."""
    _STDOUT.write(text + '\n')
    _STDOUT.flush()


def main():
    global _STDOUT
    sys.stdout.flush()
    _STDOUT = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
