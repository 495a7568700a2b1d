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
MAX_PLIES, LABEL = None, 'BASELINE configs[1]'
# --config: the default line is BASELINE configs[1] (the configuration the metric is quoted on); the others give the
# remaining self-play configurations of BASELINE.json a bench line of the same shape (same timing rules, same keys)
CONFIGS = {
    'connect4': dict(game='Connect4', search='fixed', sims=800, games=4096, prelude=45, max_plies=None, label='BASELINE configs[1]'),
    'connect4-dynamic': dict(game='Connect4', search='dynamic', sims=800, games=2048, prelude=45, max_plies=None,
                             label='BASELINE configs[2]: one GPU share of 16384 games on 8 GPUs'),
    'dragonchess': dict(game='DragonChess', search='fixed', sims=100, games=1024, prelude=4, max_plies=64,
                        label='BASELINE configs[3], 64-ply cap per game (the reference has no move cap, SURVEY.md A.20)'),
    'tictactoe': dict(game='TicTacToe', search='fixed', sims=10, games=4096, prelude=4, max_plies=None,
                      label='BASELINE configs[0]: parameters_template.yaml defaults'),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', default='connect4', choices=sorted(CONFIGS), help='workload (default: BASELINE configs[1])')
    ap.add_argument('--games', type=int, default=None, help='concurrent games per GPU (default: the config\'s)')
    ap.add_argument('--sims', type=int, default=None, help='simulations per move (playLimit; default: the config\'s)')
    ap.add_argument('--epsilon', type=float, default=0.3, help='policy noise weight of the headline loop (parameters_template.yaml: 0.3, alpha 0.2)')
    ap.add_argument('--other-epsilon', type=float, default=0.0, help='second timed loop at this noise weight (0 = the parity setting); <0: skip')
    ap.add_argument('--cpu-seconds', type=float, default=25.0, help='cpu_baseline sample length (GPU arm)')
    ap.add_argument('--ref-seconds', type=float, default=150.0, help='--impl reference: total length of warm-up + timed windows')
    ap.add_argument('--precision', default='bf16', choices=['bf16', 'fp32'], help='network kernel: bf16 tcgen05 or fp32 SIMT')
    ap.add_argument('--lanes', type=int, default=0, help='engine lanes (0 = engine default)')
    ap.add_argument('--filters', type=int, default=16, choices=[16, 64, 256], help='tower width (default: the template\'s 16; 64 / 256 run the layer-by-layer '
                    'tensor-core kernels of net_wide.cuh behind a tick-mode search: SURVEY.md 8d row 5a)')
    ap.add_argument('--contexts', type=int, default=0, help='0 (default): exact mode, the reference\'s search. W >= 1: virtual-loss mode with W simulations of a '
                    'game in flight -- NOT the reference\'s search for W > 1; the line says so in config.search_mode and is never the headline')
    ap.add_argument('--prelude', type=int, default=None, help='untimed move rounds at <= 32 sims/move that desynchronise the game slots')
    ap.add_argument('--e2e-rounds', type=int, default=4, help='e2e plays this many times --games whole games per GPU on the --games slots')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-e2e-records', action='store_true')
    args = ap.parse_args()
    global GAME, SEARCH, METRIC, MAX_PLIES, LABEL
    NET['filters'] = args.filters
    c = CONFIGS[args.config]
    GAME, SEARCH, MAX_PLIES, LABEL = c['game'], c['search'], c['max_plies'], c['label']
    METRIC = '%s_mcts_simulations_per_sec' % GAME.lower()
    args.games = args.games or c['games']
    args.sims = args.sims or c['sims']
    args.prelude = c['prelude'] if args.prelude is None else args.prelude
    if args.config != 'connect4':                  # the CPU arm and the mirror-level e2e are measured on the headline configuration
        args.no_cpu = True
        args.no_e2e_records = True
    return args


def config_dict(args, n_gpus):
    search = 'FixedMCTS(maxDepth 10, c 0.85)' if SEARCH == 'fixed' else 'DynamicMCTS(c 0.85)'
    return {'workload': '%s %s self-play, %d sims/move, %d concurrent games/GPU, '
                        'network blocks%d/filters%d/dense%d (%s)' % (GAME, search, args.sims, args.games, NET['blocks'], NET['filters'], NET['dense'], LABEL),
            'games_per_gpu': args.games, 'sims_per_move': args.sims, 'search': 'FixedMCTS' if SEARCH == 'fixed' else 'DynamicMCTS', 'max_depth': MAX_DEPTH,
            'step': 'one move round: the %d slots together play %d moves (FindMove+MoveRoot each), no barrier between moves, finished games restart in place' % (args.games, args.games),
            'prelude': '%d untimed move rounds at %d sims/move to desynchronise the slots, then the warm-up steps' % (args.prelude, min(32, args.sims)),
            'epsilon': args.epsilon, 'alpha': 0.2,
            'search_mode': ('exact: one simulation in flight per game, results identical to the reference\'s' if not getattr(args, 'contexts', 0) else
                            'virtual loss, %d simulations in flight per game%s' % (args.contexts, ' (same results as exact)' if args.contexts == 1 else
                                                                                   ' -- NON-EXACT: a different search from the reference\'s, reported separately')),
            'noise': ('policy noise ON as in parameters_template.yaml (every evaluation mixes Beta(alpha, 1-alpha) draws into the policy, NetworkFactory.py:176-182)'
                      if args.epsilon > 0 else 'policy noise off (the parity setting)'),
            'parallelism': 'games sharded over %d GPU(s), no data-path collective' % n_gpus,
            'l2_policy': 'working set (tree pool, GBs) larger than L2; no flush needed'}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference (Blackbird.GenerateTrainingSamples -> FindMove ...)
# ---------------------------------------------------------------------------------------------
def _cpu_worker(idx, shared, stop, sims, epsilon):
    """One persistent single-threaded worker (the reference is single-threaded Python): whole games, one after the
    other, until told to stop.  A sampler thread publishes the worker's cumulative simulation / game / evaluation
    counts every 50 ms, so the parent can cut time windows anywhere.  The FIRST game starts with idx-dependent cheap
    moves (32 simulations each), which spreads the workers over all game phases from the first window on -- the same
    desynchronisation the GPU arm's prelude does for its slots; every later game is a full game from the empty board."""
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass
    import numpy as np
    from oracle import network_np, search_np
    from oracle.rules_np import GAMES
    w = network_np.init_weights(3, 7, seed=0, **NET)
    ev = network_np.NetEvaluator(w, eps=epsilon, rng=np.random.RandomState(77 + idx))
    s = search_np.NetSearch(ev, SEARCH, maxDepth=MAX_DEPTH, explorationRate=C_PUCT, playLimit=sims,
                            rng=np.random.RandomState(1000 + idx))
    games = [0]
    ready = [0.0]

    def sampler():
        while not stop.value:
            shared[4 * idx + 0] = s.Stats['sims']
            shared[4 * idx + 1] = games[0]
            shared[4 * idx + 2] = s.Evals
            shared[4 * idx + 3] = ready[0]
            time.sleep(0.05)

    threading.Thread(target=sampler, daemon=True).start()
    cheap = (idx * 7) % 24                                   # plies of the first game played at 32 simulations
    first = True
    while not stop.value:
        state = GAMES[GAME]()
        s.DropRoot()
        winner, ply = None, 0
        while winner is None and not stop.value:
            lim = 32 if (first and ply < cheap) else sims
            if first and ply >= cheap:
                ready[0] = 1.0                               # from here on the worker runs the measured workload
            state, _, _ = s.FindMove(state, 1, playLimit=lim)
            s.MoveRoot(state)
            winner = state.Winner(None)
            ply += 1
        ready[0] = 1.0
        games[0] += winner is not None
        first = False


class CpuArm(object):
    """os.cpu_count() persistent worker processes playing the bench workload on the host cores (oracle port of the
    reference; numpy fp32 network at batch 1, eager child evaluation exactly like MCTS.py:134-139)."""

    def __init__(self, sims, epsilon, cores=None):
        import multiprocessing as mp
        self.cores = cores or os.cpu_count() or 1
        ctx = mp.get_context('fork')
        self.shared = ctx.Array('d', 4 * self.cores, lock=False)
        self.stop = ctx.Value('i', 0, lock=False)
        self.procs = [ctx.Process(target=_cpu_worker, args=(i, self.shared, self.stop, sims, epsilon), daemon=True)
                      for i in range(self.cores)]
        for p_ in self.procs:
            p_.start()

    def wait_ready(self, timeout=40.0):
        t0 = time.time()
        while time.time() - t0 < timeout and sum(self.shared[4 * i + 3] for i in range(self.cores)) < self.cores:
            time.sleep(0.1)

    def read(self):
        a = list(self.shared)
        return time.perf_counter(), sum(a[0::4]), sum(a[1::4]), sum(a[2::4])

    def window(self, seconds):
        t0, s0, g0, e0 = self.read()
        time.sleep(seconds)
        t1, s1, g1, e1 = self.read()
        return (s1 - s0) / (t1 - t0), g1 - g0, (e1 - e0) / (t1 - t0), t1 - t0

    def close(self):
        self.stop.value = 1
        for p_ in self.procs:
            p_.join(timeout=5)
            if p_.is_alive():
                p_.terminate()


CPU_SAMPLE = ('%d persistent single-threaded worker processes (one per host core) playing whole Connect4 FixedMCTS games back to back '
              '(oracle port of the reference: numpy fp32 network at batch 1, eager child evaluation), first game of every worker '
              'fast-forwarded by 0-23 cheap plies so that all game phases are present; %s')


def cpu_sample(seconds, sims, epsilon, cores=None):
    arm = CpuArm(sims, epsilon, cores)
    try:
        arm.wait_ready()
        arm.window(min(3.0, seconds / 4))                    # settle
        v, games, evals, wall = arm.window(seconds)
    finally:
        arm.close()
    return v, games, arm.cores, wall


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    # one pool for the whole run; steps are consecutive time windows of the same continuously running workload, sized
    # so that warm-up + timed windows cover >= 120 s and the run ends within a few minutes
    n_win = max(1, args.steps + args.warmup)
    seconds = max(1.0, min(30.0, args.ref_seconds / n_win))
    arm = CpuArm(args.sims, args.epsilon)
    try:
        arm.wait_ready()
        for _ in range(args.warmup):
            arm.window(seconds)
        t0, s0, g0, e0 = arm.read()
        per = [arm.window(seconds) for _ in range(args.steps)]
        t1, s1, g1, e1 = arm.read()
    finally:
        arm.close()
    cores = arm.cores
    value = (s1 - s0) / (t1 - t0)
    sample = CPU_SAMPLE % (cores, '%d warm-up + %d timed windows of %.1f s, %.0f s in all; %d games finished, %.1f network evaluations per simulation'
                           % (args.warmup, args.steps, seconds, n_win * seconds, g1 - g0, (e1 - e0) / max(1.0, s1 - s0)))
    print_line(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * (t1 - t0) / max(1, args.steps), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config_dict(args, args.gpus),
        'games_per_sec': (g1 - g0) / (t1 - t0), 'per_window': [round(w_[0], 1) for w_ in per],
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
    from blackbird_b200.engine import n_actions, state_bytes
    A, S = n_actions(GAME), state_bytes(GAME)
    eng = Engine(GAME, SEARCH, n_games=G, max_depth=MAX_DEPTH, exploration_rate=C_PUCT, evaluator='net',
                 play_limit_hint=P, device=local, lanes=args.lanes, max_plies=MAX_PLIES, contexts=args.contexts)
    weights = netinit.random_weights(cin, A, seed=0, **NET)

    def load_net(eps):
        eng.load_network(weights, NET['blocks'], NET['filters'], NET['dense'], epsilon=eps, alpha=0.2, seed=1 + rank, precision=args.precision)

    load_net(args.epsilon)
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
    eng.stream_moves(1, min(32, P), 1.0, uniforms)   # uploads the uniform table
    if args.prelude > 0:
        eng.stream_moves(args.prelude * G, min(32, P))
    flops = netinit.flops_per_eval(h, w, cin, A, **NET)
    names = ('sims', 'evals', 'nodes', 'expansions', 'select_steps', 'games_done', 'kernel_launches', 'moves', 'ticks', 'cache_lookups', 'cache_hits')

    def timed_loop(steps, warmup):
        """W untimed + K timed steps; CUDA events on the engine's stream, barrier + synchronize on both sides, time =
        max over ranks, counters summed over ranks.  -> (ms, counters, clocks, stream-stat deltas of this rank)"""
        for _ in range(max(1, warmup)):
            eng.stream_moves(G, P)
        eng.check_status()
        eng.counters(reset=True)
        st0 = eng.stream_stats()
        sampler = ClockSampler(local)
        sampler.start()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            eng.stream_moves(G, P)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.summary()
        eng.check_status()
        c = eng.counters(reset=True)
        st1 = eng.stream_stats()
        t = torch.tensor([ms], dtype=torch.float64, device='cuda')
        cnt = torch.tensor([c[k] for k in names], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        return float(t.item()), dict(zip(names, [float(x) for x in cnt.tolist()])), clocks, {k: st1[k] - st0[k] for k in st1}

    ms, cn, clocks, dst = timed_loop(args.steps, args.warmup)
    sims, evals, games_done, launches, moves, ticks = (cn[k] for k in ('sims', 'evals', 'games_done', 'kernel_launches', 'moves', 'ticks'))
    value = sims / (ms * 1e-3)
    streaming = eng.streaming

    # the same loop at the other noise setting (template: epsilon 0.3 = every evaluation draws Beta noise per action,
    # NetworkFactory.py:176-182; parity runs: epsilon 0), fewer steps
    other = None
    if args.other_epsilon >= 0 and args.other_epsilon != args.epsilon:
        load_net(args.other_epsilon)
        ms2, cn2, _, _ = timed_loop(max(3, args.steps // 2), max(3, args.warmup))
        other = {'epsilon': args.other_epsilon, 'value': cn2['sims'] / (ms2 * 1e-3), 'unit': UNIT, 'steps': max(3, args.steps // 2),
                 'ms_per_step': ms2 / max(3, args.steps // 2), 'evals_per_sec': cn2['evals'] / (ms2 * 1e-3),
                 'what': 'the same timed loop with the policy noise %s' % ('off (the parity setting)' if args.other_epsilon == 0 else 'on')}
        load_net(args.epsilon)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    tensor_peak = float(peaks.get('bf16_tflops_sustained', 1400.0))
    hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'measured (MEASURED_PEAKS.json, sustained bf16: the kernel runs inside a long step)' if peaks else 'fallback'
    traffic = {}
    try:                                             # ncu DRAM bytes of the same kernels in tick mode (profiles/r02_traffic.json: Connect4)
        if args.config == 'connect4':
            traffic = json.load(open(os.path.join(ROOT, 'profiles', 'r02_traffic.json')))
    except Exception:
        pass
    if streaming:
        # the two persistent kernels are resident on every SM for the whole timed region: one launch of each per
        # step, launch duration = step duration, so the roofline numbers come from the TIMED loop itself
        adv_ms = eval_ms = ms
        n_launch = args.steps
        cp = cn
        prof_ticks = max(1.0, dst['batches'])
    else:
        eng.profile(True, 16384)                     # tick mode: CUDA events around every advance / network launch of one more step
        eng.stream_moves(G, P)
        prof = eng.profile_read()
        eng.profile(False)
        cp = {k: float(v) for k, v in eng.counters(reset=True).items()}
        adv_ms, eval_ms, n_launch, prof_ticks = prof['advance_ms'], prof['eval_ms'], max(1, prof['ticks']), max(1, prof['ticks'])
    net_tflops = cp['evals'] * flops / (eval_ms * 1e-3) / 1e12 if eval_ms > 0 else 0.0
    # tree traffic model of SURVEY.md 8d with measured d, k, n_exp (Connect4: S folded out -- child states are not stored)
    d_bar = cp['select_steps'] / max(1, cp['sims'])
    nexp = cp['expansions'] / max(1, cp['sims'])
    k_bar = cp['nodes'] / max(1, cp['expansions'])
    # select: one 32-byte node per child per level; expand: k 32-byte nodes + the parent's header; evaluator
    # mailbox: state out/in + policy/value out/in; backup: 8-byte read-modify-write per path node
    b_sim = d_bar * 32 * k_bar + nexp * (k_bar * 32 + 16) + (cp['evals'] / max(1, cp['sims'])) * (2 * S + 2 * 4 * (min(A, k_bar) + 1)) + 16 * (d_bar + 1)
    tree_gbs = cp['sims'] * b_sim / (adv_ms * 1e-3) / 1e9 if adv_ms > 0 else 0.0
    per_world = max(1, world)
    if streaming:
        kname = 'k_net_forward_tc<%s> serve mode' % GAME + ' (bf16 tcgen05 tower, persistent: %d evals/launch/GPU, batches of %.2f)' % (
            cp['evals'] / n_launch / per_world, cp['evals'] / per_world / prof_ticks)
    elif args.precision == 'bf16' and NET['filters'] != 16:
        kname = 'k_wide_conv<%d> x %d layers + stage + heads (bf16 tcgen05, layer by layer; %d evals per tick avg)' % (NET['filters'], 1 + 2 * NET['blocks'], cp['evals'] / prof_ticks)
    elif args.precision == 'bf16':
        kname = 'k_net_forward_tc<%s> (bf16 tcgen05 tower, %d evals/launch avg)' % (GAME, cp['evals'] / prof_ticks)
    else:
        kname = 'k_net_forward<%s> (fp32 SIMT tower, %d evals/launch avg)' % (GAME, cp['evals'] / prof_ticks)
    net_traffic = None
    if traffic.get('net_dram_bytes_per_eval') is not None:
        net_traffic = traffic['net_dram_bytes_per_eval'] * cp['evals'] / per_world / n_launch
    roofline = {'bound': 'tensor', 'kernel': kname,
                'achieved': net_tflops / per_world, 'peak': tensor_peak, 'unit': 'TFLOP/s', 'frac': net_tflops / per_world / tensor_peak,
                'traffic': net_traffic, 'traffic_how': traffic.get('how'), 'peak_source': peak_src, 'flops_per_eval': flops,
                'algorithmic_bytes_per_eval': S + 4 * (A + 1), 'dram_bytes_per_eval_ncu': traffic.get('net_dram_bytes_per_eval'),
                'launch_ms_avg': eval_ms / n_launch, 'from': 'the timed loop (evals and CUDA-event time of the %d timed steps)' % args.steps if streaming else 'one profiled step',
                'step_share': eval_ms / max(1e-9, eval_ms + adv_ms)}
    roofline_tree = {'bound': 'hbm', 'kernel': ('k_search_stream<%s> (persistent select/expand/backup)' % GAME if streaming
                                                else 'k_advance<%s> (select/expand/backup)' % GAME), 'achieved': tree_gbs / per_world, 'peak': hbm_peak,
                     'unit': 'GB/s', 'frac': tree_gbs / per_world / hbm_peak, 'bytes_per_sim_model': b_sim, 'mean_depth': d_bar,
                     'dram_bytes_per_sim_ncu': traffic.get('search_dram_bytes_per_sim'),
                     'traffic': (traffic['search_dram_bytes_per_sim'] * cp['sims'] / per_world / n_launch) if traffic.get('search_dram_bytes_per_sim') is not None else None,
                     'expansions_per_sim': nexp, 'children_per_expansion': k_bar, 'evals_per_sim': cp['evals'] / max(1, cp['sims']),
                     'launch_ms_avg': adv_ms / n_launch}

    # the kernel with the larger share of the step is the one `roofline` reports; the other rides along
    roofline_tree['step_share'] = 1.0 - roofline['step_share']
    if streaming:
        # each persistent launch spans the step, so its share is 1 by construction; what tells the two apart is how
        # busy the network pipelines were (cycle counters kept by the kernel itself: wait for tickets / tower / heads)
        dw, dt, dh = dst['cyc_wait'], dst['cyc_tower'], dst['cyc_heads']
        nb = max(1, dst['batches'])
        roofline['step_share'] = roofline_tree['step_share'] = 1.0
        roofline['concurrent'] = roofline_tree['concurrent'] = 'both persistent kernels are resident on every SM for the whole step'
        roofline['pipeline_busy_frac'] = (dt + dh) / max(1, dw + dt + dh)
        roofline['cycles_per_batch'] = {'wait': dw / nb, 'tower': dt / nb, 'heads_publish': dh / nb}
        roofline['batches'] = nb
        roofline_tree['warp_busy_frac'] = dst['cyc_search_busy'] / max(1, dst['cyc_search_total'])
        roofline_tree['cycles_per_advance_call'] = dst['cyc_search_busy'] / max(1, dst['search_calls'])
        dominant, other_k = roofline, roofline_tree           # the search warps wait on the network (busy_frac ~ 0.9)
    else:
        dominant, other_k = (roofline, roofline_tree) if roofline['step_share'] >= 0.5 else (roofline_tree, roofline)

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

    # e2e_records: the reference-facing call itself -- the mirror's GenerateTrainingSamples (Blackbird.py:219-268): engine
    # set-up, uniforms H2D, whole games, SerializeState + labels on the device (bb_records_serialize), compact byte stream
    # D2H, Connection.PutGames into sqlite (DataManager.py:41-50); at N > 1 one process per GPU, every rank its own
    # sink, plus the all-gather of the compact streams for the trainer.  Everything inside the timed region.
    e2e_records = None
    if not args.no_e2e and not args.no_e2e_records:
        import shutil
        import tempfile
        from blackbird_b200 import Blackbird as BB
        from blackbird_b200.Connect4 import BoardState
        from blackbird_b200.parallel import generate_training_samples_sharded
        rows = G * max(1, args.e2e_rounds)
        cwd, tmp = os.getcwd(), tempfile.mkdtemp(prefix='bb_bench_rank%d_' % rank)
        os.chdir(tmp)
        try:
            cls = type('BenchModel', (BB.FixedModel,), {'DEVICE': local})
            mcfg = {'maxDepth': MAX_DEPTH, 'explorationRate': C_PUCT, 'timeLimit': None, 'playLimit': P,
                    'temperature': {'exploration': 1, 'exploitation': 0.1}}
            ncfg = {'blocks': NET['blocks'], 'filters': NET['filters'], 'eval': {'dense': NET['dense']},
                    'policy': {'dirichlet': {'alpha': 0.2, 'epsilon': args.epsilon}}}
            model = cls(BoardState, 'bench', mcfg, ncfg, {})
            model.setWeights(weights)
            np.random.seed(1234 + rank)
            barrier()
            t0 = time.perf_counter()
            written, gathered, lens = generate_training_samples_sharded(model, rows * world, 1, max_slots=G, gather=world > 1)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            lo_hi = rows                                          # games of this rank (contiguous equal blocks)
            tt = torch.tensor([dt], dtype=torch.float64, device='cuda')
            ss = torch.tensor([float(written - lo_hi) * P, float(lo_hi), float(written)], dtype=torch.float64, device='cuda')
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                dist.all_reduce(ss, op=dist.ReduceOp.SUM)
            db_bytes = os.path.getsize(os.path.join(tmp, 'data', 'blackbird.db'))
            e2e_records = {'value': float(ss[0].item()) / float(tt.item()), 'unit': UNIT, 'games_per_sec': float(ss[1].item()) / float(tt.item()),
                           'records_per_sec': float(ss[2].item()) / float(tt.item()), 'records': int(ss[2].item()), 'seconds': float(tt.item()),
                           'sqlite_bytes_rank0': int(db_bytes), 'allgather_bytes_per_rank': int(lens.max()) if lens is not None else 0,
                           'what': 'blackbird_b200.Blackbird.GenerateTrainingSamples(model, %d games/GPU, temp 1) on %d slots/GPU: engine set-up, self-play, '
                                   'device-side SerializeState + labels, compact stream D2H, Connection.PutGames (sqlite, one transaction per rank)%s'
                                   % (rows, G, ', all-gather of the compact record streams' if world > 1 else '')}
            model.Conn.Close()
        finally:
            os.chdir(cwd)
            shutil.rmtree(tmp, ignore_errors=True)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, games, cores, wall = cpu_sample(args.cpu_seconds, P, args.epsilon)
        cpu = {'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'games_finished': games,
               'sample': CPU_SAMPLE % (cores, 'one %.0f s window after a 3 s settle' % wall)}

    if rank == 0:
        print_line(json.dumps({
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms / max(1, args.steps), 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': args.precision, 'data': 'synthetic', 'config': config_dict(args, world),
            'games_per_sec': games_done / (ms * 1e-3), 'evals_per_sec': evals / (ms * 1e-3),
            'moves_per_sec': moves / (ms * 1e-3), 'ticks_per_step': ticks / max(1, args.steps) / world,
            'eval_cache': {'lookups': cn['cache_lookups'], 'hits': cn['cache_hits'], 'hit_rate': cn['cache_hits'] / max(1.0, cn['cache_lookups']),
                           'what': 'state-keyed memo of the evaluator shared by all game slots (the reference memoises GetPriors / SampleValue with lru_cache, '
                                   'Blackbird.py:350,372); evals_per_sec counts the evaluations the network computed; BB_EVAL_CACHE=0 switches it off'},
            'roofline': dominant, 'roofline_other': other_k, 'other_noise_setting': other, 'lanes': eng.lanes, 'streaming': bool(streaming), 'cpu_baseline': cpu, 'e2e': e2e, 'e2e_records': e2e_records,
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
