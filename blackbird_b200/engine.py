"""Thin Python handle over the C-ABI engine (include/bbengine.h): batched rule kernels and the
multi-game search / self-play engine.  torch is used only for device buffers and streams."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check

_STATE_BYTES = {g: lib.bb_game_state_bytes(i) for g, i in _lib.GAME_IDS.items()}
_ACTIONS = {g: lib.bb_game_actions(i) for g, i in _lib.GAME_IDS.items()}


def plane_shape(game):
    h, w, c = C.c_int(), C.c_int(), C.c_int()
    check(lib.bb_game_plane_shape(_lib.GAME_IDS[game], C.byref(h), C.byref(w), C.byref(c)))
    return h.value, w.value, c.value


def state_bytes(game):
    return _STATE_BYTES[game]


def n_actions(game):
    return _ACTIONS[game]


def _need_cuda():
    if not torch.cuda.is_available():
        raise _lib.EngineError('no CUDA device: blackbird_b200 has no CPU fallback')


def _dev(device):
    return torch.device('cuda', device)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---- batched rule kernels over packed states: uint8 tensors [n, state_bytes] on the GPU ---------
def init_states(game, n, device=0):
    _need_cuda()
    s = torch.empty((n, _STATE_BYTES[game]), dtype=torch.uint8, device=_dev(device))
    check(lib.bb_game_init(_lib.GAME_IDS[game], _ptr(s), n, _stream()))
    return s


def legal(game, states):
    out = torch.empty((states.shape[0], _ACTIONS[game]), dtype=torch.float64, device=states.device)
    check(lib.bb_game_legal(_lib.GAME_IDS[game], _ptr(states), states.shape[0], _ptr(out), _stream()))
    return out


def apply(game, states, actions):
    """In place; returns uint8 ok flags (0 where the reference raises ValueError)."""
    actions = actions.to(device=states.device, dtype=torch.int32).contiguous()
    ok = torch.empty((states.shape[0],), dtype=torch.uint8, device=states.device)
    check(lib.bb_game_apply(_lib.GAME_IDS[game], _ptr(states), _ptr(actions), states.shape[0], _ptr(ok), _stream()))
    return ok


def winner(game, states, prev_actions=None):
    if prev_actions is not None:
        prev_actions = prev_actions.to(device=states.device, dtype=torch.int32).contiguous()
    out = torch.empty((states.shape[0],), dtype=torch.int32, device=states.device)
    check(lib.bb_game_winner(_lib.GAME_IDS[game], _ptr(states), _ptr(prev_actions), states.shape[0], _ptr(out), _stream()))
    return out


def planes(game, states):
    h, w, c = plane_shape(game)
    out = torch.empty((states.shape[0], h, w, c), dtype=torch.int8, device=states.device)
    check(lib.bb_game_planes(_lib.GAME_IDS[game], _ptr(states), states.shape[0], _ptr(out), _stream()))
    return out


def equal(game, a, b):
    out = torch.empty((a.shape[0],), dtype=torch.uint8, device=a.device)
    check(lib.bb_game_equal(_lib.GAME_IDS[game], _ptr(a), _ptr(b), a.shape[0], _ptr(out), _stream()))
    return out


def pack(game, cells, player, prev=None, castle=None, device=0):
    """cells int8 [n, H*W] (0/1/2, or signed piece codes), player [n], prev [n] (0 = None)."""
    _need_cuda()
    d = _dev(device)
    cells = torch.as_tensor(np.ascontiguousarray(cells, dtype=np.int8)).to(d)
    n = cells.shape[0]
    u8 = lambda v: None if v is None else torch.as_tensor(np.ascontiguousarray(v, dtype=np.uint8)).to(d)
    player, prev, castle = u8(player), u8(prev), u8(castle)
    out = torch.empty((n, _STATE_BYTES[game]), dtype=torch.uint8, device=d)
    check(lib.bb_game_pack(_lib.GAME_IDS[game], _ptr(cells), _ptr(player), _ptr(prev), _ptr(castle), n, _ptr(out), _stream()))
    return out


def unpack(game, states):
    h, w, _ = plane_shape(game)
    n = states.shape[0]
    cells = torch.empty((n, h * w), dtype=torch.int8, device=states.device)
    pl, pv, ca = (torch.empty((n,), dtype=torch.uint8, device=states.device) for _ in range(3))
    check(lib.bb_game_unpack(_lib.GAME_IDS[game], _ptr(states), n, _ptr(cells), _ptr(pl), _ptr(pv), _ptr(ca), _stream()))
    return cells, pl, pv, ca


def kat_forward(game, states):
    n = states.shape[0]
    pol = torch.empty((n, _ACTIONS[game]), dtype=torch.float32, device=states.device)
    val = torch.empty((n,), dtype=torch.float32, device=states.device)
    check(lib.bb_kat_forward(_lib.GAME_IDS[game], _ptr(states), n, _ptr(pol), _ptr(val), _stream()))
    return pol, val


def beta_noise(alpha, seed, n, device=0):
    """n draws of the network kernels' policy-noise sampler, Beta(alpha, 1 - alpha) (NetworkFactory.py:176-180)."""
    _need_cuda()
    out = torch.empty((n,), dtype=torch.float32, device=_dev(device))
    check(lib.bb_debug_beta_noise(float(alpha), int(seed), int(n), _ptr(out), _stream()))
    return out


def kernels_concurrent(device=0):
    """False when kernel launches are being serialised (a profiler in kernel-replay mode is attached)."""
    _need_cuda()
    return bool(check(lib.bb_debug_kernels_concurrent(int(device))))


def default_node_cap(game, search, play_limit, max_depth=10):
    """Nodes per half-arena: one move's growth plus the subtree carried over MoveRoot.

    A dynamic simulation expands at most one node (<= A children); a fixed-depth one expands up to
    max_depth nodes while the tree is young and ~3 on a fresh 800-simulation root (SURVEY.md
    Appendix E).  Overflow is reported (BB_ENOMEM), never silent."""
    a = min(_ACTIONS[game], 48)
    per_sim = a * ((min(max_depth, 10) if play_limit <= 64 else 4) if search == 'fixed' else 1)
    # dynamic trees are narrow and deep: when one child keeps most of the visits move after move the carried
    # subtree approaches growth / (1 - kept fraction); 6 moves' worth of growth covers a kept fraction of 0.8
    # (2.5 overflowed in 1 of 131072 Connect4 games at 800 simulations per move, profiles/r01e_summary.md)
    moves_worth = 2.5 if search == 'fixed' else 6.0
    return int(min(1 << 22, max(1024, moves_worth * play_limit * per_sim + 64)))


_DEFAULT_PLIES = {'TicTacToe': 9, 'Connect4': 42, 'DragonChess': 512}


class Engine(object):
    """n_games concurrent searches / self-play games on one GPU."""

    def __init__(self, game, search='dynamic', n_games=1, max_depth=10, exploration_rate=0.85, evaluator='kat',
                 node_cap=None, max_plies=None, play_limit_hint=100, device=0, lanes=0, contexts=0):
        """contexts: 0 = exact mode (one simulation in flight per game: the reference's search, default); W >= 1 = virtual-loss
        mode with W simulations of a game in flight (W = 1 reproduces the exact mode; W > 1 is a different, non-exact search)."""
        _need_cuda()
        self.game, self.search, self.n_games, self.device = game, search, int(n_games), device
        self.A = _ACTIONS[game]
        self.state_bytes = _STATE_BYTES[game]
        self.max_plies = int(max_plies or _DEFAULT_PLIES[game])
        self.node_cap = int(node_cap or default_node_cap(game, search, play_limit_hint, int(max_depth or 10)))
        cfg = _lib.bb_config()
        cfg.game = _lib.GAME_IDS[game]
        cfg.search = {'fixed': _lib.FIXED, 'dynamic': _lib.DYNAMIC}[search]
        cfg.max_depth = int(max_depth or 0)
        cfg.evaluator = {'kat': _lib.EVAL_KAT, 'net': _lib.EVAL_NET, 'external': _lib.EVAL_EXTERNAL, 'rollout': _lib.EVAL_ROLLOUT}[evaluator]
        cfg.value_f64 = 1 if evaluator == 'rollout' else 0
        cfg.n_games, cfg.node_cap, cfg.max_plies, cfg.device = self.n_games, self.node_cap, self.max_plies, device
        cfg.exploration_rate = float(exploration_rate)
        cfg.lanes = int(lanes)
        cfg.contexts = int(contexts)
        self.contexts = int(contexts)
        self._h = C.c_void_p()
        check(lib.bb_engine_create(C.byref(cfg), C.byref(self._h)))
        self.evaluator = evaluator
        self.lanes = lib.bb_engine_lanes(self._h)

    @property
    def streaming(self):
        """True when searches run as the persistent search + network kernel pair (needs a loaded bf16 network)."""
        return bool(lib.bb_engine_streaming(self._h))

    def stream_stats(self):
        """Cumulative busy accounting of the streaming network kernel (batches, wait / tower / heads cycles, pipelines)."""
        out = (C.c_uint64 * 8)()
        check(lib.bb_stream_stats(self._h, out))
        return dict(zip(('batches', 'cyc_wait', 'cyc_tower', 'cyc_heads', 'pipelines', 'cyc_search_busy', 'cyc_search_total', 'search_calls'), [int(x) for x in out]))

    def stream_last_answers(self):
        """Per slot: the last state asked through the request ring and the serve-mode network's answer to it
        -> (states uint8[n, S], policy f32[n, A], value f32[n], seq uint32[n]; seq 0 = the slot never asked)."""
        n = self.n_games
        st, pol = np.empty((n, self.state_bytes), np.uint8), np.empty((n, self.A), np.float32)
        val, seq = np.empty(n, np.float32), np.empty(n, np.uint32)
        check(lib.bb_stream_last_answers(self._h, *(a.ctypes.data_as(C.c_void_p) for a in (st, pol, val, seq))))
        return st, pol, val, seq

    def close(self):
        if getattr(self, '_h', None):
            lib.bb_engine_destroy(self._h)
            self._h = None

    __del__ = close

    # -- network --------------------------------------------------------------------------------
    def load_network(self, flat_weights, blocks, filters, dense, epsilon=0.0, alpha=0.2, seed=0, precision='fp32'):
        """precision 'fp32': SIMT kernel, 1e-5 parity with the reference graph; 'bf16': tcgen05 tensor-core kernel."""
        w = np.ascontiguousarray(flat_weights, dtype=np.float32)
        d = _lib.bb_net_desc()
        d.blocks, d.filters, d.dense, d.cin = blocks, filters, dense, plane_shape(self.game)[2]
        d.epsilon, d.alpha, d.noise_seed = epsilon, alpha, seed
        d.precision = {'fp32': 0, 'bf16': 1}[precision]
        check(lib.bb_net_load(self._h, C.byref(d), w.ctypes.data_as(C.c_void_p), w.size))

    def net_forward(self, states):
        n = states.shape[0]
        pol = torch.empty((n, self.A), dtype=torch.float32, device=states.device)
        val = torch.empty((n,), dtype=torch.float32, device=states.device)
        check(lib.bb_net_forward(self._h, _ptr(states), n, _ptr(pol), _ptr(val), _stream()))
        return pol, val

    # -- search ---------------------------------------------------------------------------------
    def reset(self, states=None):
        if states is not None:
            torch.cuda.synchronize()
        check(lib.bb_search_reset(self._h, _ptr(states)))

    def drop_root(self):
        check(lib.bb_search_drop_root(self._h))

    def begin_move(self, play_limit):
        check(lib.bb_search_begin_move(self._h, int(play_limit)))

    def run(self):
        pending = C.c_int(0)
        check(lib.bb_search_run(self._h, C.byref(pending)))
        return pending.value

    def root_stats(self, kmax=None):
        kmax = int(kmax or min(self.A, 256))
        n = self.n_games
        out = dict(n_children=np.empty(n, np.int32), actions=np.empty((n, kmax), np.int32),
                   plays=np.empty((n, kmax), np.uint32), child_value=np.empty((n, kmax), np.float32),
                   child_prior=np.empty((n, kmax), np.float64), root_plays=np.empty(n, np.uint32),
                   root_value=np.empty(n, np.float64), root_has_parent=np.empty(n, np.int32))
        p = lambda k: out[k].ctypes.data_as(C.c_void_p)
        check(lib.bb_search_root_stats(self._h, kmax, p('n_children'), p('actions'), p('plays'), p('child_value'),
                                       p('child_prior'), p('root_plays'), p('root_value'), p('root_has_parent')))
        return out

    def choose(self, temp, uniforms=None):
        acts = np.empty(self.n_games, np.int32)
        u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64)
        check(lib.bb_search_choose(self._h, float(temp), None if u is None else u.ctypes.data_as(C.c_void_p),
                                   acts.ctypes.data_as(C.c_void_p)))
        return acts

    def move_root(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int32)
        check(lib.bb_search_move_root(self._h, a.ctypes.data_as(C.c_void_p)))

    def set_rng(self, words):
        """Rollout evaluator: raw 32-bit MT19937 outputs, uint32 [n_games, words_per_slot]; rewinds the slots' cursors."""
        w = np.ascontiguousarray(words, dtype=np.uint32).reshape(self.n_games, -1)
        check(lib.bb_search_set_rng(self._h, w.ctypes.data_as(C.c_void_p), w.shape[1]))

    def rng_used(self):
        used = np.empty(self.n_games, np.uint32)
        check(lib.bb_search_rng_used(self._h, used.ctypes.data_as(C.c_void_p)))
        return used

    def states(self):
        out = np.empty((self.n_games, self.state_bytes), np.uint8)
        check(lib.bb_search_states(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def phases(self):
        ph, st = np.empty(self.n_games, np.int32), np.empty(self.n_games, np.int32)
        check(lib.bb_search_phases(self._h, ph.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p)))
        return ph, st

    def check_status(self):
        ph, st = self.phases()
        bad = np.flatnonzero(ph == _lib.PH_ERROR)
        if bad.size:
            code = int(st[bad[0]])
            if code == _lib.BB_ENOMEM:
                raise MemoryError('tree arena or record pool of slot %d overflowed (node_cap=%d)' % (bad[0], self.node_cap))
            if code == _lib.BB_EASSERT:
                raise AssertionError('Selected move is legal.')
            raise _lib.EngineError('slot %d failed with status %d' % (bad[0], code))
        return ph

    def eval_pending(self):
        """-> (n, states uint8[n, S] device tensor, slots int32[n] device tensor) of the open requests"""
        n = C.c_int32()
        check(lib.bb_eval_pending(self._h, C.byref(n), None, None))
        k = n.value
        states = torch.empty((k, self.state_bytes), dtype=torch.uint8, device=_dev(self.device))
        slots = torch.empty((k,), dtype=torch.int32, device=_dev(self.device))
        if k:
            check(lib.bb_eval_fetch(self._h, _ptr(states), _ptr(slots)))
        return k, states, slots

    def eval_submit(self, policy, value):
        policy = policy.to(dtype=torch.float32).contiguous()
        value = value.to(dtype=torch.float32).contiguous()
        torch.cuda.synchronize()
        check(lib.bb_eval_submit(self._h, _ptr(policy), _ptr(value)))

    # -- whole self-play ------------------------------------------------------------------------
    def selfplay(self, play_limit, temp=1.0, uniforms=None, n_rows=None):
        """Whole games to completion.  With n_rows (or a uniforms table of more than n_games rows) the engine plays
        n_rows games on its n_games slots: a slot that ends a game starts the next unplayed one (bb_selfplay_games)."""
        u = None
        rows = self.n_games if n_rows is None else int(n_rows)
        if uniforms is not None:
            u = np.ascontiguousarray(uniforms, dtype=np.float64)
            if n_rows is None:
                rows = u.shape[0]
            assert u.shape == (rows, self.max_plies), u.shape
        up = None if u is None else u.ctypes.data_as(C.c_void_p)
        if rows == self.n_games:
            check(lib.bb_selfplay_run(self._h, int(play_limit), float(temp), up))
        else:
            check(lib.bb_selfplay_games(self._h, rows, int(play_limit), float(temp), up))
        self.check_status()

    @property
    def rec_rows(self):
        """Games covered by the record buffers (rows of the last selfplay call)."""
        return int(lib.bb_records_rows(self._h))

    def step(self, n_moves, play_limit, temp=1.0, uniforms=None):
        """Steady-state stepping: every slot plays n_moves moves; finished games restart in place."""
        u = None
        if uniforms is not None:
            u = np.ascontiguousarray(uniforms, dtype=np.float64)
            assert u.shape == (self.n_games, self.max_plies), u.shape
        check(lib.bb_selfplay_step(self._h, int(n_moves), int(play_limit), float(temp),
                                   None if u is None else u.ctypes.data_as(C.c_void_p)))

    def stream_moves(self, target_moves, play_limit, temp=1.0, uniforms=None):
        """Steady state without a per-move barrier: run until target_moves further moves were played."""
        u = None
        if uniforms is not None:
            u = np.ascontiguousarray(uniforms, dtype=np.float64)
            assert u.shape == (self.n_games, self.max_plies), u.shape
        check(lib.bb_selfplay_stream(self._h, int(target_moves), int(play_limit), float(temp),
                                     None if u is None else u.ctypes.data_as(C.c_void_p)))

    def stream(self):
        """torch view of the engine's CUDA stream (for CUDA-event timing)."""
        return torch.cuda.ExternalStream(lib.bb_engine_stream(self._h), device=_dev(self.device))

    def profile(self, enable, max_ticks=4096):
        check(lib.bb_profile_enable(self._h, int(enable), int(max_ticks)))

    def profile_read(self):
        a, b, n = C.c_double(), C.c_double(), C.c_int()
        check(lib.bb_profile_read(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return dict(advance_ms=a.value, eval_ms=b.value, ticks=n.value)

    def records_raw(self):
        n, mp, cap = self.rec_rows, self.max_plies, lib.bb_records_pool_cap(self._h)
        out = dict(n_moves=np.empty(n, np.int32), winner=np.empty(n, np.int32),
                   states=np.empty((n, mp + 1, self.state_bytes), np.uint8), rec_off=np.empty((n, mp + 2), np.uint32),
                   pool_actions=np.empty((n, cap), np.uint16), pool_plays=np.empty((n, cap), np.uint32),
                   chosen=np.empty((n, mp), np.uint16))
        p = lambda k: out[k].ctypes.data_as(C.c_void_p)
        check(lib.bb_records_export(self._h, p('n_moves'), p('winner'), p('states'), p('rec_off'), p('pool_actions'),
                                    p('pool_plays'), p('chosen')))
        return out

    def records_serialized(self, to_host=True):
        """Every record of the last self-play call as final protobuf bytes, packed on the device (bb_records_serialize).
        -> (bytes: numpy uint8 [n_bytes] (or the device tensor with to_host=False), game_off uint64 [rows+1],
        rec_bytes uint32 [rows]); record r of game g is bytes[game_off[g] + r * rec_bytes[g] :][: rec_bytes[g]]."""
        rows = self.rec_rows
        n_rec, n_bytes = C.c_uint64(), C.c_uint64()
        off, sz = np.empty(rows + 1, np.uint64), np.empty(rows, np.uint32)
        check(lib.bb_records_serialize(self._h, None, 0, C.byref(n_rec), C.byref(n_bytes), off.ctypes.data_as(C.c_void_p), sz.ctypes.data_as(C.c_void_p)))
        buf = torch.empty((int(n_bytes.value),), dtype=torch.uint8, device=_dev(self.device))
        check(lib.bb_records_serialize(self._h, _ptr(buf), buf.numel(), None, None, None, None))
        if not to_host:
            return buf, off, sz
        host = torch.empty((buf.numel(),), dtype=torch.uint8, pin_memory=True)
        host.copy_(buf)
        return host.numpy(), off, sz

    def counters(self, reset=False):
        c = _lib.bb_counters()
        check(lib.bb_counters_read(self._h, C.byref(c), int(reset)))
        return {n: int(getattr(c, n)) for n, _ in c._fields_}
