"""ctypes binding of libbbengine.so (include/bbengine.h).

There is no CPU fallback: importing this module fails loudly when the shared library has not
been built (``python -c 'import __graft_entry__ as g; g.build()'``), and every compute entry
point returns BB_ECUDA without a GPU.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('BB_ENGINE_LIB') or os.path.join(HERE, 'libbbengine.so')   # override: debug builds (-DBB_DEBUG)

BB_OK, BB_EINVAL, BB_EILLEGAL, BB_EASSERT, BB_ENOMEM, BB_ECUDA, BB_ESTATE = 0, -1, -2, -3, -4, -5, -6
TICTACTOE, CONNECT4, DRAGONCHESS = 0, 1, 2
FIXED, DYNAMIC = 0, 1
EVAL_KAT, EVAL_NET, EVAL_EXTERNAL, EVAL_ROLLOUT = 0, 1, 2, 3
PH_IDLE, PH_RUN, PH_WAIT_EXPAND, PH_WAIT_LEAF, PH_MOVE_READY, PH_GAME_OVER, PH_ERROR, PH_COMPACT = range(8)
GAME_IDS = {'TicTacToe': TICTACTOE, 'Connect4': CONNECT4, 'DragonChess': DRAGONCHESS}


class bb_config(C.Structure):
    _fields_ = [('game', C.c_int32), ('search', C.c_int32), ('max_depth', C.c_int32), ('evaluator', C.c_int32),
                ('n_games', C.c_int32), ('node_cap', C.c_int32), ('max_plies', C.c_int32), ('device', C.c_int32),
                ('exploration_rate', C.c_double), ('value_f64', C.c_int32), ('lanes', C.c_int32), ('contexts', C.c_int32), ('reserved', C.c_int32 * 5)]


class bb_net_desc(C.Structure):
    _fields_ = [('blocks', C.c_int32), ('filters', C.c_int32), ('dense', C.c_int32), ('cin', C.c_int32),
                ('epsilon', C.c_float), ('alpha', C.c_float), ('noise_seed', C.c_uint64), ('precision', C.c_int32), ('reserved', C.c_int32 * 3)]


class bb_train_desc(C.Structure):
    _fields_ = [('game', C.c_int32), ('blocks', C.c_int32), ('filters', C.c_int32), ('dense', C.c_int32), ('optimizer', C.c_int32),
                ('momentum', C.c_float), ('epsilon', C.c_float), ('alpha', C.c_float), ('noise_seed', C.c_uint64),
                ('device', C.c_int32), ('reserved', C.c_int32 * 5)]


class bb_counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ('sims', 'evals', 'nodes', 'expansions', 'select_steps', 'ticks',
                                          'kernel_launches', 'moves', 'games_done', 'cache_lookups', 'cache_hits')]


VP, I32, I = C.c_void_p, C.c_int32, C.c_int
SIGNATURES = {
    'bb_last_error': (C.c_char_p, []),
    'bb_abi_version': (I, []),
    'bb_game_state_bytes': (I, [I]),
    'bb_game_actions': (I, [I]),
    'bb_game_plane_shape': (I, [I, C.POINTER(I), C.POINTER(I), C.POINTER(I)]),
    'bb_game_init': (I, [I, VP, I, VP]),
    'bb_game_legal': (I, [I, VP, I, VP, VP]),
    'bb_game_apply': (I, [I, VP, VP, I, VP, VP]),
    'bb_game_winner': (I, [I, VP, VP, I, VP, VP]),
    'bb_game_planes': (I, [I, VP, I, VP, VP]),
    'bb_game_equal': (I, [I, VP, VP, I, VP, VP]),
    'bb_game_pack': (I, [I, VP, VP, VP, VP, I, VP, VP]),
    'bb_game_unpack': (I, [I, VP, I, VP, VP, VP, VP, VP]),
    'bb_engine_create': (I, [C.POINTER(bb_config), C.POINTER(VP)]),
    'bb_engine_destroy': (I, [VP]),
    'bb_net_load': (I, [VP, C.POINTER(bb_net_desc), VP, C.c_size_t]),
    'bb_net_forward': (I, [VP, VP, I, VP, VP, VP]),
    'bb_kat_forward': (I, [I, VP, I, VP, VP, VP]),
    'bb_debug_beta_noise': (I, [C.c_float, C.c_uint64, I, VP, VP]),
    'bb_search_reset': (I, [VP, VP]),
    'bb_search_drop_root': (I, [VP]),
    'bb_search_begin_move': (I, [VP, I]),
    'bb_search_run': (I, [VP, C.POINTER(I)]),
    'bb_search_root_stats': (I, [VP, I, VP, VP, VP, VP, VP, VP, VP, VP]),
    'bb_search_choose': (I, [VP, C.c_double, VP, VP]),
    'bb_search_move_root': (I, [VP, VP]),
    'bb_search_set_rng': (I, [VP, VP, C.c_uint32]),
    'bb_search_rng_used': (I, [VP, VP]),
    'bb_search_states': (I, [VP, VP]),
    'bb_search_phases': (I, [VP, VP, VP]),
    'bb_eval_pending': (I, [VP, C.POINTER(I32), C.POINTER(VP), C.POINTER(VP)]),
    'bb_eval_fetch': (I, [VP, VP, VP]),
    'bb_eval_submit': (I, [VP, VP, VP]),
    'bb_selfplay_run': (I, [VP, I, C.c_double, VP]),
    'bb_selfplay_games': (I, [VP, I, I, C.c_double, VP]),
    'bb_records_rows': (I, [VP]),
    'bb_selfplay_step': (I, [VP, I, I, C.c_double, VP]),
    'bb_selfplay_stream': (I, [VP, C.c_longlong, I, C.c_double, VP]),
    'bb_engine_stream': (VP, [VP]),
    'bb_profile_enable': (I, [VP, I, I]),
    'bb_profile_read': (I, [VP, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(I)]),
    'bb_records_pool_cap': (I, [VP]),
    'bb_records_export': (I, [VP, VP, VP, VP, VP, VP, VP, VP]),
    'bb_records_device': (I, [VP] + [C.POINTER(VP)] * 7),
    'bb_records_serialize': (I, [VP, VP, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), VP, VP]),
    'bb_trainer_create': (I, [C.POINTER(bb_train_desc), VP, C.c_size_t, C.POINTER(VP)]),
    'bb_trainer_destroy': (I, [VP]),
    'bb_trainer_step': (I, [VP, VP, VP, VP, I, C.c_float, VP]),
    'bb_trainer_step_stream': (I, [VP, VP, VP, I, C.c_float, VP]),
    'bb_trainer_losses': (I, [VP, C.POINTER(C.c_float)]),
    'bb_trainer_weights': (I, [VP, VP, C.c_size_t]),
    'bb_trainer_set_weights': (I, [VP, VP, C.c_size_t]),
    'bb_trainer_grads': (I, [VP, VP, C.c_size_t]),
    'bb_trainer_steps': (C.c_uint64, [VP]),
    'bb_records_serialize_static': (I, [VP, VP, C.c_uint64, VP, VP, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    'bb_records_static_max_rows': (I, []),
    'bb_counters_read': (I, [VP, C.POINTER(bb_counters), I]),
    'bb_engine_lanes': (I, [VP]),
    'bb_engine_streaming': (I, [VP]),
    'bb_stream_stats': (I, [VP, C.POINTER(C.c_uint64)]),
    'bb_stream_last_answers': (I, [VP, VP, VP, VP, VP]),
    'bb_debug_kernels_concurrent': (I, [I]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError('%s is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
                          '(the engine has no CPU fallback)' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header and library out of sync
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


class EngineError(RuntimeError):
    pass


def check(rc):
    """Map a status code to the exception the reference raises for the same condition."""
    if rc >= 0:
        return rc
    msg = (lib.bb_last_error() or b'').decode()
    if rc == BB_EILLEGAL:
        raise ValueError('Tried to make an illegal move.')
    if rc == BB_EASSERT:
        raise AssertionError(msg or 'Selected move is legal.')
    if rc == BB_EINVAL:
        raise ValueError(msg)
    if rc == BB_ENOMEM:
        raise MemoryError(msg)
    raise EngineError('libbbengine error %d: %s' % (rc, msg))
