"""CPU: the C-ABI library loads and exports every symbol include/bbengine.h declares; the host-side
mirrors import without a GPU and refuse to compute without one; record wire format and the sqlite
sink behave like the reference's (golden bytes from the unmodified reference)."""
import ctypes
import os
import re
import sqlite3

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    import __graft_entry__ as ge
    ge.build()
    from blackbird_b200 import _lib
    return _lib


def header_symbols():
    text = open(os.path.join(ROOT, 'include', 'bbengine.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(bb_[a-z0-9_]+)\s*\(', text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = header_symbols()
    assert len(names) >= 40
    raw = ctypes.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), 'declared in bbengine.h but not exported: ' + n
        assert n in lib.SIGNATURES, 'declared in bbengine.h but not bound in _lib.py: ' + n
    assert sorted(lib.SIGNATURES) == names, 'bound but not declared'
    assert lib.lib.bb_abi_version() >= 1


def test_game_description_needs_no_gpu(lib):
    assert [lib.lib.bb_game_state_bytes(g) for g in range(3)] == [4, 16, 40]
    assert [lib.lib.bb_game_actions(g) for g in range(3)] == [9, 7, 4032]
    from blackbird_b200 import engine
    assert engine.plane_shape('TicTacToe') == (3, 3, 3)
    assert engine.plane_shape('Connect4') == (6, 7, 3)
    assert engine.plane_shape('DragonChess') == (8, 8, 17)


def test_struct_layouts_match_header(lib):
    assert ctypes.sizeof(lib.bb_config) == 8 * 4 + 8 + 2 * 4 + 6 * 4
    assert ctypes.sizeof(lib.bb_net_desc) == 4 * 4 + 2 * 4 + 8 + 4 + 3 * 4
    assert ctypes.sizeof(lib.bb_counters) == 11 * 8


def test_no_cpu_fallback(lib):
    """Without a CUDA device engine creation fails loudly (BB_ECUDA), it never computes on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    cfg = lib.bb_config()
    cfg.game, cfg.search, cfg.max_depth, cfg.evaluator = lib.CONNECT4, lib.FIXED, 10, lib.EVAL_KAT
    cfg.n_games, cfg.node_cap, cfg.max_plies, cfg.exploration_rate = 4, 1024, 42, 0.85
    h = ctypes.c_void_p()
    rc = lib.lib.bb_engine_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == lib.BB_ECUDA and b'no CPU fallback' in lib.lib.bb_last_error()
    from blackbird_b200 import Connect4
    with pytest.raises(lib.EngineError):
        Connect4.BoardState()
    from blackbird_b200.engine import Engine
    with pytest.raises(lib.EngineError):
        Engine('Connect4', 'fixed', n_games=4)


def test_invalid_configs_are_rejected(lib):
    h = ctypes.c_void_p()
    cfg = lib.bb_config()
    cfg.game, cfg.search, cfg.max_depth, cfg.evaluator = lib.CONNECT4, lib.FIXED, 0, lib.EVAL_KAT
    cfg.n_games, cfg.node_cap, cfg.max_plies = 4, 1024, 42
    assert lib.lib.bb_engine_create(ctypes.byref(cfg), ctypes.byref(h)) == lib.BB_EINVAL
    assert b'MaxDepth for MCTS must be > 0.' in lib.lib.bb_last_error()          # FixedMCTS.py:15-16
    cfg.max_depth, cfg.n_games = 10, 0
    assert lib.lib.bb_engine_create(ctypes.byref(cfg), ctypes.byref(h)) == lib.BB_EINVAL
    with pytest.raises(ValueError):
        lib.check(lib.BB_EILLEGAL)
    with pytest.raises(AssertionError):
        lib.check(lib.BB_EASSERT)
    with pytest.raises(MemoryError):
        lib.check(lib.BB_ENOMEM)


def test_record_wire_format_against_reference_bytes(lib, golden):
    """ExampleState.SerializeState / FromSerialized (Blackbird.py:44-81) on the golden records."""
    from blackbird_b200.Blackbird import ExampleState
    n = 0
    for case in golden('records.json'):
        for hexrec in case['records']:
            raw = bytes.fromhex(hexrec)
            ex = ExampleState.FromSerialized(raw)
            assert ex.Board.dtype == np.int8 and ex.Board.ndim == 4 and ex.Board.shape[0] == 1
            assert ExampleState(ex.MctsEval[0], ex.MctsPolicy, ex.Board).SerializeState() == raw
            n += 1
    assert n > 100
    # Connect4 record = 200 bytes with a non-zero label
    ex = ExampleState(1, np.full(7, 1 / 7), np.zeros((1, 6, 7, 3), np.int8))
    b = ex.SerializeState()
    assert len(b) == 200 and b[:5] == bytes.fromhex('0d0000803f') and b[-9:] == bytes.fromhex('22040106070' + '32a0107')


def test_search_mirror_argument_checks_without_gpu(lib):
    from blackbird_b200.FixedMCTS import FixedMCTS
    from blackbird_b200.DynamicMCTS import DynamicMCTS
    from blackbird_b200.MCTS import MCTS
    with pytest.raises(ValueError, match='MaxDepth for MCTS must be > 0.'):
        FixedMCTS(maxDepth=-1, explorationRate=0.85)
    m = DynamicMCTS(explorationRate=0.85, timeLimit=None, playLimit=10, maxDepth=10, temperature={'exploration': 1})
    assert (m.ExplorationRate, m.TimeLimit, m.PlayLimit, m.Root) == (0.85, None, 10, None)
    with pytest.raises(TypeError, match='State not of type GameState'):
        m.FindMove('not a state')
    m.DropRoot()
    m.MoveRoot(None)                         # Root is None: no-op (MCTS.py:274-275)
    m.ResetRoot()
    with pytest.raises(NotImplementedError):
        MCTS(0.85)._findLeaf(None, 1)


def test_datamanager_schema_and_roundtrip(lib, tmp_path, monkeypatch):
    """DataManager.py:6-125: same tables, GetLastVersion registers version 1, PutGames/GetGames by (name, version)."""
    monkeypatch.chdir(tmp_path)
    from blackbird_b200.DataManager import Connection
    c = Connection()
    assert c.GetLastVersion('Connect4', 'bb') == 1
    assert c.GetLastVersion('Connect4', 'bb') == 1
    c.PutGames('bb', 1, 'Connect4', [b'\x01\x02', b'\x03'])
    c.PutModel('Connect4', 'bb', 2)
    c.PutGames('bb', 2, 'Connect4', [b'\x09'])
    assert c.GetGames('bb', 1) == [b'\x01\x02', b'\x03'] and c.GetGames('bb', 2) == [b'\x09']
    assert c.GetLastVersion('Connect4', 'bb') == 2
    c.PutTrainingStatistic(1, 'bb', 2, 'RANDOM')
    c.DumpToZip('bb', 1)
    import gzip
    assert gzip.open(tmp_path / 'data' / 'states_bb_v1.gz').read() == b'\x01\x02\x07\x07\x07\x03\x07\x07\x07'
    with pytest.raises(sqlite3.IntegrityError):
        c.PutGames('unknown', 1, 'Connect4', [b'x'])
    c.Close()
    db = sqlite3.connect(str(tmp_path / 'data' / 'blackbird.db'))
    tables = {r[0] for r in db.execute("SELECT name FROM sqlite_master WHERE type='table'")}
    assert {'ModelDim', 'ConfigurationDim', 'TrainingStatisticsFact', 'JobQueue', 'GameStateFact'} <= tables
    assert db.execute("SELECT Name, Version FROM ModelDim ORDER BY ModelKey LIMIT 2").fetchall() == [('RANDOM', 0), ('MCTS', 0)]
    cols = [r[1] for r in db.execute('PRAGMA table_info(GameStateFact)')]
    assert cols == ['GameStateKey', 'ModelKey', 'GameType', 'State']


def test_product_does_not_import_oracle():
    """oracle/ is test infrastructure: no module of the product package may import or execute it."""
    pkg = os.path.join(ROOT, 'blackbird_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
    import subprocess
    import sys
    code = ('import sys; sys.path.insert(0, %r); import blackbird_b200.Blackbird, blackbird_b200.records, blackbird_b200.parallel; '
            'assert not [m for m in sys.modules if m == "oracle" or m.startswith("oracle.")]') % ROOT
    subprocess.check_call([sys.executable, '-c', code])
