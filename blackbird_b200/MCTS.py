"""MCTS mirror: the reference's search API (MCTS.py:101-399) over the CUDA engine.

One MCTS object owns ONE engine slot (``Engine(n_games=1)``) and answers FindMove / MoveRoot /
DropRoot / Root with the same argument meaning, return values and exceptions as the reference, so
main.py and Blackbird.TestModels keep working when the import is switched.  The tree lives in the
engine's pool on the GPU; ``Root`` is a read-only snapshot (bb_search_root_stats).  This
per-object path exists for drop-in compatibility; bulk self-play goes through
Blackbird.GenerateTrainingSamples, which runs thousands of slots in one engine.

Evaluator hook.  The reference lets subclasses override GetPriors / SampleValue
(MCTS.py:346-383); Blackbird.Model implements both through Network.getPolicy / getEvaluation
(Blackbird.py:350-389).  The engine reproduces Model's arithmetic on the device
(priors = f32 policy * mask / pairwise f64 sum; v = (value + 1) * 0.5 in f32, flipped by side),
so the hook it honours sits one level lower: a subclass that defines ``getPolicy(planes)`` and
``getEvaluation(planes)`` in Python is served through the engine's external-evaluator mailbox
(bb_eval_fetch / bb_eval_submit), one call per requested state exactly like the reference's
batch-1 sess.run.  A class with neither (plain FixedMCTS / DynamicMCTS: the 'good' opponent of
Blackbird.TestGood) gets the base class' own evaluator on the device (BB_EVAL_ROLLOUT): all-ones
priors and a fresh uniformly random rollout on every leaf visit (MCTS.py:346-383), values exact.
Its rollout draws are np.random.choice calls on numpy's GLOBAL stream in the reference; the mirror
hands the engine the next raw MT19937 outputs of that very stream and afterwards advances the
stream by what the device consumed, so games replay the reference draw for draw.
"""
from time import time

import numpy as np
import torch

from . import _lib
from . import engine as _engine
from .GameState import GameState, DeviceBoard


class Node(object):
    """Read-only view of a tree node (MCTS.py:7-98).  Statistics are a snapshot taken after the
    last FindMove / MoveRoot; Children holds one level (views with Children = None)."""

    def __init__(self, state, legalActions, priors, plays=0, value=0, children=None, parent=None):
        self.State = state
        self.Value = value
        self.Plays = plays
        self.LegalActions = np.array(legalActions)
        self.Children = children
        self.Parent = parent
        self.Priors = np.multiply(priors, legalActions) if priors is not None else None
        n = len(self.LegalActions)
        self._childWinRates = np.zeros(n)
        self._childPlays = np.zeros(n, dtype=float)

    def WinRate(self):
        return self.Value / self.Plays if self.Plays > 0 else 0          # MCTS.py:53

    def ChildProbability(self):
        plays = self.ChildPlays()                                        # MCTS.py:66-68
        total = sum(plays)
        return plays / total if total > 0 else np.zeros(len(plays), dtype=float)

    def ChildWinRates(self):
        for i, kid in enumerate(self.Children or ()):                    # MCTS.py:80-83
            if kid is not None:
                self._childWinRates[i] = kid.WinRate()
        return self._childWinRates

    def ChildPlays(self):
        for i, kid in enumerate(self.Children or ()):                    # MCTS.py:95-98
            if kid is not None:
                self._childPlays[i] = kid.Plays
        return self._childPlays


class MCTS(object):
    """MCTS.py:101-399.  ``_KIND`` selects the engine's _findLeaf variant (FixedMCTS / DynamicMCTS)."""
    _KIND = None
    DEVICE = 0
    _TIME_MODE_SIMS = 4096           # wall-clock searches stop here at the latest: the tree arena is sized for it

    def __init__(self, explorationRate, timeLimit=None, playLimit=None, **kwargs):
        self.TimeLimit = timeLimit
        self.PlayLimit = playLimit
        self.ExplorationRate = explorationRate
        self._eng = None
        self._has_root = False
        self._root_view = None

    # -- engine plumbing --------------------------------------------------------------------------
    def _evaluator_kind(self):
        if getattr(self, '_net_loaded_weights', None) is not None:
            return 'net'
        if callable(getattr(self, 'getPolicy', None)) and callable(getattr(self, 'getEvaluation', None)):
            return 'external'
        return 'rollout'                                                 # the base class' own GetPriors / SampleValue

    def _engine_for(self, state, playLimit):
        if self._KIND is None:
            raise NotImplementedError                                    # MCTS._findLeaf (MCTS.py:385-391)
        game = state.GameType
        hint = int(playLimit or self.PlayLimit or self._TIME_MODE_SIMS)
        if self._eng is not None and (self._eng.game != game or hint > self._eng_hint):
            if self._has_root:
                raise _lib.EngineError('playLimit grew beyond the tree arena of a live Root; DropRoot() first')
            self._eng.close()
            self._eng = None
        if self._eng is None:
            kind = self._evaluator_kind()
            self._eng = _engine.Engine(game, self._KIND, n_games=1, max_depth=getattr(self, 'MaxDepth', None) or 10,
                                       exploration_rate=self.ExplorationRate, evaluator=kind, play_limit_hint=hint,
                                       device=self.DEVICE, lanes=1)
            self._eng_hint = hint
            if kind == 'net':
                self._load_network_into(self._eng)
        return self._eng

    def _serve_requests(self, eng):
        """External evaluator: answer every open request with self.getPolicy / self.getEvaluation,
        one state at a time (Network.py:48-64 is batch 1 too)."""
        k, states, _ = eng.eval_pending()
        planes = _engine.planes(eng.game, states).cpu().numpy()
        pol = np.empty((k, eng.A), np.float32)
        val = np.empty((k,), np.float32)
        for i in range(k):
            pol[i] = self.getPolicy(planes[i:i + 1])
            val[i] = self.getEvaluation(planes[i:i + 1])
        dev = torch.device('cuda', eng.device)
        eng.eval_submit(torch.as_tensor(pol).to(dev), torch.as_tensor(val).to(dev))

    # raw MT19937 words a rollout simulation may consume, by game (a draw per rollout ply, ~1.4 words each; DragonChess
    # playouts run a few hundred plies); running out is reported (BB_ESTATE), never silent
    _ROLLOUT_WORDS = {'TicTacToe': 32, 'Connect4': 160, 'DragonChess': 4096}

    def _simulate(self, eng, n):
        if eng.evaluator == 'rollout':
            budget = int(n) * self._ROLLOUT_WORDS[eng.game] + 256
            peek = np.random.RandomState()
            peek.set_state(np.random.get_state())                        # the words np.random would hand out next
            eng.set_rng(np.frombuffer(peek.bytes(4 * budget), dtype='<u4'))
            eng.begin_move(n)
            eng.run()
            eng.check_status()
            used = int(eng.rng_used()[0])
            if used:
                np.random.bytes(4 * used)                                # the global stream moves by exactly what the rollouts drew
            return
        eng.begin_move(n)
        if eng.evaluator == 'external':
            while eng.run():
                self._serve_requests(eng)
        else:
            eng.run()
        eng.check_status()

    def _snapshot(self, state_cls):
        """Root view from bb_search_root_stats + bb_search_states."""
        eng = self._eng
        st = eng.root_stats()
        root_state = state_cls.FromPacked(eng.states()[0], device=eng.device)
        legal = root_state.LegalActions()
        A = eng.A
        k = int(st['n_children'][0])
        plays = int(st['root_plays'][0])
        # Node.Value stays the int 0 until a parent exists (A.13), then accumulates in f32 (A.9)
        ro = eng.evaluator == 'rollout'                                  # exact Python numbers instead of np.float32 (A.9)
        value = (float(st['root_value'][0]) if ro else np.float32(st['root_value'][0])) if int(st['root_has_parent'][0]) else 0
        priors = np.zeros(A)
        children = None
        if k >= 0:
            children = [None] * A
            for j in range(k):
                a = int(st['actions'][0, j])
                priors[a] = st['child_prior'][0, j]
                kid = Node(None, np.zeros(A), None, plays=int(st['plays'][0, j]),
                           value=float(st['child_value'][0, j]) if ro else np.float32(st['child_value'][0, j]))
                children[a] = kid
        root = Node(root_state, legal, priors, plays=plays, value=value, children=children)
        for kid in (children or ()):
            if kid is not None:
                kid.Parent = root
        self._root_view = root
        return root

    @property
    def Root(self):
        return self._root_view if self._has_root else None

    @Root.setter
    def Root(self, value):
        if value is not None:
            raise AttributeError('Root lives in the engine; only Root = None (DropRoot) can be assigned')
        self.DropRoot()

    # -- reference API ----------------------------------------------------------------------------
    def AddChildren(self, node):
        raise NotImplementedError('expansion happens inside the engine (MCTS.py:122-139 -> csrc/search_dev.cuh expand_node)')

    def DropRoot(self):
        """MCTS.py:141-144."""
        self._has_root = False
        self._root_view = None

    def FindMove(self, state, temp=0.1, moveTime=None, playLimit=None):
        """MCTS.py:146-199: (next state, Root.WinRate(), Root.ChildProbability())."""
        if not isinstance(state, GameState):
            raise TypeError('State not of type GameState')
        endTime = None
        if moveTime is None:
            moveTime = self.TimeLimit
        if moveTime is not None:
            endTime = time() + moveTime
        if playLimit is None:
            playLimit = self.PlayLimit
        if endTime is None and playLimit is None:
            raise ValueError('Not enough information to decide a stop time.')
        if not isinstance(state, DeviceBoard):
            raise TypeError('State not of type GameState')               # host-only states cannot enter the engine
        eng = self._engine_for(state, playLimit)
        if not self._has_root:
            eng.reset(state._st.to(torch.device('cuda', eng.device)))    # Root = Node(state, ...) happens in begin_move
            self._has_root = True
        else:
            assert self._root_view.State == state, 'Primed for the correct input state.'
        # _runMCTS (MCTS.py:284-303)
        if endTime is None:
            self._simulate(eng, int(playLimit))
        else:
            done = 0
            chunk = 8                                                    # wall-clock mode is non-deterministic anyway (A.25)
            while True:
                have_children = done > 0
                if not ((time() < endTime or not have_children) and (playLimit is None or done < playLimit)):
                    break
                if done + chunk > self._eng_hint:
                    break
                n = chunk if playLimit is None else min(chunk, int(playLimit) - done)
                self._simulate(eng, n)
                done += n
        root = self._snapshot(type(state))
        action = self._selectAction(root, temp, exploring=False)
        self._last_action = int(action)
        return (self._applyAction(state, action), root.WinRate(), root.ChildProbability())

    def MoveRoot(self, state):
        self._moveRoot(state)

    def ResetRoot(self):
        """MCTS.py:201-213 walks back to the game's first root.  The engine keeps only the current
        root's subtree (half-arena copy at MoveRoot), and nothing in the reference calls ResetRoot."""
        if not self._has_root:
            return
        raise NotImplementedError('ancestors of Root are not kept by the engine (DESIGN.md, tree pool)')

    def _applyAction(self, state, action):
        s = state.Copy()
        s.ApplyAction(action)
        return s

    def _moveRoot(self, state):
        """MCTS.py:260-282: Root <- the first child whose State == state; Root <- None if unexpanded."""
        if not self._has_root:
            return
        root = self._root_view
        if root is None or root.Children is None:
            self.DropRoot()
            return
        eng = self._eng
        actions = [a for a, kid in enumerate(root.Children) if kid is not None]
        if not actions:
            return
        # children's states in one batched rule-kernel call; Copy() semantics per game (A.6)
        reps = root.State._st.repeat(len(actions), 1)
        if not root.State.COPY_KEEPS_PREV:
            cells, pl, pv, ca = _engine.unpack(eng.game, reps)
            reps = _engine.pack(eng.game, cells.cpu().numpy(), pl.cpu().numpy(), np.zeros(len(actions), np.uint8),
                                ca.cpu().numpy(), device=eng.device)
        _engine.apply(eng.game, reps, torch.tensor(actions, dtype=torch.int32))
        same = _engine.equal(eng.game, reps, state._st.to(reps.device).repeat(len(actions), 1)).cpu().numpy()
        hit = np.flatnonzero(same)
        if hit.size == 0:
            return                                                       # the reference leaves Root where it is
        eng.move_root(np.array([actions[int(hit[0])]], dtype=np.int32))
        ph, _ = eng.phases()
        self._snapshot(type(state))

    def _runMCTS(self, temp, endTime=None, nPlays=None):
        raise NotImplementedError('the simulation loop runs inside the engine (csrc/search_kernels.cuh k_advance)')

    def _selectAction(self, root, temp, exploring=True):
        """MCTS.py:305-342 on the exported root statistics, numpy expression for numpy expression
        (same dtype chain, same draws from the global np.random stream)."""
        assert root.Children is not None, 'The node has children to select.'
        if exploring or temp == 0:
            allPlays = sum(root.ChildPlays())
            upperConfidence = (root.ChildWinRates()
                               + (self.ExplorationRate * root.Priors * np.sqrt(1.0 + allPlays))
                               / (1.0 + root.ChildPlays()))
            choice = np.argmax(upperConfidence)
        else:
            allPlays = sum([p ** (1 / temp) for p in root.ChildPlays()])
            p = [c ** (1 / temp) / allPlays for c in root.ChildPlays()]
            choice = np.random.choice(len(root.ChildPlays()), p=p)
        assert root.LegalActions[choice] == 1, 'Selected move is legal.'
        return choice

    # -- overridables of the reference (kept for interface completeness) ---------------------------
    def GetPriors(self, state):
        """MCTS.py:346-358 default: ones."""
        return np.array([1] * len(state.LegalActions()))

    def SampleValue(self, state, player):
        """MCTS.py:360-383 on the host, for callers; the search plays its rollouts on the device (rollout_value)."""
        rolloutState = state
        winner = rolloutState.Winner()
        while winner is None:
            actions = np.where(rolloutState.LegalActions() == 1)[0]
            action = np.random.choice(actions)
            rolloutState = self._applyAction(rolloutState, action)
            winner = rolloutState.Winner(action)
        return 0.5 if winner == 0 else int(player == winner)

    def _findLeaf(self, node, temp):
        raise NotImplementedError

    def __del__(self):
        eng = getattr(self, '_eng', None)
        if eng is not None:
            eng.close()
