"""Oracle tree search: numpy restatement of MCTS.py / FixedMCTS.py / DynamicMCTS.py and of
Blackbird.Model's two evaluator hooks.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The arithmetic keeps the reference's numpy expressions verbatim where the dtype
chain matters: the oracle of record is "reference code +
numpy 2.x (NEP 50)".  Everything else (tree bookkeeping) is restated iteratively.
"""
from time import time

import numpy as np

np.seterr(divide='ignore', invalid='ignore')      # Blackbird.py:17


class TreeNode(object):
    """MCTS.py:7-98."""
    __slots__ = ('State', 'Value', 'Plays', 'LegalActions', 'Children', 'Parent', 'Priors',
                 '_wr', '_pl')

    def __init__(self, state, legal, priors):
        self.State = state
        self.Value = 0                              # int 0; becomes np.float32 on first add (A.9)
        self.Plays = 0
        self.LegalActions = np.array(legal)
        self.Children = None
        self.Parent = None
        self.Priors = np.multiply(priors, legal)    # MCTS.py:38 (A.12)
        self._wr = np.zeros(len(legal))
        self._pl = np.zeros(len(legal), dtype=float)

    def WinRate(self):
        return self.Value / self.Plays if self.Plays > 0 else 0          # MCTS.py:53

    def ChildPlays(self):
        for i, kid in enumerate(self.Children):                          # MCTS.py:95-98
            if kid is not None:
                self._pl[i] = kid.Plays
        return self._pl

    def ChildWinRates(self):
        for i, kid in enumerate(self.Children):                          # MCTS.py:80-83
            if kid is not None:
                self._wr[i] = kid.WinRate()
        return self._wr

    def ChildProbability(self):
        plays = self.ChildPlays()                                        # MCTS.py:66-68
        total = sum(plays)
        return plays / total if total > 0 else np.zeros(len(plays), dtype=float)


class Search(object):
    """MCTS.py:101-399 (base) with both _findLeaf variants selected by ``kind``.

    kind='dynamic' -> DynamicMCTS.py:14-34; kind='fixed' -> FixedMCTS.py:21-34.
    Counters (sims, expansions, nodes, depth) feed SURVEY.md 8d's byte model.
    """

    def __init__(self, kind, explorationRate, playLimit=None, timeLimit=None, maxDepth=None,
                 rng=None, **_ignored):
        if kind == 'fixed' and maxDepth <= 0:
            raise ValueError('MaxDepth for MCTS must be > 0.')          # FixedMCTS.py:15-16
        self.Kind = kind
        self.MaxDepth = maxDepth
        self.TimeLimit = timeLimit
        self.PlayLimit = playLimit
        self.ExplorationRate = explorationRate
        self.Root = None
        self.Rng = rng if rng is not None else np.random   # reference uses the global stream (MCTS.py:338)
        self.Stats = dict(sims=0, expansions=0, nodes=0, depth=0)

    # -- hooks (the defaults of MCTS.py:346-383 are RolloutSearch below)
    def GetPriors(self, state):
        raise NotImplementedError

    def SampleValue(self, state, player):
        raise NotImplementedError

    # -- tree growth ------------------------------------------------------------
    def _child_state(self, state, action):
        s = state.Copy()                                                 # MCTS.py:227-236
        s.ApplyAction(action)
        return s

    def _new_node(self, state):
        self.Stats['nodes'] += 1
        return TreeNode(state, state.LegalActions(), self.GetPriors(state))

    def AddChildren(self, node):
        self.Stats['expansions'] += 1                                    # MCTS.py:122-139
        node.Children = [None] * len(node.LegalActions)
        for a in range(len(node.LegalActions)):
            if node.LegalActions[a] == 1:
                kid = self._new_node(self._child_state(node.State, a))
                kid.Parent = node
                node.Children[a] = kid

    def DropRoot(self):
        self.Root = None

    def MoveRoot(self, state):
        # MCTS.py:260-282
        if self.Root is None:
            return
        if self.Root.Children is None:
            self.Root = None
            return
        for kid in self.Root.Children:
            if kid is not None and kid.State == state:
                self.Root = kid
                break

    # -- one move -----------------------------------------------------------------
    def FindMove(self, state, temp=0.1, moveTime=None, playLimit=None):
        # MCTS.py:146-199
        endTime = None
        if moveTime is None:
            moveTime = self.TimeLimit
        if moveTime is not None:
            endTime = time() + moveTime
        if playLimit is None:
            playLimit = self.PlayLimit
        if endTime is None and playLimit is None:
            raise ValueError('Not enough information to decide a stop time.')
        if self.Root is None:
            self.Root = self._new_node(state)
        assert self.Root.State == state, 'Primed for the correct input state.'
        self._run(temp, endTime, playLimit)
        action = self._select(self.Root, temp, exploring=False)
        self.LastAction = int(action)
        return (self._child_state(state, action), self.Root.WinRate(), self.Root.ChildProbability())

    def _run(self, temp, endTime, nPlays):
        # MCTS.py:284-303
        endPlays = self.Root.Plays + (nPlays if nPlays is not None else 0)
        while ((endTime is None or (time() < endTime or self.Root.Children is None))
               and (nPlays is None or self.Root.Plays < endPlays)):
            leaf = self._find_leaf(self.Root, temp)
            val = self.SampleValue(leaf.State, leaf.State.PreviousPlayer)
            self._backup(leaf, val, leaf.State.PreviousPlayer)
            self.Stats['sims'] += 1

    def _find_leaf(self, node, temp):
        last = None
        depth = 0
        while True:
            if self.Kind == 'fixed' and depth == self.MaxDepth:          # FixedMCTS.py:23
                break
            if node.Children is None:
                if node.State.Winner(last) is not None:                  # :24-26 in both files
                    break
                self.AddChildren(node)
                if self.Kind == 'dynamic':                               # DynamicMCTS.py:27-28
                    break
            if np.sum(node.LegalActions) == 0:
                break
            last = self._select(node, temp)
            node = node.Children[last]
            depth += 1
        if self.Kind == 'fixed':
            assert last is not None, 'There is at least one legal option.'   # FixedMCTS.py:32
        self.Stats['depth'] += depth
        return node

    def _backup(self, leaf, value, player):
        # MCTS.py:238-258, iteratively; climbs past Root to the top of the chain (A.13)
        node = leaf
        while True:
            node.Plays += 1
            if node.Parent is None:
                break
            if node.Parent.State.Player == player:
                node.Value += value
            else:
                node.Value += 1 - value
            node = node.Parent

    def _select(self, root, temp, exploring=True):
        # MCTS.py:305-342 -- expressions kept verbatim for the dtype chain (A.8, A.15)
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
            choice = self.Rng.choice(len(root.ChildPlays()), p=p)
        assert root.LegalActions[choice] == 1, 'Selected move is legal.'
        return choice


class NetSearch(Search):
    """Search + the evaluator hooks of Blackbird.Model (Blackbird.py:350-389).

    ``evaluator`` exposes getPolicy(planes)->f32[A] and getEvaluation(planes)->f32 scalar
    (Network.py:48-64).  The reference wraps both hooks in lru_cache(4096); with a pure
    evaluator the cache is transparent (A.26) and is left out.
    """

    def __init__(self, evaluator, kind, **mctsConfig):
        super().__init__(kind, **mctsConfig)
        self.Evaluator = evaluator
        self.Evals = 0

    def SampleValue(self, state, player):
        value = self.Evaluator.getEvaluation(state.AsInputArray())      # Blackbird.py:365
        value = (value + 1) * 0.5                                        # f32 under NEP 50 (A.9)
        if state.Player != player:
            value = 1 - value
        assert value >= 0, 'Value: {}'.format(value)
        return value

    def GetPriors(self, state):
        self.Evals += 1
        policy = self.Evaluator.getPolicy(state.AsInputArray()) * state.LegalActions()   # :386
        policy /= np.sum(policy)                                         # pairwise f64 sum (A.11)
        return policy


class RolloutSearch(Search):
    """The base-class evaluator of MCTS.py:346-383: un-normalised all-ones priors and a fresh uniformly random rollout
    on EVERY visit of a leaf.  This is the 'good' opponent of Blackbird.TestGood (FixedMCTS(maxDepth=10,
    explorationRate=0.85, timeLimit=1), Blackbird.py:164).  Values are Python 0 / 1 / 0.5, so Node.Value accumulates
    exactly (int / float, A.9) and WinRate is an f64 division.  Draws: rollouts use rng.choice(actions) -- one masked
    32-bit rejection draw of the MT19937 stream per step (none when a single action is legal)."""

    def GetPriors(self, state):
        return np.array([1] * len(state.LegalActions()))                 # MCTS.py:346-358

    def SampleValue(self, state, player):
        rolloutState = state                                             # MCTS.py:360-383
        winner = rolloutState.Winner()
        while winner is None:
            actions = np.where(rolloutState.LegalActions() == 1)[0]
            action = self.Rng.choice(actions)
            rolloutState = self._child_state(rolloutState, action)
            winner = rolloutState.Winner(action)
        return 0.5 if winner == 0 else int(player == winner)


class KatEvaluator(object):
    """Deterministic integer evaluator of SURVEY.md Appendix B.2 (not in the reference;
    it is the parity vehicle both sides share)."""

    def __init__(self, nActions):
        self.A = nActions

    @staticmethod
    def _h(planes):
        x = planes.astype(np.int64).ravel()
        return int((np.arange(1, x.size + 1, dtype=np.int64) * x).sum() % 1000003)

    def getPolicy(self, planes):
        h = self._h(planes)
        n = np.array([1 + (h * (i + 1)) % 97 for i in range(self.A)], dtype=np.int64)
        return n.astype(np.float32) / np.float32(n.sum())

    def getEvaluation(self, planes):
        h = self._h(planes)
        return np.float32((h % 1009) * 2 - 1008) / np.float32(1009)
