"""Host-side mirror of the reference's ``class MCTS`` (src/mcts.py:11-136) over the CUDA engine.

The reference object is ``MCTS(game, nnet, args)`` with ``getActionProb(canonicalBoard, temp)``
running ``args.numMCTSSims`` x ``search`` and returning the visit-count policy.  Here one object
drives the search trees of ALL resident games of a ``LockstepGames`` engine at once (one tree per
game on the GPU); names and argument meaning follow the reference:

    args.numMCTSSims, args.cpuct      (mcts.py:37, 112)
    getActionProb(temp=1)             (mcts.py:28-53)  -> moves, probs, counts  (batched)
    search()                          (mcts.py:55-136) -> runs the batched simulations only

Differences that the reference leaves undefined and this build defines (SURVEY.md §A.11,
DESIGN.md): the default evaluator is the fixed stub (uniform priors over the legal moves, value
0) fused into one kernel; a real evaluator plugs in as ``nnet`` (see :class:`MCTS`); action
index = position in the canonical legal-move list of the position; tree (not transposition-merged)
node identity; temp == 0 picks the FIRST maximum instead of a random one among ties.
"""
from types import SimpleNamespace

import numpy as np

EPS = 1e-8   # mcts.py:6


class MCTS:
    def __init__(self, games, args=None, nnet=None):
        """games: engine.LockstepGames; args: object with numMCTSSims and cpuct (mcts.py:16-26).

        nnet: None -> the stub evaluator, whole search in one fused kernel.  Otherwise the batched
        form of ``nnet.predict`` (mcts.py:85): an object with ``predict_batch(batch)`` that, for
        the ``_abi.LeafBatch`` of device pointers it is given, enqueues work on ``batch.stream``
        that writes ``d_priors`` [n, MAX_MOVES] f32 (action a = a-th canonical legal move of the
        leaf, ``d_moves``) and ``d_values`` [n] f32 for every row with ``d_counts > 0``; or one of
        the built-in device evaluators as ``("stub",)`` / ``("hash", seed)``."""
        self.nnet = nnet
        self.games = games
        self.args = args or SimpleNamespace(numMCTSSims=800, cpuct=1.0)
        self.draw_value = getattr(self.args, "drawValue", 1e-4)
        games.mcts_configure(int(self.args.numMCTSSims), float(self.args.cpuct), float(self.draw_value),
                             int(getattr(self.args, "edgeCapacity", 0)))

    def search(self, sync=True):
        """numMCTSSims simulations from a fresh tree rooted at every game's current state."""
        from . import _abi as A
        if self.nnet is None:
            self.games.mcts_search(sync=sync)
        elif isinstance(self.nnet, tuple):
            kind = {"stub": A.EVAL_STUB, "hash": A.EVAL_HASH}[self.nnet[0]]
            self.games.mcts_search_builtin(kind, int(self.nnet[1]) if len(self.nnet) > 1 else 0)
        else:
            self.games.mcts_search_with(self.nnet.predict_batch)

    def getActionProb(self, temp=1):  # noqa: N802  (reference name)
        """Returns (moves [n, MAX_MOVES] u16, probs [n, MAX_MOVES] f64, counts [n]): for game i the
        first counts[i] entries are its legal moves in canonical order and their probabilities
        (mcts.py:40-53).  Games that are over have counts == 0."""
        self.search()
        probs = self.games.mcts_policy(temp)   # mcts.py:43-53 on the device (temp 0: first max, no random tie choice)
        r = self.games.mcts_results()
        self.last = r
        return r["moves"], probs, r["counts"]

    def choose(self, temp=1, seed=0):
        """Draw the move to play from the temp policy of the last search (the Coach's
        np.random.choice(p=pi); build-defined counter-based draw, see include/tafl.h)."""
        self.games.mcts_policy(temp, choose=True, seed=seed, want_probs=False)

    def play_best(self):
        """Apply the temp == 0 choice of the last search in every Ongoing game; returns statuses."""
        return self.games.mcts_play_best(True)

    def stats(self):
        return self.games.mcts_stats().as_dict()
