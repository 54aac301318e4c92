#!/usr/bin/env python
"""ncu target for the evaluator hook: one search of 4096 Copenhagen trees x 200 rounds, built-in stub evaluator."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402

with E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, 4096, 0, 128) as g:
    g.step_random(20261104, 20, auto_reset=True)
    g.mcts_configure(200, 1.0, 1e-4)
    g.mcts_search_builtin(A.EVAL_STUB)
    print("mcts", g.mcts_stats().as_dict())
