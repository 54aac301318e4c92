#!/usr/bin/env python
"""The short program ncu profiles: a few launches of each hot kernel at BASELINE sizes.

    python scripts/ncu_target.py [step] [legal] [mcts] [hook]

step : 65 536 Copenhagen games, 300 warm-up plies (steady state) + 8 single-ply launches of k_step_random
legal: 4 launches of k_legal_moves_g on those games
mcts : one search of 4096 trees x 800 sims (k_mcts_sims_g) from 20-ply positions
hook : the same search through the evaluator hook with the built-in stub evaluator
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402

what = set(sys.argv[1:]) or {"step", "legal", "mcts"}
rules, fen = E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN
if what & {"step", "legal"}:
    with E.LockstepGames(rules, fen, 65536, 0, 128) as g:
        g.step_random(1, 300, auto_reset=True, steps_per_launch=100)     # 3 multi-ply launches: steady-state positions
        if "step" in what:
            for _ in range(8):
                g.step_random(1, 1, auto_reset=True)
        if "legal" in what:
            for _ in range(4):
                g.legal_moves_device()
        print("stats", g.stats().as_dict()["steps"])
if what & {"mcts", "hook"}:
    with E.LockstepGames(rules, fen, 4096, 0, 128) as g:
        g.step_random(20261104, 20, auto_reset=True)
        g.mcts_configure(800, 1.0, 1e-4)
        if "mcts" in what:
            g.mcts_search()
        if "hook" in what:
            g.mcts_search_builtin(A.EVAL_STUB)
        print("mcts", g.mcts_stats().as_dict())
