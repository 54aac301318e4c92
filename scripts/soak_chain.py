#!/usr/bin/env python
"""Soak of the CTA-chained step graph at the BASELINE size: 65 536 Copenhagen games x 2048 lockstep plies (32 graph
launches of 64 chained one-ply kernels, auto-restart) on the GPU against the CPU oracle on all host threads — final
states byte for byte, statistics field by field."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import oracle_ffi  # noqa: E402
from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402

n, plies, seed = 65536, int(os.environ.get("SOAK_PLIES", "2048")), 99
orc = oracle_ffi.oracle()
rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
with E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, 0, 128) as g:
    t0 = time.perf_counter()
    for _ in range(plies // 64):
        g.step_random(seed, 64, auto_reset=True, steps_per_launch=1, sync=False)
    got = g.get_states()
    t_gpu = time.perf_counter() - t0
    stats = g.stats().as_dict()
t0 = time.perf_counter()
exp, exp_stats = orc.random_playout(rules, fen, 128, seed, n, plies, auto_reset=True, n_threads=max(1, orc.hw_threads()))
t_cpu = time.perf_counter() - t0
same = got.tobytes() == exp.tobytes()
print("chained graph soak: %d games x %d plies: states %s, statistics %s; GPU %.2f s, oracle %.1f s; %d games finished" % (
    n, plies, "identical" if same else "DIFFER", "identical" if stats == exp_stats.as_dict() else "DIFFER", t_gpu, t_cpu,
    stats["games_finished"]))
sys.exit(0 if same and stats == exp_stats.as_dict() else 1)
