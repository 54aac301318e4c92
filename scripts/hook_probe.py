#!/usr/bin/env python
"""Probe of the evaluator-hook rounds (select leaves -> evaluate -> expand + backup): sims/s for the
built-in stub / hash evaluators; run under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from alphazeroforhnefatafl_b200 import engine as E, selfplay as SP  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sims = int(sys.argv[2]) if len(sys.argv) > 2 else 800
moves = int(sys.argv[3]) if len(sys.argv) > 3 else 2
for ev in (("stub",), ("hash", 5)):
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, 0, 128)
    g.step_random(1, 20, auto_reset=True)
    sp = SP.SelfPlay(g, num_sims=sims, cpuct=1.0, evaluator=ev)
    sp.move(gather=False)
    s0 = g.mcts_stats().as_dict()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(sp.stream)
    for _ in range(moves):
        sp.move(gather=False)
    b.record(sp.stream)
    torch.cuda.synchronize()
    s1 = g.mcts_stats().as_dict()
    ms = a.elapsed_time(b)
    print("%-10s %d trees x %d sims: %.3e sims/s  (%.1f us per round of %d leaves)" %
          (ev[0], n, sims, (s1["sims"] - s0["sims"]) / (ms * 1e-3), ms * 1e3 / (moves * sims), n), flush=True)
    g.close()
