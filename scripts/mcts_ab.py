#!/usr/bin/env python
"""A/B probe for the fused MCTS search: sims/s of 5 searches of 4096 Copenhagen trees x 800 sims from 20-ply roots,
for the library and knobs selected by the environment (TAFL_LIB, TAFL_MCTS_G, ...).
    TAFL_LIB=.../libtafl_b200_x.so python scripts/mcts_ab.py [label] [preset]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402

label = sys.argv[1] if len(sys.argv) > 1 else "run"
preset = sys.argv[2] if len(sys.argv) > 2 else "COPENHAGEN"
hook = os.environ.get("AB_HOOK", "0") == "1"
g = E.LockstepGames(getattr(E.preset.rules, preset), getattr(E.preset.boards, preset), 4096, 0, 128)
g.step_random(20261104, 20, auto_reset=True)
g.mcts_configure(800, 1.0, 1e-4, vl_width=int(os.environ.get("AB_VL", "0")))
st = torch.cuda.ExternalStream(g.stream, device=0)
run = (lambda sync=True: g.mcts_search_builtin(A.EVAL_STUB)) if hook else g.mcts_search
run()
s0 = g.mcts_stats().as_dict()
reps = 2 if hook else 5
with torch.cuda.stream(st):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(reps):
        run(sync=False) if not hook else run()
    b.record(st)
    torch.cuda.synchronize()
s1 = g.mcts_stats().as_dict()
ms = a.elapsed_time(b)
d = {k: s1[k] - s0[k] for k in s0}
print("%-28s %.3e sims/s  %.2f ms per search  mean depth %.3f" % (label, d["sims"] / (ms * 1e-3), ms / reps, d["depth_sum"] / d["sims"]), flush=True)
g.close()
