#!/usr/bin/env python
"""A/B probe for the env-step and legal-move kernels: device time per ply / per launch for the library
and knobs selected by the environment (TAFL_LIB, TAFL_STEP_SINGLE, TAFL_STEP_MINB, TAFL_LM_G ...).

    TAFL_LIB=.../libtafl_b200_v1.so python scripts/step_ab.py [label]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402
from alphazeroforhnefatafl_b200._lib import lib  # noqa: E402

label = sys.argv[1] if len(sys.argv) > 1 else "run"
n = int(os.environ.get("AB_GAMES", "65536"))
P = int(os.environ.get("AB_PLIES", "64"))          # plies per graph launch
g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, 0, 128)
st = torch.cuda.ExternalStream(g.stream, device=0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
L = lib()
with torch.cuda.stream(st):
    for _ in range(6):
        g.step_random(20261104, P, auto_reset=True, steps_per_launch=1, sync=False)
    best = []
    for rep in range(5):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(1024 // P):
            g.step_random(20261104, P, auto_reset=True, steps_per_launch=1, sync=False)
        b.record(st)
        torch.cuda.synchronize()
        best.append(1e3 * a.elapsed_time(b) / (1024 // P * P))
    g.legal_moves_device()
    lm = []
    for rep in range(20):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        L.tafl_legal_moves_async(g.handle, None, None)
        b.record(st)
        torch.cuda.synchronize()
        lm.append(1e3 * a.elapsed_time(b))
cs = g.stats().state_checksum
print("%-28s step us/ply min %.3f med %.3f | legal_moves us min %.2f med %.2f | checksum %016x" % (
    label, min(best), sorted(best)[2], min(lm), sorted(lm)[10], cs), flush=True)
g.close()
