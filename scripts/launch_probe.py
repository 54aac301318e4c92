#!/usr/bin/env python
"""Host time to enqueue one kernel through the C ABI: a launch that carries the 1.4 KB rule context as a
__grid_constant__ parameter (tafl_legal_moves_async) against one that does not (tafl_reset_finished), on a
64-game engine whose kernels take ~2 us, 4000 calls back to back without a sync."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402
from alphazeroforhnefatafl_b200._lib import lib  # noqa: E402

g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, 64, 0, 128)
L, h = lib(), g.handle
g.legal_moves_device()
for name, fn in (("with the rule context (k_legal_moves_g)", lambda: L.tafl_legal_moves_async(h, None, None)),
                 ("without (k_reset_finished)", lambda: L.tafl_reset_finished(h, None)),
                 ("python call overhead only (tafl_launch_count)", lambda: L.tafl_launch_count(h))):
    for _ in range(200):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4000):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("%-50s %.2f us per call to enqueue, %.2f us per call until the GPU is done" % (name, 1e6 * (t1 - t0) / 4000, 1e6 * (t2 - t0) / 4000))
g.close()
