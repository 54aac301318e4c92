#!/usr/bin/env python
"""Tuning probe: tafl_step_random_host (by-value states, 64 B per game each way) under TAFL_HOST_CHUNKS."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402


def run(chunks, n=65536, steps=60):
    os.environ["TAFL_HOST_CHUNKS"] = str(chunks)
    from alphazeroforhnefatafl_b200 import engine as E
    from alphazeroforhnefatafl_b200._lib import lib
    L = lib()
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, 0, 128)
    a = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    b = torch.empty(n * 64, dtype=torch.uint8).pin_memory()
    a.numpy()[:] = np.frombuffer(g.get_states().tobytes(), np.uint8)
    best = 1e9
    for r in range(4):
        t0 = time.perf_counter()
        for _ in range(steps):
            assert L.tafl_step_random_host(g.handle, a.data_ptr(), b.data_ptr(), 3, 1, 1) == 0
            a, b = b, a
        best = min(best, time.perf_counter() - t0)
    print("chunks=%2d  %7.1f us/step  %.3e steps/s" % (chunks, best * 1e6 / steps, n * steps / best), flush=True)
    g.close()


if __name__ == "__main__":
    for c in (1, 2, 3, 4, 6, 8, 12, 16, 32):
        run(c)
