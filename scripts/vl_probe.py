#!/usr/bin/env python
"""sims/s of the fused stub search against the virtual-loss width W (4096 Copenhagen trees x 800 sims, 20-ply roots).
W = 0: k_mcts_sims_g (sequential, two trees per warp); W >= 1: k_mcts_sims_vl."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sims = int(sys.argv[2]) if len(sys.argv) > 2 else 800
preset = sys.argv[3] if len(sys.argv) > 3 else "COPENHAGEN"
for W in (0, 1, 2, 4, 8, 16):
    g = E.LockstepGames(getattr(E.preset.rules, preset), getattr(E.preset.boards, preset), n, 0, 128)
    g.step_random(20261104, 20, auto_reset=True)
    g.mcts_configure(sims, 1.0, 1e-4, vl_width=W)
    st = torch.cuda.ExternalStream(g.stream, device=0)
    g.mcts_search()
    s0 = g.mcts_stats().as_dict()
    with torch.cuda.stream(st):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(5):
            g.mcts_search(sync=False)
        b.record(st)
        torch.cuda.synchronize()
    s1 = g.mcts_stats().as_dict()
    ms = a.elapsed_time(b)
    d = {k: s1[k] - s0[k] for k in s0}
    print("W=%2d  %.3e sims/s  %.2f ms per search  nodes/sim %.3f  mean depth %.3f  terminal %.4f" % (
        W, d["sims"] / (ms * 1e-3), ms / 5, d["nodes"] / d["sims"], d["depth_sum"] / d["sims"], d["terminal_hits"] / d["sims"]), flush=True)
    g.close()
