#!/usr/bin/env python
"""Tuning probe: microseconds per 65 536-game ply of k_step_random under different launch modes
(plain launches, CUDA graph, graph + PDL edges, and several plies per launch = no relaunch at all)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def run(label, env, spl, plies=64, reps=60):
    for k in ("TAFL_STEP_GRAPH", "TAFL_STEP_PDL", "TAFL_STEP_MINB"):
        os.environ.pop(k, None)
    os.environ.update(env)
    from alphazeroforhnefatafl_b200 import engine as E
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, 65536, 0, 128)
    st = torch.cuda.ExternalStream(g.stream, device=0)
    for _ in range(5):
        g.step_random(1, plies, steps_per_launch=spl, sync=False)
    g.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    t0 = time.perf_counter()
    for _ in range(reps):
        g.step_random(1, plies, steps_per_launch=spl, sync=False)
    t1 = time.perf_counter()
    b.record(st)
    g.sync()
    us = a.elapsed_time(b) * 1e3 / (reps * plies)
    print("%-44s %7.2f us/ply   (host enqueue %.2f us/ply)" % (label, us, (t1 - t0) * 1e6 / (reps * plies)), flush=True)
    g.close()


if __name__ == "__main__":
    run("plain launches", {"TAFL_STEP_GRAPH": "0"}, 1)
    run("graph, full-serialisation edges", {"TAFL_STEP_GRAPH": "1", "TAFL_STEP_PDL": "0"}, 1)
    run("graph + PDL edges", {"TAFL_STEP_GRAPH": "1", "TAFL_STEP_PDL": "1"}, 1)
    run("graph + PDL, MINB=4", {"TAFL_STEP_GRAPH": "1", "TAFL_STEP_PDL": "1", "TAFL_STEP_MINB": "4"}, 1)
    for spl in (2, 4, 8, 64):
        run("plain, %d plies per launch" % spl, {"TAFL_STEP_GRAPH": "0"}, spl)
