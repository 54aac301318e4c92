#!/usr/bin/env python
"""Multi-rank probe of the self-play sweep: ms per move with the record gather on / off (torchrun)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E, selfplay as SP  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, sims, moves = 4096, 800, int(os.environ.get("MOVES", "60"))
for mode in sys.argv[1:] or ["gather", "nogather"]:
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, local, 128)
    g.set_game_id_base(rank * n)
    g.step_random(20261104, 20, auto_reset=True)
    kw = {}
    if mode.startswith("every"):
        kw["gather_every"] = int(mode[5:])
    sp = SP.SelfPlay(g, num_sims=sims, overlap=True, restart_finished=True, **kw)
    for _ in range(4):
        sp.move(gather=mode != "nogather")
    sp.flush(gather=mode != "nogather")
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record(sp.stream)
    rows = 0
    for _ in range(moves):
        out = sp.move(gather=mode != "nogather")
        rows += 0 if out is None else out.shape[0]
    out = sp.flush(gather=mode != "nogather")
    rows += 0 if out is None else out.shape[0]
    b.record(sp.stream)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    t = torch.tensor([a.elapsed_time(b) / moves, 1e3 * wall / moves], device="cuda")
    tmax = t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("%-10s world %d: ms/move device min %.3f max %.3f | wall min %.3f max %.3f | rows on rank 0: %d" % (
            mode, world, t[0].item(), tmax[0].item(), t[1].item(), tmax[1].item(), rows), flush=True)
    sp.close()
    g.close()
if world > 1:
    dist.destroy_process_group()
