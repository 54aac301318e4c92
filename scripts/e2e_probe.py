#!/usr/bin/env python
"""Per-rank time of the host-buffer do_play call (tafl_step_status_restart, pinned buffers) for each TAFL_HOST_IO mode
(bit 0: the kernel reads the plays in place over PCIe, bit 1: it writes status / error bytes in place; 0 = copy engine both
ways).  Run alone or under torch.distributed.run with one rank per GPU; every rank prints its own line.
    python -m torch.distributed.run --nproc-per-node 8 ... scripts/e2e_probe.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402
from alphazeroforhnefatafl_b200._lib import lib  # noqa: E402

rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def physical_cores():
    """logical CPUs grouped by physical core (thread_siblings_list), within what this process may use"""
    allowed, cores, seen = os.sched_getaffinity(0), [], set()
    for c in sorted(allowed):
        if c in seen:
            continue
        try:
            with open("/sys/devices/system/cpu/cpu%d/topology/thread_siblings_list" % c) as f:
                sib = set()
                for part in f.read().strip().split(","):
                    lo, _, hi = part.partition("-")
                    sib.update(range(int(lo), int(hi or lo) + 1))
        except OSError:
            sib = {c}
        sib &= allowed
        seen |= sib
        cores.append(sorted(sib))
    return cores


if os.environ.get("PIN", "0") == "1" and world > 1:   # whole physical cores per rank: no two ranks share a core's hyperthreads
    cores = physical_cores()
    per = max(1, len(cores) // world)
    mine = [c for core in cores[local * per:(local + 1) * per] for c in core]
    if mine:
        os.sched_setaffinity(0, mine)
    if rank == 0:
        print("pinning: %d physical cores, %d per rank, rank 0 on CPUs %s" % (len(cores), per, mine), flush=True)
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, plies = 65536, 128
L = lib()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


for mode in os.environ.get("MODES", "3,0,1,2").split(","):
    os.environ["TAFL_HOST_IO"] = mode
    g = E.LockstepGames(E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN, n, local, 128)
    g.set_game_id_base(rank * n)
    g.step_random(7, 200, auto_reset=True, steps_per_launch=50)
    steady = g.get_states()
    rec = torch.empty((plies, n, 4), dtype=torch.uint8).pin_memory()
    st = torch.empty((plies, n), dtype=torch.uint8).pin_memory()
    er = torch.empty((plies, n), dtype=torch.uint8).pin_memory()
    for p in range(plies):
        g.step_random(8, 1, auto_reset=True)
        moves, _, _ = g.last_actions()
        rec[p].numpy()[:] = E.moves_to_plays(moves, 128).view(np.uint8).reshape(n, 4)
    ptrs = [(rec[p].data_ptr(), st[p].data_ptr(), er[p].data_ptr()) for p in range(plies)]
    best = []
    for rep in range(4):
        g.set_states(steady)
        barrier()
        t0 = time.perf_counter()
        for a, b, c in ptrs:
            rc = L.tafl_step_status_restart(g.handle, a, b, c)
            assert rc == 0, rc
        dt = time.perf_counter() - t0
        barrier()
        best.append(1e6 * dt / plies)
    print("rank %d host_io %s: us per ply %s" % (rank, mode, " ".join("%.1f" % x for x in best)), flush=True)
    g.close()
    barrier()
if world > 1:
    dist.destroy_process_group()
