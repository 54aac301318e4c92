#!/usr/bin/env python
"""N-rank check of tafl_gather_records under torchrun: every rank runs self-play moves on its shard,
the records are gathered twice — through the C ABI (NCCL by dlopen) and through torch.distributed —
and rank 0 compares them row for row.  Ranks hold different numbers of rows (finished games drop out).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/nccl_gather_check.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from alphazeroforhnefatafl_b200 import engine as E  # noqa: E402
from alphazeroforhnefatafl_b200 import selfplay as SP  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = 512 + 64 * rank                                    # ragged shards
    g = E.LockstepGames(E.preset.rules.BRANDUBH, E.preset.boards.BRANDUBH, n, local, 64)
    g.set_game_id_base(rank * 100000)
    g.step_random(3, 10 + 5 * rank, auto_reset=False)
    from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
    sp = SP.SelfPlay(g, num_sims=100, native_gather=False)
    uid = [SP.NcclGather.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ng = SP.NcclGather(g, uid[0], world, rank)
    recv = torch.empty((world * (512 + 64 * world), A.REC_STRIDE), dtype=torch.uint8, device="cuda") if rank == 0 else None
    ok = True
    seen = []
    for mv in range(12):
        rows = sp.move(gather=False).contiguous()
        torch.cuda.synchronize()
        a, counts = ng.gather(rows, recv)
        b = SP.gather_records(rows, 0)
        if rank == 0:
            same = a.shape == b.shape and torch.equal(a.cpu(), b.cpu())
            ok = ok and same
            seen.append(b.cpu().clone())
            print("move %2d rows per rank %s  c-abi == torch.distributed: %s" % (mv, list(counts), same), flush=True)
    ng.close()
    sp.close()
    g.close()
    # The product path: SelfPlay with the native gather on its side stream, collection deferred by one move
    # (overlap=True).  The same games from the same start must yield the same rows, move for move.
    g = E.LockstepGames(E.preset.rules.BRANDUBH, E.preset.boards.BRANDUBH, n, local, 64)
    g.set_game_id_base(rank * 100000)
    g.step_random(3, 10 + 5 * rank, auto_reset=False)
    sp = SP.SelfPlay(g, num_sims=100, overlap=True)
    assert sp.ng is not None
    got = []
    for mv in range(12):
        out = sp.move()
        if mv > 0 and rank == 0:
            got.append(out.cpu().clone())
    out = sp.flush()
    if rank == 0:
        got.append(out.cpu().clone())
        same = len(got) == len(seen) and all(x.shape == y.shape and torch.equal(x, y) for x, y in zip(got, seen))
        ok = ok and same
        print("SelfPlay(overlap=True, native side-stream gather) == torch path over 12 moves: %s" % same, flush=True)
    sp.close()
    g.close()
    dist.barrier()
    if rank == 0:
        print("NCCL GATHER CHECK", "OK" if ok else "FAILED", "world", world, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
