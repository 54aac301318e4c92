"""CPU tests of the multi-GPU host logic: record gather with torch.distributed (gloo, world size 2
and 3, uneven shard sizes), outcome back-fill, the rolling replay window, and the MCTS policy
post-processing (temperature) of the mcts.py mirror."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import selfplay as SP


def _fake_records(rank, n):
    recs = np.zeros(n, SP.REC_DTYPE)
    recs["h"]["game_id"] = 1000 * rank + np.arange(n)
    recs["h"]["ply"] = 7 + rank
    recs["h"]["n_moves"] = 5
    recs["moves"][:, :5] = np.arange(5) + rank
    recs["visits"][:, :5] = 10 * (rank + 1)
    return recs


def _worker(rank, world, port, counts, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        local = torch.from_numpy(_fake_records(rank, counts[rank]).view(np.uint8).reshape(counts[rank], A.REC_STRIDE).copy())
        out = SP.gather_records(local, dst=0)
        if rank == 0:
            q.put(out.numpy().copy())
        else:
            assert out is None
        # a second round with different sizes (an all-empty round too)
        out = SP.gather_records(local[:0], dst=0)
        if rank == 0:
            assert out.shape[0] == 0
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(world, counts):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, counts, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    exp = np.concatenate([_fake_records(r, counts[r]) for r in range(world)])
    assert got.shape == (sum(counts), A.REC_STRIDE)
    assert got.tobytes() == exp.tobytes()   # rank order == game-id order, padding trimmed


def test_gather_records_world2_uneven():
    _run(2, [5, 3])


def test_gather_records_world3_with_empty_rank():
    _run(3, [4, 0, 6])


def test_gather_single_process_passthrough():
    t = torch.zeros((3, A.REC_STRIDE), dtype=torch.uint8)
    assert SP.gather_records(t) is t


def test_backfill_outcomes():
    recs = np.zeros(6, SP.REC_DTYPE)
    h = recs["h"]
    h["game_id"] = [1, 1, 1, 2, 2, 3]
    h["side"] = [A.ATTACKER, A.DEFENDER, A.ATTACKER, A.ATTACKER, A.DEFENDER, A.ATTACKER]
    h["status"] = [0, 0, A.status_win(A.WIN_KING_CAPTURED, False), 0, A.status_draw(A.DRAW_REPETITION), 0]
    z = SP.backfill_outcomes(recs, draw_value=0.0)
    assert list(z[:3]) == [1.0, -1.0, 1.0]
    assert list(z[3:5]) == [0.0, 0.0]
    assert np.isnan(z[5])
    # one game id hosts several episodes (tafl_reset_finished): the result of episode 0 must not leak into episode 1
    recs = np.zeros(5, SP.REC_DTYPE)
    h = recs["h"]
    h["game_id"] = 9
    h["episode"] = [0, 0, 1, 1, 2]
    h["side"] = [A.ATTACKER, A.DEFENDER, A.ATTACKER, A.DEFENDER, A.ATTACKER]
    h["status"] = [0, A.status_win(A.WIN_KING_ESCAPED, True), 0, A.status_win(A.WIN_KING_CAPTURED, False), 0]
    z = SP.backfill_outcomes(recs)
    assert list(z[:4]) == [-1.0, 1.0, 1.0, -1.0] and np.isnan(z[4])


def test_replay_buffer_rolling_window():
    rb = SP.ReplayBuffer(10)
    for k in range(5):
        r = np.zeros(4, SP.REC_DTYPE)
        r["h"]["ply"] = 4 * k + np.arange(4)
        rb.extend(r)
    assert len(rb) == 10
    assert list(rb.as_array()["h"]["ply"]) == list(range(10, 20))   # the oldest entries were dropped


def test_mcts_policy_temperature():
    """MCTS.getActionProb plumbing (mcts.py:43-53) on a fake engine whose mcts_policy is the
    reference formula (the device kernel is checked against the reference in the GPU tests)."""
    from alphazeroforhnefatafl_b200.mcts import MCTS

    class Fake:
        def mcts_configure(self, *a):
            self.cfg = a

        def mcts_search(self, sync=True):
            pass

        def mcts_policy(self, temp=1.0, choose=False, seed=0, want_probs=True):
            r = self.mcts_results()
            v = r["visits"].astype(np.float64)
            p = np.zeros_like(v)
            if temp == 0:
                has = r["counts"] > 0
                p[np.arange(len(v))[has], r["best"][has]] = 1.0
            else:
                c = v ** (1.0 / temp)
                s = c.sum(1, keepdims=True)
                np.divide(c, s, out=p, where=s > 0)
            return p

        def mcts_results(self):
            v = np.zeros((2, A.MAX_MOVES), np.uint32)
            v[0, :3] = [1, 3, 3]
            return dict(moves=np.zeros((2, A.MAX_MOVES), np.uint16), visits=v, q=np.zeros((2, A.MAX_MOVES)),
                        counts=np.array([3, 0], np.uint32), best=np.array([1, 0], np.uint32))

    m = MCTS(Fake())
    _, p, c = m.getActionProb(temp=1)
    assert np.allclose(p[0, :3], [1 / 7, 3 / 7, 3 / 7]) and p[1].sum() == 0 and list(c) == [3, 0]
    _, p, _ = m.getActionProb(temp=0)
    assert list(p[0, :3]) == [0.0, 1.0, 0.0] and p[1].sum() == 0    # first maximum, no random tie choice
    _, p, _ = m.getActionProb(temp=0.5)
    assert np.allclose(p[0, :3], np.array([1, 9, 9]) / 19)
