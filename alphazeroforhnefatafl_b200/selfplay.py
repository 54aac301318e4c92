"""Self-play generation over game-sharded GPUs (BASELINE.json configs 4-5).

Games shard by id range across ranks (one process per GPU); searching and stepping need no
communication.  The only exchange is the per-move gather of fixed-stride replay records
(``tafl_pack_records``: root position, legal moves, visit counts, chosen move, status) to the
replay-buffer rank with ``torch.distributed`` (NCCL over NVLink on GPUs, gloo in the CPU tests).

Nearest reference analogue: ``write_to_file`` (game/main.rs:86-132), a rolling window of the
newest ``max_entries`` positions -> :class:`ReplayBuffer`.
"""
import collections

import numpy as np
import torch
import torch.distributed as dist

from . import _abi as A

REC_HEADER = np.dtype([("game_id", "<u8"), ("ply", "<u4"), ("n_moves", "<u2"), ("side", "u1"), ("status", "u1"),
                       ("att", "<u8", (2,)), ("def", "<u8", (2,)), ("best", "<u2"), ("_pad", "<u2"), ("episode", "<u4")])
assert REC_HEADER.itemsize == 56
REC_DTYPE = np.dtype([("h", REC_HEADER), ("moves", "<u2", (A.REC_MOVES,)), ("visits", "<u4", (A.REC_MOVES,))])
assert REC_DTYPE.itemsize == A.REC_STRIDE


def gather_records(local, dst=0, group=None):
    """Gather [n_local, REC_STRIDE] uint8 record tensors of every rank to rank ``dst``.

    Ranks may hold different numbers of rows: counts are all-gathered first, rows are padded to
    the maximum for the fixed-size gather and trimmed on arrival.  Returns the concatenated
    [sum n, REC_STRIDE] tensor on ``dst`` (rank order == game-id order), None elsewhere."""
    assert local.dtype == torch.uint8 and local.dim() == 2
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(counts)
    if m == 0:
        return local[:0] if rank == dst else None
    send = local
    if local.shape[0] < m:
        send = torch.zeros((m, local.shape[1]), dtype=torch.uint8, device=local.device)
        send[:local.shape[0]] = local
    send = send.contiguous()
    if rank == dst:
        bufs = [torch.empty_like(send) for _ in range(world)]
        dist.gather(send, bufs, dst=dst, group=group)
        return torch.cat([b[:c] for b, c in zip(bufs, counts)], 0)
    dist.gather(send, None, dst=dst, group=group)
    return None


class NcclGather:
    """The same gather through the C ABI alone (``tafl_gather_records``: NCCL resolved with dlopen,
    counts all-gathered, rows sent straight to their offset on the root) — what a host without
    PyTorch binds.  ``uid`` is the 128-byte ncclUniqueId made on rank 0 (``NcclGather.unique_id()``)
    and shipped to the other ranks by any means."""

    def __init__(self, games, uid, n_ranks, rank, root=0):
        import ctypes as C
        from ._lib import lib
        self.g, self.n_ranks, self.rank, self.root = games, n_ranks, rank, root
        self.comm = C.c_void_p()
        buf = (C.c_char * 128).from_buffer_copy(bytes(uid))
        games._ck(lib().tafl_nccl_comm_create(games._h, buf, n_ranks, rank, C.byref(self.comm)))

    @staticmethod
    def unique_id():
        import ctypes as C
        from ._lib import lib
        buf = (C.c_char * 128)()
        rc = lib().tafl_nccl_unique_id(buf)
        if rc != 0:
            raise RuntimeError("tafl_nccl_unique_id failed (%d): is libnccl.so.2 loadable?" % rc)
        return bytes(buf)

    def gather(self, local, recv=None):
        """local: [n_local, REC_STRIDE] uint8 CUDA tensor; returns (rows tensor on the root / None, counts).
        recv (root): [capacity, REC_STRIDE] uint8 CUDA tensor; default capacity = n_games * n_ranks rows,
        pass a larger one when the shards are ragged."""
        from ._lib import lib
        assert local.dtype == torch.uint8 and local.dim() == 2 and local.shape[1] == A.REC_STRIDE and local.is_contiguous()
        counts = np.zeros(self.n_ranks, np.uint32)
        cap = 0
        if self.rank == self.root:
            if recv is None:
                recv = torch.empty((self.g.n_games * self.n_ranks, A.REC_STRIDE), dtype=torch.uint8, device=local.device)
            cap = recv.shape[0]
        self.g._ck(lib().tafl_gather_records(self.g._h, self.comm, self.n_ranks, self.rank, self.root,
                                             local.data_ptr() if local.shape[0] else None, local.shape[0],
                                             recv.data_ptr() if recv is not None else None, cap, A.ptr(counts)))
        self.g.sync()
        return (recv[:int(counts.sum())] if self.rank == self.root else None), counts

    def close(self):
        from ._lib import lib
        if self.comm:
            lib().tafl_nccl_comm_destroy(self.comm)
            self.comm = None


def records_to_numpy(t):
    """[n, REC_STRIDE] uint8 tensor -> structured REC_DTYPE array (host)."""
    return np.ascontiguousarray(t.cpu().numpy()).view(REC_DTYPE).reshape(-1)


def backfill_outcomes(recs, draw_value=0.0):
    """z for every record: +1 if the side to move at the record's position went on to win its game,
    -1 if it lost, draw_value for a draw, NaN while the game is still unfinished (mcts.py callers
    assign the episode result to every position of the episode)."""
    z = np.full(len(recs), np.nan)
    h = recs["h"]
    final = {}   # one game id hosts many episodes (tafl_reset_finished): the key is (game_id, episode)
    for gid, ep, st in zip(h["game_id"], h["episode"], h["status"]):
        if st & 0x80:
            final[(int(gid), int(ep))] = int(st)
    for i in range(len(recs)):
        st = final.get((int(h["game_id"][i]), int(h["episode"][i])))
        if st is None:
            continue
        if (st & 0xC0) == 0xC0:
            z[i] = draw_value
        else:
            winner = A.DEFENDER if st & 1 else A.ATTACKER
            z[i] = 1.0 if winner == int(h["side"][i]) else -1.0
    return z


class ReplayBuffer:
    """Rolling window of the newest ``max_entries`` records (write_to_file, game/main.rs:98-106:
    when the window is full the oldest entry is dropped)."""

    def __init__(self, max_entries):
        self.max_entries = max_entries
        self.q = collections.deque()
        self.n = 0

    def extend(self, recs):
        if len(recs) == 0:
            return
        self.q.append(recs)
        self.n += len(recs)
        while self.n > self.max_entries:
            over = self.n - self.max_entries
            if len(self.q[0]) <= over:
                self.n -= len(self.q.popleft())
            else:
                self.q[0] = self.q[0][over:]
                self.n -= over

    def __len__(self):
        return self.n

    def as_array(self):
        return np.concatenate(list(self.q)) if self.q else np.zeros(0, REC_DTYPE)


class SelfPlay:
    """MCTS self-play of this rank's shard of games; records of every move are gathered to
    ``dst``.  ``games`` is an engine.LockstepGames whose game-id base is already the shard's."""

    def __init__(self, games, num_sims=800, cpuct=1.0, draw_value=1e-4, dst=0, group=None, edge_capacity=0,
                 tree_reuse=False, evaluator=None, root_noise=None):
        """tree_reuse: one MCTS object per episode as in the reference (mcts.py:20-26) — the subtree
        below the move played is kept.  evaluator: None = stub fused into the search kernel;
        ("stub",) / ("hash", seed) = built-in device evaluators through the evaluator hook; or an
        object with predict_batch(batch) (see mcts.MCTS).  root_noise: (alpha, eps) = Dirichlet(alpha)
        exploration noise mixed into each root's priors with weight eps (the reference's TODO at
        src/mcts.rs:53; evaluator searches only), drawn per move from numpy Generator(seed of move())."""
        self.g, self.dst, self.group = games, dst, group
        self.tree_reuse, self.evaluator = tree_reuse, evaluator
        if root_noise is not None and evaluator is None:
            raise ValueError("root_noise needs an evaluator (the fused stub search has implicit uniform priors)")
        self.root_noise, self.moves_done = root_noise, 0
        games.mcts_configure(num_sims, cpuct, draw_value, edge_capacity, tree_reuse=tree_reuse)
        self.stream = torch.cuda.ExternalStream(games.stream, device=games.device)
        with torch.cuda.device(games.device):
            self.rec = torch.zeros((games.n_games, A.REC_STRIDE), dtype=torch.uint8, device="cuda")

    def move(self, gather=True, only_ongoing=True, temp=0, seed=0):
        """One self-play move for every Ongoing game: search, choose the move from the ``temp``
        policy (mcts.py:43-53; temp 0 = first max, temp 1 = visit-proportional draw keyed by
        (seed, game id, ply)), play it, pack the records, gather them.  Returns the gathered
        uint8 tensor on ``dst`` (None elsewhere)."""
        g = self.g
        with torch.cuda.stream(self.stream):
            ev = self.evaluator
            if self.root_noise is not None:
                alpha, eps = self.root_noise
                g.mcts_root_noise(g.dirichlet_root_noise(alpha, (int(seed), g.game_id_base, self.moves_done)), eps)
            self.moves_done += 1
            if ev is None:
                g.mcts_search(sync=False)
            elif isinstance(ev, tuple):
                g.mcts_search_builtin({"stub": A.EVAL_STUB, "hash": A.EVAL_HASH}[ev[0]], int(ev[1]) if len(ev) > 1 else 0)
            else:
                g.mcts_search_with(ev.predict_batch)
            if temp != 0:
                g.mcts_policy(temp, choose=True, seed=seed, want_probs=False)
            g.mcts_play_best(False)
            g.pack_records(self.rec.data_ptr())
            if self.tree_reuse:
                g.mcts_reroot()
            out = self.rec
            if only_ongoing:   # rows of games that were already over carry n_moves == 0
                n_moves = self.rec[:, 12].to(torch.int32) | (self.rec[:, 13].to(torch.int32) << 8)
                out = self.rec[n_moves > 0]
            if gather:
                out = gather_records(out, self.dst, self.group)
        return out
