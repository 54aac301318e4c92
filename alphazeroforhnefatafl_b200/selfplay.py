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


def _wait(stream):
    """host wait for a stream through a blocking-sync event: the waiting thread sleeps instead of spinning on a core
    that another rank's launch thread or an NCCL proxy thread could use (eight ranks share one host)"""
    ev = torch.cuda.Event(blocking=True)
    ev.record(stream)
    ev.synchronize()


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

    def gather(self, local, recv=None, stream=None, wait=True):
        """local: [n_local, REC_STRIDE] uint8 CUDA tensor; returns (rows tensor on the root / None, counts).
        recv (root): [capacity, REC_STRIDE] uint8 CUDA tensor; default capacity = n_games * n_ranks rows,
        pass a larger one when the shards are ragged.  stream: a torch.cuda.Stream to run the exchange on
        (None = the engine's stream); the call returns when the rows have landed.  wait=False (ranks other than the
        root): only enqueue — no row counts come back, the host does not block."""
        from ._lib import lib
        assert local.dtype == torch.uint8 and local.dim() == 2 and local.shape[1] == A.REC_STRIDE and local.is_contiguous()
        counts = np.zeros(self.n_ranks, np.uint32)
        nowait = not wait and self.rank != self.root
        cap = 0
        if self.rank == self.root:
            if recv is None:
                recv = torch.empty((self.g.n_games * self.n_ranks, A.REC_STRIDE), dtype=torch.uint8, device=local.device)
            cap = recv.shape[0]
        self.g._ck(lib().tafl_gather_records_on(self.g._h, self.comm, self.n_ranks, self.rank, self.root,
                                                local.data_ptr() if local.shape[0] else None, local.shape[0],
                                                recv.data_ptr() if recv is not None else None, cap, None if nowait else A.ptr(counts),
                                                stream.cuda_stream if stream is not None else None))
        if nowait:
            return None, None
        if stream is not None:
            _wait(stream)
        else:
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


class DeviceReplayBuffer:
    """The same rolling window kept in HBM on the replay-buffer rank (180 GB: millions of rows fit): a ring
    of ``max_entries`` rows; gathered rows are inserted with device-to-device copies on the gather's
    side stream, so nothing crosses PCIe per move.  ``as_array()`` brings the window to the host, oldest first."""

    def __init__(self, max_entries, device):
        self.max_entries, self.head, self.n = int(max_entries), 0, 0
        with torch.cuda.device(device):
            self.buf = torch.empty((self.max_entries, A.REC_STRIDE), dtype=torch.uint8, device="cuda")

    def extend(self, rows, stream=None):
        k = int(rows.shape[0])
        if k == 0:
            return
        if k > self.max_entries:
            rows, k = rows[k - self.max_entries:], self.max_entries
        ctx = torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream())
        with ctx:
            first = min(k, self.max_entries - self.head)
            self.buf[self.head:self.head + first].copy_(rows[:first])
            if k > first:
                self.buf[:k - first].copy_(rows[first:])
        if stream is not None:
            _wait(stream)            # the source rows may be overwritten two moves from now
        self.head = (self.head + k) % self.max_entries
        self.n = min(self.max_entries, self.n + k)

    def __len__(self):
        return self.n

    def as_array(self):
        torch.cuda.synchronize()
        if self.n < self.max_entries:
            t = self.buf[:self.n]
        else:
            t = torch.cat([self.buf[self.head:], self.buf[:self.head]], 0)
        return records_to_numpy(t)


class SelfPlay:
    """MCTS self-play of this rank's shard of games; records of every move are gathered to
    ``dst``.  ``games`` is an engine.LockstepGames whose game-id base is already the shard's.

    Record path: ``tafl_pack_records_compact`` compacts the rows of the searched games
    on the device (stable, game order) into one of two row buffers and writes the row count and the
    search's overflow flags into pinned host words — no host synchronisation, no eager tensor ops on
    the engine's stream.  On several GPUs the rows travel with ``tafl_gather_records_on`` (NCCL
    through the C ABI, ragged, no padding) on a SIDE stream ordered behind the pack by an event.
    With ``overlap=True`` the collection of move k is deferred until move k+1 has been enqueued, so
    the gather runs under the next search and the one host wait it needs never idles the GPU;
    ``move()`` then returns the rows of the PREVIOUS move and ``flush()`` the last ones.
    """

    def __init__(self, games, num_sims=800, cpuct=1.0, draw_value=1e-4, dst=0, group=None, edge_capacity=0,
                 tree_reuse=False, evaluator=None, root_noise=None, overlap=False, restart_finished=False,
                 native_gather=None):
        """tree_reuse: one MCTS object per episode as in the reference (mcts.py:20-26) — the subtree
        below the move played is kept.  evaluator: None = stub fused into the search kernel;
        ("stub",) / ("hash", seed) = built-in device evaluators through the evaluator hook; or an
        object with predict_batch(batch) (see mcts.MCTS).  root_noise: (alpha, eps) = Dirichlet(alpha)
        exploration noise mixed into each root's priors with weight eps (the reference's TODO at
        src/mcts.rs:53; evaluator searches only), drawn per move from numpy Generator(seed of move()).
        restart_finished: finished games start over after their last record (tafl_reset_finished; the
        record header's episode counter tells the episodes of one game id apart).
        native_gather: None = the C-ABI NCCL gather when torch.distributed runs on NCCL, the
        torch.distributed gather otherwise (gloo in the CPU tests); True / False force one."""
        from . import engine as E
        self.g, self.dst, self.group = games, dst, group
        self.tree_reuse, self.evaluator = tree_reuse, evaluator
        if root_noise is not None and evaluator is None:
            raise ValueError("root_noise needs an evaluator (the fused stub search has implicit uniform priors)")
        self.root_noise, self.moves_done = root_noise, 0
        self.overlap, self.restart_finished = overlap, restart_finished
        games.mcts_configure(num_sims, cpuct, draw_value, edge_capacity, tree_reuse=tree_reuse)
        self.stream = torch.cuda.ExternalStream(games.stream, device=games.device)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if native_gather is None:
            native_gather = self.world > 1 and dist.get_backend(group) == "nccl"
        with torch.cuda.device(games.device):
            self.rec = [torch.zeros((games.n_games, A.REC_STRIDE), dtype=torch.uint8, device="cuda") for _ in range(2)]
            self.ev_pack = [torch.cuda.Event(blocking=True) for _ in range(2)]
            self.ev_sent = [None, None]                  # side-stream event behind the last send of each row buffer
            self.side = torch.cuda.Stream()
        self.cnt = E.PinnedArray((2, 2), np.uint32)      # [buffer][rows, overflow flags], written by the device
        self.pending = None                              # buffer whose rows have not been collected yet
        self.buf = 0
        self.ng, self.recv = None, None
        if native_gather and self.world > 1:
            uid = [NcclGather.unique_id() if self.rank == 0 else None]
            dist.broadcast_object_list(uid, src=0, group=group)
            self.ng = NcclGather(games, uid[0], self.world, self.rank, root=dst)
            n_all = torch.tensor([games.n_games], dtype=torch.int64, device="cuda")
            dist.all_reduce(n_all, group=group)          # rows the root must be able to hold: every rank's games
            if self.rank == dst:
                with torch.cuda.device(games.device):
                    self.recv = [torch.empty((int(n_all.item()), A.REC_STRIDE), dtype=torch.uint8, device="cuda") for _ in range(2)]

    def close(self):
        if self.ng is not None:
            self.side.synchronize()      # sends enqueued without waiting are out before the communicator goes
            self.ng.close()
            self.ng = None
        if self.cnt is not None:
            self.g.sync()
            self.cnt.close()
            self.cnt = None

    def move(self, gather=True, only_ongoing=True, temp=0, seed=0):
        """One self-play move for every Ongoing game: search, choose the move from the ``temp``
        policy (mcts.py:43-53; temp 0 = first max, temp 1 = visit-proportional draw keyed by
        (seed, game id, ply)), play it, pack the records, gather them.  Returns the gathered
        uint8 tensor on ``dst`` (None elsewhere) — with ``overlap=True`` that of the previous move
        (None on the first call)."""
        g, b = self.g, self.buf
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
            if self.ev_sent[b] is not None:   # the rows this buffer held two moves ago have left (device-side wait only)
                self.stream.wait_event(self.ev_sent[b])
                self.ev_sent[b] = None
            if only_ongoing:   # rows of the searched games only, compacted on the device
                g.pack_records_compact(self.rec[b].data_ptr(), A.ptr(self.cnt.a[b]))
            else:
                g.pack_records(self.rec[b].data_ptr())
                self.cnt.a[b, 0] = 0xFFFFFFFF     # "all rows"; overflow is asked for when the rows are collected
            if self.tree_reuse:
                g.mcts_reroot()
            if self.restart_finished:
                g.reset_finished()
            self.ev_pack[b].record(self.stream)
        self.buf ^= 1
        if not self.overlap:
            return self._collect(b, gather)
        prev, self.pending = self.pending, b
        return None if prev is None else self._collect(prev, gather)

    def flush(self, gather=True):
        """overlap mode: the rows of the last move (None when nothing is pending)"""
        prev, self.pending = self.pending, None
        return None if prev is None else self._collect(prev, gather)

    def _collect(self, b, gather):
        """rows of buffer ``b`` once its pack has run: checks the search's overflow flags, then gathers"""
        self.ev_pack[b].synchronize()
        n_rows, flags = int(self.cnt.a[b, 0]), int(self.cnt.a[b, 1])
        if n_rows == 0xFFFFFFFF:
            n_rows, flags = self.g.n_games, self.g.mcts_overflow()
        if flags:
            from ._lib import TaflError
            raise TaflError(A.ERR_CAPACITY, "MCTS search stopped early in at least one tree (overflow flags %d: 1 edge pool, "
                                            "2 node pool, 4 piece table): raise the pool capacities" % flags)
        rows = self.rec[b][:n_rows]
        if not gather or self.world == 1:
            return rows
        if self.ng is None:
            with torch.cuda.stream(self.stream):
                return gather_records(rows, self.dst, self.group)
        self.side.wait_event(self.ev_pack[b])
        # the root waits for the row counts and the rows; every other rank only enqueues its send and carries on
        out, _ = self.ng.gather(rows, self.recv[b] if self.recv is not None else None, stream=self.side, wait=self.rank == self.dst)
        ev = torch.cuda.Event()
        ev.record(self.side)
        self.ev_sent[b] = ev
        return out
