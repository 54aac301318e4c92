"""Host-side mirror of the reference's game API over the C ABI (``include/tafl.h``).

Names follow the reference crate (paths relative to the reference tree, ``game/``):

* ``preset.rules.COPENHAGEN`` ... / ``preset.boards.COPENHAGEN`` ...   — preset.rs:12-134
* ``Play``                                                             — play.rs:23-92
* ``GameLogic`` (by-value batch methods: validate_play, do_play, iter_plays, side_can_play,
  get_captures, detect_shieldwall, find_enclosure, enclosure_secure, detect_exit_fort)
                                                                       — game/logic.rs:61-880
* ``LockstepGames``: ``n_games`` resident ``Game``s advanced together  — game/mod.rs:76-116

Every array argument is a batch: the GPU advances tens of thousands of games per launch.  All
compute happens in libtafl_b200.so on the device; nothing here re-implements rules on the CPU.
"""
import ctypes as C

import numpy as np

from . import _abi as A
from ._lib import TaflError, lib


def _check(rc, eng=None):
    if rc != 0:
        msg = lib().tafl_last_error(eng).decode() if rc == A.ERR_CUDA or eng else ""
        raise TaflError(rc, msg)


def preset_rules(preset_id):
    r = A.Rules()
    _check(lib().tafl_preset_rules(preset_id, C.byref(r)))
    return r


def preset_board(preset_id):
    p = C.c_char_p()
    _check(lib().tafl_preset_board_fen(preset_id, C.byref(p)))
    return p.value.decode()


class _Namespace:
    def __init__(self, fn, names):
        self._fn, self._names = fn, names

    def __getattr__(self, name):
        if name in self._names:
            return self._fn(self._names[name])
        raise AttributeError(name)


class preset:  # noqa: N801  (mirrors the reference module name `preset`)
    rules = _Namespace(preset_rules, dict(COPENHAGEN=A.PRESET_COPENHAGEN, BRANDUBH=A.PRESET_BRANDUBH,
                                          MAGPIE=A.PRESET_MAGPIE, TABLUT=A.PRESET_TABLUT, KOCH=A.PRESET_KOCH))
    boards = _Namespace(preset_board, dict(COPENHAGEN=A.PRESET_COPENHAGEN, BRANDUBH=A.PRESET_BRANDUBH,
                                           MAGPIE=A.PRESET_MAGPIE, TABLUT=A.PRESET_TABLUT))


def state_from_fen(fen, width=0, side=A.ATTACKER):
    """``GameState::new(fen, side)`` (game/state.rs:136-145) -> 1-element STATE_DTYPE array."""
    if width == 0:
        width = A.width_for(_fen_width(fen))
    st = np.zeros(1, A.STATE_DTYPE)
    _check(lib().tafl_state_from_fen(fen.encode(), width, side, A.ptr(st)))
    return st


def _fen_width(fen):
    n, num = 0, ""
    for ch in fen.split("/")[0]:
        if ch.isdigit():
            num += ch
        else:
            n += int(num) if num else 0
            num = ""
            n += 1
    return n + (int(num) if num else 0)


def state_to_fen(st, width):
    buf = C.create_string_buffer(512)
    _check(lib().tafl_state_to_fen(A.ptr(np.ascontiguousarray(st)), width, buf, 512))
    return buf.value.decode()


def board_to_matrix(st, width):
    """board_to_matrix (game/main.rs:55-83) of ONE packed state: [side_len, side_len] uint8 with
    corners 20, throne 30, +1 per soldier, +5 for the king (sides are not distinguished)."""
    st = np.ascontiguousarray(st, A.STATE_DTYPE).reshape(-1)[:1]
    n = int(st["side_len"][0])
    out = np.zeros((n, n), np.uint8)
    _check(lib().tafl_board_to_matrix(A.ptr(st), width, A.ptr(out), n * n))
    return out


def play_from_str(s):
    p = A.Play()
    _check(lib().tafl_play_from_str(s.encode(), C.byref(p)))
    return p


def play_to_str(p):
    buf = C.create_string_buffer(64)
    _check(lib().tafl_play_to_str(C.byref(p), buf, 64))
    return buf.value.decode()


def plays_array(plays):
    """list of A.Play / (from_row, from_col, axis, disp) tuples -> PLAY_DTYPE array"""
    out = np.zeros(len(plays), A.PLAY_DTYPE)
    for i, p in enumerate(plays):
        if isinstance(p, A.Play):
            out[i] = (p.from_row, p.from_col, p.axis, p.disp)
        else:
            out[i] = tuple(p)
    return out


def moves_to_plays(moves, width):
    """compact tafl_move array -> PLAY_DTYPE array (Play::from_tiles, play.rs:36-49)"""
    moves = np.asarray(moves, np.uint16)
    s = A.stride_for(width)
    f, t = (moves & 127).astype(np.int32), ((moves >> 7) & 127).astype(np.int32)
    fr, fc, tr, tc = f // s, f % s, t // s, t % s
    out = np.zeros(moves.shape, A.PLAY_DTYPE)
    hz = fr == tr
    out["from_row"], out["from_col"] = fr, fc
    out["axis"] = np.where(hz, A.HORIZONTAL, A.VERTICAL)
    out["disp"] = np.where(hz, tc - fc, tr - fr)
    return out


class PinnedArray:
    """numpy view of page-locked host memory from tafl_host_alloc: buffers of this kind let the HOST-buffer
    calls (tafl_step_status) run without copy-engine staging.  Keep the object alive while `.a` is in use."""

    def __init__(self, shape, dtype):
        dtype = np.dtype(dtype)
        nbytes = max(1, int(np.prod(shape)) * dtype.itemsize)
        self._p = C.c_void_p()
        _check(lib().tafl_host_alloc(nbytes, C.byref(self._p)))
        buf = (C.c_uint8 * nbytes).from_address(self._p.value)
        self.a = np.frombuffer(buf, np.uint8)[:int(np.prod(shape)) * dtype.itemsize].view(dtype).reshape(shape)

    def close(self):
        if self._p:
            self.a = None
            lib().tafl_host_free(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001  interpreter shutdown
            pass


class LockstepGames:
    """``n_games`` resident games on one GPU, advanced in lockstep (Game, game/mod.rs:76-116)."""

    def __init__(self, rules, start_fen, n_games, device=0, width=0):
        self._h = C.c_void_p()
        self.rules = rules
        rc = lib().tafl_engine_create(C.byref(rules), start_fen.encode(), width, n_games, device, C.byref(self._h))
        if rc != 0:
            raise TaflError(rc, lib().tafl_last_error(None).decode())
        n, w, sl, dev, sm = C.c_uint32(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _check(lib().tafl_engine_info(self._h, C.byref(n), C.byref(w), C.byref(sl), C.byref(dev), C.byref(sm)))
        self.n_games, self.width, self.side_len, self.device, self.sm_count = n.value, w.value, sl.value, dev.value, sm.value
        self.start_fen = start_fen
        self.game_id_base = 0
        self._pin = None

    # -- lifetime -------------------------------------------------------------------------------
    def close(self):
        if self._h:
            lib().tafl_engine_destroy(self._h)
            self._h = C.c_void_p()
        if self._pin is not None:
            for p in self._pin:
                p.close()
            self._pin = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        _check(rc, self._h)

    @property
    def handle(self):
        return self._h

    @property
    def stream(self):
        return lib().tafl_engine_stream(self._h)

    @property
    def launches(self):
        return int(lib().tafl_launch_count(self._h))

    def sync(self):
        self._ck(lib().tafl_engine_sync(self._h))

    @property
    def spec(self):
        """kernel specialisation in use: 0 generic, 1 Copenhagen 11x11, 2 Tablut 9x9, 3 Brandubh 7x7/u64"""
        return int(lib().tafl_engine_spec(self._h))

    def force_generic(self, on=True):
        self._ck(lib().tafl_engine_force_generic(self._h, int(on)))

    # -- resident state -------------------------------------------------------------------------
    def reset(self, game_ids=None):
        if game_ids is None:
            self._ck(lib().tafl_reset(self._h, None, 0))
        else:
            ids = np.ascontiguousarray(game_ids, np.uint32)
            self._ck(lib().tafl_reset(self._h, A.ptr(ids), len(ids)))

    def get_states(self, first=0, n=None):
        n = self.n_games - first if n is None else n
        out = np.zeros(n, A.STATE_DTYPE)
        self._ck(lib().tafl_get_states(self._h, A.ptr(out), first, n))
        return out

    def set_states(self, states, first=0):
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        self._ck(lib().tafl_set_states(self._h, A.ptr(states), first, len(states)))

    def states_device_ptr(self):
        return lib().tafl_states_device(self._h)

    # -- legal moves (get_all_possible_moves, game/main.rs:33-43) ---------------------------------
    def legal_moves(self, first=0, n=None):
        n = self.n_games - first if n is None else n
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        counts = np.zeros(n, np.uint32)
        self._ck(lib().tafl_legal_moves_host(self._h, A.ptr(moves), A.ptr(counts), first, n))
        return moves, counts

    def legal_moves_device(self):
        """enqueue movegen into the engine-owned device buffers; returns (d_moves, d_counts) pointers"""
        self._ck(lib().tafl_legal_moves(self._h, None, None))
        dm, dc = C.c_void_p(), C.c_void_p()
        self._ck(lib().tafl_moves_device(self._h, C.byref(dm), C.byref(dc)))
        return dm.value, dc.value

    # -- stepping -------------------------------------------------------------------------------
    def step(self, plays, want_effects=True):
        """Game::do_play for every resident game; plays is a PLAY_DTYPE array [n_games]."""
        plays = np.ascontiguousarray(plays, A.PLAY_DTYPE)
        assert len(plays) == self.n_games
        fx = np.zeros(self.n_games, A.EFFECTS_DTYPE) if want_effects else None
        self._ck(lib().tafl_step(self._h, A.ptr(plays), A.ptr(fx) if want_effects else None))
        return fx

    def game_do_play(self, plays, restart_finished=False):
        """``Game::do_play`` (game/mod.rs:95-101) for every resident game with the reference's return
        type: (status[n] uint8 packed GameStatus, err[n] uint8 PlayInvalid code).  restart_finished: games that
        had ended are first replaced by new ones, in the same launch (tafl_step_status_restart)."""
        plays = np.ascontiguousarray(plays, A.PLAY_DTYPE)
        assert len(plays) == self.n_games
        if self._pin is None:   # page-locked staging: the kernel reads / writes these in place (tafl_step_status)
            self._pin = (PinnedArray(self.n_games, A.PLAY_DTYPE), PinnedArray(self.n_games, np.uint8), PinnedArray(self.n_games, np.uint8))
        p_pl, p_st, p_er = self._pin
        p_pl.a[:] = plays
        fn = lib().tafl_step_status_restart if restart_finished else lib().tafl_step_status
        self._ck(fn(self._h, A.ptr(p_pl.a), A.ptr(p_st.a), A.ptr(p_er.a)))
        return p_st.a.copy(), p_er.a.copy()

    def last_actions(self):
        """(moves uint16, status uint8, stepped bool) of the most recent random ply (tafl_actions_host)."""
        a = np.empty(self.n_games, np.uint32)
        self._ck(lib().tafl_actions_host(self._h, A.ptr(a), 0, self.n_games))
        return (a & 0xFFFF).astype(np.uint16), ((a >> 16) & 0xFF).astype(np.uint8), ((a >> 24) & 1).astype(bool)

    def step_random(self, seed, n_steps, auto_reset=True, steps_per_launch=1, sync=True):
        fn = lib().tafl_step_random if sync else lib().tafl_step_random_async
        self._ck(fn(self._h, seed, n_steps, int(auto_reset), steps_per_launch))

    def step_random_host(self, states, seed, n_steps=1, auto_reset=True, out=None):
        """By-value random plies with HOST arrays (tafl_step_random_host): returns the successor states."""
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        assert len(states) == self.n_games
        if out is None:
            out = np.empty_like(states)
        self._ck(lib().tafl_step_random_host(self._h, A.ptr(states), A.ptr(out), seed, n_steps, int(auto_reset)))
        return out

    # -- Game<T> history / undo (game/mod.rs:77-108) ------------------------------------------------
    def history_enable(self, capacity):
        """keep (state, PlayRecord) of the last `capacity` valid plays of do_play per game"""
        self._ck(lib().tafl_history_enable(self._h, capacity))

    def undo_last_play(self, mask=None):
        """Game::undo_last_play in every game (or those with mask != 0); no-op where nothing is recorded"""
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self._ck(lib().tafl_undo_last_play(self._h, None if m is None else A.ptr(m)))

    def history_len(self):
        out = np.zeros(self.n_games, np.uint32)
        self._ck(lib().tafl_history_len(self._h, A.ptr(out)))
        return out

    def play_history(self, game, cap=1024):
        out = np.zeros(cap, A.PLAY_RECORD_DTYPE)
        n = C.c_uint32(0)
        self._ck(lib().tafl_play_history(self._h, game, A.ptr(out), cap, C.byref(n)))
        return out[:n.value]

    def set_game_id_base(self, v):
        self._ck(lib().tafl_set_game_id_base(self._h, v))
        self.game_id_base = int(v)

    def set_random_ply(self, v):
        self._ck(lib().tafl_set_random_ply(self._h, v))

    @property
    def random_ply(self):
        return int(lib().tafl_random_ply(self._h))

    def stats_reset(self):
        self._ck(lib().tafl_stats_reset(self._h))

    def stats(self):
        s = A.Stats()
        self._ck(lib().tafl_stats_get(self._h, C.byref(s)))
        return s

    # -- MCTS (src/mcts.py) -----------------------------------------------------------------------
    def mcts_configure(self, num_sims, cpuct=1.0, draw_value=1e-4, edge_capacity=0, tree_reuse=False, node_capacity=0, vl_width=0):
        """vl_width > 1: virtual loss, that many simulations in flight per tree — in the fused stub search and through the
        evaluator hook, where the leaf batch then has n_games * vl_width rows (row = game * vl_width + simulation)"""
        self._hook_width = vl_width if vl_width > 1 else 1
        cfg = A.MctsConfig(num_sims=num_sims, cpuct=cpuct, draw_value=draw_value, edge_capacity=edge_capacity,
                           flags=A.MCTS_TREE_REUSE if tree_reuse else 0, node_capacity=node_capacity, vl_width=vl_width)
        self._ck(lib().tafl_mcts_configure(self._h, C.byref(cfg)))

    def mcts_search(self, sync=True):
        self._ck((lib().tafl_mcts_search if sync else lib().tafl_mcts_search_async)(self._h))

    def mcts_results(self, first=0, n=None):
        n = self.n_games - first if n is None else n
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        visits = np.zeros((n, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, A.MAX_MOVES), np.float64)
        counts = np.zeros(n, np.uint32)
        best = np.zeros(n, np.uint32)
        self._ck(lib().tafl_mcts_results_host(self._h, first, n, A.ptr(moves), A.ptr(visits), A.ptr(q), A.ptr(counts),
                                              A.ptr(best)))
        return dict(moves=moves, visits=visits, q=q, counts=counts, best=best)

    def mcts_policy(self, temp=1.0, choose=False, seed=0, want_probs=True):
        """getActionProb's policy (mcts.py:40-53) on the device; choose=True also draws the move
        tafl_mcts_play_best will play from it.  Returns probs [n, MAX_MOVES] f64 (or None)."""
        probs = np.zeros((self.n_games, A.MAX_MOVES), np.float64) if want_probs else None
        self._ck(lib().tafl_mcts_policy(self._h, float(temp), int(choose), int(seed), A.ptr(probs) if want_probs else None))
        return probs

    def mcts_play_best(self, want_status=True):
        st = np.zeros(self.n_games, np.uint8) if want_status else None
        self._ck(lib().tafl_mcts_play_best(self._h, A.ptr(st) if want_status else None))
        return st

    def mcts_advance(self, want_status=True):
        """play the chosen move and re-root every tree at it (tree reuse, mcts.py:20-26)"""
        st = np.zeros(self.n_games, np.uint8) if want_status else None
        self._ck(lib().tafl_mcts_advance(self._h, A.ptr(st) if want_status else None))
        return st

    def mcts_reroot(self):
        self._ck(lib().tafl_mcts_reroot(self._h))

    def mcts_root_noise(self, noise, eps):
        """Root exploration noise for the evaluator-path searches that follow (tafl_mcts_root_noise):
        priors of each root become (1 - eps) * P + eps * noise[game]; noise None / eps 0 switches it off."""
        if noise is None or eps == 0:
            self._ck(lib().tafl_mcts_root_noise(self._h, None, 0.0))
            return
        noise = np.ascontiguousarray(noise, np.float64)
        assert noise.shape == (self.n_games, A.MAX_MOVES)
        self._ck(lib().tafl_mcts_root_noise(self._h, A.ptr(noise), float(eps)))

    def dirichlet_root_noise(self, alpha, seed):
        """Dirichlet(alpha) rows over each game's legal root moves (zeros beyond), numpy Generator(seed)."""
        counts = self.legal_moves()[1]
        rng = np.random.default_rng(seed)
        noise = np.zeros((self.n_games, A.MAX_MOVES), np.float64)
        for g, L in enumerate(counts):
            if L:
                noise[g, :L] = rng.dirichlet(np.full(int(L), float(alpha)))
        return noise

    def mcts_stats(self):
        s = A.MctsStats()
        self._ck(lib().tafl_mcts_stats_get(self._h, C.byref(s)))
        return s

    # -- evaluator hook: nnet.predict (mcts.py:85) as a batched device boundary ---------------------
    def mcts_begin(self):
        self._ck(lib().tafl_mcts_begin(self._h))

    def mcts_select_leaves(self):
        """Phase 1 of one simulation in every tree; returns the LeafBatch of device pointers."""
        b = A.LeafBatch()
        self._ck(lib().tafl_mcts_select_leaves(self._h, C.byref(b)))
        return b

    def mcts_expand_backup(self):
        self._ck(lib().tafl_mcts_expand_backup(self._h))

    def mcts_finish(self):
        self._ck(lib().tafl_mcts_finish(self._h))

    def eval_builtin(self, evaluator, seed=0):
        self._ck(lib().tafl_eval_builtin(self._h, evaluator, seed))

    def mcts_search_builtin(self, evaluator, seed=0):
        self._ck(lib().tafl_mcts_search_builtin(self._h, evaluator, seed))

    def mcts_search_with(self, fn):
        """fn(batch: LeafBatch) enqueues the evaluator on batch.stream; called once per round (num_sims times, or
        1 + ceil((num_sims - 1) / vl_width) with virtual loss)."""
        err = []

        def tramp(_user, bp):
            try:
                fn(bp.contents)
                return 0
            except Exception as ex:   # noqa: BLE001  (must not unwind through the C frames)
                err.append(ex)
                return -1
        cb = A.EVALUATOR_FN(tramp)
        rc = lib().tafl_mcts_search_with(self._h, cb, None)
        if err:
            raise err[0]
        self._ck(rc)

    @property
    def leaf_rows(self):
        """rows of the evaluator hook's leaf batch"""
        return self.n_games * getattr(self, "_hook_width", 1)

    def leaf_batch_host(self):
        n, cells = self.leaf_rows, self.side_len * self.side_len
        st = np.zeros(n, A.STATE_DTYPE)
        planes = np.zeros((n, cells), np.uint8)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        counts = np.zeros(n, np.uint32)
        self._ck(lib().tafl_leaf_batch_host(self._h, A.ptr(st), A.ptr(planes), A.ptr(moves), A.ptr(counts)))
        return dict(states=st, planes=planes, moves=moves, counts=counts)

    def leaf_eval_host(self, priors, values):
        priors = np.ascontiguousarray(priors, np.float32)
        values = np.ascontiguousarray(values, np.float32)
        assert priors.shape == (self.leaf_rows, A.MAX_MOVES) and values.shape == (self.leaf_rows,)
        self._ck(lib().tafl_leaf_eval_host(self._h, A.ptr(priors), A.ptr(values)))

    def pack_records(self, d_out):
        self._ck(lib().tafl_pack_records(self._h, d_out))

    def pack_records_compact(self, d_out, h_count=None, d_count=None):
        """rows of the searched games only, compacted in game order on the device; h_count: pointer to two
        pinned uint32 (rows, overflow flags of the search), written when the enqueued work has run"""
        self._ck(lib().tafl_pack_records_compact(self._h, d_out, d_count, h_count))

    def reset_finished(self, want_count=False):
        """restart every finished game from the start position (episode counter + 1)"""
        n = C.c_uint32(0)
        self._ck(lib().tafl_reset_finished(self._h, C.byref(n) if want_count else None))
        return n.value if want_count else None

    def mcts_overflow(self):
        """overflow flags of the last search (0 = complete; 1 edge pool, 2 node pool, 4 piece table); syncs"""
        f = C.c_uint32(0)
        self._ck(lib().tafl_mcts_overflow(self._h, C.byref(f)))
        return f.value

    # -- GameLogic by-value batch API (logic.rs) --------------------------------------------------
    def validate_play(self, states, plays):
        states, plays = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_validate_plays(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(out)))
        return out

    def do_play(self, states, plays):
        states, plays = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        out = np.zeros(len(states), A.STATE_DTYPE)
        fx = np.zeros(len(states), A.EFFECTS_DTYPE)
        self._ck(lib().tafl_do_plays(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(out), A.ptr(fx)))
        return out, fx

    def legal_moves_of(self, states):
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        moves = np.zeros((len(states), A.MAX_MOVES), np.uint16)
        counts = np.zeros(len(states), np.uint32)
        self._ck(lib().tafl_legal_moves_of(self._h, A.ptr(states), len(states), A.ptr(moves), A.ptr(counts)))
        return moves, counts

    def legal_moves_of_static(self, states):
        """legal_moves_of through tafl_legal_moves_of_static: device staging in the library's static array, no device
        allocation; at most lib().tafl_static_max_games() states per call (TAFL_ERR_CAPACITY beyond)"""
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        moves = np.zeros((len(states), A.MAX_MOVES), np.uint16)
        counts = np.zeros(len(states), np.uint32)
        self._ck(lib().tafl_legal_moves_of_static(self._h, A.ptr(states), len(states), A.ptr(moves), A.ptr(counts)))
        return moves, counts

    def iter_plays(self, states, tiles_rc):
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        tiles_rc = np.ascontiguousarray(tiles_rc, np.uint8)
        moves = np.zeros((len(states), 32), np.uint16)
        counts = np.zeros(len(states), np.int32)
        self._ck(lib().tafl_piece_plays(self._h, A.ptr(states), A.ptr(tiles_rc), len(states), A.ptr(moves), A.ptr(counts)))
        return moves, counts

    def get_captures(self, states, plays):
        states, plays = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        fx = np.zeros(len(states), A.EFFECTS_DTYPE)
        self._ck(lib().tafl_get_captures(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(fx)))
        return fx

    def detect_shieldwall(self, states, plays):
        states, plays = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        fx = np.zeros(len(states), A.EFFECTS_DTYPE)
        self._ck(lib().tafl_detect_shieldwall(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(fx)))
        return fx

    def find_enclosure(self, states, queries):
        states, queries = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(queries, A.ENCL_QUERY_DTYPE)
        out = np.zeros(len(states), A.ENCLOSURE_DTYPE)
        self._ck(lib().tafl_find_enclosure(self._h, A.ptr(states), A.ptr(queries), len(states), A.ptr(out)))
        return out

    def enclosure_secure(self, states, queries):
        states, queries = np.ascontiguousarray(states, A.STATE_DTYPE), np.ascontiguousarray(queries, A.ENCL_QUERY_DTYPE)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_enclosure_secure(self._h, A.ptr(states), A.ptr(queries), len(states), A.ptr(out)))
        return out

    def detect_exit_fort(self, states):
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_detect_exit_fort(self._h, A.ptr(states), len(states), A.ptr(out)))
        return out

    def side_can_play(self, states, sides):
        states = np.ascontiguousarray(states, A.STATE_DTYPE)
        sides = np.ascontiguousarray(sides, np.uint8)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_side_can_play(self._h, A.ptr(states), A.ptr(sides), len(states), A.ptr(out)))
        return out


class WideLogic:
    """``GameLogic`` over the large bitfields — ``BitfieldBoardState<U256>`` (up to 15x15) and ``<U512>`` (up to
    21x21), game/board/state.rs:337-340 — by value: every method takes a batch of ``_abi.STATE_W_DTYPE`` states and
    runs one CUDA kernel over it (``tafl_w_*``, include/tafl.h "Large boards").  Move lists are ``PLAY_DTYPE`` arrays
    (a tile index of a 21x21 board does not fit the compact ``tafl_move``)."""

    def __init__(self, rules, side_len, width=0, device=0):
        self._h = C.c_void_p()
        self.rules, self.side_len = rules, side_len
        self.width = width or A.width_for(side_len)
        rc = lib().tafl_wide_create(C.byref(rules), side_len, self.width, device, C.byref(self._h))
        if rc != 0:
            raise TaflError(rc, lib().tafl_wide_last_error(None).decode())

    def close(self):
        if self._h:
            lib().tafl_wide_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def _ck(self, rc):
        if rc != 0:
            raise TaflError(rc, lib().tafl_wide_last_error(self._h).decode())

    @staticmethod
    def state_from_fen(fen, width, side=A.ATTACKER):
        st = np.zeros(1, A.STATE_W_DTYPE)
        _check(lib().tafl_w_state_from_fen(fen.encode(), width, side, A.ptr(st)))
        return st

    @staticmethod
    def state_to_fen(st, width):
        buf = C.create_string_buffer(1024)
        _check(lib().tafl_w_state_to_fen(A.ptr(np.ascontiguousarray(st, A.STATE_W_DTYPE)), width, buf, 1024))
        return buf.value.decode()

    @staticmethod
    def move_piece(st, width, frm, to):
        _check(lib().tafl_w_board_move_piece(A.ptr(st), width, frm[0], frm[1], to[0], to[1]))

    def _st(self, states):
        return np.ascontiguousarray(states, A.STATE_W_DTYPE)

    def validate_plays(self, states, plays):
        states, plays = self._st(states), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_w_validate_plays(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(out)))
        return out

    def do_plays(self, states, plays):
        states, plays = self._st(states), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        out, fx = np.zeros(len(states), A.STATE_W_DTYPE), np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._ck(lib().tafl_w_do_plays(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(out), A.ptr(fx)))
        return out, fx

    def legal_plays_of(self, states):
        states = self._st(states)
        plays = np.zeros((len(states), A.W_MAX_PLAYS), A.PLAY_DTYPE)
        counts = np.zeros(len(states), np.uint32)
        self._ck(lib().tafl_w_legal_plays_of(self._h, A.ptr(states), len(states), A.ptr(plays), A.ptr(counts)))
        return plays, counts

    def piece_plays(self, states, tiles_rc):
        states, tiles_rc = self._st(states), np.ascontiguousarray(tiles_rc, np.uint8)
        plays = np.zeros((len(states), 64), A.PLAY_DTYPE)
        counts = np.zeros(len(states), np.int32)
        self._ck(lib().tafl_w_piece_plays(self._h, A.ptr(states), A.ptr(tiles_rc), len(states), A.ptr(plays), A.ptr(counts)))
        return plays, counts

    def get_captures(self, states, plays):
        states, plays = self._st(states), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        fx = np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._ck(lib().tafl_w_get_captures(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(fx)))
        return fx

    def detect_shieldwall(self, states, plays):
        states, plays = self._st(states), np.ascontiguousarray(plays, A.PLAY_DTYPE)
        fx = np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._ck(lib().tafl_w_detect_shieldwall(self._h, A.ptr(states), A.ptr(plays), len(states), A.ptr(fx)))
        return fx

    def find_enclosure(self, states, queries):
        states, queries = self._st(states), np.ascontiguousarray(queries, A.ENCL_QUERY_DTYPE)
        out = np.zeros(len(states), A.ENCLOSURE_W_DTYPE)
        self._ck(lib().tafl_w_find_enclosure(self._h, A.ptr(states), A.ptr(queries), len(states), A.ptr(out)))
        return out

    def enclosure_secure(self, states, queries):
        states, queries = self._st(states), np.ascontiguousarray(queries, A.ENCL_QUERY_DTYPE)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_w_enclosure_secure(self._h, A.ptr(states), A.ptr(queries), len(states), A.ptr(out)))
        return out

    def detect_exit_fort(self, states):
        states = self._st(states)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_w_detect_exit_fort(self._h, A.ptr(states), len(states), A.ptr(out)))
        return out

    def side_can_play(self, states, sides):
        states, sides = self._st(states), np.ascontiguousarray(sides, np.uint8)
        out = np.zeros(len(states), np.uint8)
        self._ck(lib().tafl_w_side_can_play(self._h, A.ptr(states), A.ptr(sides), len(states), A.ptr(out)))
        return out
