"""Test-only loaders for the CPU oracle (oracle/liboracle.so) and for the host build of the
kernel math (tests/hostsim/libhostsim.so).  Nothing in the product package imports this file."""
import ctypes as C
import os
import subprocess

import numpy as np

from alphazeroforhnefatafl_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
HOSTSIM_DIR = os.path.join(ROOT, "tests", "hostsim")
CSRC = os.path.join(ROOT, "alphazeroforhnefatafl_b200", "csrc")


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    deps = [os.path.join(ORACLE_DIR, f) for f in ("oracle_capi.cpp", "tafl_oracle.hpp", "mcts_oracle.hpp")]
    deps.append(os.path.join(ROOT, "include", "tafl.h"))
    if _newer(so, deps) or _newer(os.path.join(ORACLE_DIR, "oracle_kat"), deps + [os.path.join(ORACLE_DIR, "kat_main.cpp")]):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "all"])
    return so


SAN_FLAGS = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"]


def build_hostsim(sanitize=None):
    """sanitize=True (or TAFL_HOSTSIM_SAN=1): the same harness built with ASan + UBSan
    (libhostsim_san.so; the loading process needs libasan preloaded, see test_hostsim.py)."""
    if sanitize is None:
        sanitize = os.environ.get("TAFL_HOSTSIM_SAN") == "1"
    so = os.path.join(HOSTSIM_DIR, "libhostsim_san.so" if sanitize else "libhostsim.so")
    deps = [os.path.join(HOSTSIM_DIR, "hostsim.cpp"), os.path.join(ROOT, "include", "tafl.h")]
    deps += [os.path.join(CSRC, f) for f in ("tafl_bb.cuh", "tafl_core.cuh", "tafl_select.cuh", "tafl_ctx.hpp", "tafl_wide.cuh")]
    if _newer(so, deps):
        opt = SAN_FLAGS if sanitize else ["-O2"]
        subprocess.check_call(["g++"] + opt + ["-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-Wno-unknown-pragmas", "-o", so, os.path.join(HOSTSIM_DIR, "hostsim.cpp")])
    return so


def sanitizer_runtime():
    """path of g++'s libasan.so (to LD_PRELOAD into a python that loads libhostsim_san.so), or None"""
    try:
        p = subprocess.check_output(["g++", "-print-file-name=libasan.so"], text=True).strip()
    except Exception:  # noqa: BLE001
        return None
    return p if os.path.isabs(p) and os.path.exists(p) else None


class ByValueImpl:
    """The functional (by-value) batch API, identical for oracle (orc_) and hostsim (sim_)."""

    def __init__(self, lib, prefix, rules, width):
        self.lib, self.prefix, self.rules, self.width = lib, prefix, rules, width

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def validate_plays(self, states, plays):
        n = len(states)
        out = np.zeros(n, np.uint8)
        rc = self._f("validate_plays")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(plays), n, A.ptr(out))
        assert rc == 0, rc
        return out

    def do_plays(self, states, plays):
        n = len(states)
        out = np.zeros(n, A.STATE_DTYPE)
        fx = np.zeros(n, A.EFFECTS_DTYPE)
        rc = self._f("do_plays")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(plays), n, A.ptr(out), A.ptr(fx))
        assert rc == 0, rc
        return out, fx

    def legal_moves_of(self, states):
        n = len(states)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        counts = np.zeros(n, np.uint32)
        rc = self._f("legal_moves_of")(C.byref(self.rules), self.width, A.ptr(states), n, A.ptr(moves), A.ptr(counts))
        assert rc == 0, rc
        return moves, counts

    def piece_plays(self, states, tiles_rc):
        n = len(states)
        tiles_rc = np.ascontiguousarray(tiles_rc, np.uint8)
        moves = np.zeros((n, 32), np.uint16)
        counts = np.zeros(n, np.int32)
        rc = self._f("piece_plays")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(tiles_rc), n, A.ptr(moves),
                                    A.ptr(counts))
        assert rc == 0, rc
        return moves, counts

    def get_captures(self, states, plays):
        n = len(states)
        fx = np.zeros(n, A.EFFECTS_DTYPE)
        rc = self._f("get_captures")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(plays), n, A.ptr(fx))
        assert rc == 0, rc
        return fx

    def detect_shieldwall(self, states, plays):
        n = len(states)
        fx = np.zeros(n, A.EFFECTS_DTYPE)
        rc = self._f("detect_shieldwall")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(plays), n, A.ptr(fx))
        assert rc == 0, rc
        return fx

    def find_enclosure(self, states, queries):
        n = len(states)
        out = np.zeros(n, A.ENCLOSURE_DTYPE)
        rc = self._f("find_enclosure")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(queries), n, A.ptr(out))
        assert rc == 0, rc
        return out

    def enclosure_secure(self, states, queries):
        n = len(states)
        out = np.zeros(n, np.uint8)
        rc = self._f("enclosure_secure")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(queries), n, A.ptr(out))
        assert rc == 0, rc
        return out

    def detect_exit_fort(self, states):
        n = len(states)
        out = np.zeros(n, np.uint8)
        rc = self._f("detect_exit_fort")(C.byref(self.rules), self.width, A.ptr(states), n, A.ptr(out))
        assert rc == 0, rc
        return out

    def side_can_play(self, states, sides):
        n = len(states)
        sides = np.ascontiguousarray(sides, np.uint8)
        out = np.zeros(n, np.uint8)
        rc = self._f("side_can_play")(C.byref(self.rules), self.width, A.ptr(states), A.ptr(sides), n, A.ptr(out))
        assert rc == 0, rc
        return out


class ByValueImplW:
    """The by-value API over the large bitfields (widths 256 / 512, tafl_state_w): oracle (orc_w_) and hostsim (sim_w_)."""

    def __init__(self, lib, prefix, rules, width):
        self.lib, self.prefix, self.rules, self.width = lib, prefix, rules, width

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _call(self, name, *args):
        rc = self._f(name)(C.byref(self.rules), self.width, *args)
        assert rc == 0, (name, rc)

    def validate_plays(self, states, plays):
        out = np.zeros(len(states), np.uint8)
        self._call("validate_plays", A.ptr(states), A.ptr(plays), len(states), A.ptr(out))
        return out

    def do_plays(self, states, plays):
        out, fx = np.zeros(len(states), A.STATE_W_DTYPE), np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._call("do_plays", A.ptr(states), A.ptr(plays), len(states), A.ptr(out), A.ptr(fx))
        return out, fx

    def legal_plays_of(self, states):
        plays, counts = np.zeros((len(states), A.W_MAX_PLAYS), A.PLAY_DTYPE), np.zeros(len(states), np.uint32)
        self._call("legal_plays_of", A.ptr(states), len(states), A.ptr(plays), A.ptr(counts))
        return plays, counts

    def piece_plays(self, states, tiles_rc):
        tiles_rc = np.ascontiguousarray(tiles_rc, np.uint8)
        plays, counts = np.zeros((len(states), 64), A.PLAY_DTYPE), np.zeros(len(states), np.int32)
        self._call("piece_plays", A.ptr(states), A.ptr(tiles_rc), len(states), A.ptr(plays), A.ptr(counts))
        return plays, counts

    def get_captures(self, states, plays):
        fx = np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._call("get_captures", A.ptr(states), A.ptr(plays), len(states), A.ptr(fx))
        return fx

    def detect_shieldwall(self, states, plays):
        fx = np.zeros(len(states), A.EFFECTS_W_DTYPE)
        self._call("detect_shieldwall", A.ptr(states), A.ptr(plays), len(states), A.ptr(fx))
        return fx

    def find_enclosure(self, states, queries):
        out = np.zeros(len(states), A.ENCLOSURE_W_DTYPE)
        self._call("find_enclosure", A.ptr(states), A.ptr(queries), len(states), A.ptr(out))
        return out

    def enclosure_secure(self, states, queries):
        out = np.zeros(len(states), np.uint8)
        self._call("enclosure_secure", A.ptr(states), A.ptr(queries), len(states), A.ptr(out))
        return out

    def detect_exit_fort(self, states):
        out = np.zeros(len(states), np.uint8)
        self._call("detect_exit_fort", A.ptr(states), len(states), A.ptr(out))
        return out

    def side_can_play(self, states, sides):
        sides = np.ascontiguousarray(sides, np.uint8)
        out = np.zeros(len(states), np.uint8)
        self._call("side_can_play", A.ptr(states), A.ptr(sides), len(states), A.ptr(out))
        return out


class Oracle:
    """ctypes face of oracle/liboracle.so."""

    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.orc_preset_board_fen.restype = C.c_char_p
        for name in ("orc_random_playout", "orc_mcts_search"):
            getattr(L, name).restype = C.c_int
        L.orc_random_playout.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_uint64, C.c_uint64, C.c_uint32,
                                         C.c_uint64, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_mcts_search.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_double, C.c_double,
                                      C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]

    def preset_rules(self, preset):
        r = A.Rules()
        assert self.lib.orc_preset_rules(preset, C.byref(r)) == 0
        return r

    def preset_board(self, preset):
        return self.lib.orc_preset_board_fen(preset).decode()

    def state_from_fen(self, fen, width, side=A.ATTACKER):
        if A.is_wide(width):
            st = np.zeros(1, A.STATE_W_DTYPE)
            rc = self.lib.orc_w_state_from_fen(fen.encode(), width, 1 if side == A.DEFENDER else 0, A.ptr(st))
            return rc, st
        st = np.zeros(1, A.STATE_DTYPE)
        rc = self.lib.orc_state_from_fen(fen.encode(), width, 1 if side == A.DEFENDER else 0, A.ptr(st))
        return rc, st

    def state_from_display_str(self, s, width, side=A.ATTACKER):
        st = np.zeros(1, A.STATE_DTYPE)
        rc = self.lib.orc_state_from_display_str(s.encode(), width, 1 if side == A.DEFENDER else 0, A.ptr(st))
        return rc, st

    def state_to_fen(self, st, width):
        buf = C.create_string_buffer(512)
        assert self.lib.orc_state_to_fen(A.ptr(st), width, buf, C.c_size_t(512)) == 0
        return buf.value.decode()

    def board_op(self, st, width, op, row=0, col=0, a=0, b=0):
        out = (C.c_int * 2)()
        rc = self.lib.orc_board_op(A.ptr(st), width, op, row, col, a, b, out)
        return rc, (out[0], out[1])

    def iter_occupied(self, st, width, side):
        buf = np.zeros(2 * 128, np.uint8)
        n = self.lib.orc_iter_occupied(A.ptr(st), width, 1 if side == A.DEFENDER else 0, A.ptr(buf))
        return [(int(buf[2 * i]), int(buf[2 * i + 1])) for i in range(n)]

    def play_from_str(self, s):
        p = A.Play()
        rc = self.lib.orc_play_from_str(s.encode(), C.byref(p))
        return rc, p

    def play_to_str(self, p):
        buf = C.create_string_buffer(64)
        assert self.lib.orc_play_to_str(C.byref(p), buf, C.c_size_t(64)) == 0
        return buf.value.decode()

    def tile_from_str(self, s):
        r, c = C.c_uint8(), C.c_uint8()
        rc = self.lib.orc_tile_from_str(s.encode(), C.byref(r), C.byref(c))
        return rc, (r.value, c.value)

    def track_play(self, st, width, side, play, captures):
        assert self.lib.orc_track_play(A.ptr(st), width, 1 if side == A.DEFENDER else 0, C.byref(play), int(captures)) == 0

    def by_value(self, rules, width):
        if A.is_wide(width):
            return ByValueImplW(self.lib, "orc_w_", rules, width)
        return ByValueImpl(self.lib, "orc_", rules, width)

    def random_playout(self, rules, fen, width, seed, n_games, n_steps, auto_reset=True, game_id_base=0, ply0=0,
                       n_threads=1, states=None):
        if states is None:
            rc, st0 = self.state_from_fen(fen, width, A.DEFENDER if rules.starting_side else A.ATTACKER)
            assert rc == 0
            states = np.repeat(st0, n_games)
        states = np.ascontiguousarray(states)
        stats = A.Stats()
        rc = self.lib.orc_random_playout(C.addressof(rules), fen.encode(), width, seed, game_id_base, n_games, ply0,
                                         n_steps, int(auto_reset), n_threads, states.ctypes.data, C.addressof(stats))
        assert rc == 0, rc
        return states, stats

    def mcts_search(self, rules, width, roots, num_sims, cpuct=1.0, draw_value=1e-4, n_threads=1):
        roots = np.ascontiguousarray(roots)
        n = len(roots)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        visits = np.zeros((n, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, A.MAX_MOVES), np.float64)
        counts = np.zeros(n, np.uint32)
        best = np.zeros(n, np.uint32)
        stats = A.MctsStats()
        rc = self.lib.orc_mcts_search(C.addressof(rules), width, roots.ctypes.data, n, num_sims, cpuct, draw_value,
                                      n_threads, moves.ctypes.data, visits.ctypes.data, q.ctypes.data,
                                      counts.ctypes.data, best.ctypes.data, C.addressof(stats))
        assert rc == 0, rc
        return dict(moves=moves, visits=visits, q=q, counts=counts, best=best, stats=stats)

    def mcts_search_eval(self, rules, width, roots, num_sims, eval_seed, cpuct=1.0, draw_value=1e-4, n_threads=1):
        """mcts_search with the TAFL_EVAL_HASH test evaluator instead of the stub."""
        roots = np.ascontiguousarray(roots)
        n = len(roots)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        visits = np.zeros((n, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, A.MAX_MOVES), np.float64)
        counts = np.zeros(n, np.uint32)
        best = np.zeros(n, np.uint32)
        stats = A.MctsStats()
        f = self.lib.orc_mcts_search_eval
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_uint64,
                      C.c_int] + [C.c_void_p] * 6
        f.restype = C.c_int
        rc = f(C.addressof(rules), width, roots.ctypes.data, n, num_sims, cpuct, draw_value, eval_seed, n_threads,
               moves.ctypes.data, visits.ctypes.data, q.ctypes.data, counts.ctypes.data, best.ctypes.data,
               C.addressof(stats))
        assert rc == 0, rc
        return dict(moves=moves, visits=visits, q=q, counts=counts, best=best, stats=stats)

    def mcts_episode(self, rules, width, roots, num_sims, n_moves, reuse, cpuct=1.0, draw_value=1e-4, eval_seed=None,
                     n_threads=1, noise=None, eps=0.0):
        """n_moves self-play moves per root with one MCTS object per game (reuse) or a fresh tree per move;
        noise [n][n_moves][MAX_MOVES] f64 = root exploration noise mixed in with weight eps."""
        roots = np.ascontiguousarray(roots)
        n = len(roots)
        visits = np.zeros((n, n_moves, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, n_moves, A.MAX_MOVES), np.float64)
        counts = np.zeros((n, n_moves), np.uint32)
        best = np.zeros((n, n_moves), np.uint32)
        final = np.zeros(n, A.STATE_DTYPE)
        stats = A.MctsStats()
        if noise is not None:
            noise = np.ascontiguousarray(noise, np.float64)
            assert noise.shape == (n, n_moves, A.MAX_MOVES)
        f = self.lib.orc_mcts_episode_noise
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_double, C.c_double,
                      C.c_int, C.c_uint64, C.c_int] + [C.c_void_p] * 7 + [C.c_double]
        f.restype = C.c_int
        rc = f(C.addressof(rules), width, roots.ctypes.data, n, num_sims, n_moves, int(reuse), cpuct, draw_value,
               int(eval_seed is not None), int(eval_seed or 0), n_threads, visits.ctypes.data, q.ctypes.data,
               counts.ctypes.data, best.ctypes.data, final.ctypes.data, C.addressof(stats),
               None if noise is None else noise.ctypes.data, float(eps))
        assert rc == 0, rc
        return dict(visits=visits, q=q, counts=counts, best=best, final=final, stats=stats)

    def mcts_episode_vl(self, rules, width, roots, num_sims, n_moves, reuse, vl_width, cpuct=1.0, draw_value=1e-4, n_threads=1):
        """the stub-evaluator episode with vl_width simulations in flight per tree (virtual loss)"""
        roots = np.ascontiguousarray(roots)
        n = len(roots)
        visits = np.zeros((n, n_moves, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, n_moves, A.MAX_MOVES), np.float64)
        counts = np.zeros((n, n_moves), np.uint32)
        best = np.zeros((n, n_moves), np.uint32)
        final = np.zeros(n, A.STATE_DTYPE)
        stats = A.MctsStats()
        f = self.lib.orc_mcts_episode_vl
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_double, C.c_double,
                      C.c_uint32, C.c_int] + [C.c_void_p] * 6
        f.restype = C.c_int
        rc = f(C.addressof(rules), width, roots.ctypes.data, n, num_sims, n_moves, int(reuse), cpuct, draw_value, int(vl_width),
               n_threads, visits.ctypes.data, q.ctypes.data, counts.ctypes.data, best.ctypes.data, final.ctypes.data,
               C.addressof(stats))
        assert rc == 0, rc
        return dict(visits=visits, q=q, counts=counts, best=best, final=final, stats=stats)


    def mcts_episode_vl_eval(self, rules, width, roots, num_sims, n_moves, reuse, vl_width, eval_seed=None, cpuct=1.0,
                             draw_value=1e-4, n_threads=1, noise=None, eps=0.0):
        """virtual loss with an evaluator: eval_seed None = stub, else the TAFL_EVAL_HASH test evaluator (+ optional root noise)"""
        roots = np.ascontiguousarray(roots)
        n = len(roots)
        visits = np.zeros((n, n_moves, A.MAX_MOVES), np.uint32)
        q = np.zeros((n, n_moves, A.MAX_MOVES), np.float64)
        counts = np.zeros((n, n_moves), np.uint32)
        best = np.zeros((n, n_moves), np.uint32)
        final = np.zeros(n, A.STATE_DTYPE)
        stats = A.MctsStats()
        if noise is not None:
            noise = np.ascontiguousarray(noise, np.float64)
            assert noise.shape == (n, n_moves, A.MAX_MOVES)
        f = self.lib.orc_mcts_episode_vl_eval
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_double, C.c_double,
                      C.c_uint32, C.c_int, C.c_uint64, C.c_int] + [C.c_void_p] * 7 + [C.c_double]
        f.restype = C.c_int
        rc = f(C.addressof(rules), width, roots.ctypes.data, n, num_sims, n_moves, int(reuse), cpuct, draw_value, int(vl_width),
               0 if eval_seed is None else 1, 0 if eval_seed is None else int(eval_seed), n_threads, visits.ctypes.data,
               q.ctypes.data, counts.ctypes.data, best.ctypes.data, final.ctypes.data, C.addressof(stats),
               None if noise is None else noise.ctypes.data, float(eps))
        assert rc == 0, rc
        return dict(visits=visits, q=q, counts=counts, best=best, final=final, stats=stats)
    def hash_eval(self, states, seed):
        states = np.ascontiguousarray(states)
        n = len(states)
        P = np.zeros((n, A.MAX_MOVES), np.float32)
        v = np.zeros(n, np.float32)
        f = self.lib.orc_hash_eval
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p]
        f.restype = C.c_int
        assert f(states.ctypes.data, n, seed, P.ctypes.data, v.ctypes.data) == 0
        return P, v

    def np_sum(self, a):
        a = np.ascontiguousarray(a, np.float64)
        f = self.lib.orc_np_sum
        f.argtypes = [C.c_void_p, C.c_uint64]
        f.restype = C.c_double
        return f(a.ctypes.data, len(a))

    def hw_threads(self):
        return self.lib.orc_hw_threads()


class HostSim:
    """ctypes face of tests/hostsim/libhostsim.so (the kernels' math compiled for the CPU)."""

    def __init__(self):
        self.lib = C.CDLL(build_hostsim())
        self.lib.sim_random_playout.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_uint32,
                                                C.c_uint64, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]

    def by_value(self, rules, width):
        if A.is_wide(width):
            return ByValueImplW(self.lib, "sim_w_", rules, width)
        return ByValueImpl(self.lib, "sim_", rules, width)

    def force_generic(self, on):
        """1: always use the generic (run-time rule flags) code; 0: preset specialisations when they match."""
        self.lib.sim_force_generic(int(on))

    def spec_used(self, spec):
        return self.lib.sim_spec_used(int(spec))

    def random_playout(self, rules, start_state, width, seed, n_games, n_steps, auto_reset=True, game_id_base=0, ply0=0,
                       states=None):
        if states is None:
            states = np.repeat(start_state, n_games)
        states = np.ascontiguousarray(states)
        start_state = np.ascontiguousarray(start_state)
        stats = A.Stats()
        rc = self.lib.sim_random_playout(C.addressof(rules), start_state.ctypes.data, width, seed, game_id_base, n_games,
                                         ply0, n_steps, int(auto_reset), states.ctypes.data, C.addressof(stats))
        assert rc == 0, rc
        return states, stats

    def piece_table_moves(self, rules, width, states):
        n = len(states)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        counts = np.zeros(n, np.uint32)
        rc = self.lib.sim_piece_table_moves(C.byref(rules), width, A.ptr(states), n, A.ptr(moves), A.ptr(counts))
        assert rc == 0, rc
        return moves, counts

    def emit_moves(self, rules, width, states):
        n = len(states)
        moves = np.zeros((n, A.MAX_MOVES), np.uint16)
        counts = np.zeros(n, np.uint32)
        rc = self.lib.sim_emit_moves(C.byref(rules), width, A.ptr(states), n, A.ptr(moves), A.ptr(counts))
        assert rc == 0, rc
        return moves, counts

    def select_moves(self, rules, width, states, ks):
        n = len(states)
        ks = np.ascontiguousarray(ks, np.uint32)
        moves = np.zeros(n, np.uint16)
        counts = np.zeros(n, np.uint32)
        rc = self.lib.sim_select_moves(C.byref(rules), width, A.ptr(states), A.ptr(ks), n, A.ptr(moves), A.ptr(counts))
        assert rc == 0, rc
        return moves, counts


_ORACLE = None
_HOSTSIM = None


def oracle():
    global _ORACLE
    if _ORACLE is None:
        _ORACLE = Oracle()
    return _ORACLE


def hostsim():
    global _HOSTSIM
    if _HOSTSIM is None:
        _HOSTSIM = HostSim()
    return _HOSTSIM


class OracleHost:
    """Text/board helpers for tests/kat_runner.py, served by the oracle."""

    def __init__(self):
        self.o = oracle()

    def preset_rules(self, pid):
        return self.o.preset_rules(pid)

    def preset_board(self, pid):
        return self.o.preset_board(pid)

    def state_from_fen(self, fen, width, side):
        return self.o.state_from_fen(fen, width, side)

    def move_piece(self, st, width, frm, to):
        if A.is_wide(width):
            rc = self.o.lib.orc_w_move_piece(A.ptr(st), width, frm[0], frm[1], to[0], to[1])
        else:
            rc, _ = self.o.board_op(st, width, 6, frm[0], frm[1], to[0], to[1])
        assert rc == 0, rc

    def play_from_str(self, s):
        return self.o.play_from_str(s)
