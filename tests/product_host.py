"""Adapters that present the PRODUCT library (libtafl_b200.so, C ABI) to tests/kat_runner.py."""
import ctypes as C

import numpy as np

from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import engine as E
from alphazeroforhnefatafl_b200._lib import lib


class ProductHost:
    """Host-only helpers of the C ABI (no device needed)."""

    def preset_rules(self, pid):
        return E.preset_rules(pid)

    def preset_board(self, pid):
        return E.preset_board(pid)

    def state_from_fen(self, fen, width, side):
        if A.is_wide(width):
            st = np.zeros(1, A.STATE_W_DTYPE)
            rc = lib().tafl_w_state_from_fen(fen.encode(), width, side, A.ptr(st))
            return rc, st
        st = np.zeros(1, A.STATE_DTYPE)
        rc = lib().tafl_state_from_fen(fen.encode(), width, side, A.ptr(st))
        return rc, st

    def move_piece(self, st, width, frm, to):
        f = lib().tafl_w_board_move_piece if A.is_wide(width) else lib().tafl_board_move_piece
        rc = f(A.ptr(st), width, frm[0], frm[1], to[0], to[1])
        assert rc == 0, rc

    def play_from_str(self, s):
        p = A.Play()
        rc = lib().tafl_play_from_str(s.encode(), C.byref(p))
        return rc, p


class GpuByValue:
    """ByValueImpl-compatible face of a LockstepGames engine (the CUDA path)."""

    def __init__(self, eng):
        self.e = eng

    def validate_plays(self, st, pl):
        return self.e.validate_play(st, pl)

    def do_plays(self, st, pl):
        return self.e.do_play(st, pl)

    def legal_moves_of(self, st):
        return self.e.legal_moves_of(st)

    def piece_plays(self, st, tiles):
        return self.e.iter_plays(st, tiles)

    def get_captures(self, st, pl):
        return self.e.get_captures(st, pl)

    def detect_shieldwall(self, st, pl):
        return self.e.detect_shieldwall(st, pl)

    def find_enclosure(self, st, q):
        return self.e.find_enclosure(st, q)

    def enclosure_secure(self, st, q):
        return self.e.enclosure_secure(st, q)

    def detect_exit_fort(self, st):
        return self.e.detect_exit_fort(st)

    def side_can_play(self, st, sides):
        return self.e.side_can_play(st, sides)


class GpuImplFactory:
    """(rules, width, side_len) -> GpuByValue, engines cached per configuration."""

    def __init__(self, device=0):
        self.device, self.cache = device, {}

    def __call__(self, rules, width, side_len):
        key = (bytes(rules), width, side_len)
        if key not in self.cache:
            if A.is_wide(width):   # large bitfields: the by-value WideLogic is the implementation itself
                self.cache[key] = E.WideLogic(rules.copy(), side_len, width, self.device)
            else:
                fen = "/".join([str(side_len)] * side_len)
                self.cache[key] = GpuByValue(E.LockstepGames(rules.copy(), fen, 1, self.device, width))
        return self.cache[key]

    def close(self):
        for v in self.cache.values():
            (v if isinstance(v, E.WideLogic) else v.e).close()
        self.cache.clear()
