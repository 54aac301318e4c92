"""CPU tests of the product library's C ABI: it loads, exports every symbol include/tafl.h
declares, its host-only helpers (FEN / notation / board accessors — no device needed) reproduce
the reference's board/state.rs, tiles.rs and game/state.rs unit tests, and it refuses to run
without a CUDA device (no CPU fallback)."""
import ctypes as C
import subprocess

import numpy as np
import pytest
import torch

import kat_runner
from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import _lib, engine as E

HOST = kat_runner.load()["host"]


@pytest.fixture(scope="module")
def L(built_lib):
    return _lib.lib()


def st_from(L, fen, width, side=A.ATTACKER):
    if fen.startswith("@"):
        fen = E.preset_board(kat_runner.PRESETS[fen[1:]])
    st = np.zeros(1, A.STATE_DTYPE)
    assert L.tafl_state_from_fen(fen.encode(), width, side, A.ptr(st)) == 0
    return st


def test_exports_every_declared_symbol(built_lib, L):
    declared = _lib.declared_symbols()
    assert len(declared) >= 60
    out = subprocess.run(["nm", "-D", "--defined-only", built_lib], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    missing = [s for s in declared if s not in exported]
    assert not missing, missing
    assert set(L._tafl_signatures) == set(declared)   # the ctypes binding covers the whole header


def test_pod_sizes_match_header(L):
    assert C.sizeof(A.State) == 64 and A.STATE_DTYPE.itemsize == 64
    assert C.sizeof(A.Rules) == 32 and C.sizeof(A.Play) == 4 and C.sizeof(A.PlayEffects) == 24


def test_presets_match_oracle(L, orc):
    for pid in range(5):
        assert bytes(E.preset_rules(pid)) == bytes(orc.preset_rules(pid))
    for pid in range(4):
        assert E.preset_board(pid) == orc.preset_board(pid)
    r = E.preset.rules.COPENHAGEN
    assert r.shieldwall_enabled and r.exit_fort and r.king_strength == A.KING_STRONG
    assert E.preset.boards.BRANDUBH == "3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3"


def test_from_str(L):
    t = HOST["test_from_str"]
    for width in (64, 128):
        a = st_from(L, t["fen"], width)
        b = np.zeros(1, A.STATE_DTYPE)
        assert L.tafl_state_from_display_str(t["display"].encode(), width, A.ATTACKER, A.ptr(b)) == 0
        assert a.tobytes() == b.tobytes()
        buf = C.create_string_buffer(512)
        assert L.tafl_state_to_display_str(A.ptr(a), width, buf, 512) == 0
        assert buf.value.decode().strip("\n") == t["display"]


def test_piece_movement(L):
    t = HOST["test_piece_movement"]
    for width in (64, 128):
        st = st_from(L, t["fen"], width)
        r, c = C.c_uint8(), C.c_uint8()
        assert L.tafl_board_get_king(A.ptr(st), width, C.byref(r), C.byref(c)) == 0
        assert [r.value, c.value] == t["king"]
        for row, col, typ, side in t["set"]:
            assert L.tafl_board_set_piece(A.ptr(st), width, row, col, typ, side) == 0
        (fr, fc), (tr, tc) = t["move"]
        assert L.tafl_board_move_piece(A.ptr(st), width, fr, fc, tr, tc) == 0
        assert L.tafl_board_get_king(A.ptr(st), width, C.byref(r), C.byref(c)) == 0
        assert [r.value, c.value] == t["king_after"]
        assert L.tafl_board_is_king(A.ptr(st), width, tr, tc) == 1
        assert E.state_to_fen(st, width) == t["fen_after"]
        for row, col in t["occupied"]:
            assert L.tafl_board_tile_occupied(A.ptr(st), width, row, col) == 1
        for row, col in t["empty"]:
            assert L.tafl_board_tile_occupied(A.ptr(st), width, row, col) == 0
        # move_piece on an empty tile: the reference panics "No piece to move." (board/state.rs:219)
        assert L.tafl_board_move_piece(A.ptr(st), width, 1, 1, 1, 2) == A.ERR_NO_PIECE


def test_iter_occupied(L):
    t = HOST["test_iter_occupied"]
    for width in (64, 128):
        st = st_from(L, t["fen"], width)
        for side, key in ((A.ATTACKER, "attackers"), (A.DEFENDER, "defenders")):
            buf = np.zeros(256, np.uint8)
            n = L.tafl_board_iter_occupied(A.ptr(st), width, side, A.ptr(buf))
            got = [[int(buf[2 * i]), int(buf[2 * i + 1])] for i in range(n)]
            assert sorted(got) == sorted(t[key])
            assert got == sorted(got)   # ascending bit order == row-major (board/state.rs:94-101)


def test_swap_pieces(L):
    t = HOST["test_swap_pieces"]
    st = st_from(L, t["fen"], 64)
    (ar, ac), (br, bc) = t["a"], t["b"]
    typ, side = C.c_int(), C.c_int()
    L.tafl_board_get_piece(A.ptr(st), 64, ar, ac, C.byref(typ), C.byref(side))
    assert (typ.value, side.value) == (A.KING, A.DEFENDER)
    L.tafl_board_get_piece(A.ptr(st), 64, br, bc, C.byref(typ), C.byref(side))
    assert (typ.value, side.value) == (A.SOLDIER, A.ATTACKER)
    assert L.tafl_board_swap_pieces(A.ptr(st), 64, ar, ac, br, bc) == 0
    L.tafl_board_get_piece(A.ptr(st), 64, br, bc, C.byref(typ), C.byref(side))
    assert (typ.value, side.value) == (A.KING, A.DEFENDER)
    L.tafl_board_get_piece(A.ptr(st), 64, ar, ac, C.byref(typ), C.byref(side))
    assert (typ.value, side.value) == (A.SOLDIER, A.ATTACKER)
    r, c = C.c_uint8(), C.c_uint8()
    L.tafl_board_get_king(A.ptr(st), 64, C.byref(r), C.byref(c))
    assert [r.value, c.value] == t["b"]


def test_count_pieces(L):
    t = HOST["test_count_pieces"]
    st = st_from(L, t["fen"], 128)
    assert L.tafl_board_count_pieces(A.ptr(st), 128, A.ATTACKER) == t["attackers"]
    assert L.tafl_board_count_pieces(A.ptr(st), 128, A.DEFENDER) == t["defenders"]


def test_moves(L):
    t = HOST["test_moves"]
    for m in t["ok"]:
        p = A.Play()
        assert L.tafl_play_from_tiles(*m["frm"], *m["to"], C.byref(p)) == 0
        assert [p.from_row, p.from_col] == m["frm"]
        assert p.axis == (A.HORIZONTAL if m["axis"] == "H" else A.VERTICAL) and p.disp == m["disp"]
        assert list(p.to()) == m["to"]
    for m in t["err"]:
        assert L.tafl_play_from_tiles(*m["frm"], *m["to"], C.byref(A.Play())) == A.ERR_PARSE_BAD_PLAY


ERR = dict(EmptyString=A.ERR_PARSE_EMPTY, BadChar=A.ERR_PARSE_BAD_CHAR, BadInt=A.ERR_PARSE_BAD_INT,
           BadPlay=A.ERR_PARSE_BAD_PLAY, BadString=A.ERR_PARSE_BAD_STRING)


def test_parsing_tiles(L):
    t = HOST["test_parsing_tiles"]
    for s, rc_ in t["ok"]:
        r, c = C.c_uint8(), C.c_uint8()
        assert L.tafl_tile_from_str(s.encode(), C.byref(r), C.byref(c)) == 0
        assert [r.value, c.value] == rc_
        buf = C.create_string_buffer(16)
        assert L.tafl_tile_to_str(rc_[0], rc_[1], buf, 16) == 0 and buf.value.decode() == s
    for s, err in t["err"]:
        assert L.tafl_tile_from_str(s.encode(), C.byref(C.c_uint8()), C.byref(C.c_uint8())) == ERR[err], s


def test_parsing_moves(L):
    t = HOST["test_parsing_moves"]
    for s, frm, to in t["ok"]:
        p = E.play_from_str(s)
        assert [p.from_row, p.from_col] == frm and list(p.to()) == to
        assert E.play_to_str(p) == s
    for s, err in t["err"]:
        assert L.tafl_play_from_str(s.encode(), C.byref(A.Play())) == ERR[err], s


def test_repetition_tracker(L):
    t = HOST["test_repetition_tracker"]
    st = np.zeros(1, A.STATE_DTYPE)
    for block in t["blocks"]:
        for i in range(t["rounds"]):
            for j, s in enumerate(block):
                side = A.DEFENDER if j % 2 else A.ATTACKER
                p = E.play_from_str(s)
                assert L.tafl_track_play(A.ptr(st), side, C.byref(p), 0) == 0
                reps = int(st["def_reps"][0] if j % 2 else st["att_reps"][0])
                assert reps == i, (block, i, j)


def test_fen_errors(L):
    st = np.zeros(1, A.STATE_DTYPE)
    assert L.tafl_state_from_fen(b"3t3/3t2", 64, A.ATTACKER, A.ptr(st)) == A.ERR_PARSE_BAD_LINE_LEN
    assert L.tafl_state_from_fen(b"3x3", 64, A.ATTACKER, A.ptr(st)) == A.ERR_PARSE_BAD_CHAR
    assert L.tafl_state_from_fen(b"11/11", 128, A.ATTACKER, A.ptr(st)) == 0 and int(st["side_len"][0]) == 11


def test_board_to_matrix(L):
    """board_to_matrix (game/main.rs:55-83): corners 20, throne 30, +1 soldier, +5 king."""
    st = st_from(L, "@BRANDUBH", 64)
    out = np.zeros(49, np.uint8)
    assert L.tafl_board_to_matrix(A.ptr(st), 64, A.ptr(out), 49) == 0
    m = out.reshape(7, 7)
    assert m[0, 0] == 20 and m[6, 6] == 20 and m[3, 3] == 35 and m[0, 3] == 1 and m[2, 3] == 1 and m[1, 1] == 0


def test_move_to_play_roundtrip(L):
    for width in (64, 128):
        s = A.stride_for(width)
        for (fr, fc, tr, tc) in ((0, 3, 0, 1), (5, 2, 1, 2), (6, 6, 6, 0)):
            m = (fr * s + fc) | ((tr * s + tc) << 7)
            p = A.Play()
            assert L.tafl_move_to_play(m, width, C.byref(p)) == 0
            assert (p.from_row, p.from_col) == (fr, fc) and p.to() == (tr, tc)
            arr = E.moves_to_plays(np.array([m], np.uint16), width)
            assert (int(arr["from_row"][0]), int(arr["from_col"][0]), int(arr["axis"][0]), int(arr["disp"][0])) == \
                (p.from_row, p.from_col, p.axis, p.disp)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(L):
    """Without a CUDA device the engine cannot be created: the product has no CPU path."""
    with pytest.raises(_lib.TaflError) as ei:
        E.LockstepGames(E.preset.rules.BRANDUBH, E.preset.boards.BRANDUBH, 4)
    assert ei.value.code == A.ERR_CUDA
    assert "no CPU fallback" in str(ei.value)


def test_large_bitfield_host_helpers_match_oracle(L, orc):
    """tafl_w_state_from_fen / tafl_w_state_to_fen / tafl_w_board_move_piece (BitfieldBoardState<U256> / <U512>, no device
    needed) against the oracle's restatement of board/state.rs at those widths; the creation call refuses without a GPU."""
    from parity_cases import wide_random_boards
    rng = np.random.default_rng(5)
    for width, N in ((256, 15), (256, 9), (512, 21), (512, 17)):
        for st in wide_random_boards(orc, rng, N, width, 25):
            st = st.reshape(1)
            fen = E.WideLogic.state_to_fen(st, width)
            side = A.DEFENDER if st["flags"][0] & A.FLAG_DEF_TO_PLAY else A.ATTACKER
            mine = E.WideLogic.state_from_fen(fen, width, side)
            rc, theirs = orc.state_from_fen(fen, width, side)
            assert rc == 0 and mine.tobytes() == theirs.tobytes()
            # the king's row / column live in four bits each (board/state.rs:136-147): a king beyond row or column 15 of a
            # 21x21 board reads back as a soldier, in the reference as here, so only smaller boards round-trip exactly
            if N <= 16:
                assert mine.tobytes() == st.tobytes()
            plays, counts = orc.by_value(orc.preset_rules(A.PRESET_COPENHAGEN), width).legal_plays_of(st)
            if counts[0]:
                p = plays[0, rng.integers(counts[0])]
                frm = (int(p["from_row"]), int(p["from_col"]))
                to = (frm[0], frm[1] + int(p["disp"])) if p["axis"] == A.HORIZONTAL else (frm[0] + int(p["disp"]), frm[1])
                a, b = st.copy(), st.copy()
                E.WideLogic.move_piece(a, width, frm, to)
                assert orc.lib.orc_w_move_piece(A.ptr(b), width, frm[0], frm[1], to[0], to[1]) == 0
                assert a.tobytes() == b.tobytes()
    bad = np.zeros(1, A.STATE_W_DTYPE)
    assert L.tafl_w_state_from_fen(b"3t3/8", 256, A.ATTACKER, A.ptr(bad)) == A.ERR_PARSE_BAD_LINE_LEN
    assert L.tafl_w_state_from_fen(b"3t3", 128, A.ATTACKER, A.ptr(bad)) == A.ERR_INVALID_ARG
    if not torch.cuda.is_available():
        with pytest.raises(E.TaflError):
            E.WideLogic(orc.preset_rules(A.PRESET_COPENHAGEN), 15, 256)
