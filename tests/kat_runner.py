"""Executes tests/golden/reference_kats.json (the reference's own unit tests) against any
implementation of the by-value GameLogic API.

An implementation is (host, make_impl):
  host       text/board helpers: state_from_fen(fen, width, side) -> (rc, st), move_piece(st, width,
             frm, to), play_from_str(s) -> (rc, Play), preset_rules(id), preset_board(id)
  make_impl  (rules, width, side_len) -> object with validate_plays / do_plays / piece_plays /
             get_captures / detect_shieldwall / find_enclosure / enclosure_secure /
             detect_exit_fort / side_can_play  (batch methods over numpy arrays)
"""
import json
import os

import numpy as np

from alphazeroforhnefatafl_b200 import _abi as A

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_kats.json")

PRESETS = dict(COPENHAGEN=A.PRESET_COPENHAGEN, BRANDUBH=A.PRESET_BRANDUBH, MAGPIE=A.PRESET_MAGPIE,
               TABLUT=A.PRESET_TABLUT, KOCH=A.PRESET_KOCH)
PLAY_INVALID = dict(Ok=A.PLAY_OK, WrongPlayer=A.WRONG_PLAYER, NoPiece=A.NO_PIECE, OutOfBounds=A.OUT_OF_BOUNDS,
                    NoCommonAxis=A.NO_COMMON_AXIS, BlockedByPiece=A.BLOCKED_BY_PIECE,
                    MoveThroughBlockedTile=A.MOVE_THROUGH_BLOCKED_TILE, MoveOntoBlockedTile=A.MOVE_ONTO_BLOCKED_TILE,
                    TooFar=A.TOO_FAR, GameOver=A.GAME_OVER)
WIN_REASON = dict(KingEscaped=0, ExitFort=1, KingCaptured=2, AllCaptured=3, Enclosed=4, NoPlays=5, Repetition=6)
DRAW_REASON = dict(Repetition=0, NoPlays=1)


def load():
    with open(GOLDEN) as f:
        return json.load(f)


def status_code(sym):
    if sym == "Ongoing":
        return A.STATUS_ONGOING
    if "win" in sym:
        return A.status_win(WIN_REASON[sym["win"][0]], sym["win"][1] == "Defender")
    return A.status_draw(DRAW_REASON[sym["draw"]])


def side_code(s):
    return A.DEFENDER if s == "D" else A.ATTACKER


def make_play(frm, to):
    """Play::from_tiles (play.rs:36-49) without the common-axis check (tests only pass aligned tiles)."""
    p = np.zeros(1, A.PLAY_DTYPE)
    p["from_row"], p["from_col"] = frm
    if frm[0] == to[0]:
        p["axis"], p["disp"] = A.HORIZONTAL, to[1] - frm[1]
    else:
        assert frm[1] == to[1]
        p["axis"], p["disp"] = A.VERTICAL, to[0] - frm[0]
    return p


def tiles_of(width, words):
    return sorted(map(list, A.bb_to_tiles(width, words)))


def run_case(case, width, host, make_impl):
    rules = host.preset_rules(PRESETS[case["rules"]])
    for k, v in case["override"].items():
        setattr(rules, k, v)
    impl = make_impl(rules, width, case["side_len"])
    st = None
    n_checks = 0
    where = lambda i: "%s [%s] width=%d step %d" % (case["name"], case["ref"], width, i)
    for i, s in enumerate(case["steps"]):
        op = s["op"]
        if op == "new":
            fen = s["fen"]
            if fen.startswith("@"):
                fen = host.preset_board(PRESETS[fen[1:]])
            rc, st = host.state_from_fen(fen, width, side_code(s["side"]))
            assert rc == 0, where(i)
            assert int(st["side_len"][0]) == case["side_len"], where(i)
        elif op == "set_side":
            if s["side"] == "D":
                st["flags"] |= A.FLAG_DEF_TO_PLAY
            else:
                st["flags"] &= ~np.uint8(A.FLAG_DEF_TO_PLAY)
        elif op == "move_piece":
            host.move_piece(st, width, s["frm"], s["to"])
        elif op == "validate":
            err = impl.validate_plays(st, make_play(s["frm"], s["to"]))
            assert int(err[0]) == PLAY_INVALID[s["expect"]], where(i) + " got %d" % err[0]
            n_checks += 1
        elif op in ("do_play", "do_play_str"):
            if op == "do_play_str":
                rc, p = host.play_from_str(s["play"])
                assert rc == 0, where(i)
                pl = np.zeros(1, A.PLAY_DTYPE)
                pl[0] = (p.from_row, p.from_col, p.axis, p.disp)
            else:
                pl = make_play(s["frm"], s["to"])
            new_st, fx = impl.do_plays(st, pl)
            assert int(fx["err"][0]) == 0, where(i) + " err %d" % fx["err"][0]
            if "status" in s:
                assert int(new_st["status"][0]) == status_code(s["status"]), where(i) + " status %#x" % new_st["status"][0]
                assert int(fx["status"][0]) == status_code(s["status"]), where(i)
                n_checks += 1
            if "captures" in s:
                assert tiles_of(width, fx["captures"][0]) == sorted(s["captures"]), where(i)
                assert int(fx["n_captures"][0]) == len(s["captures"]), where(i)
                n_checks += 1
            st = new_st
        elif op == "iter_plays":
            moves, counts = impl.piece_plays(st, np.array([s["tile"]], np.uint8))
            if s["expect"] == "err":
                assert int(counts[0]) == -1, where(i)
            else:
                n = int(counts[0])
                assert n >= 0, where(i)
                got = []
                for m in moves[0, :n]:
                    if A.is_wide(width):   # wide boards list plays, not compact moves
                        frm = (int(m["from_row"]), int(m["from_col"]))
                        to = (frm[0], frm[1] + int(m["disp"])) if m["axis"] == A.HORIZONTAL else (frm[0] + int(m["disp"]), frm[1])
                    else:
                        frm, to = A.move_from_to(width, m)
                    assert list(frm) == s["tile"], where(i)
                    got.append(list(to))
                assert len(set(map(tuple, got))) == n, where(i)
                assert sorted(got) == sorted(s["expect"]), where(i) + " got %r" % got
            n_checks += 1
        elif op == "get_captures":
            fx = impl.get_captures(st, make_play(s["frm"], s["to"]))
            assert tiles_of(width, fx["captures"][0]) == sorted(s["expect"]), where(i)
            n_checks += 1
        elif op == "detect_shieldwall":
            fx = impl.detect_shieldwall(st, make_play(s["frm"], s["to"]))
            if s["expect"] is None:
                assert int(fx["n_captures"][0]) == 0xFF, where(i)
            else:
                assert int(fx["n_captures"][0]) == len(s["expect"]), where(i)
                assert tiles_of(width, fx["captures"][0]) == sorted(s["expect"]), where(i)
            n_checks += 1
        elif op in ("find_enclosure", "enclosure_secure"):
            q = np.zeros(1, A.ENCL_QUERY_DTYPE)
            q["row"], q["col"] = s["tile"]
            q["abort_on_edge"], q["abort_on_corner"] = int(s["abort_edge"]), int(s["abort_corner"])
            q["enclosed_pieces"], q["enclosing_pieces"] = s["enclosed"], s["enclosing"]
            if op == "enclosure_secure":
                q["inside_safe"], q["outside_safe"] = int(s["inside_safe"]), int(s["outside_safe"])
                out = impl.enclosure_secure(st, q)
                assert int(out[0]) == (1 if s["expect"] else 0), where(i) + " got %d" % out[0]
            else:
                e = impl.find_enclosure(st, q)
                if s["expect"] is None:
                    assert int(e["found"][0]) == 0, where(i)
                else:
                    assert int(e["found"][0]) == 1, where(i)
                    if s["expect"] != "some":
                        for k in ("occupied", "unoccupied", "boundary"):
                            assert tiles_of(width, e[k][0]) == sorted(s["expect"][k]), where(i) + " " + k
            n_checks += 1
        elif op == "exit_fort":
            out = impl.detect_exit_fort(st)
            assert bool(out[0]) == s["expect"], where(i)
            n_checks += 1
        elif op == "can_play":
            out = impl.side_can_play(st, np.array([1 if s["side"] == "D" else 0], np.uint8))
            assert bool(out[0]) == s["expect"], where(i)
            n_checks += 1
        else:
            raise ValueError(op)
    return n_checks


def run_all(host, make_impl, widths=(64, 128)):
    """widths 256 / 512: every case also runs on the large bitfields (the reference instantiates its width-generic
    tests on all four, game/game/logic.rs:1015-1021,1081-1087; a 7x7 or 9x9 position sits in any wider row stride)"""
    data = load()
    total = 0
    for case in data["cases"]:
        for w in widths:
            if w in case["widths"] or (A.is_wide(w) and case["side_len"] <= (15 if w == 256 else 21)):
                total += run_case(case, w, host, make_impl)
    return total
