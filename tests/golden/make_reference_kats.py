#!/usr/bin/env python
"""Writes tests/golden/reference_kats.json: the reference crate's own unit tests for the hot path,
transcribed VERBATIM (FEN strings, tiles, rule overrides, expected values) from the Rust sources
under /root/reference/game into an implementation-neutral op list.

The Rust crate cannot be compiled in this image (no cargo/rustc), so these vectors are hand
transcriptions, not generated outputs; every case cites the reference test it carries over
(paths relative to the reference tree).  tests/kat_runner.py executes the op list against any
implementation of the by-value API: the CPU oracle, the host build of the kernel math, and the
CUDA library through the C ABI.

    python tests/golden/make_reference_kats.py
"""
import json
import os

K, S = 1, 2                      # PieceType::King, Soldier (pieces.rs:31-38)
A, D = 0, 8                      # Side::Attacker, Defender (pieces.rs:13-17)


def ps_type(t):                  # PieceSet::from(PieceType)  (pieces.rs:198-202)
    return (t | (t << 8)) & 0xFFFF


def ps_piece(t, s):              # PieceSet::from(Piece)      (pieces.rs:192-196)
    return (t << s) & 0xFFFF


def ps_side(s):                  # PieceSet::from(Side)       (pieces.rs:204-208)
    return (0x00FF << s) & 0xFFFF


WIN = lambda reason, side: {"win": [reason, side]}
ONGOING = "Ongoing"

cases = []


def case(name, ref, rules, side_len, steps, widths=(64, 128), override=None):
    cases.append(dict(name=name, ref=ref, rules=rules, override=override or {}, side_len=side_len,
                      widths=list(widths), steps=steps))


# rules.rs / logic.rs:904-908  TEST_RULES = { slow_pieces: from_piece_type(King), throne_movement: NoPass, ..BRANDUBH }
TEST_RULES = dict(slow_pieces=ps_type(K), throne_movement=1)

# ------------------------------------------------------------------------------------------------
# game/mod.rs:137-207  test_iter_plays
# ------------------------------------------------------------------------------------------------
case("test_iter_plays/start", "game/game/mod.rs:137-184", "BRANDUBH", 7, [
    dict(op="new", fen="@BRANDUBH", side="A"),
    dict(op="iter_plays", tile=[0, 0], expect="err"),
    dict(op="iter_plays", tile=[1, 0], expect="err"),
    dict(op="iter_plays", tile=[0, 3], expect=[[0, 1], [0, 2], [0, 4], [0, 5]]),
    dict(op="iter_plays", tile=[1, 3], expect=[[1, 0], [1, 1], [1, 2], [1, 4], [1, 5], [1, 6]]),
    dict(op="iter_plays", tile=[2, 3], expect=[[2, 0], [2, 1], [2, 2], [2, 4], [2, 5], [2, 6]]),
    dict(op="iter_plays", tile=[3, 3], expect=[]),
])
case("test_iter_plays/through_throne", "game/game/mod.rs:185-206", "BRANDUBH", 7, [
    dict(op="new", fen="1T5/7/7/1t3K1/7/7/7", side="A"),
    dict(op="iter_plays", tile=[3, 1],
         expect=[[1, 1], [2, 1], [4, 1], [5, 1], [6, 1], [3, 0], [3, 2], [3, 4]]),
])

# ------------------------------------------------------------------------------------------------
# logic.rs:923-1021  test_play_validity (generic over the bitfield widths; u64 and u128 here)
# ------------------------------------------------------------------------------------------------
case("test_play_validity/brandubh", "game/game/logic.rs:923-990", "BRANDUBH", 7, [
    dict(op="new", fen="@BRANDUBH", side="A"),
    dict(op="validate", frm=[3, 1], to=[4, 1], expect="Ok"),
    dict(op="validate", frm=[0, 3], to=[0, 0], expect="MoveOntoBlockedTile"),
    dict(op="validate", frm=[1, 1], to=[2, 1], expect="NoPiece"),
    dict(op="validate", frm=[0, 3], to=[0, 7], expect="OutOfBounds"),
    dict(op="validate", frm=[0, 3], to=[2, 3], expect="BlockedByPiece"),
    dict(op="do_play", frm=[3, 1], to=[4, 1]),
    dict(op="validate", frm=[3, 3], to=[3, 2], expect="BlockedByPiece"),
    dict(op="move_piece", frm=[3, 2], to=[4, 2]),
    dict(op="move_piece", frm=[3, 3], to=[3, 2]),
    dict(op="validate", frm=[2, 3], to=[3, 3], expect="MoveOntoBlockedTile"),
    dict(op="validate", frm=[3, 2], to=[3, 3], expect="Ok"),
])
case("test_play_validity/test_rules", "game/game/logic.rs:992-1013", "BRANDUBH", 7, [
    dict(op="new", fen="7/5Tt/2T4/2t2t1/Tt4T/2t4/2T2K1", side="D"),
    dict(op="validate", frm=[6, 5], to=[6, 3], expect="TooFar"),
    dict(op="validate", frm=[6, 5], to=[6, 4], expect="Ok"),
    dict(op="set_side", side="A"),
    dict(op="validate", frm=[3, 2], to=[3, 4], expect="MoveThroughBlockedTile"),
], override=TEST_RULES)

# ------------------------------------------------------------------------------------------------
# logic.rs:1023-1087  test_play_outcome
# ------------------------------------------------------------------------------------------------
PROTO = "4t2/5Tt/2T4/2t2t1/Tt4T/2t4/2T2K1"
case("test_play_outcome/king_captured", "game/game/logic.rs:1029-1042", "BRANDUBH", 7, [
    dict(op="new", fen=PROTO, side="A"),
    dict(op="get_captures", frm=[0, 4], to=[6, 4], expect=[[6, 5]]),
    dict(op="do_play", frm=[0, 4], to=[6, 4], status=WIN("KingCaptured", "Attacker")),
], override=TEST_RULES)
case("test_play_outcome/triple_capture", "game/game/logic.rs:1044-1058", "BRANDUBH", 7, [
    dict(op="new", fen=PROTO, side="D"),
    dict(op="get_captures", frm=[4, 6], to=[4, 2], expect=[[4, 1], [3, 2], [5, 2]]),
    dict(op="do_play", frm=[4, 6], to=[4, 2], status=ONGOING, captures=[[4, 1], [3, 2], [5, 2]]),
], override=TEST_RULES)
case("test_play_outcome/king_escaped", "game/game/logic.rs:1060-1069", "BRANDUBH", 7, [
    dict(op="new", fen=PROTO, side="D"),
    dict(op="get_captures", frm=[6, 5], to=[6, 6], expect=[]),
    dict(op="do_play", frm=[6, 5], to=[6, 6], status=WIN("KingEscaped", "Defender")),
], override=TEST_RULES)
case("test_play_outcome/ongoing", "game/game/logic.rs:1071-1079", "BRANDUBH", 7, [
    dict(op="new", fen=PROTO, side="D"),
    dict(op="get_captures", frm=[6, 5], to=[5, 5], expect=[]),
    dict(op="do_play", frm=[6, 5], to=[5, 5], status=ONGOING),
], override=TEST_RULES)

# ------------------------------------------------------------------------------------------------
# logic.rs:1090-1169  test_shieldwalls (MediumBasicBoardState = u128)
# ------------------------------------------------------------------------------------------------
NO_CORNER = dict(shieldwall_enabled=1, sw_corners_may_close=0, sw_captures=ps_type(S))
KING_CAP = dict(shieldwall_enabled=1, sw_corners_may_close=0, sw_captures=0xFFFF)
corner_sw = "9/9/9/9/6t2/7tT/7tT/7tT/9"
regular_sw = "9/9/9/6t2/7tT/7tT/7tT/8t/9"
regular_sw_king = "9/9/9/6t2/7tT/7tK/7tT/8t/9"
no_sw_gap = "9/9/9/6t2/7tT/8T/7tT/8t/9"
no_sw_friend = "9/9/9/6t2/7tT/6tTT/7tT/8t/9"
no_sw_small = "9/9/9/6t2/7tT/8t/9/9/9"
CM, M, N_ = ([4, 6], [4, 8]), ([3, 6], [3, 8]), ([3, 6], [3, 7])


def sw(play, expect):
    return dict(op="detect_shieldwall", frm=play[0], to=play[1], expect=expect)


case("test_shieldwalls/corner", "game/game/logic.rs:1131-1138", "COPENHAGEN", 9, [
    dict(op="new", fen=corner_sw, side="A"), sw(N_, None), sw(CM, [[5, 8], [6, 8], [7, 8]])], widths=(128,))
case("test_shieldwalls/no_corner_rules", "game/game/logic.rs:1140-1141", "COPENHAGEN", 9, [
    dict(op="new", fen=corner_sw, side="A"), sw(M, None)], widths=(128,), override=NO_CORNER)
case("test_shieldwalls/regular", "game/game/logic.rs:1143-1149", "COPENHAGEN", 9, [
    dict(op="new", fen=regular_sw, side="A"), sw(M, [[4, 8], [5, 8], [6, 8]])], widths=(128,), override=NO_CORNER)
case("test_shieldwalls/king_not_captured", "game/game/logic.rs:1151-1155", "COPENHAGEN", 9, [
    dict(op="new", fen=regular_sw_king, side="A"), sw(M, [[4, 8], [6, 8]])], widths=(128,), override=NO_CORNER)
case("test_shieldwalls/king_captured", "game/game/logic.rs:1157-1162", "COPENHAGEN", 9, [
    dict(op="new", fen=regular_sw_king, side="A"), sw(M, [[4, 8], [5, 8], [6, 8]])], widths=(128,), override=KING_CAP)
for nm, fen, ref in (("gap", no_sw_gap, "1164-1165"), ("friend", no_sw_friend, "1167-1168"),
                     ("small", no_sw_small, "1170-1171")):
    case("test_shieldwalls/" + nm, "game/game/logic.rs:" + ref, "COPENHAGEN", 9, [
        dict(op="new", fen=fen, side="A"), sw(M, None)], widths=(128,), override=NO_CORNER)

# ------------------------------------------------------------------------------------------------
# logic.rs:1172-1213  test_encl_secure (SmallBasicBoardState = u64, Copenhagen rules on 7 wide)
# ------------------------------------------------------------------------------------------------
setup = ["7/2ttt2/1t1K1t1/2ttt2/7", "7/1tttt2/1t1K1t1/2tttt1/7", "2t1t2/1t1t1t1/1t1K1t1/2ttt2/7",
         "2t2t1/1t3t1/1t1K1t1/2ttt2/7"]
SAFE_CORNERS = dict(hostility_corners=0, hostility_edge=0, hostility_throne=0)
for i, (s, ins, outs, secure, ov) in enumerate([
        (0, False, True, True, None), (0, False, False, False, None), (1, False, True, True, None),
        (1, True, False, True, None), (2, False, True, False, None), (3, False, True, False, None),
        (3, False, True, True, SAFE_CORNERS), (3, True, False, True, None)]):
    case("test_encl_secure/%d" % i, "game/game/logic.rs:1188-1211", "COPENHAGEN", 7, [
        dict(op="new", fen=setup[s], side="A"),
        dict(op="enclosure_secure", tile=[2, 3], enclosed=ps_type(K), enclosing=ps_piece(S, A), abort_edge=False,
             abort_corner=False, inside_safe=ins, outside_safe=outs, expect=secure)], override=ov)

# ------------------------------------------------------------------------------------------------
# logic.rs:1216-1233  test_exit_forts (u128, Copenhagen, 9x9)
# ------------------------------------------------------------------------------------------------
for nm, fen, exp in (("flat", "9/9/8t/7tT/7T1/6tT1/7TK/7tT/9", True), ("bulge", "9/9/9/9/9/5TTTT/5T2K/6TTT/9", True),
                     ("enemy", "9/9/9/8T/7Tt/7T1/7TK/8T/9", False), ("unfree", "9/9/9/8T/7TT/7TT/7TK/8T/9", False),
                     ("gap", "9/9/9/8T/9/4t2T1/7TK/8T/9", False), ("vuln", "9/9/9/9/9/6TTT/5T2K/6TTT/9", False)):
    case("test_exit_forts/" + nm, "game/game/logic.rs:1216-1233", "COPENHAGEN", 9, [
        dict(op="new", fen=fen, side="A"), dict(op="exit_fort", expect=exp)], widths=(128,))

# ------------------------------------------------------------------------------------------------
# logic.rs:1236-1385  test_enclosures (u64, Brandubh)
# ------------------------------------------------------------------------------------------------
BOUND8 = [[0, 2], [0, 3], [0, 4], [1, 1], [1, 5], [2, 2], [2, 3], [2, 4]]


def encl(tile, enclosed, enclosing, ae, ac, expect):
    return dict(op="find_enclosure", tile=tile, enclosed=enclosed, enclosing=enclosing, abort_edge=ae,
                abort_corner=ac, expect=expect)


case("test_enclosures/full", "game/game/logic.rs:1243-1266", "BRANDUBH", 7, [
    dict(op="new", fen="2ttt2/1t1K1t1/2tttt1/7/7/7/7", side="A"),
    encl([1, 3], ps_type(K), ps_type(S), True, True,
         dict(occupied=[[1, 3]], unoccupied=[[1, 2], [1, 4]], boundary=BOUND8))])
case("test_enclosures/with_edge", "game/game/logic.rs:1268-1302", "BRANDUBH", 7, [
    dict(op="new", fen="2t1t2/1t1K1t1/2tttt1/7/7/7/7", side="A"),
    encl([1, 3], ps_type(K), ps_type(S), True, True, None),
    encl([1, 3], ps_type(K), ps_type(S), False, True,
         dict(occupied=[[1, 3]], unoccupied=[[0, 3], [1, 2], [1, 4]],
              boundary=[[0, 2], [0, 4], [1, 1], [1, 5], [2, 2], [2, 3], [2, 4]]))])
case("test_enclosures/with_corner", "game/game/logic.rs:1304-1337", "BRANDUBH", 7, [
    dict(op="new", fen="5t1/4tK1/4ttt/7/7/7/7", side="A"),
    encl([1, 3], ps_type(K), ps_type(S), False, True, None),
    encl([1, 5], ps_type(K), ps_type(S), False, False,
         dict(occupied=[[1, 5]], unoccupied=[[0, 6], [1, 6]], boundary=[[0, 5], [1, 4], [2, 5], [2, 6]]))])
case("test_enclosures/with_soldier", "game/game/logic.rs:1339-1373", "BRANDUBH", 7, [
    dict(op="new", fen="2ttt2/1t1KTt1/2tttt1/7/7/7/7", side="A"),
    encl([1, 3], ps_type(K), ps_piece(S, A), True, True, None),
    encl([1, 3], ps_piece(K, D) | ps_piece(S, D), ps_piece(S, A), True, True,
         dict(occupied=[[1, 3], [1, 4]], unoccupied=[[1, 2]], boundary=BOUND8))])
case("test_enclosures/edge_2", "game/game/logic.rs:1375-1385", "BRANDUBH", 7, [
    dict(op="new", fen="1t2t2/1t1K1t1/2tttt1/7/7/7/7", side="A"),
    encl([1, 3], ps_type(K), ps_piece(S, A), False, False, "some")])

# ------------------------------------------------------------------------------------------------
# logic.rs:1388-1403  test_can_play
# ------------------------------------------------------------------------------------------------
for i, fen in enumerate(("2tt3/1tTKt2/2tt3/7/7/7/7", "2tKt2/3t3/7/7/7/7/7")):
    case("test_can_play/%d" % i, "game/game/logic.rs:1388-1403", "BRANDUBH", 7, [
        dict(op="new", fen=fen, side="A"),
        dict(op="can_play", side="A", expect=True), dict(op="can_play", side="D", expect=False)])

# ------------------------------------------------------------------------------------------------
# logic.rs:1406-1421  test_repetitions
# ------------------------------------------------------------------------------------------------
rep_steps = [dict(op="new", fen="@BRANDUBH", side="A")]
for _ in range(3):
    for p in ("d6-f6", "d5-f5", "f6-d6", "f5-d5"):
        rep_steps.append(dict(op="do_play_str", play=p, status=ONGOING))
rep_steps.append(dict(op="do_play_str", play="d6-f6", status=WIN("Repetition", "Defender")))
case("test_repetitions", "game/game/logic.rs:1406-1421", "BRANDUBH", 7, rep_steps)

# ------------------------------------------------------------------------------------------------
# logic.rs:1424-1462  test_strong_king_capture (Brandubh, u64)
# ------------------------------------------------------------------------------------------------
for i, (fen, frm, to, caps, st) in enumerate([
        ("1T5/7/4t2/4K1t/4t2/7/7", [3, 6], [3, 5], [[3, 4]], WIN("KingCaptured", "Attacker")),
        ("1T5/4t2/7/4Kt1/4t2/7/7", [1, 4], [2, 4], [[3, 4]], WIN("KingCaptured", "Attacker")),
        ("1T5/7/7/4K1t/4t2/7/7", [3, 6], [3, 5], [], ONGOING),
        ("1T5/4t2/7/4K2/4t2/7/7", [1, 4], [2, 4], [], ONGOING)]):
    case("test_strong_king_capture/%d" % i, "game/game/logic.rs:1424-1462", "BRANDUBH", 7, [
        dict(op="new", fen=fen, side="A"), dict(op="do_play", frm=frm, to=to, status=st, captures=caps)])

# ------------------------------------------------------------------------------------------------
# logic.rs:1465-1482  test_linnaean_capture (Tablut, u128)
# ------------------------------------------------------------------------------------------------
case("test_linnaean_capture", "game/game/logic.rs:1465-1482", "TABLUT", 9, [
    dict(op="new", fen="tT7/9/9/4t4/t2TKt3/4t4/9/9/9", side="A"),
    dict(op="do_play", frm=[4, 0], to=[4, 2], captures=[[4, 3]])], widths=(128,))

# ------------------------------------------------------------------------------------------------
# Host-side (no rules engine) known answers: board/state.rs, tiles.rs, game/state.rs tests
# ------------------------------------------------------------------------------------------------
host = dict(
    test_from_str=dict(ref="game/board/state.rs:354-372", fen="3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3",
                       display="...t...\n...t...\n...T...\nttTKTtt\n...T...\n...t...\n...t..."),
    test_piece_movement=dict(ref="game/board/state.rs:375-404", fen="3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3",
                             king=[3, 3], set=[[1, 5, S, A], [4, 1, S, D]], move=[[3, 3], [0, 4]], king_after=[0, 4],
                             fen_after="3tK2/3t1t1/3T3/ttT1Ttt/1T1T3/3t3/3t3",
                             occupied=[[0, 3], [2, 3], [0, 4]], empty=[[3, 3], [5, 4], [1, 1]]),
    test_iter_occupied=dict(ref="game/board/state.rs:408-431", fen="3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3",
                            attackers=[[0, 3], [1, 3], [5, 3], [6, 3], [3, 0], [3, 1], [3, 5], [3, 6]],
                            defenders=[[2, 3], [3, 3], [4, 3], [3, 2], [3, 4]]),
    test_swap_pieces=dict(ref="game/board/state.rs:434-444", fen="5/1K3/5/5/3t1", a=[1, 1], b=[4, 3]),
    test_count_pieces=dict(ref="game/board/state.rs:447-451", fen="@COPENHAGEN", attackers=24, defenders=13),
    test_moves=dict(ref="game/tiles.rs:237-273", ok=[
        dict(frm=[2, 4], to=[2, 6], axis="H", disp=2), dict(frm=[2, 3], to=[5, 3], axis="V", disp=3),
        dict(frm=[1, 4], to=[1, 1], axis="H", disp=-3), dict(frm=[7, 5], to=[0, 5], axis="V", disp=-7)],
        err=[dict(frm=[2, 3], to=[3, 6])]),
    test_parsing_tiles=dict(ref="game/tiles.rs:276-292", ok=[["a8", [7, 0]], ["f14", [13, 5]]],
                            err=[["", "EmptyString"], ["[53", "BadChar"], ["a!!", "BadInt"]]),
    test_parsing_moves=dict(ref="game/tiles.rs:295-325", ok=[["a8-a11", [7, 0], [10, 0]], ["f5-d5", [4, 5], [4, 3]]],
                            err=[["f5-d6", "BadPlay"], ["f5-d7-d6", "BadString"], ["f5-d]", "BadInt"],
                                 ["!5-d5", "BadChar"]]),
    test_repetition_tracker=dict(ref="game/game/state.rs:165-187",
                                 blocks=[["a1-b1", "a2-b2", "b1-a1", "b2-a2"], ["f1-g1", "f2-g2", "g1-f1", "g2-f2"]],
                                 rounds=5),
)

out = dict(
    provenance="hand-transcribed from the reference crate's #[cfg(test)] modules (Rust toolchain unavailable); "
               "see tests/golden/make_reference_kats.py",
    cases=cases, host=host)
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_kats.json")
with open(path, "w") as f:
    json.dump(out, f, indent=1, sort_keys=True)
    f.write("\n")
print("wrote", path, len(cases), "cases")
