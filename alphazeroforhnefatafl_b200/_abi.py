"""ctypes mirrors of the POD types in ``include/tafl.h`` (one-to-one, same field order).

Shared by the product binding (``_lib.py``) and by the test-only oracle / hostsim loaders.
"""
import ctypes as C

import numpy as np

MAX_MOVES = 512
REC_MOVES = MAX_MOVES
REC_STRIDE = 56 + 2 * REC_MOVES + 4 * REC_MOVES

ATTACKER, DEFENDER = 0, 8
KING, SOLDIER, KNIGHT, COMMANDER, GUARD, MERCENARY = 1, 2, 4, 8, 16, 32
VERTICAL, HORIZONTAL = 0, 0x80

THRONE_NOTHRONE, THRONE_NOPASS, THRONE_KINGPASS, THRONE_NOENTRY, THRONE_KINGENTRY = range(5)
KING_STRONG, KING_STRONG_BY_THRONE, KING_WEAK = range(3)
KING_ARMED, KING_ANVIL, KING_HAMMER = range(3)
ENCL_NONE, ENCL_WITH_EDGE_ACCESS, ENCL_WITHOUT_EDGE_ACCESS = range(3)

(PLAY_OK, WRONG_PLAYER, NO_PIECE, OUT_OF_BOUNDS, NO_COMMON_AXIS, BLOCKED_BY_PIECE,
 MOVE_THROUGH_BLOCKED_TILE, MOVE_ONTO_BLOCKED_TILE, TOO_FAR, GAME_OVER) = range(10)

(WIN_KING_ESCAPED, WIN_EXIT_FORT, WIN_KING_CAPTURED, WIN_ALL_CAPTURED, WIN_ENCLOSED, WIN_NO_PLAYS,
 WIN_REPETITION) = range(7)
DRAW_REPETITION, DRAW_NO_PLAYS = range(2)

PRESET_COPENHAGEN, PRESET_BRANDUBH, PRESET_MAGPIE, PRESET_TABLUT, PRESET_KOCH = range(5)

(OK, ERR_INVALID_ARG, ERR_CUDA, ERR_PARSE_BAD_LINE_LEN, ERR_PARSE_BAD_CHAR, ERR_PARSE_EMPTY,
 ERR_PARSE_BAD_INT, ERR_PARSE_BAD_PLAY, ERR_PARSE_BAD_STRING, ERR_UNSUPPORTED, ERR_CAPACITY,
 ERR_NO_PIECE) = (0, -1, -2, -3, -4, -5, -6, -7, -8, -9, -10, -11)

FLAG_ATT_MID, FLAG_DEF_MID, FLAG_DEF_TO_PLAY = 1, 2, 4


def pieceset_type(t):
    return (t | (t << 8)) & 0xFFFF


def pieceset_side(s):
    return (0x00FF << s) & 0xFFFF


def pieceset_piece(t, s):
    return (t << s) & 0xFFFF


PIECESET_NONE, PIECESET_ALL = 0, 0xFFFF


def status_win(reason, winner_is_defender):
    return 0x80 | (reason << 1) | (1 if winner_is_defender else 0)


def status_draw(reason):
    return 0xC0 | (reason << 1)


STATUS_ONGOING = 0


class Play(C.Structure):
    _fields_ = [("from_row", C.c_uint8), ("from_col", C.c_uint8), ("axis", C.c_uint8), ("disp", C.c_int8)]

    def to(self):
        if self.axis == HORIZONTAL:
            return self.from_row, self.from_col + self.disp
        return self.from_row + self.disp, self.from_col

    def __repr__(self):
        return f"Play(({self.from_row},{self.from_col})->{self.to()})"


class Rules(C.Structure):
    _fields_ = [
        ("edge_escape", C.c_uint8), ("king_strength", C.c_uint8), ("king_attack", C.c_uint8),
        ("shieldwall_enabled", C.c_uint8), ("sw_corners_may_close", C.c_uint8), ("exit_fort", C.c_uint8),
        ("throne_movement", C.c_uint8), ("starting_side", C.c_uint8),
        ("sw_captures", C.c_uint16), ("may_enter_corners", C.c_uint16), ("hostility_throne", C.c_uint16),
        ("hostility_corners", C.c_uint16), ("hostility_edge", C.c_uint16), ("slow_pieces", C.c_uint16),
        ("enclosure_win", C.c_uint8), ("repetition_enabled", C.c_uint8), ("repetition_is_loss", C.c_uint8),
        ("draw_on_no_plays", C.c_uint8), ("linnaean_capture", C.c_uint8), ("_pad", C.c_uint8 * 3),
        ("n_repetitions", C.c_uint32),
    ]

    def copy(self, **overrides):
        r = Rules.from_buffer_copy(bytes(self))
        for k, v in overrides.items():
            setattr(r, k, v)
        return r


class State(C.Structure):
    _fields_ = [
        ("att", C.c_uint64 * 2), ("def_", C.c_uint64 * 2), ("recent", C.c_uint32 * 4),
        ("att_reps", C.c_uint8), ("def_reps", C.c_uint8), ("flags", C.c_uint8), ("status", C.c_uint8),
        ("turn", C.c_uint32), ("plays_since_capture", C.c_uint32),
        ("ring_head", C.c_uint8), ("side_len", C.c_uint8), ("_pad", C.c_uint8 * 2),
    ]


class PlayEffects(C.Structure):
    _fields_ = [("captures", C.c_uint64 * 2), ("status", C.c_uint8), ("err", C.c_uint8),
                ("n_captures", C.c_uint8), ("_pad", C.c_uint8 * 5)]


class Enclosure(C.Structure):
    _fields_ = [("occupied", C.c_uint64 * 2), ("unoccupied", C.c_uint64 * 2), ("boundary", C.c_uint64 * 2),
                ("found", C.c_uint8), ("_pad", C.c_uint8 * 7)]


class EnclosureQuery(C.Structure):
    _fields_ = [("row", C.c_uint8), ("col", C.c_uint8), ("abort_on_edge", C.c_uint8),
                ("abort_on_corner", C.c_uint8), ("enclosed_pieces", C.c_uint16),
                ("enclosing_pieces", C.c_uint16), ("inside_safe", C.c_uint8), ("outside_safe", C.c_uint8),
                ("_pad", C.c_uint8 * 6)]


class Stats(C.Structure):
    _fields_ = [("steps", C.c_uint64), ("legal_moves", C.c_uint64), ("captures", C.c_uint64),
                ("games_finished", C.c_uint64), ("finished_plies", C.c_uint64),
                ("win_reason", (C.c_uint64 * 8) * 2), ("draw_reason", C.c_uint64 * 2),
                ("state_checksum", C.c_uint64)]

    def as_dict(self):
        return {
            "steps": self.steps, "legal_moves": self.legal_moves, "captures": self.captures,
            "games_finished": self.games_finished, "finished_plies": self.finished_plies,
            "win_reason": [[self.win_reason[i][j] for j in range(8)] for i in range(2)],
            "draw_reason": [self.draw_reason[j] for j in range(2)],
        }


class MctsConfig(C.Structure):
    _fields_ = [("num_sims", C.c_uint32), ("cpuct", C.c_double), ("draw_value", C.c_double),
                ("edge_capacity", C.c_uint32), ("flags", C.c_uint32), ("node_capacity", C.c_uint32), ("vl_width", C.c_uint32)]


MCTS_TREE_REUSE = 1


class MctsStats(C.Structure):
    _fields_ = [("sims", C.c_uint64), ("nodes", C.c_uint64), ("depth_sum", C.c_uint64), ("edges", C.c_uint64),
                ("terminal_hits", C.c_uint64), ("max_edges_used", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class LeafBatch(C.Structure):
    """tafl_leaf_batch: device pointers of the pending evaluator batch (include/tafl.h)."""
    _fields_ = [("d_states", C.c_void_p), ("d_planes", C.c_void_p), ("d_moves", C.c_void_p), ("d_counts", C.c_void_p),
                ("d_priors", C.c_void_p), ("d_values", C.c_void_p), ("n_games", C.c_uint32), ("side_len", C.c_uint32),
                ("stream", C.c_void_p)]


EVALUATOR_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(LeafBatch))
EVAL_STUB, EVAL_HASH = 0, 1

assert C.sizeof(LeafBatch) == 64
assert C.sizeof(State) == 64 and C.sizeof(Play) == 4 and C.sizeof(PlayEffects) == 24
assert C.sizeof(Enclosure) == 56 and C.sizeof(EnclosureQuery) == 16 and C.sizeof(Rules) == 32

STATE_DTYPE = np.dtype([
    ("att", "<u8", (2,)), ("def", "<u8", (2,)), ("recent", "<u4", (4,)),
    ("att_reps", "u1"), ("def_reps", "u1"), ("flags", "u1"), ("status", "u1"),
    ("turn", "<u4"), ("plays_since_capture", "<u4"), ("ring_head", "u1"), ("side_len", "u1"), ("_pad", "u1", (2,)),
])
PLAY_DTYPE = np.dtype([("from_row", "u1"), ("from_col", "u1"), ("axis", "u1"), ("disp", "i1")])
EFFECTS_DTYPE = np.dtype([("captures", "<u8", (2,)), ("status", "u1"), ("err", "u1"), ("n_captures", "u1"),
                          ("_pad", "u1", (5,))])
PLAY_RECORD_DTYPE = np.dtype([("play", PLAY_DTYPE), ("side", "u1"), ("_pad", "u1", (3,)), ("effects", EFFECTS_DTYPE)])
assert PLAY_RECORD_DTYPE.itemsize == 32
ENCLOSURE_DTYPE = np.dtype([("occupied", "<u8", (2,)), ("unoccupied", "<u8", (2,)), ("boundary", "<u8", (2,)),
                            ("found", "u1"), ("_pad", "u1", (7,))])
ENCL_QUERY_DTYPE = np.dtype([("row", "u1"), ("col", "u1"), ("abort_on_edge", "u1"), ("abort_on_corner", "u1"),
                             ("enclosed_pieces", "<u2"), ("enclosing_pieces", "<u2"), ("inside_safe", "u1"),
                             ("outside_safe", "u1"), ("_pad", "u1", (6,))])
assert STATE_DTYPE.itemsize == 64 and PLAY_DTYPE.itemsize == 4 and EFFECTS_DTYPE.itemsize == 24
assert ENCLOSURE_DTYPE.itemsize == 56 and ENCL_QUERY_DTYPE.itemsize == 16


# large boards (BitfieldBoardState<U256> / <U512>): tafl_state_w, tafl_play_effects_w, tafl_enclosure_w
W_MAX_PLAYS = 2048
STATE_W_DTYPE = np.dtype([
    ("att", "<u8", (8,)), ("def", "<u8", (8,)), ("recent", "<u4", (4,)),
    ("att_reps", "u1"), ("def_reps", "u1"), ("flags", "u1"), ("status", "u1"),
    ("turn", "<u4"), ("plays_since_capture", "<u4"), ("ring_head", "u1"), ("side_len", "u1"), ("_pad", "u1", (2,)),
])
EFFECTS_W_DTYPE = np.dtype([("captures", "<u8", (8,)), ("status", "u1"), ("err", "u1"), ("n_captures", "u1"), ("_pad", "u1", (5,))])
ENCLOSURE_W_DTYPE = np.dtype([("occupied", "<u8", (8,)), ("unoccupied", "<u8", (8,)), ("boundary", "<u8", (8,)),
                              ("found", "u1"), ("_pad", "u1", (7,))])
assert STATE_W_DTYPE.itemsize == 160 and EFFECTS_W_DTYPE.itemsize == 72 and ENCLOSURE_W_DTYPE.itemsize == 200


def is_wide(width):
    return width in (256, 512)


def state_dtype(width):
    return STATE_W_DTYPE if is_wide(width) else STATE_DTYPE


def ptr(arr, ctype=None):
    """ctypes pointer to a contiguous numpy array's buffer."""
    assert arr.flags["C_CONTIGUOUS"]
    p = arr.ctypes.data_as(C.c_void_p)
    return C.cast(p, C.POINTER(ctype)) if ctype is not None else p


def width_for(side_len):
    """Smallest bitfield width that fits the board (board/state.rs:331-334)."""
    if side_len <= 7:
        return 64
    if side_len <= 11:
        return 128
    if side_len <= 15:
        return 256
    if side_len <= 21:
        return 512
    raise ValueError("no bitfield holds a board larger than 21x21")


def stride_for(width):
    return {64: 7, 128: 11, 256: 15, 512: 21}[width]   # bitfield.rs:178-181


def tile_bit(width, row, col):
    return row * stride_for(width) + col


def bb_to_tiles(width, words):
    """tile bitboard (two u64 words; eight for the wide widths) -> sorted list of (row, col)."""
    v = 0
    for k, x in enumerate(words):
        v |= int(x) << (64 * k)
    s = stride_for(width)
    out = []
    i = 0
    while v:
        if v & 1:
            out.append((i // s, i % s))
        v >>= 1
        i += 1
    return out


def move_from_to(width, m):
    s = stride_for(width)
    f, t = int(m) & 127, (int(m) >> 7) & 127
    return (f // s, f % s), (t // s, t % s)
