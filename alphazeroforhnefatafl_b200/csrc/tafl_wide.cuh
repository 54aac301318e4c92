// tafl_wide.cuh — the by-value GameLogic queries for the large bitfields (BB256 / BB512, tafl_state_w).
//
// The rules themselves are the width-generic functions of tafl_core.cuh (the reference is generic over
// BitfieldBoardState<T> in the same way: game/board/state.rs:337-340, game/bitfield.rs:180-181, tests
// instantiated at game/game/logic.rs:1015-1021,1081-1087); this header only adds what depends on the wide POD
// types: move lists as tafl_play (a 21x21 tile index does not fit a tafl_move), effects / enclosures with
// eight-word tile sets.  __host__ __device__, so tests/hostsim checks the same code against the oracle on the CPU.
#pragma once
#include "tafl_core.cuh"

namespace tafl {

template <class BB> TAFL_HD void put_words(uint64_t (&o)[8], BB b) {
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = 2 * i < BB::WORDS ? ((uint64_t)b.word(2 * i < BB::WORDS ? 2 * i : 0) | ((uint64_t)b.word(2 * i + 1 < BB::WORDS ? 2 * i + 1 : 0) << 32)) : 0ull;
}
TAFL_HD tafl_play make_play(int row, int col, int horizontal, int disp) {
    tafl_play p;
    p.from_row = (uint8_t)row;
    p.from_col = (uint8_t)col;
    p.axis = horizontal ? TAFL_HORIZONTAL : TAFL_VERTICAL;
    p.disp = (int8_t)disp;
    return p;
}
// the plays of one piece in ValidPlayIterator order (play.rs:189-225): V+, V-, H+, H-, distance ascending
template <class BB, int SP>
TAFL_HD int emit_plays_of(const Ctx<BB, SP>& cx, BB occ, int kind, int row, int col, tafl_play* out, int n, int cap) {
    const Rays r = piece_rays(cx, occ, kind, row, col);
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        uint32_t m = r.m[d];
        while (m) {
            int pos;
            if ((d & 1) == 0) { pos = ctz32(m); m &= m - 1; }
            else { pos = 31 - clz32(m); m &= ~(1u << pos); }
            if (n < cap) out[n] = make_play(row, col, d >= 2, d < 2 ? pos - row : pos - col);
            ++n;
        }
    }
    return n;
}
// get_all_possible_moves (game/main.rs:33-43): pieces ascending bit index
template <class BB, int SP>
TAFL_HD int gen_plays(const Ctx<BB, SP>& cx, const tafl_state_w& st, tafl_play* out, int cap) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;  // logic.rs:165-167
    const Pos<BB> p = load_pos(cx, st);
    const int def_side = (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    BB own = p.side(def_side);
    const BB occ = p.occ();
    int n = 0;
    while (own.any()) {
        const int s = own.ctz();
        own = own & ~BB::bit(s);
        const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
        const int k = def_side ? ((BB::bit(s) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
        n = emit_plays_of(cx, occ, k, row, col, out, n, cap);
    }
    return n;
}
// GameLogic::iter_plays(tile) (logic.rs:850-856): -1 == Err(NoPiece)
template <class BB, int SP>
TAFL_HD int piece_plays_w(const Ctx<BB, SP>& cx, const tafl_state_w& st, int row, int col, tafl_play* out, int cap) {
    const Pos<BB> p = load_pos(cx, st);
    const int s = idx_of<BB>(row, col);
    int kind;
    if (Nib<BB>::raw_def(st, s)) kind = (row == p.kr && col == p.kc) ? KIND_K : KIND_DS;
    else if (Nib<BB>::raw_att(st, s)) kind = KIND_AS;
    else return -1;
    if (st.status != TAFL_STATUS_ONGOING) return 0;
    if (row >= cx.n() || col >= cx.n()) return 0;
    return emit_plays_of(cx, p.occ(), kind, row, col, out, 0, cap);
}
TAFL_HD tafl_play_effects_w blank_effects_w() {
    tafl_play_effects_w fx;
    for (int i = 0; i < 8; ++i) fx.captures[i] = 0;
    fx.status = 0; fx.err = 0; fx.n_captures = 0;
    for (int i = 0; i < 5; ++i) fx._pad[i] = 0;
    return fx;
}
// Game::do_play by value (logic.rs:822-834): validate, apply, report
template <class BB, int SP>
TAFL_HD tafl_play_effects_w do_play_w(const Ctx<BB, SP>& cx, tafl_state_w& s, const tafl_play& pl) {
    tafl_play_effects_w fx = blank_effects_w();
    const int err = validate_play(cx, s, pl, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
    fx.err = (uint8_t)err;
    if (!err) {
        const BB caps = do_valid_play(cx, s, pl.from_row, pl.from_col, pl.axis == TAFL_HORIZONTAL, pl.disp);
        put_words(fx.captures, caps);
        fx.n_captures = (uint8_t)caps.popc();
    }
    fx.status = s.status;
    return fx;
}
// GameLogic::get_captures (logic.rs:604-699) after moving the piece as do_valid_play does
template <class BB, int SP>
TAFL_HD tafl_play_effects_w query_captures_w(const Ctx<BB, SP>& cx, const tafl_state_w& st, const tafl_play& pl) {
    tafl_play_effects_w fx = blank_effects_w();
    Pos<BB> p = load_pos(cx, st);
    const int fr = pl.from_row, fc = pl.from_col, hz = pl.axis == TAFL_HORIZONTAL;
    const int tr = hz ? fr : fr + pl.disp, tc = hz ? fc + pl.disp : fc;
    const BB fb = (fr < cx.n() && fc < cx.n()) ? BB::bit(idx_of<BB>(fr, fc)) : BB::zero();
    const int pk = kind_at(p, fb);
    if (pk == KIND_NONE || tr < 0 || tc < 0 || tr >= cx.n() || tc >= cx.n()) { fx.err = TAFL_NO_PIECE; return fx; }
    const BB tb = BB::bit(idx_of<BB>(tr, tc));
    const int piece_def = pk != KIND_AS;
    if (piece_def) { p.def = p.def | tb; p.att = p.att & ~tb; } else { p.att = p.att | tb; p.def = p.def & ~tb; }
    if (pk == KIND_K) { p.kr = tr; p.kc = tc; p.kbit = tb; }
    p.att = p.att & ~fb; p.def = p.def & ~fb;
    const BB caps = get_captures(cx, p, piece_def, pk, (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, tr, tc);
    put_words(fx.captures, caps);
    fx.n_captures = (uint8_t)caps.popc();
    return fx;
}
// GameLogic::detect_shieldwall (logic.rs:535-569) on the board as given; n_captures == 0xFF: None
template <class BB, int SP>
TAFL_HD tafl_play_effects_w query_shieldwall_w(const Ctx<BB, SP>& cx, const tafl_state_w& st, const tafl_play& pl) {
    tafl_play_effects_w fx = blank_effects_w();
    fx.n_captures = 0xFF;
    const Pos<BB> p = load_pos(cx, st);
    const int fr = pl.from_row, fc = pl.from_col, hz = pl.axis == TAFL_HORIZONTAL;
    const int tr = hz ? fr : fr + pl.disp, tc = hz ? fc + pl.disp : fc;
    BB caps;
    if (detect_shieldwall(cx, p, (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, tr, tc, &caps)) {
        put_words(fx.captures, caps);
        fx.n_captures = (uint8_t)caps.popc();
    }
    return fx;
}
// GameLogic::find_enclosure (logic.rs:309-401) [+ enclosure_secure logic.rs:408-463 -> *secure, 0xFF: no enclosure]
template <class BB, int SP>
TAFL_HD tafl_enclosure_w query_enclosure_w(const Ctx<BB, SP>& cx, const tafl_state_w& st, const tafl_enclosure_query& q, uint8_t* secure) {
    tafl_enclosure_w o;
    for (int i = 0; i < 8; ++i) o.occupied[i] = o.unoccupied[i] = o.boundary[i] = 0;
    o.found = 0;
    for (int i = 0; i < 7; ++i) o._pad[i] = 0;
    const Pos<BB> p = load_pos(cx, st);
    const BB enclosed = pieceset_bb(p, q.enclosed_pieces);
    const BB enclosing = pieceset_bb(p, q.enclosing_pieces) & ~enclosed;
    const Encl<BB> e = find_enclosure(cx, p, q.row, q.col, enclosed, enclosing, q.abort_on_edge != 0, q.abort_on_corner != 0);
    *secure = 0xFF;
    if (!e.found) return o;
    o.found = 1;
    put_words(o.occupied, e.region & p.occ());
    put_words(o.unoccupied, e.region & ~p.occ());
    put_words(o.boundary, e.boundary);
    *secure = enclosure_secure(cx, p, e, q.inside_safe != 0, q.outside_safe != 0) ? 1 : 0;
    return o;
}
template <class BB, int SP>
TAFL_HD uint8_t query_can_play_w(const Ctx<BB, SP>& cx, const tafl_state_w& st, int def_side) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;  // every play validates as GameOver (logic.rs:165-167)
    return side_can_play(cx, load_pos(cx, st), def_side) ? 1 : 0;
}

}  // namespace tafl
