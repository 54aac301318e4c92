// tafl_ctx.hpp — host-side construction of Ctx<BB> (`GameLogic::new`, logic.rs:70-72 +
// `SpecialTiles::from`, board/geometry.rs:13-24): rules + side length -> constant masks.
#pragma once
#include <cstring>
#include <string>

#include "tafl_core.cuh"

namespace tafl {

inline bool host_set_has(uint16_t set, int kind) { return pieceset_has(set, kind); }

template <class BB>
inline Ctx<BB> make_ctx(const tafl_rules& r, int N) {
    Ctx<BB> cx;
    std::memset(&cx, 0, sizeof(cx));
    cx.rules = r;
    cx.N = N;
    cx.mid = N / 2;
    auto B = [&](int row, int col) { return BB::bit(row * BB::STRIDE + col); };
    BB board = BB::zero(), edge = BB::zero(), row0 = BB::zero(), rowL = BB::zero(), col0 = BB::zero(), colL = BB::zero();
    for (int row = 0; row < N; ++row)
        for (int col = 0; col < N; ++col) {
            BB b = B(row, col);
            board = board | b;
            if (row == 0) row0 = row0 | b;
            if (row == N - 1) rowL = rowL | b;
            if (col == 0) col0 = col0 | b;
            if (col == N - 1) colL = colL | b;
            if (row == 0 || row == N - 1 || col == 0 || col == N - 1) edge = edge | b;
        }
    cx.board = board; cx.edge = edge; cx.row0 = row0; cx.rowL = rowL; cx.col0 = col0; cx.colL = colL;
    cx.not_col0 = board & ~col0;
    cx.not_colL = board & ~colL;
    cx.corners = N > 0 ? (B(0, 0) | B(0, N - 1) | B(N - 1, N - 1) | B(N - 1, 0)) : BB::zero();
    cx.throne = N > 0 ? B(N / 2, N / 2) : BB::zero();
    {
        BB t = cx.throne;
        BB adj = t.template shr<BB::STRIDE>() | (t.template shl<BB::STRIDE>() & board) |
                 (t.template shr<1>() & cx.not_colL) | (t.template shl<1>() & cx.not_col0);
        cx.throne_adj = adj;
    }
    const int tm = r.throne_movement;
    for (int k = 0; k < 3; ++k) {
        const bool is_king = (k == KIND_K);
        cx.sh[k] = (host_set_has(r.hostility_throne, k) ? cx.throne : BB::zero()) |
                   (host_set_has(r.hostility_corners, k) ? cx.corners : BB::zero());
        cx.edge_hostile[k] = host_set_has(r.hostility_edge, k);
        cx.pass_block[k] = (tm == TAFL_THRONE_NOPASS) || (tm == TAFL_THRONE_KINGPASS && !is_king);
        cx.noentry[k] = (tm == TAFL_THRONE_NOENTRY) || (tm == TAFL_THRONE_KINGENTRY && !is_king);
        cx.may_corner[k] = host_set_has(r.may_enter_corners, k);
        cx.slow[k] = host_set_has(r.slow_pieces, k);
        cx.sw_caps[k] = host_set_has(r.sw_captures, k);
        cx.dest_block[k] = (cx.noentry[k] ? cx.throne : BB::zero()) | (cx.may_corner[k] ? BB::zero() : cx.corners);
    }
    cx.king_hostile = (r.king_attack == TAFL_KING_ARMED || r.king_attack == TAFL_KING_ANVIL);
    cx.king_captures = (r.king_attack == TAFL_KING_ARMED || r.king_attack == TAFL_KING_HAMMER);
    cx.sbt_special = (r.king_strength == TAFL_KING_STRONG_BY_THRONE) &&
                     (tm == TAFL_THRONE_NOENTRY || tm == TAFL_THRONE_KINGENTRY);
    return cx;
}

}  // namespace tafl
