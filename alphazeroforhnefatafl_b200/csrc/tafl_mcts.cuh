// tafl_mcts.cuh — device side of the batched MCTS (`class MCTS`, reference src/mcts.py:11-136):
// one search tree per resident game, ONE WARP per tree, structure-of-arrays node / edge pools.
//
//   select  (mcts.py:104-119)  warp-parallel PUCT argmax over the node's edges, f64, the exact
//                              operation order of the oracle (oracle/mcts_oracle.hpp), strict '>'
//                              first-max tie-break;
//   expand  (mcts.py:83-102)   the new child's state is produced by do_valid_move (tafl_core.cuh),
//                              its legal moves by a warp-cooperative generator (one piece per
//                              lane, shuffle prefix sum to compact in canonical order);
//   backup  (mcts.py:125-136)  walk the parent links, incremental-mean Q, sign flip per ply.
//
// Build-defined semantics (the reference has no Game/NNet adapter): tree (node =
// path), action index = position in the canonical move list, stub evaluator P = 1/L, v = 0,
// terminal value +-1 / draw_value from the point of view of the side to move.
#pragma once
#include "tafl_core.cuh"

namespace tafl {

// node header (16 B): x = edge_base, y = Ns, z = parent node (0xFFFFFFFF for the root),
// w = n_edges[0:10) | parent_edge[10:20) | expanded[20] | es[21:23)
enum { ES_ONGOING = 0, ES_WIN = 1, ES_LOSS = 2, ES_DRAW = 3 };   // for the side to move at the node
#define MCTS_NO_PARENT 0xFFFFFFFFu
__device__ __forceinline__ uint32_t hdr_pack(uint32_t n_edges, uint32_t pedge, uint32_t expanded, uint32_t es) {
    return (n_edges & 1023u) | ((pedge & 1023u) << 10) | (expanded << 20) | (es << 21);
}
__device__ __forceinline__ uint32_t hdr_n_edges(uint32_t w) { return w & 1023u; }
__device__ __forceinline__ uint32_t hdr_pedge(uint32_t w) { return (w >> 10) & 1023u; }
__device__ __forceinline__ uint32_t hdr_expanded(uint32_t w) { return (w >> 20) & 1u; }
__device__ __forceinline__ uint32_t hdr_es(uint32_t w) { return (w >> 21) & 3u; }

// game.getGameEnded(board, 1) as a 2-bit code (oracle: Mcts::game_ended)
__host__ __device__ __forceinline__ uint32_t es_code(const tafl_state& s) {
    if (s.status == TAFL_STATUS_ONGOING) return ES_ONGOING;
    if (TAFL_STATUS_IS_DRAW(s.status)) return ES_DRAW;
    const bool winner_def = TAFL_STATUS_WINNER_IS_DEFENDER(s.status);
    const bool to_play_def = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    return winner_def == to_play_def ? ES_WIN : ES_LOSS;
}

struct MctsView {   // raw device pointers, passed to kernels by value
    tafl_state* n_state;   // [trees][node_cap]
    uint4* n_hdr;          // [trees][node_cap]
    uint32_t* e_N;         // [trees][edge_cap]   Nsa
    double* e_Q;           // [trees][edge_cap]   Qsa
    uint32_t* e_child;     // [trees][edge_cap]   valid iff Nsa > 0
    tafl_move* e_move;     // [trees][edge_cap]
    uint32_t* t_nodes;     // [trees] nodes in use
    uint32_t* t_edges;     // [trees] edges in use
    uint32_t* t_overflow;  // [trees] 1 = edge pool exhausted (search of that tree stopped)
    unsigned long long* stats;   // sims, nodes, depth_sum, edges, terminal_hits, max_edges_used
    uint32_t node_cap, edge_cap;
};

#ifdef __CUDACC__

// n-th (0-based) set bit of a board; caller guarantees n < popc
__device__ __forceinline__ int nth_set_bit(BB64 b, int n) {
    const uint32_t w0 = (uint32_t)b.v, w1 = (uint32_t)(b.v >> 32);
    const int c0 = __popc(w0);
    return n < c0 ? (int)__fns(w0, 0, n + 1) : 32 + (int)__fns(w1, 0, n - c0 + 1);
}
__device__ __forceinline__ int nth_set_bit(BB128 b, int n) {
    const int c = __popcll(b.lo);
    return n < c ? nth_set_bit(BB64{b.lo}, n) : 64 + nth_set_bit(BB64{b.hi}, n - c);
}

// get_all_possible_moves (game/main.rs:33-43) by one warp: lane i takes the i-th piece (ascending
// bit index), counts its plays with piece_rays, an exclusive shuffle scan gives its slot in the
// canonical list.  Returns the list length (same value on every lane).  `out` may hold `cap` moves.
template <class BB, int SP>
__device__ __forceinline__ int warp_gen_moves(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side, tafl_move* out, int cap,
                                              int lane) {
    const BB own = p.side(def_side), occ = p.occ();
    const int n_pieces = own.popc();
    int total = 0;
    for (int base = 0; base < n_pieces; base += 32) {
        const int idx = base + lane;
        const bool active = idx < n_pieces;
        int s = 0, row = 0, col = 0, c = 0;
        Rays r;
        r.m[0] = r.m[1] = r.m[2] = r.m[3] = 0;
        if (active) {
            s = nth_set_bit(own, idx);
            row = s / BB::STRIDE;
            col = s - row * BB::STRIDE;
            const int k = def_side ? ((BB::bit(s) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
            r = piece_rays(cx, occ, k, row, col);
            c = r.cnt[0] + r.cnt[1] + r.cnt[2] + r.cnt[3];
        }
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        int off = total + incl - c;
        total += __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (active) {
#pragma unroll
            for (int d = 0; d < 4; ++d) {
                uint32_t m = r.m[d];
                while (m) {
                    int pos;
                    if ((d & 1) == 0) { pos = __ffs(m) - 1; m &= m - 1; }
                    else { pos = 31 - __clz(m); m &= ~(1u << pos); }
                    const int to = (d < 2) ? idx_of<BB>(pos, col) : idx_of<BB>(row, pos);
                    if (off < cap) out[off] = (tafl_move)(s | (to << 7));
                    ++off;
                }
            }
        }
    }
    return total;
}

#endif  // __CUDACC__

}  // namespace tafl
