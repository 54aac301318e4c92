// tafl_mcts.cuh — device side of the batched MCTS (`class MCTS`, reference src/mcts.py:11-136):
// one search tree per resident game, ONE WARP per tree, structure-of-arrays node / edge pools.
//
//   select  (mcts.py:104-119)  warp-parallel PUCT argmax over the node's edges, f64, the exact
//                              operation order of the oracle (oracle/mcts_oracle.hpp), strict '>'
//                              first-max tie-break;
//   expand  (mcts.py:83-102)   the new child's state is produced by do_valid_move (tafl_core.cuh);
//                              its legal moves are kept as a PIECE TABLE (tafl_select.cuh: one
//                              32-bit entry per piece = tile + plays per direction, one piece per
//                              lane, one coalesced store) instead of an explicit list — the move
//                              behind an edge is decoded the one time the edge is first followed;
//   backup  (mcts.py:125-136)  walk the parent links, incremental-mean Q, sign flip per ply.
//
// Build-defined semantics (the reference has no Game/NNet adapter): tree (node =
// path), action index = position in the canonical move list, stub evaluator P = 1/L, v = 0,
// terminal value +-1 / draw_value from the point of view of the side to move.
#pragma once
#include "tafl_select.cuh"

namespace tafl {

// node header (16 B): x = edge_base, y = Ns, z = parent node (0xFFFFFFFF for the root),
// w = n_edges[0:10) | parent_edge[10:20) | expanded[20] | es[21:23) | n_pieces[23:30)
enum { ES_ONGOING = 0, ES_WIN = 1, ES_LOSS = 2, ES_DRAW = 3 };   // for the side to move at the node
#define MCTS_NO_PARENT 0xFFFFFFFFu
__host__ __device__ __forceinline__ uint32_t hdr_pack(uint32_t n_edges, uint32_t pedge, uint32_t expanded, uint32_t es,
                                                      uint32_t n_pieces = 0) {
    return (n_edges & 1023u) | ((pedge & 1023u) << 10) | (expanded << 20) | (es << 21) | ((n_pieces & 127u) << 23);
}
__host__ __device__ __forceinline__ uint32_t hdr_n_edges(uint32_t w) { return w & 1023u; }
__host__ __device__ __forceinline__ uint32_t hdr_pedge(uint32_t w) { return (w >> 10) & 1023u; }
__host__ __device__ __forceinline__ uint32_t hdr_expanded(uint32_t w) { return (w >> 20) & 1u; }
__host__ __device__ __forceinline__ uint32_t hdr_es(uint32_t w) { return (w >> 21) & 3u; }
__host__ __device__ __forceinline__ uint32_t hdr_n_pieces(uint32_t w) { return (w >> 23) & 127u; }

// game.getGameEnded(board, 1) as a 2-bit code (oracle: Mcts::game_ended)
__host__ __device__ __forceinline__ uint32_t es_code(const tafl_state& s) {
    if (s.status == TAFL_STATUS_ONGOING) return ES_ONGOING;
    if (TAFL_STATUS_IS_DRAW(s.status)) return ES_DRAW;
    const bool winner_def = TAFL_STATUS_WINNER_IS_DEFENDER(s.status);
    const bool to_play_def = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    return winner_def == to_play_def ? ES_WIN : ES_LOSS;
}

// one edge (s, a): 16 bytes, one vector load per lane in select
struct __align__(16) MctsEdge {
    double Q;         // Qsa
    uint32_t N;       // Nsa; 0 == "(s,a) not in Qsa"
    uint32_t child;   // node reached by (s,a); valid iff N > 0
};
#define MCTS_PT_MAX 64   // piece-table entries per node (pieces of the side to move)

struct MctsView {   // raw device pointers, passed to kernels by value
    tafl_state* n_state;   // [trees][node_cap]
    uint4* n_hdr;          // [trees][node_cap]
    uint32_t* n_ptab;      // [trees][node_cap][MCTS_PT_MAX]   piece table of the node's legal moves
    MctsEdge* edges;       // [trees][edge_cap]
    uint32_t* t_nodes;     // [trees] nodes in use
    uint32_t* t_edges;     // [trees] edges in use
    uint32_t* t_overflow;  // [trees] 1 = edge pool exhausted (search of that tree stopped)
    unsigned long long* stats;   // sims, nodes, depth_sum, edges, terminal_hits, max_edges_used
    double* e_prior;       // [trees][edge_cap]  Ps[s][a] (evaluator-hook searches only, else nullptr)
    uint32_t* t_pending;   // [trees] leaf waiting for its evaluation between the two phases (or MCTS_NO_PARENT)
    uint32_t node_cap, edge_cap;
};

#ifdef __CUDACC__

// n-th (0-based) set bit of a board; caller guarantees n < popc
__device__ __forceinline__ int nth_set_bit(BB64 b, int n) {
    const uint32_t w0 = (uint32_t)b.v, w1 = (uint32_t)(b.v >> 32);
    const int c0 = __popc(w0);
    return n < c0 ? nth_bit32(w0, n) : 32 + nth_bit32(w1, n - c0);
}
__device__ __forceinline__ int nth_set_bit(BB128 b, int n) {
    const int c = __popcll(b.lo);
    return n < c ? nth_set_bit(BB64{b.lo}, n) : 64 + nth_set_bit(BB64{b.hi}, n - c);
}

// Expansion by one warp: lane i takes the i-th piece of the side to move (ascending bit index),
// piece_rays gives its plays per direction, the entry goes to ptab[i].  Returns the number of legal
// moves L (same on every lane); *n_pieces_out = pieces written (> MCTS_PT_MAX: table overflow).
template <class BB, int SP>
__device__ __forceinline__ int warp_build_ptab(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side,
                                               uint32_t* __restrict__ ptab, int lane, int* n_pieces_out) {
    const BB own = p.side(def_side), occ = p.occ();
    const int n_pieces = own.popc();
    *n_pieces_out = n_pieces;
    int total = 0;
    for (int base = 0; base < n_pieces && base < MCTS_PT_MAX; base += 32) {
        const int idx = base + lane;
        uint32_t e = 0;
        if (idx < n_pieces) {
            const int s = nth_set_bit(own, idx);
            const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
            const bool is_king = def_side && BB::bit(s) == p.kbit;
            e = pt_pack(s, is_king, piece_rays(cx, occ, is_king ? KIND_K : (def_side ? KIND_DS : KIND_AS), row, col));
            ptab[idx] = e;
        }
        total += (int)__reduce_add_sync(0xFFFFFFFFu, (unsigned)pt_total(e));
    }
    return total;
}
// the move behind canonical action index a of a node with `n_pieces` table entries (a < L)
template <class BB, int SP>
__device__ __forceinline__ tafl_move warp_decode_edge(const Ctx<BB, SP>& cx, const uint32_t* __restrict__ ptab,
                                                      int n_pieces, int def_side, int a, int lane) {
    int before = 0;
    tafl_move m = 0;
    for (int base = 0; base < n_pieces; base += 32) {
        const int idx = base + lane;
        const uint32_t e = idx < n_pieces ? ptab[idx] : 0u;
        const int c = pt_total(e);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        const unsigned owns = __ballot_sync(0xFFFFFFFFu, a < before + incl);
        if (owns) {   // the first lane whose running total exceeds a holds the piece
            const int src = __ffs(owns) - 1;
            const uint32_t es = __shfl_sync(0xFFFFFFFFu, e, src);
            const int excl = __shfl_sync(0xFFFFFFFFu, before + incl - c, src);
            m = pt_decode(cx, es, def_side, a - excl);
            break;
        }
        before += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    return m;
}

// ---------------------------------------------------------------------------------------------
// The same cooperative helpers for a GROUP of G lanes (G = 32, 16, 8) owning one tree, so that one
// warp can carry 32/G trees: `gmask` = the group's lanes, `gl` = lane index inside the group.
// ---------------------------------------------------------------------------------------------
template <int G> __device__ __forceinline__ unsigned grp_ballot(unsigned gmask, int gshift, bool pred) {
    const unsigned b = __ballot_sync(gmask, pred) >> gshift;
    return G == 32 ? b : (b & ((1u << (G & 31)) - 1u));
}
template <int G, class BB, int SP>
__device__ __forceinline__ int grp_build_ptab(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side, uint32_t* __restrict__ ptab,
                                              unsigned gmask, int gl, int* n_pieces_out) {
    const BB own = p.side(def_side), occ = p.occ();
    const int n_pieces = own.popc();
    *n_pieces_out = n_pieces;
    int total = 0;
    for (int base = 0; base < n_pieces && base < MCTS_PT_MAX; base += G) {
        const int idx = base + gl;
        uint32_t e = 0;
        if (idx < n_pieces && idx < MCTS_PT_MAX) {
            const int s = nth_set_bit(own, idx);
            const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
            const bool is_king = def_side && BB::bit(s) == p.kbit;
            e = pt_pack(s, is_king, piece_rays(cx, occ, is_king ? KIND_K : (def_side ? KIND_DS : KIND_AS), row, col));
            ptab[idx] = e;
        }
        total += (int)__reduce_add_sync(gmask, (unsigned)pt_total(e));
    }
    return total;
}
template <int G, class BB, int SP>
__device__ __forceinline__ tafl_move grp_decode_edge(const Ctx<BB, SP>& cx, const uint32_t* __restrict__ ptab, int n_pieces,
                                                     int def_side, int a, unsigned gmask, int gshift, int gl) {
    int before = 0;
    tafl_move m = 0;
    for (int base = 0; base < n_pieces; base += G) {
        const int idx = base + gl;
        const uint32_t e = idx < n_pieces ? ptab[idx] : 0u;
        const int c = pt_total(e);
        int incl = c;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) {
            const int t = __shfl_up_sync(gmask, incl, o, G);
            if (gl >= o) incl += t;
        }
        const unsigned owns = grp_ballot<G>(gmask, gshift, a < before + incl);
        if (owns) {   // the first lane whose running total exceeds a holds the piece
            const int src = __ffs(owns) - 1;
            const uint32_t es = __shfl_sync(gmask, e, src, G);
            const int excl = __shfl_sync(gmask, before + incl - c, src, G);
            m = pt_decode(cx, es, def_side, a - excl);
            break;
        }
        before += __shfl_sync(gmask, incl, G - 1, G);
    }
    return m;
}

#endif  // __CUDACC__

}  // namespace tafl
