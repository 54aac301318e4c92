// tafl_mcts.cu — batched MCTS kernels (K5 select, K6 expand, K7 backup, K8 root policy / play,
// K9 pack_records) and their C ABI (include/tafl.h, "MCTS" and "Self-play records" sections).
// Reference: src/mcts.py:11-136 (`class MCTS`), restated for the CPU in oracle/mcts_oracle.hpp.
#include <cstdio>
#include <cstring>

#define TAFL_OUTLINE_RARE 1
#include "tafl_engine.hpp"
#include "tafl_mcts.cuh"
#include "tafl_select.cuh"

using namespace tafl;

namespace tafl {

struct MctsPools {
    tafl_mcts_config cfg{};
    MctsView v{};
    // root results, [n_games][TAFL_MAX_MOVES] (+ [n_games] for counts / best)
    tafl_move* r_moves = nullptr;
    uint32_t* r_visits = nullptr;
    double* r_q = nullptr;
    uint32_t* r_counts = nullptr;
    uint32_t* r_best = nullptr;
    double* r_probs = nullptr;        // [n_games][TAFL_MAX_MOVES], allocated by the first tafl_mcts_policy
    uint32_t* remap = nullptr;        // [n_games][node_cap] scratch of k_mcts_reroot (tree reuse only)
    uint32_t* d_any_overflow = nullptr;
    uint32_t* row_of = nullptr;       // [n_games + 1] compacted record row of each game (+ the row count), tafl_pack_records_compact
    bool searched = false;
    // evaluator hook (tafl_mcts_begin / select_leaves / expand_backup): the leaf batch of one round
    tafl_state* l_state = nullptr;    // [n_games]
    uint8_t* l_planes = nullptr;      // [n_games][side_len^2]  board_to_matrix
    tafl_move* l_moves = nullptr;     // [n_games][TAFL_MAX_MOVES]
    uint32_t* l_counts = nullptr;     // [n_games]
    float* l_priors = nullptr;        // [n_games][TAFL_MAX_MOVES]  evaluator output
    float* l_values = nullptr;        // [n_games]                  evaluator output
    // virtual loss through the hook (cfg.vl_width = W > 1): the leaf batch has n_games x W rows, row = game * W + simulation
    uint32_t l_width = 1;             // W the leaf buffers are sized for
    uint32_t* b_leaf = nullptr;       // [n_games][16] leaf of each simulation of the pending batch
    uint8_t* b_kind = nullptr;        // [n_games][16]
    uint32_t* b_w = nullptr;          // [n_games] simulations in the pending batch
    uint32_t* b_done = nullptr;       // [n_games] simulations finished since tafl_mcts_begin
    double* d_noise = nullptr;        // [n_games][TAFL_MAX_MOVES] root exploration noise (tafl_mcts_root_noise), eps below
    double noise_eps = 0.0;
    uint8_t* t_noised = nullptr;      // [n_games] this search's root already carries the noise
    double rg_eps = 0.0;
    cudaGraphExec_t round_graph = nullptr;   // num_sims rounds with a built-in evaluator (tafl_mcts_search_builtin)
    int rg_kind = -1, rg_spec = -1, rg_gen = -1;
    uint64_t rg_launches = 0;
    uint64_t rg_seed = 0;
    uint32_t rg_sims = 0;
    double rg_cpuct = 0, rg_draw = 0;
    int tree_mode = 0;                // 0: every edge of an expanded node is initialised (k_mcts_sims, evaluator hook);
                                      // 1: k_mcts_sims_g trees, edges written lazily up to index Ns (uniform-prior prefix)
    int phase = 0;                    // 0 idle, 1 begun / expanded, 2 leaves selected (evaluation pending)
    uint32_t rounds = 0;              // simulations run since tafl_mcts_begin
};

}  // namespace tafl

namespace {

constexpr int MCTS_WARPS = 4;                 // trees per CTA
constexpr double MCTS_EPS = 1e-8;             // mcts.py:6
enum { ST_SIMS = 0, ST_NODES, ST_DEPTH, ST_EDGES, ST_TERMINAL, ST_MAXEDGES, ST_COUNT };
// why a tree stopped (t_overflow[tree], OR-ed into the engine-wide word that the host reads)
enum { OVF_EDGES = 1u, OVF_NODES = 2u, OVF_PIECES = 4u };

// fresh tree per game, rooted at the game's current state (MCTS.__init__ + first getActionProb)
__global__ void k_mcts_init(const tafl_state* __restrict__ states, uint32_t n, MctsView mv) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const tafl_state s = load_state(states + g);
    store_state(mv.n_state + (size_t)g * mv.node_cap, s);
    mv.n_hdr[(size_t)g * mv.node_cap] = make_uint4(0u, 0u, MCTS_NO_PARENT, hdr_pack(0, 0, 0, es_code(s)));
    mv.t_nodes[g] = 1;
    mv.t_edges[g] = 0;
    mv.t_overflow[g] = 0;
    if (mv.t_pending) mv.t_pending[g] = MCTS_NO_PARENT;
    atomicAdd(&mv.stats[ST_NODES], 1ull);   // the root (oracle: Mcts::reset -> new_node)
}

// PUCT argmax over the L edges of one node by one warp (mcts.py:104-119): f64, the oracle's exact
// operation order, strict '>' per lane, then a shuffle argmax that breaks ties toward the lower
// action index = mcts.py's first-max.  PRIORS = false: uniform P = 1/L (stub evaluator);
// PRIORS = true: per-action P read from `prior` (evaluator hook).
template <bool PRIORS>
__device__ __forceinline__ void warp_puct_select(const MctsEdge* __restrict__ ed, const double* __restrict__ prior, uint32_t L,
                                                 uint32_t Ns, double cpuct, int lane, uint32_t& out_a, uint32_t& out_N,
                                                 uint32_t& out_child) {
    const double sq_vis = sqrt((double)Ns), sq_new = sqrt((double)Ns + MCTS_EPS);
    const double c_uni = cpuct * (1.0 / (double)L);          // uniform prior over the valid moves
    const double t_vis = c_uni * sq_vis;                      // numerator for visited edges
    const double u_new = c_uni * sq_new;
    double best_u = -INFINITY;
    uint32_t best_a = 0xFFFFFFFFu, best_N = 0, best_child = 0;
    // four edges per lane in flight: the loads of a group are issued before any is used
    for (uint32_t a0 = lane; a0 < L; a0 += 128) {
        MctsEdge e4[4];
        double p4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + 32u * j;
            if (a < L) {
                e4[j] = ed[a];
                if (PRIORS) p4[j] = prior[a];
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + 32u * j;
            if (a < L) {
                double u;
                if (PRIORS) {
                    const double c = cpuct * p4[j];
                    if (e4[j].N > 0) u = e4[j].Q + (c * sq_vis) / (double)(1u + e4[j].N);
                    else u = c * sq_new;
                } else {
                    if (e4[j].N > 0) u = e4[j].Q + t_vis / (double)(1u + e4[j].N);
                    else u = u_new;
                }
                if (u > best_u) { best_u = u; best_a = a; best_N = e4[j].N; best_child = e4[j].child; }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ou = __shfl_xor_sync(0xFFFFFFFFu, best_u, o);
        const uint32_t oa = __shfl_xor_sync(0xFFFFFFFFu, best_a, o);
        const uint32_t on = __shfl_xor_sync(0xFFFFFFFFu, best_N, o);
        const uint32_t oc = __shfl_xor_sync(0xFFFFFFFFu, best_child, o);
        if (ou > best_u || (ou == best_u && oa < best_a)) { best_u = ou; best_a = oa; best_N = on; best_child = oc; }
    }
    out_a = best_a; out_N = best_N; out_child = best_child;
}

// backup (mcts.py:125-136): walk the parent links from `node`, incremental-mean Q, sign flip per ply
__device__ __forceinline__ void warp_backup(uint4* __restrict__ n_hdr, MctsEdge* __restrict__ edges, uint32_t node, double v, int lane) {
    while (true) {
        const uint4 h = n_hdr[node];
        if (h.z == MCTS_NO_PARENT) break;
        const uint32_t parent = h.z, a = hdr_pedge(h.w);
        uint4 ph = n_hdr[parent];
        MctsEdge* const ep = edges + ph.x + a;
        const MctsEdge e = *ep;
        double q;
        if (e.N > 0) q = ((double)e.N * e.Q + v) / (double)(e.N + 1u);
        else q = v;
        __syncwarp();
        if (lane == 0) {
            ep->Q = q;
            ep->N = e.N + 1u;
            ph.y += 1u;   // Ns[s] += 1
            n_hdr[parent] = ph;
        }
        __syncwarp();
        v = -v;
        node = parent;
    }
}

// `sims` x MCTS.search (mcts.py:55-136) for every tree; warp w of the grid owns tree w.
template <class BB, int SP, int MINB>
__global__ void __launch_bounds__(MCTS_WARPS * 32, MINB)
k_mcts_sims(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, uint32_t sims, double cpuct, double draw_value,
            uint32_t* __restrict__ any_overflow) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    uint32_t n_nodes = mv.t_nodes[tree], n_edges_used = mv.t_edges[tree];
    uint32_t overflow = mv.t_overflow[tree];
    unsigned long long c_sims = 0, c_nodes = 0, c_depth = 0, c_edges = 0, c_term = 0;

    for (uint32_t sim = 0; sim < sims && !overflow; ++sim) {
        uint32_t node = 0, depth = 0;
        double v;   // value returned by search(node) to its caller
        // ---- descend ---------------------------------------------------------------------------
        while (true) {
            uint4 h = n_hdr[node];
            const uint32_t es = hdr_es(h.w);
            if (es != ES_ONGOING) {   // mcts.py:77-81  terminal: return -Es[s]
                const double Es = es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value);
                v = -Es;
                c_term += 1;
                c_depth += depth;
                break;
            }
            if (!hdr_expanded(h.w)) {   // mcts.py:83-102  leaf: expand, Ns = 0, return -v with v = 0
                const tafl_state s = load_state(n_state + node);
                const Pos<BB> p = load_pos(cx, s);
                int n_pieces;
                const int L = warp_build_ptab(cx, p, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0,
                                              n_ptab + (size_t)node * MCTS_PT_MAX, lane, &n_pieces);
                if ((uint32_t)L > mv.edge_cap - n_edges_used || n_pieces > MCTS_PT_MAX) {
                    overflow = n_pieces > MCTS_PT_MAX ? OVF_PIECES : OVF_EDGES;
                    break;
                }
                MctsEdge z;
                z.Q = 0.0; z.N = 0u; z.child = 0u;   // Nsa absent == 0
                for (int a = lane; a < L; a += 32) edges[n_edges_used + a] = z;
                if (lane == 0)
                    n_hdr[node] = make_uint4(n_edges_used, 0u, h.z, hdr_pack((uint32_t)L, hdr_pedge(h.w), 1u, es, (uint32_t)n_pieces));
                n_edges_used += (uint32_t)L;
                c_edges += (unsigned long long)L;
                c_depth += depth;
                __syncwarp();
                v = -0.0;
                break;
            }
            // ---- select (mcts.py:104-119) --------------------------------------------------------
            const uint32_t L = hdr_n_edges(h.w), base = h.x, Ns = h.y;
            if (L == 0) { v = 0.0; break; }   // Ongoing node without plays: oracle returns 0.0 unnegated
            uint32_t best_a, best_N, best_child;
            warp_puct_select<false>(edges + base, nullptr, L, Ns, cpuct, lane, best_a, best_N, best_child);
            const uint32_t a = best_a;
            uint32_t child;
            if (best_N == 0) {   // first traversal: getNextState (mcts.py:122-123)
                tafl_state s = load_state(n_state + node);
                const tafl_move m = warp_decode_edge(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                                     (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)a, lane);
                if (n_nodes >= mv.node_cap) { overflow = OVF_NODES; break; }   // only reachable with tree reuse
                do_valid_move(cx, s, m);
                child = n_nodes++;
                c_nodes += 1;
                if (lane == 0) {
                    store_state(n_state + child, s);
                    n_hdr[child] = make_uint4(0u, 0u, node, hdr_pack(0, a, 0, es_code(s)));
                    edges[base + a].child = child;
                }
                __syncwarp();
            } else {
                child = best_child;
            }
            node = child;
            depth += 1;
        }
        if (overflow) break;
        // ---- backup (mcts.py:125-136) ----------------------------------------------------------------
        warp_backup(n_hdr, edges, node, v, lane);
        c_sims += 1;
    }
    if (lane == 0) {
        mv.t_nodes[tree] = n_nodes;
        mv.t_edges[tree] = n_edges_used;
        if (overflow) { mv.t_overflow[tree] = overflow; atomicOr(any_overflow, overflow); }
        atomicAdd(&mv.stats[ST_SIMS], c_sims);
        atomicAdd(&mv.stats[ST_NODES], c_nodes);
        atomicAdd(&mv.stats[ST_DEPTH], c_depth);
        atomicAdd(&mv.stats[ST_EDGES], c_edges);
        atomicAdd(&mv.stats[ST_TERMINAL], c_term);
        atomicMax(&mv.stats[ST_MAXEDGES], (unsigned long long)n_edges_used);
    }
}

// ---------------------------------------------------------------------------------------------
// k_mcts_sims_g: the same search with G lanes per tree (32/G trees per warp).  Everything up to the
// choice of the edge to open is done per group; the expensive part of a simulation — do_valid_move
// for the new child and the piece table of its expansion, identical straight-line code for every
// tree — sits behind one full-warp convergence point, so it is issued ONCE for all trees of the
// warp instead of once per tree.  The child is created and expanded from registers (no store ->
// load round trip through its header and state).  Results are bit-identical to k_mcts_sims.
// ---------------------------------------------------------------------------------------------
template <int G>
__device__ __forceinline__ void grp_puct_select(const MctsEdge* __restrict__ ed, uint32_t Lall, uint32_t Ns, double cpuct, unsigned gmask,
                                                int gl, uint32_t& out_a, uint32_t& out_N, uint32_t& out_child) {
    const double sq_vis = sqrt((double)Ns), sq_new = sqrt((double)Ns + MCTS_EPS);
    const double c_uni = cpuct * (1.0 / (double)Lall);
    const double t_vis = c_uni * sq_vis, u_new = c_uni * sq_new;
    // Under uniform priors every unvisited edge has the same u, and the first maximum takes the
    // lowest index: edges are opened in index order, the visited ones are a prefix of length
    // k <= Ns.  Edges behind index Ns cannot win against edge k, so only [0, min(L, Ns + 1)) is read.
    const uint32_t L = Lall < Ns + 1u ? Lall : Ns + 1u;
    double best_u = -INFINITY;
    uint32_t best_a = 0xFFFFFFFFu, best_N = 0, best_child = 0;
    for (uint32_t a0 = gl; a0 < L; a0 += 4 * G) {
        MctsEdge e4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) e4[j] = ed[a];
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) {
                double u;
                if (e4[j].N > 0) u = e4[j].Q + t_vis / (double)(1u + e4[j].N);
                else u = u_new;
                if (u > best_u) { best_u = u; best_a = a; best_N = e4[j].N; best_child = e4[j].child; }
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        const double ou = __shfl_xor_sync(gmask, best_u, o, G);
        const uint32_t oa = __shfl_xor_sync(gmask, best_a, o, G);
        const uint32_t on = __shfl_xor_sync(gmask, best_N, o, G);
        const uint32_t oc = __shfl_xor_sync(gmask, best_child, o, G);
        if (ou > best_u || (ou == best_u && oa < best_a)) { best_u = ou; best_a = oa; best_N = on; best_child = oc; }
    }
    out_a = best_a; out_N = best_N; out_child = best_child;
}

template <bool LAZY = true>
__device__ __forceinline__ void grp_backup(uint4* __restrict__ n_hdr, MctsEdge* __restrict__ edges, uint32_t node, double v,
                                           unsigned gmask, int gl) {
    while (true) {
        const uint4 h = n_hdr[node];
        if (h.z == MCTS_NO_PARENT) break;
        const uint32_t parent = h.z, a = hdr_pedge(h.w);
        uint4 ph = n_hdr[parent];
        MctsEdge* const ep = edges + ph.x + a;
        const MctsEdge e = *ep;
        double q;
        if (e.N > 0) q = ((double)e.N * e.Q + v) / (double)(e.N + 1u);
        else q = v;
        __syncwarp(gmask);
        if (gl == 0) {
            ep->Q = q;
            ep->N = e.N + 1u;
            ph.y += 1u;
            n_hdr[parent] = ph;
            // edges are zeroed lazily: select reads [0, min(L, Ns + 1)), so edge index Ns becomes
            // readable now (it is unvisited: the visited prefix is never longer than Ns).  Trees with
            // per-action priors (evaluator hook) initialise every edge at expansion: LAZY = false.
            if (LAZY && ph.y < hdr_n_edges(ph.w)) {
                MctsEdge z;
                z.Q = 0.0; z.N = 0u; z.child = 0u;
                edges[ph.x + ph.y] = z;
            }
        }
        __syncwarp(gmask);
        v = -v;
        node = parent;
    }
}

#if !defined(TAFL_SIMS_WARPS)
#define TAFL_SIMS_WARPS 4
#endif
constexpr int SIMS_WARPS = TAFL_SIMS_WARPS;   // warps per CTA of k_mcts_sims_g (they enter the create step together): 4 -> 4.09, 2 -> 3.92, 8 -> 3.84 x 10^8 sims/s
template <class BB, int SP, int G, int MINB>
__global__ void __launch_bounds__(SIMS_WARPS * 32, (MINB * 4 / SIMS_WARPS > 0 ? MINB * 4 / SIMS_WARPS : 1))
k_mcts_sims_g(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, uint32_t sims, double cpuct, double draw_value,
              uint32_t* __restrict__ any_overflow) {
    constexpr int TPW = 32 / G;
    enum { M_NONE = 0, M_BACKUP, M_CREATE, M_EXPAND };
    const int lane = threadIdx.x & 31, gl = lane % G, gshift = (lane / G) * G;
    const unsigned gmask = G == 32 ? 0xFFFFFFFFu : (((1u << (G & 31)) - 1u) << gshift);
    const uint32_t tree_raw = (blockIdx.x * SIMS_WARPS + (threadIdx.x >> 5)) * TPW + (uint32_t)(lane / G);
    const bool live = tree_raw < n_trees;
    const uint32_t tree = live ? tree_raw : 0u;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    uint32_t n_nodes = mv.t_nodes[tree], n_edges_used = mv.t_edges[tree];
    uint32_t overflow = mv.t_overflow[tree];
    bool stop = !live || overflow != 0;   // a stopped group only keeps the warp-wide convergence points
    unsigned long long c_sims = 0, c_nodes = 0, c_depth = 0, c_edges = 0, c_term = 0;

    for (uint32_t sim = 0; sim < sims; ++sim) {
#if defined(TAFL_MCTS_PHASE_SYNC)   // A/B: keep the warps of a CTA in phase (instruction-cache locality) at the price of waiting for the slowest
        if ((sim % TAFL_MCTS_PHASE_SYNC) == 0) __syncthreads();
#endif
        int mode = M_NONE;
        uint32_t node = 0, depth = 0, sel_a = 0, sel_base = 0;
        uint4 h = make_uint4(0u, 0u, 0u, 0u);
        double v = 0.0;
        if (!stop) {
            while (true) {   // ---- descend to the edge to open (mcts.py:104-123) or to a terminal node
                h = n_hdr[node];
                const uint32_t es = hdr_es(h.w);
                if (es != ES_ONGOING) {   // mcts.py:77-81
                    v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                    c_term += 1;
                    c_depth += depth;
                    mode = M_BACKUP;
                    break;
                }
                if (!hdr_expanded(h.w)) { mode = M_EXPAND; break; }   // the root of a fresh tree
                const uint32_t L = hdr_n_edges(h.w);
                if (L == 0) { v = 0.0; mode = M_BACKUP; break; }
                // the node's state and piece table are needed only if the selected edge is new — most
                // simulations end that way one or two levels down — so they are prefetched into L1 here,
                // under the edge loads and the PUCT arithmetic (registers are at the 128 cap that keeps
                // all 2048 warps resident; a register prefetch cost a second wave)
                if (gl < 3) {   // one 64-byte state line, two 128-byte piece-table lines
                    const void* pf = gl == 0 ? (const void*)(n_state + node)
                                             : (const void*)(n_ptab + (size_t)node * MCTS_PT_MAX + (gl - 1) * 32);
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
                }
                uint32_t a, bN, bchild;
                grp_puct_select<G>(edges + h.x, L, h.y, cpuct, gmask, gl, a, bN, bchild);
                if (bN == 0) { mode = M_CREATE; sel_a = a; sel_base = h.x; break; }
                node = bchild;
                depth += 1;
            }
        }
        // Every tree of the CTA opens its new node together: the create step is the longest stretch of straight-line code
        // in the kernel, and four warps that enter it in phase fetch it once (3.92 -> 4.09 x 10^8 sims/s; a barrier at the
        // top of the simulation instead gives 4.02, both 4.05).  Safe: every warp runs all `sims` iterations.
#if defined(TAFL_MCTS_NO_PHASE_SYNC)
        __syncwarp();
#else
        __syncthreads();
#endif
        if (mode == M_CREATE || mode == M_EXPAND) {
            tafl_state s = load_state(n_state + node);
            uint32_t target = node, parent = h.z, pedge = hdr_pedge(h.w), es = ES_ONGOING;
            bool expand = true;
            if (mode == M_CREATE) {   // getNextState (mcts.py:122-123)
                if (n_nodes >= mv.node_cap) {   // only reachable with tree reuse
                    overflow = OVF_NODES;
                    expand = false;
                } else {
                    const tafl_move m = grp_decode_edge<G>(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                                           (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)sel_a, gmask, gshift, gl);
                    do_valid_move(cx, s, m);
                    target = n_nodes++;
                    c_nodes += 1;
                    parent = node;
                    pedge = sel_a;
                    depth += 1;
                    es = es_code(s);
                    if (gl == 0) {
                        store_state(n_state + target, s);
                        edges[sel_base + sel_a].child = target;
                    }
                    if (es != ES_ONGOING) {   // the new node is terminal (mcts.py:77-81)
                        if (gl == 0) n_hdr[target] = make_uint4(0u, 0u, parent, hdr_pack(0, pedge, 0, es));
                        v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                        c_term += 1;
                        c_depth += depth;
                        expand = false;
                    }
                }
            }
            if (expand) {   // mcts.py:83-102: leaf, Ns = 0, return -v with v = 0
                const Pos<BB> p = load_pos(cx, s);
                int n_pieces;
                const int L = grp_build_ptab<G>(cx, p, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, n_ptab + (size_t)target * MCTS_PT_MAX, gmask,
                                                gl, &n_pieces);
                if ((uint32_t)L > mv.edge_cap - n_edges_used || n_pieces > MCTS_PT_MAX) {
                    overflow = n_pieces > MCTS_PT_MAX ? OVF_PIECES : OVF_EDGES;
                    if (mode == M_CREATE && gl == 0) n_hdr[target] = make_uint4(0u, 0u, parent, hdr_pack(0, pedge, 0, es));
                } else {
                    if (gl == 0) {
                        MctsEdge z;
                        z.Q = 0.0; z.N = 0u; z.child = 0u;
                        if (L > 0) edges[n_edges_used] = z;   // Ns = 0: edge 0 is the only readable one (see grp_backup)
                        n_hdr[target] = make_uint4(n_edges_used, 0u, parent, hdr_pack((uint32_t)L, pedge, 1u, es, (uint32_t)n_pieces));
                    }
                    n_edges_used += (uint32_t)L;
                    c_edges += (unsigned long long)L;
                    c_depth += depth;
                    v = -0.0;
                }
            }
            node = target;
            __syncwarp(gmask);
        }
        if (overflow) stop = true;
        if (!stop && mode != M_NONE) {   // ---- backup (mcts.py:125-136)
            grp_backup(n_hdr, edges, node, v, gmask, gl);
            c_sims += 1;
        }
    }
    if (gl == 0 && live) {
        mv.t_nodes[tree] = n_nodes;
        mv.t_edges[tree] = n_edges_used;
        if (overflow) { mv.t_overflow[tree] = overflow; atomicOr(any_overflow, overflow); }
        atomicAdd(&mv.stats[ST_SIMS], c_sims);
        atomicAdd(&mv.stats[ST_NODES], c_nodes);
        atomicAdd(&mv.stats[ST_DEPTH], c_depth);
        atomicAdd(&mv.stats[ST_EDGES], c_edges);
        atomicAdd(&mv.stats[ST_TERMINAL], c_term);
        atomicMax(&mv.stats[ST_MAXEDGES], (unsigned long long)n_edges_used);
    }
}

// ---------------------------------------------------------------------------------------------
// k_mcts_sims_vl: the fused stub search with VIRTUAL LOSS — up to W <= G simulations in flight per tree
// (north star: "select (PUCT with virtual loss)"; intent at src/mcts.rs:49-55; semantics defined and
// pinned by oracle/mcts_oracle.hpp::run_from_root_vl; W = 1 is mcts.py's sequential search).
// One batch of w simulations of a tree, by its group of G lanes:
//   select   the w descents run one after the other (each sees the virtual losses of the earlier ones);
//            lane i keeps the outcome of simulation i: the leaf, the edge to open and its decoded move;
//   create   do_valid_move for ALL new children of the batch at once, one lane per child — the
//            straight-line part that k_mcts_sims_g issues once per simulation with every lane of the
//            group computing the same child is issued once per batch with every lane computing its own;
//   expand   piece tables of the new children, group-cooperative, one child after the other;
//   backup   in simulation order: virtual losses removed, real update (mcts.py:125-136).
// The virtual-loss counts live in the top byte of the words they qualify: edge.N = Nsa | vl << 24,
// node header .y = Ns | vl(node) << 24 (a search leaves them zero).  The uniform-prior prefix rule of
// k_mcts_sims_g carries over with Ns_eff = Ns + vl(node): edges are first touched in index order, so
// select reads [0, min(L, Ns_eff + 1)) and the edge that becomes readable is zeroed when Ns_eff grows.
// ---------------------------------------------------------------------------------------------
#define MCTS_VL_ONE (1u << 24)
#define MCTS_VL_MASK 0x00FFFFFFu
template <int G>
__device__ __forceinline__ void grp_puct_select_vl(const MctsEdge* __restrict__ ed, uint32_t Lall, uint32_t ns_raw, double cpuct,
                                                   unsigned gmask, int gl, uint32_t& out_a, uint32_t& out_Nraw, uint32_t& out_child) {
    const uint32_t ns_eff = (ns_raw & MCTS_VL_MASK) + (ns_raw >> 24);
    const double sq_vis = sqrt((double)ns_eff), sq_new = sqrt((double)ns_eff + MCTS_EPS);
    const double c_uni = cpuct * (1.0 / (double)Lall);
    const double t_vis = c_uni * sq_vis, u_new = c_uni * sq_new;
    const uint32_t L = Lall < ns_eff + 1u ? Lall : ns_eff + 1u;
    double best_u = -INFINITY;
    uint32_t best_a = 0xFFFFFFFFu, best_N = 0, best_child = 0;
    for (uint32_t a0 = gl; a0 < L; a0 += 4 * G) {
        MctsEdge e4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) e4[j] = ed[a];
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) {
                const uint32_t N = e4[j].N & MCTS_VL_MASK, vl = e4[j].N >> 24, nn = N + vl;
                double u;
                if (nn > 0) {
                    const double qv = vl ? ((double)N * e4[j].Q - (double)vl) / (double)nn : e4[j].Q;
                    u = qv + t_vis / (double)(1u + nn);
                } else {
                    u = u_new;
                }
                if (u > best_u) { best_u = u; best_a = a; best_N = e4[j].N; best_child = e4[j].child; }
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        const double ou = __shfl_xor_sync(gmask, best_u, o, G);
        const uint32_t oa = __shfl_xor_sync(gmask, best_a, o, G);
        const uint32_t on = __shfl_xor_sync(gmask, best_N, o, G);
        const uint32_t oc = __shfl_xor_sync(gmask, best_child, o, G);
        if (ou > best_u || (ou == best_u && oa < best_a)) { best_u = ou; best_a = oa; best_N = on; best_child = oc; }
    }
    out_a = best_a; out_Nraw = best_N; out_child = best_child;
}

// backup of one simulation of a batch: its virtual losses come off, the real update goes on
__device__ __forceinline__ void grp_backup_vl(uint4* __restrict__ n_hdr, MctsEdge* __restrict__ edges, uint32_t node, double v,
                                              unsigned gmask, int gl) {
    while (true) {
        const uint4 h = n_hdr[node];
        if (h.z == MCTS_NO_PARENT) break;
        const uint32_t parent = h.z, a = hdr_pedge(h.w);
        uint4 ph = n_hdr[parent];
        MctsEdge* const ep = edges + ph.x + a;
        const MctsEdge e = *ep;
        const uint32_t N = e.N & MCTS_VL_MASK, vl = e.N >> 24;
        double q;
        if (N > 0) q = ((double)N * e.Q + v) / (double)(N + 1u);
        else q = v;
        __syncwarp(gmask);
        if (gl == 0) {
            ep->Q = q;
            ep->N = (N + 1u) | ((vl - 1u) << 24);
            ph.y = ((ph.y & MCTS_VL_MASK) + 1u) | (((ph.y >> 24) - 1u) << 24);   // Ns += 1, one virtual loss less: Ns_eff unchanged
            n_hdr[parent] = ph;
        }
        __syncwarp(gmask);
        v = -v;
        node = parent;
    }
}

template <class BB, int SP, int G>
__global__ void __launch_bounds__(MCTS_WARPS * 32, 4)
k_mcts_sims_vl(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, uint32_t sims, uint32_t W, double cpuct,
               double draw_value, uint32_t* __restrict__ any_overflow) {
    constexpr int TPW = 32 / G;
    enum { K_NONE = 0, K_TERM, K_ROOT, K_EMPTY, K_NEW, K_COLL };
    const int lane = threadIdx.x & 31, gl = lane % G, gshift = (lane / G) * G;
    const unsigned gmask = G == 32 ? 0xFFFFFFFFu : (((1u << (G & 31)) - 1u) << gshift);
    const uint32_t tree_raw = (blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5)) * TPW + (uint32_t)(lane / G);
    const bool live = tree_raw < n_trees;
    const uint32_t tree = live ? tree_raw : 0u;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    uint32_t n_nodes = mv.t_nodes[tree], n_edges_used = mv.t_edges[tree];
    uint32_t overflow = mv.t_overflow[tree];
    bool stop = !live || overflow != 0;
    unsigned long long c_sims = 0, c_nodes = 0, c_depth = 0, c_edges = 0, c_term = 0;
    if (W > (uint32_t)G) W = (uint32_t)G;
    if (W < 1u) W = 1u;
    uint32_t done = 0;
    // every group of the warp takes part in the warp-wide convergence points until the slowest one is through
    while (__any_sync(0xFFFFFFFFu, !stop && done < sims)) {
        const bool active = !stop && done < sims;
        uint32_t w = 0;
        if (active) {
            const uint4 h0 = n_hdr[0];
            w = sims - done < W ? sims - done : W;
            if (!hdr_expanded(h0.w) && hdr_es(h0.w) == ES_ONGOING) w = 1;   // one simulation per batch while the root is unexpanded
        }
        // ---- select: simulation i of the batch is remembered by lane i of the group
        int my_kind = K_NONE;
        uint32_t my_leaf = 0, my_parent = 0, my_pedge = 0, my_depth = 0;
        tafl_move my_move = 0;
        double my_v = 0.0;
        for (uint32_t i = 0; i < w; ++i) {
            int kind = K_NONE;
            uint32_t node = 0, depth = 0, leaf = 0, parent = 0, pedge = 0;
            tafl_move mvv = 0;
            double v = 0.0;
            while (true) {
                uint4 h = n_hdr[node];
                const uint32_t es = hdr_es(h.w);
                if (es != ES_ONGOING) {   // mcts.py:77-81
                    v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                    kind = K_TERM; leaf = node;
                    break;
                }
                if (!hdr_expanded(h.w)) { kind = K_ROOT; leaf = node; break; }
                const uint32_t L = hdr_n_edges(h.w);
                if (L == 0) { kind = K_EMPTY; leaf = node; v = 0.0; break; }
                uint32_t a, bNraw, bchild;
                grp_puct_select_vl<G>(edges + h.x, L, h.y, cpuct, gmask, gl, a, bNraw, bchild);
                depth += 1;
                const bool fresh = (bNraw & MCTS_VL_MASK) == 0;        // (s,a) not in Qsa
                const bool collide = fresh && (bNraw >> 24) != 0;      // ... and an earlier simulation of this batch is opening it
                uint32_t child = bchild;
                if (fresh && !collide) {
                    if (n_nodes >= mv.node_cap) { overflow = OVF_NODES; break; }   // only reachable with tree reuse
                    child = n_nodes++;
                    c_nodes += 1;
                    const tafl_state ps = load_state(n_state + node);
                    mvv = grp_decode_edge<G>(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                             (ps.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)a, gmask, gshift, gl);
                }
                __syncwarp(gmask);
                if (gl == 0) {   // the virtual loss of this simulation on (s,a) and on s
                    MctsEdge* const ep = edges + h.x + a;
                    ep->N = bNraw + MCTS_VL_ONE;
                    if (fresh && !collide) ep->child = child;
                    const uint32_t ns_eff = (h.y & MCTS_VL_MASK) + (h.y >> 24) + 1u;
                    reinterpret_cast<uint32_t*>(n_hdr + node)[1] = h.y + MCTS_VL_ONE;
                    if (ns_eff < L) {   // the edge that becomes readable (untouched so far)
                        MctsEdge z;
                        z.Q = 0.0; z.N = 0u; z.child = 0u;
                        edges[h.x + ns_eff] = z;
                    }
                }
                __syncwarp(gmask);
                if (fresh) {
                    kind = collide ? K_COLL : K_NEW;
                    leaf = child; parent = node; pedge = a;
                    break;
                }
                node = child;
            }
            if (overflow) break;
            if (gl == (int)i) { my_kind = kind; my_leaf = leaf; my_parent = parent; my_pedge = pedge; my_depth = depth; my_move = mvv; my_v = v; }
        }
        if (overflow) { stop = true; w = 0; my_kind = K_NONE; }
        __syncwarp();   // ---- create: every new child of every tree of the warp, one lane each
        if (my_kind == K_NEW) {
            tafl_state s = load_state(n_state + my_parent);
            do_valid_move(cx, s, my_move);
            store_state(n_state + my_leaf, s);
            n_hdr[my_leaf] = make_uint4(0u, 0u, my_parent, hdr_pack(0, my_pedge, 0, es_code(s)));
        }
        __syncwarp();
        // ---- expand (mcts.py:83-102, stub evaluator), one leaf after the other, group-cooperative
        for (uint32_t i = 0; i < w; ++i) {
            const int kind = __shfl_sync(gmask, my_kind, (int)i, G);
            const uint32_t leaf = __shfl_sync(gmask, my_leaf, (int)i, G);
            const uint32_t depth = __shfl_sync(gmask, my_depth, (int)i, G);
            if (kind == K_NONE || kind == K_EMPTY) continue;
            c_depth += depth;
            if (kind == K_TERM) { c_term += 1; continue; }
            const uint4 h = n_hdr[leaf];
            const uint32_t es = hdr_es(h.w);
            double v = -0.0;
            if (es != ES_ONGOING) {   // the new (or collided) child is terminal
                v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                c_term += 1;
            } else if (kind != K_COLL) {
                const tafl_state s = load_state(n_state + leaf);
                const Pos<BB> p = load_pos(cx, s);
                int n_pieces;
                const int L = grp_build_ptab<G>(cx, p, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, n_ptab + (size_t)leaf * MCTS_PT_MAX, gmask,
                                                gl, &n_pieces);
                if ((uint32_t)L > mv.edge_cap - n_edges_used || n_pieces > MCTS_PT_MAX) {
                    overflow = n_pieces > MCTS_PT_MAX ? OVF_PIECES : OVF_EDGES;
                    break;
                }
                if (gl == 0) {
                    MctsEdge z;
                    z.Q = 0.0; z.N = 0u; z.child = 0u;
                    if (L > 0) edges[n_edges_used] = z;   // Ns_eff = 0: edge 0 is the only readable one
                    n_hdr[leaf] = make_uint4(n_edges_used, 0u, h.z, hdr_pack((uint32_t)L, hdr_pedge(h.w), 1u, es, (uint32_t)n_pieces));
                }
                n_edges_used += (uint32_t)L;
                c_edges += (unsigned long long)L;
            }
            if (gl == (int)i) my_v = v;
        }
        if (overflow) { stop = true; w = 0; }
        __syncwarp(gmask);
        // ---- backup, in simulation order
        for (uint32_t i = 0; i < w; ++i) {
            const int kind = __shfl_sync(gmask, my_kind, (int)i, G);
            const uint32_t leaf = __shfl_sync(gmask, my_leaf, (int)i, G);
            const double v = __shfl_sync(gmask, my_v, (int)i, G);
            if (kind == K_NONE) continue;
            grp_backup_vl(n_hdr, edges, leaf, v, gmask, gl);
            c_sims += 1;
        }
        done += w;
        if (active && w == 0) stop = true;
    }
    if (gl == 0 && live) {
        mv.t_nodes[tree] = n_nodes;
        mv.t_edges[tree] = n_edges_used;
        if (overflow) { mv.t_overflow[tree] = overflow; atomicOr(any_overflow, overflow); }
        atomicAdd(&mv.stats[ST_SIMS], c_sims);
        atomicAdd(&mv.stats[ST_NODES], c_nodes);
        atomicAdd(&mv.stats[ST_DEPTH], c_depth);
        atomicAdd(&mv.stats[ST_EDGES], c_edges);
        atomicAdd(&mv.stats[ST_TERMINAL], c_term);
        atomicMax(&mv.stats[ST_MAXEDGES], (unsigned long long)n_edges_used);
    }
}

// MCTS.getActionProb (mcts.py:40-47), temp = 0 with first-max: root visit counts, Q, moves, argmax
template <class BB, int SP>
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_results(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, tafl_move* __restrict__ r_moves,
               uint32_t* __restrict__ r_visits, double* __restrict__ r_q, uint32_t* __restrict__ r_counts,
               uint32_t* __restrict__ r_best, int prefix) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    const uint4 h = mv.n_hdr[(size_t)tree * mv.node_cap];
    const uint32_t L = hdr_expanded(h.w) ? hdr_n_edges(h.w) : 0u, base = h.x;
    const int n_pieces = (int)hdr_n_pieces(h.w);
    const MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap + base;
    const uint32_t* const ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;   // root = node 0
    const tafl_state root = load_state(mv.n_state + (size_t)tree * mv.node_cap);
    const int def_side = (root.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    const size_t ro = (size_t)tree * TAFL_MAX_MOVES;
    uint32_t best_n = 0, best_a = 0xFFFFFFFFu;
    for (uint32_t a0 = 0; a0 < TAFL_MAX_MOVES; a0 += 32) {
        const uint32_t a = a0 + lane;
        uint32_t N = 0;
        double q = 0.0;
        tafl_move m = 0;
        if (a0 < L) {   // warp-uniform: the decode is a warp-wide operation
            // every lane decodes its own edge: run the warp-cooperative decode once per lane index
            for (int src = 0; src < 32; ++src) {
                const int aa = (int)(a0 + src);
                if ((uint32_t)aa >= L) break;
                const tafl_move mm = warp_decode_edge(cx, ptab, n_pieces, def_side, aa, lane);
                if (src == lane) m = mm;
            }
            if (a < L && (!prefix || a <= h.y)) {   // prefix trees: edges behind index Ns were never written
                const MctsEdge e = edges[a];
                N = e.N;
                if (N > 0) q = e.Q;
            }
        }
        r_moves[ro + a] = m;
        r_visits[ro + a] = N;
        r_q[ro + a] = q;
        if (N > best_n) { best_n = N; best_a = a; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t on = __shfl_xor_sync(0xFFFFFFFFu, best_n, o), oa = __shfl_xor_sync(0xFFFFFFFFu, best_a, o);
        if (on > best_n || (on == best_n && oa < best_a)) { best_n = on; best_a = oa; }
    }
    if (lane == 0) {
        r_counts[tree] = L;
        r_best[tree] = best_n > 0 ? best_a : 0u;
    }
}

// MCTS.getActionProb's policy (mcts.py:40-53) and the move choice made from it, one warp per tree.
//   temp == 0 : one-hot on the FIRST maximum (mcts.py:43-48 without the random tie choice)
//   temp == 1 : probs = N / sum(N) — exact (x ** 1.0 == float(x), integer sum, one division)
//   else      : probs = N^(1/temp) / sum(N^(1/temp))  (CUDA pow: within 1e-12 relative of CPython's)
// The chosen action (what a Coach would draw with np.random.choice(p=probs); not in the reference,
// build-defined): temp == 0 -> first max; temp == 1 -> the visit-proportional draw
// k = pick_index(seed, game_id, ply, sum N), first a with cumsum(N) > k (integer, reproducible on
// the host); else the same inverse-CDF walk over the f64 weights with u = 53 random bits.
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_policy(MctsView mv, uint32_t n_trees, uint64_t gid_base, const uint32_t* __restrict__ r_visits,
              const uint32_t* __restrict__ r_counts, double temp, uint64_t seed, int choose, double* __restrict__ r_probs,
              uint32_t* __restrict__ r_best) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    const uint32_t L = r_counts[tree];
    const size_t ro = (size_t)tree * TAFL_MAX_MOVES;
    // first maximum of the visit counts (r_best may hold an earlier draw)
    uint32_t bn = 0, best = 0xFFFFFFFFu;
    for (uint32_t a = lane; a < L; a += 32) {
        const uint32_t N = r_visits[ro + a];
        if (N > bn) { bn = N; best = a; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t on = __shfl_xor_sync(0xFFFFFFFFu, bn, o), oa = __shfl_xor_sync(0xFFFFFFFFu, best, o);
        if (on > bn || (on == bn && oa < best)) { bn = on; best = oa; }
    }
    if (bn == 0) best = 0u;
    if (temp == 0.0) {
        if (r_probs) for (uint32_t a = lane; a < TAFL_MAX_MOVES; a += 32) r_probs[ro + a] = (a < L && a == best) ? 1.0 : 0.0;
        if (choose && lane == 0) r_best[tree] = best;
        return;
    }
    const bool t1 = temp == 1.0;
    const double inv_t = 1.0 / temp;
    // per-lane weights of the 16 actions a = 32*i + lane
    double w[TAFL_MAX_MOVES / 32];
    double part = 0.0;
    unsigned long long ipart = 0;
#pragma unroll
    for (int i = 0; i < TAFL_MAX_MOVES / 32; ++i) {
        const uint32_t a = 32u * i + lane;
        const uint32_t N = a < L ? r_visits[ro + a] : 0u;
        w[i] = t1 ? (double)N : (N ? pow((double)N, inv_t) : 0.0);
        part += w[i];
        ipart += N;
    }
    double sum = part;
    unsigned long long isum = ipart;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
        isum += __shfl_xor_sync(0xFFFFFFFFu, isum, o);
    }
    if (t1) sum = (double)isum;
    if (r_probs) {
#pragma unroll
        for (int i = 0; i < TAFL_MAX_MOVES / 32; ++i) r_probs[ro + 32u * i + lane] = sum > 0 ? w[i] / sum : 0.0;
    }
    if (!choose || L == 0 || !(sum > 0)) return;
    const tafl_state root = load_state(mv.n_state + (size_t)tree * mv.node_cap);
    const uint64_t x = splitmix64(seed ^ ((gid_base + tree) * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)root.turn * 0xBF58476D1CE4E5B9ull));
    const unsigned long long k = mulhi64(x, isum);                                  // temp == 1
    const double u = (double)(x >> 11) * 1.1102230246251565e-16 * sum;              // other temperatures
    // inverse CDF in action order: rows of 32 consecutive actions, inclusive scan across the lanes
    unsigned long long ibefore = 0;
    double before = 0.0;
    uint32_t pick = best;
    for (int i = 0; i < TAFL_MAX_MOVES / 32; ++i) {
        if (32u * i >= L) break;
        const uint32_t a = 32u * i + lane;
        unsigned long long iinc = a < L ? r_visits[ro + a] : 0u;
        double inc = w[i];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long it = __shfl_up_sync(0xFFFFFFFFu, iinc, o);
            const double ft = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) { iinc += it; inc += ft; }
        }
        const bool hit = t1 ? (ibefore + iinc > k) : (before + inc > u);
        const unsigned m = __ballot_sync(0xFFFFFFFFu, hit && a < L);
        if (m) { pick = 32u * i + (uint32_t)(__ffs(m) - 1); break; }
        ibefore += __shfl_sync(0xFFFFFFFFu, iinc, 31);
        before += __shfl_sync(0xFFFFFFFFu, inc, 31);
    }
    if (lane == 0) r_best[tree] = pick;
}

// ---------------------------------------------------------------------------------------------
// Tree reuse (TAFL_MCTS_TREE_REUSE): the reference keeps one MCTS object per episode, so the
// subtree below the move played carries its Ns / Nsa / Qsa / Ps into the next getActionProb
// (mcts.py:20-26 — the dicts are never cleared).  k_mcts_reroot makes the child reached by the
// chosen root edge the new node 0 and compacts its subtree IN PLACE: nodes are created in
// increasing index order with parent < child, and edge blocks are handed out in the same
// order, so every kept node / edge block moves to a lower-or-equal address and an ascending
// sweep never overwrites something it still has to read.  One warp per tree.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_reroot(MctsView mv, uint32_t n_trees, const uint32_t* __restrict__ r_counts, const uint32_t* __restrict__ r_best,
              uint32_t* __restrict__ remap_all, int prefix) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    double* const prior = mv.e_prior ? mv.e_prior + (size_t)tree * mv.edge_cap : nullptr;
    uint32_t* const remap = remap_all + (size_t)tree * mv.node_cap;
    const uint32_t n_nodes = mv.t_nodes[tree];
    const uint4 h0 = n_hdr[0];
    uint32_t new_root = MCTS_NO_PARENT;
    if (n_nodes > 0 && mv.t_overflow[tree] == 0 && hdr_es(h0.w) != ES_ONGOING) return;   // finished game: no move was played
    if (n_nodes > 0 && mv.t_overflow[tree] == 0 && hdr_expanded(h0.w) && r_counts[tree] > 0) {
        const MctsEdge e = edges[h0.x + r_best[tree]];
        if (e.N > 0) new_root = e.child;
    }
    if (new_root == MCTS_NO_PARENT) {   // nothing to keep: the next search starts a fresh tree
        if (lane == 0) { mv.t_nodes[tree] = 0; mv.t_edges[tree] = 0; }
        return;
    }
    // pass 1: remap[i] = new index of node i, or NONE if it is outside the kept subtree
    uint32_t kept = 0;
    for (uint32_t base = 0; base < n_nodes; base += 32) {
        const uint32_t i = base + lane;
        const bool valid = i < n_nodes;
        const uint32_t parent = valid ? n_hdr[i].z : MCTS_NO_PARENT;
        bool keep = valid && i == new_root;
        const bool has_parent = valid && i > new_root && parent != MCTS_NO_PARENT;
        const bool in_chunk = has_parent && parent >= base;
        if (has_parent && !in_chunk) keep = parent >= new_root && remap[parent] != MCTS_NO_PARENT;
        while (true) {   // parents inside this chunk: propagate until stable
            const bool pk = __shfl_sync(0xFFFFFFFFu, (int)keep, in_chunk ? (int)(parent - base) : 0) != 0;
            const bool nk = keep || (in_chunk && pk);
            const unsigned changed = __ballot_sync(0xFFFFFFFFu, nk != keep);
            keep = nk;
            if (!changed) break;
        }
        const unsigned km = __ballot_sync(0xFFFFFFFFu, keep);
        if (valid) remap[i] = keep ? kept + (uint32_t)__popc(km & ((1u << lane) - 1u)) : MCTS_NO_PARENT;
        kept += (uint32_t)__popc(km);
        __syncwarp();
    }
    // pass 2: move the kept nodes and their edge blocks down, ascending
    uint32_t e_used = 0;
    for (uint32_t i = new_root; i < n_nodes; ++i) {
        const uint32_t j = remap[i];
        if (j == MCTS_NO_PARENT) continue;
        uint4 h = n_hdr[i];
        const bool expanded = hdr_expanded(h.w) != 0;
        const uint32_t L = expanded ? hdr_n_edges(h.w) : 0u, old_base = h.x;
        if (j != i) {
            const uint32_t w0 = n_ptab[(size_t)i * MCTS_PT_MAX + lane], w1 = n_ptab[(size_t)i * MCTS_PT_MAX + 32 + lane];
            tafl_state s;
            if (lane == 0) s = load_state(n_state + i);
            __syncwarp();
            n_ptab[(size_t)j * MCTS_PT_MAX + lane] = w0;
            n_ptab[(size_t)j * MCTS_PT_MAX + 32 + lane] = w1;
            if (lane == 0) store_state(n_state + j, s);
        }
        const uint32_t Lc = (prefix && h.y + 1u < L) ? h.y + 1u : L;   // prefix trees: only [0, Ns] was ever written
        for (uint32_t off = 0; off < Lc; off += 32) {
            const uint32_t a = off + lane;
            MctsEdge e;
            double p = 0.0;
            if (a < Lc) {
                e = edges[old_base + a];
                if (prior) p = prior[old_base + a];
                e.child = e.N > 0 ? remap[e.child] : 0u;
            }
            __syncwarp();
            if (a < Lc) {
                edges[e_used + a] = e;
                if (prior) prior[e_used + a] = p;
            }
            __syncwarp();
        }
        if (lane == 0) {
            h.x = expanded ? e_used : 0u;
            h.z = i == new_root ? MCTS_NO_PARENT : remap[h.z];
            n_hdr[j] = h;
        }
        e_used += L;
        __syncwarp();
    }
    if (lane == 0) { mv.t_nodes[tree] = kept; mv.t_edges[tree] = e_used; }
}

// start of a search with tree reuse: keep a carried tree whose root IS the game's current state,
// start a fresh one otherwise (first move, game reset / states replaced by the caller, no child)
__global__ void k_mcts_init_carry(const tafl_state* __restrict__ states, uint32_t n, MctsView mv) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const tafl_state s = load_state(states + g);
    if (mv.t_nodes[g] > 0 && mv.t_overflow[g] == 0) {
        const uint4* const p = reinterpret_cast<const uint4*>(states + g);
        const uint4* const q = reinterpret_cast<const uint4*>(mv.n_state + (size_t)g * mv.node_cap);
        bool same = true;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 x = p[k], y = q[k];
            same = same && x.x == y.x && x.y == y.y && x.z == y.z && x.w == y.w;
        }
        if (same) {
            if (mv.t_pending) mv.t_pending[g] = MCTS_NO_PARENT;
            return;
        }
    }
    store_state(mv.n_state + (size_t)g * mv.node_cap, s);
    mv.n_hdr[(size_t)g * mv.node_cap] = make_uint4(0u, 0u, MCTS_NO_PARENT, hdr_pack(0, 0, 0, es_code(s)));
    mv.t_nodes[g] = 1;
    mv.t_edges[g] = 0;
    mv.t_overflow[g] = 0;
    if (mv.t_pending) mv.t_pending[g] = MCTS_NO_PARENT;
    atomicAdd(&mv.stats[ST_NODES], 1ull);
}

// play the chosen root move in every Ongoing game (do_valid_play, logic.rs:782-820)
template <class BB, int SP>
__global__ void __launch_bounds__(64)
k_mcts_play_best(const __grid_constant__ Ctx<BB, SP> cx, tafl_state* __restrict__ states, uint32_t n,
                 const tafl_move* __restrict__ r_moves, const uint32_t* __restrict__ r_counts,
                 const uint32_t* __restrict__ r_best, uint8_t* __restrict__ status_out) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    tafl_state s = load_state(states + g);
    if (s.status == TAFL_STATUS_ONGOING && r_counts[g] > 0) {
        do_valid_move(cx, s, r_moves[(size_t)g * TAFL_MAX_MOVES + r_best[g]]);
        store_state(states + g, s);
    }
    if (status_out) status_out[g] = s.status;
}

// one replay-buffer row per game: header (root position, side, chosen move, status after it) +
// root moves + visit counts, fixed stride (include/tafl.h TAFL_REC_STRIDE); one warp per game.
// `row_of` (may be null): compacted row index of each game, 0xFFFFFFFF = the game emits no row.
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_pack_records(MctsView mv, const tafl_state* __restrict__ states, uint32_t n, uint64_t gid_base,
               const tafl_move* __restrict__ r_moves, const uint32_t* __restrict__ r_visits,
               const uint32_t* __restrict__ r_counts, const uint32_t* __restrict__ r_best,
               const uint32_t* __restrict__ episode, const uint32_t* __restrict__ row_of, uint8_t* __restrict__ out) {
    static_assert(TAFL_REC_MOVES == TAFL_MAX_MOVES, "a record row holds every legal move of its root");
    static_assert(sizeof(tafl_record_header) == 56 && TAFL_REC_STRIDE % 8 == 0, "record layout");
    const int lane = threadIdx.x & 31;
    const uint32_t g = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (g >= n) return;
    const uint32_t row = row_of ? row_of[g] : g;
    if (row == 0xFFFFFFFFu) return;
    uint8_t* rec = out + (size_t)row * TAFL_REC_STRIDE;
    const uint32_t cnt = r_counts[g];
    if (lane == 0) {
        const tafl_state root = load_state(mv.n_state + (size_t)g * mv.node_cap);
        const tafl_state cur = load_state(states + g);
        tafl_record_header h;
        h.game_id = gid_base + g;
        h.ply = root.turn;
        h.n_moves = (uint16_t)cnt;
        h.side_to_play = (root.flags & TAFL_FLAG_DEF_TO_PLAY) ? TAFL_DEFENDER : TAFL_ATTACKER;
        h.status = cur.status;
        h.att[0] = root.att[0]; h.att[1] = root.att[1]; h.def[0] = root.def[0]; h.def[1] = root.def[1];
        h.best = (uint16_t)r_best[g];
        h._pad = 0;
        h.episode = episode[g];
        *reinterpret_cast<tafl_record_header*>(rec) = h;
    }
    // 8-byte vectors: 4 moves / 2 visit counts per lane per store (rows are 8-byte aligned: 56 + 2*512 + 4*512)
    const uint2* const ms = reinterpret_cast<const uint2*>(r_moves + (size_t)g * TAFL_MAX_MOVES);
    const uint2* const vs = reinterpret_cast<const uint2*>(r_visits + (size_t)g * TAFL_MAX_MOVES);
    uint2* const mo = reinterpret_cast<uint2*>(rec + sizeof(tafl_record_header));
    uint2* const vo = reinterpret_cast<uint2*>(rec + sizeof(tafl_record_header) + 2 * TAFL_REC_MOVES);
    const uint2 zero = make_uint2(0u, 0u);
    for (uint32_t i = lane; i < TAFL_REC_MOVES / 4; i += 32) {   // k_mcts_results zero-fills the rows behind cnt
        mo[i] = 4u * i < cnt ? ms[i] : zero;
    }
    for (uint32_t i = lane; i < TAFL_REC_MOVES / 2; i += 32) {
        vo[i] = 2u * i < cnt ? vs[i] : zero;
    }
}

// stable compaction of the searched games (r_counts > 0) in game order: row_of[g] = number of searched
// games before g, one CTA walks the array in chunks of its size (n_games is a few thousand trees)
__global__ void __launch_bounds__(1024)
k_record_rows(const uint32_t* __restrict__ r_counts, uint32_t n, uint32_t* __restrict__ row_of, uint32_t* __restrict__ d_count,
              uint32_t* __restrict__ h_count, const uint32_t* __restrict__ any_overflow) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t base_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (uint32_t g0 = 0; g0 < n; g0 += blockDim.x) {
        const uint32_t g = g0 + threadIdx.x;
        const bool keep = g < n && r_counts[g] > 0;
        const unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
        if (lane == 0) warp_tot[warp] = (uint32_t)__popc(m);
        __syncthreads();
        uint32_t before = base_s;
        for (int w = 0; w < warp; ++w) before += warp_tot[w];
        if (g < n) row_of[g] = keep ? before + (uint32_t)__popc(m & ((1u << lane) - 1u)) : 0xFFFFFFFFu;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t t = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += warp_tot[w];
            base_s += t;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (d_count) *d_count = base_s;
        if (h_count) { h_count[0] = base_s; h_count[1] = *any_overflow; }   // device alias of the caller's pinned words
    }
}


// ---------------------------------------------------------------------------------------------
// Evaluator hook: MCTS.search (mcts.py:55-136) cut at `nnet.predict` (mcts.py:85).
// One round = one simulation in EVERY tree: k_mcts_select_leaf descends to the leaf and emits the
// leaf batch, the evaluator (any code enqueued on the engine's stream) fills priors and values,
// k_mcts_expand_backup masks / renormalises the priors exactly as mcts.py:87-99 does and backs the
// value up.  A simulation that meets a terminal node completes inside the first phase.
// ---------------------------------------------------------------------------------------------

// board_to_matrix (game/main.rs:55-83): corners 20, throne 30, + 1 per soldier, + 5 for the king
template <class BB, int SP>
__device__ __forceinline__ void warp_board_planes(const Ctx<BB, SP>& cx, const Pos<BB>& p, uint8_t* __restrict__ out, int lane) {
    const int n = cx.n(), mid = n / 2;
    const BB occ = p.occ(), king = p.king();
    for (int t = lane; t < n * n; t += 32) {
        const int r = t / n, c = t - r * n, b = idx_of<BB>(r, c);
        int v = 0;
        if ((r == 0 || r == n - 1) && (c == 0 || c == n - 1)) v = 20;
        if (r == mid && c == mid) v = 30;
        if (occ.test(b)) v += king.test(b) ? 5 : 1;
        out[t] = (uint8_t)v;
    }
}

template <class BB, int SP>
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_select_leaf(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, double cpuct, double draw_value,
                   tafl_state* __restrict__ l_state, uint8_t* __restrict__ l_planes, tafl_move* __restrict__ l_moves,
                   uint32_t* __restrict__ l_counts, uint32_t* __restrict__ any_overflow) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    const double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap;
    uint32_t want = 0, pending = MCTS_NO_PARENT;
    if (mv.t_overflow[tree] == 0) {
        uint32_t node = 0, depth = 0;
        while (true) {
            const uint4 h = n_hdr[node];
            const uint32_t es = hdr_es(h.w);
            if (es != ES_ONGOING) {   // mcts.py:77-81: no evaluation, the simulation completes here
                const double Es = es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value);
                warp_backup(n_hdr, edges, node, -Es, lane);
                if (lane == 0) {
                    atomicAdd(&mv.stats[ST_TERMINAL], 1ull);
                    atomicAdd(&mv.stats[ST_DEPTH], (unsigned long long)depth);
                    atomicAdd(&mv.stats[ST_SIMS], 1ull);
                }
                break;
            }
            if (!hdr_expanded(h.w)) {   // mcts.py:83-86: leaf, ask the evaluator
                const tafl_state s = load_state(n_state + node);
                const Pos<BB> p = load_pos(cx, s);
                const int def_side = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
                uint32_t* const ptab = n_ptab + (size_t)node * MCTS_PT_MAX;
                int n_pieces;
                const int L = warp_build_ptab(cx, p, def_side, ptab, lane, &n_pieces);
                if ((uint32_t)L > mv.edge_cap - mv.t_edges[tree] || n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) {
                    const uint32_t why = (n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) ? OVF_PIECES : OVF_EDGES;
                    if (lane == 0) { mv.t_overflow[tree] = why; atomicOr(any_overflow, why); }
                    break;
                }
                if (L == 0) {   // Ongoing position without plays: unreachable through do_play; expand empty, v = 0
                    if (lane == 0) n_hdr[node] = make_uint4(mv.t_edges[tree], 0u, h.z, hdr_pack(0u, hdr_pedge(h.w), 1u, es, (uint32_t)n_pieces));
                    __syncwarp();
                    warp_backup(n_hdr, edges, node, -0.0, lane);
                    if (lane == 0) atomicAdd(&mv.stats[ST_SIMS], 1ull);
                    break;
                }
                __syncwarp();
                // the leaf batch row: position, planes, canonical move list (action index -> move)
                if (lane == 0) {
                    store_state(l_state + tree, s);
                    n_hdr[node] = make_uint4(0u, 0u, h.z, hdr_pack((uint32_t)L, hdr_pedge(h.w), 0u, es, (uint32_t)n_pieces));
                }
                warp_board_planes(cx, p, l_planes + (size_t)tree * (size_t)(cx.n() * cx.n()), lane);
                tafl_move* const mo = l_moves + (size_t)tree * TAFL_MAX_MOVES;
                int before = 0;
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
                    const int off = before + incl - c;
                    for (int j = 0; j < c; ++j) mo[off + j] = pt_decode(cx, e, def_side, j);
                    before += __shfl_sync(0xFFFFFFFFu, incl, 31);
                }
                want = (uint32_t)L;
                pending = node;
                if (lane == 0) atomicAdd(&mv.stats[ST_DEPTH], (unsigned long long)depth);
                break;
            }
            const uint32_t L = hdr_n_edges(h.w), base = h.x, Ns = h.y;
            if (L == 0) {
                warp_backup(n_hdr, edges, node, 0.0, lane);
                if (lane == 0) atomicAdd(&mv.stats[ST_SIMS], 1ull);
                break;
            }
            uint32_t a, best_N, best_child;
            warp_puct_select<true>(edges + base, prior + base, L, Ns, cpuct, lane, a, best_N, best_child);
            uint32_t child;
            if (best_N == 0) {   // first traversal: getNextState (mcts.py:122-123)
                tafl_state s = load_state(n_state + node);
                const tafl_move m = warp_decode_edge(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                                     (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)a, lane);
                do_valid_move(cx, s, m);
                child = mv.t_nodes[tree];
                __syncwarp();
                if (child >= mv.node_cap) {
                    if (lane == 0) { mv.t_overflow[tree] = OVF_NODES; atomicOr(any_overflow, OVF_NODES); }
                    break;
                }
                if (lane == 0) {
                    mv.t_nodes[tree] = child + 1u;
                    store_state(n_state + child, s);
                    n_hdr[child] = make_uint4(0u, 0u, node, hdr_pack(0, a, 0, es_code(s)));
                    edges[base + a].child = child;
                    atomicAdd(&mv.stats[ST_NODES], 1ull);
                }
                __syncwarp();
            } else {
                child = best_child;
            }
            node = child;
            depth += 1;
        }
    }
    if (lane == 0) {
        l_counts[tree] = want;
        mv.t_pending[tree] = pending;
    }
}

// mcts.py:87-102 for the pending leaf of every tree, then the backup of -v (mcts.py:125-136).
// np.sum (mcts.py:88,99) is numpy's pairwise add-reduce over the action_size = 512 vector: four
// 128-element blocks of eight stride-8 accumulators — lane = 8*block + accumulator reproduces
// its association order exactly (see oracle/mcts_oracle.hpp np_pairwise_sum).
__device__ __forceinline__ double warp_np_sum512(const double (&x)[16], int lane) {
    double r = x[0];
#pragma unroll
    for (int i = 1; i < 16; ++i) r += x[i];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) r += __shfl_xor_sync(0xFFFFFFFFu, r, o);
    (void)lane;
    return 0.0 + r;
}

// Root exploration noise (the reference's TODO at src/mcts.rs:53; oracle: Mcts::noise_root): once per
// search, P[a] <- (1 - eps) * P[a] + eps * noise[a] on the root's priors as soon as the root is
// expanded (first round of a fresh tree, or before the first round of a carried one).
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_root_noise(MctsView mv, uint32_t n_trees, const double* __restrict__ noise, double eps, uint8_t* __restrict__ noised) {
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees || noised[tree]) return;
    const uint4 h = mv.n_hdr[(size_t)tree * mv.node_cap];
    if (!hdr_expanded(h.w)) return;
    double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap + h.x;
    const double* const nz = noise + (size_t)tree * TAFL_MAX_MOVES;
    const uint32_t L = hdr_n_edges(h.w);
    for (uint32_t a = lane; a < L; a += 32) prior[a] = (1.0 - eps) * prior[a] + eps * nz[a];
    __syncwarp();
    if (lane == 0) noised[tree] = 1;
}

__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_mcts_expand_backup(MctsView mv, uint32_t n_trees, const float* __restrict__ l_priors, const float* __restrict__ l_values) {
    static_assert(TAFL_MAX_MOVES == 512, "np.sum restatement is laid out for action_size == 512");
    const int lane = threadIdx.x & 31;
    const uint32_t tree = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (tree >= n_trees) return;
    const uint32_t node = mv.t_pending[tree];
    if (node == MCTS_NO_PARENT) return;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap;
    const uint4 h = n_hdr[node];
    const uint32_t L = hdr_n_edges(h.w), base = mv.t_edges[tree];
    const float* const pf = l_priors + (size_t)tree * TAFL_MAX_MOVES;
    const int blk = lane >> 3, acc = lane & 7;
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t a = (uint32_t)(blk * 128 + acc + 8 * i);
        x[i] = (double)pf[a] * (a < L ? 1.0 : 0.0);   // self.Ps[s] * valids  (mcts.py:87)
    }
    double sum = warp_np_sum512(x, lane);                 // mcts.py:88
    if (!(sum > 0)) {                                      // mcts.py:91-99: all valid moves were masked
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const uint32_t a = (uint32_t)(blk * 128 + acc + 8 * i);
            x[i] = x[i] + (a < L ? 1.0 : 0.0);
        }
        sum = warp_np_sum512(x, lane);
    }
    MctsEdge z;
    z.Q = 0.0; z.N = 0u; z.child = 0u;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t a = (uint32_t)(blk * 128 + acc + 8 * i);
        if (a < L) {
            prior[base + a] = x[i] / sum;                  // mcts.py:90 / :99
            edges[base + a] = z;
        }
    }
    const double v = (double)l_values[tree];
    if (lane == 0) {
        n_hdr[node] = make_uint4(base, 0u, h.z, h.w | (1u << 20));   // expanded, Ns = 0 (mcts.py:101)
        mv.t_edges[tree] = base + L;
        mv.t_pending[tree] = MCTS_NO_PARENT;
        atomicAdd(&mv.stats[ST_EDGES], (unsigned long long)L);
        atomicAdd(&mv.stats[ST_SIMS], 1ull);
        atomicMax(&mv.stats[ST_MAXEDGES], (unsigned long long)(base + L));
    }
    __syncwarp();
    warp_backup(n_hdr, edges, node, -v, lane);             // mcts.py:102: return -v
}

// ---------------------------------------------------------------------------------------------
// Second-generation evaluator-hook kernel: G lanes per tree (32/G trees per warp) and BOTH halves of a
// round in one launch — phase 1 finishes the previous round (mask / renormalise the evaluator's priors,
// mcts.py:87-102, and back its value up), phase 2 runs the next descent to a leaf and emits the leaf
// batch row.  A search of `sims` rounds is then  select | eval | (expand+select | eval) x (sims-1) | expand:
// two launches per simulation instead of three, and the descent's create step (do_valid_move + the piece
// table of the new child, the expensive straight-line part) is issued once per warp for all its trees,
// as in k_mcts_sims_g.  The split C-ABI calls use the same kernel with one phase switched off.
// Results are bit-identical to k_mcts_select_leaf + k_mcts_expand_backup (tested).
// ---------------------------------------------------------------------------------------------
template <int G>
__device__ __forceinline__ void grp_puct_select_priors(const MctsEdge* __restrict__ ed, const double* __restrict__ prior, uint32_t L,
                                                       uint32_t Ns, double cpuct, unsigned gmask, int gl, uint32_t& out_a,
                                                       uint32_t& out_N, uint32_t& out_child) {
    const double sq_vis = sqrt((double)Ns), sq_new = sqrt((double)Ns + MCTS_EPS);
    double best_u = -INFINITY;
    uint32_t best_a = 0xFFFFFFFFu, best_N = 0, best_child = 0;
    for (uint32_t a0 = gl; a0 < L; a0 += 4 * G) {
        MctsEdge e4[4];
        double p4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) { e4[j] = ed[a]; p4[j] = prior[a]; }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) {
                const double c = cpuct * p4[j];
                double u;
                if (e4[j].N > 0) u = e4[j].Q + (c * sq_vis) / (double)(1u + e4[j].N);
                else u = c * sq_new;
                if (u > best_u) { best_u = u; best_a = a; best_N = e4[j].N; best_child = e4[j].child; }
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        const double ou = __shfl_xor_sync(gmask, best_u, o, G);
        const uint32_t oa = __shfl_xor_sync(gmask, best_a, o, G);
        const uint32_t on = __shfl_xor_sync(gmask, best_N, o, G);
        const uint32_t oc = __shfl_xor_sync(gmask, best_child, o, G);
        if (ou > best_u || (ou == best_u && oa < best_a)) { best_u = ou; best_a = oa; best_N = on; best_child = oc; }
    }
    out_a = best_a; out_N = best_N; out_child = best_child;
}

// mcts.py:87-101 for the pending leaf `node` of ONE tree by the full warp (np.sum needs all 32 lanes: lane = 8 * block +
// accumulator holds the 16 evaluator outputs pfv[i] = priors[block * 128 + accumulator + 8 i], loaded by the caller so
// that the loads of all the warp's trees are in flight together).  The backup of -v is left to the tree's own group.
__device__ __forceinline__ void warp_expand_leaf(const MctsView& mv, uint32_t tree, uint32_t node, const float (&pfv)[16],
                                                 const float* __restrict__ pf, int lane, bool single) {
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap;
    const uint4 h = n_hdr[node];
    const uint32_t L = hdr_n_edges(h.w), base = mv.t_edges[tree];
    const int blk = lane >> 3, acc = lane & 7;
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t a = (uint32_t)(blk * 128 + acc + 8 * i);
        x[i] = (double)pfv[i] * (a < L ? 1.0 : 0.0);   // self.Ps[s] * valids  (mcts.py:87)
    }
    double sum = warp_np_sum512(x, lane);                 // mcts.py:88
    const bool fallback = !(sum > 0);                      // mcts.py:91-99: all valid moves were masked
    if (fallback) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const uint32_t a = (uint32_t)(blk * 128 + acc + 8 * i);
            x[i] = x[i] + (a < L ? 1.0 : 0.0);
        }
        sum = warp_np_sum512(x, lane);
    }
    // Ps /= sum for the valid actions, one action per lane (the np.sum layout above would leave 24 of the 32 lanes idle
    // here): the numerator is recomputed from the evaluator output with the same operations, so the quotient is the same
    MctsEdge z;
    z.Q = 0.0; z.N = 0u; z.child = 0u;
    for (uint32_t a = (uint32_t)lane; a < L; a += 32u) {
        double xa = (double)pf[a] * 1.0;
        if (fallback) xa = xa + 1.0;
        prior[base + a] = xa / sum;                        // mcts.py:90 / :99
        edges[base + a] = z;
    }
    __syncwarp();
    if (lane == 0) {
        n_hdr[node] = make_uint4(base, 0u, h.z, h.w | (1u << 20));   // expanded, Ns = 0 (mcts.py:101)
        mv.t_edges[tree] = base + L;
        atomicAdd(&mv.stats[ST_EDGES], (unsigned long long)L);
        atomicMax(&mv.stats[ST_MAXEDGES], (unsigned long long)(base + L));
        if (single) {   // one leaf per tree and round (k_mcts_hook_round): the simulation is counted here
            mv.t_pending[tree] = MCTS_NO_PARENT;
            atomicAdd(&mv.stats[ST_SIMS], 1ull);
        }
    }
}
__device__ __forceinline__ void warp_expand_tree(const MctsView& mv, uint32_t tree, uint32_t node, const float (&pfv)[16],
                                                 const float* __restrict__ l_priors, int lane) {
    warp_expand_leaf(mv, tree, node, pfv, l_priors + (size_t)tree * TAFL_MAX_MOVES, lane, true);
}

// one row of the leaf batch: the position, its board_to_matrix planes and its canonical move list, by the tree's group
template <class BB, int SP, int G>
__device__ __forceinline__ void grp_emit_leaf_row(const Ctx<BB, SP>& cx, const tafl_state& s, const Pos<BB>& p, int def_side,
                                                  const uint32_t* __restrict__ ptab, int n_pieces, tafl_state* __restrict__ o_state,
                                                  uint8_t* __restrict__ o_planes, tafl_move* __restrict__ mo, unsigned gmask, int gl) {
    if (gl == 0) store_state(o_state, s);
    {   // board_to_matrix (game/main.rs:55-83)
        const int n = cx.n(), mid = n / 2;
        const BB occ = p.occ(), king = p.king();
        for (int t = gl; t < n * n; t += G) {
            const int r = t / n, c = t - r * n, b = idx_of<BB>(r, c);
            int val = 0;
            if ((r == 0 || r == n - 1) && (c == 0 || c == n - 1)) val = 20;
            if (r == mid && c == mid) val = 30;
            if (occ.test(b)) val += king.test(b) ? 5 : 1;
            o_planes[t] = (uint8_t)val;
        }
    }
    int before = 0;
    for (int base = 0; base < n_pieces; base += G) {
        const int idx = base + gl;
        const uint32_t e = idx < n_pieces ? ptab[idx] : 0u;   // written by this same lane in grp_build_ptab
        const int c = pt_total(e);
        int incl = c;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) {
            const int t = __shfl_up_sync(gmask, incl, o, G);
            if (gl >= o) incl += t;
        }
        if (c > 0) {   // the piece's plays by tile arithmetic (emit_piece_plays == pt_decode for j = 0 .. c-1)
            const int tile = (int)(e & 127u), row = tile / BB::STRIDE;
            Rays r;
            r.cnt[0] = (int)((e >> 8) & 15u); r.cnt[1] = (int)((e >> 12) & 15u);
            r.cnt[2] = (int)((e >> 16) & 15u); r.cnt[3] = (int)((e >> 20) & 15u);
            emit_piece_plays(cx, tile, row, tile - row * BB::STRIDE, (e & 0x80u) ? KIND_K : (def_side ? KIND_DS : KIND_AS), r, mo,
                             before + incl - c, TAFL_MAX_MOVES);
        }
        before += __shfl_sync(gmask, incl, G - 1, G);
    }
}

template <class BB, int SP, int G>
__global__ void __launch_bounds__(MCTS_WARPS * 32, 4)
k_mcts_hook_round(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, double cpuct, double draw_value,
                  tafl_state* __restrict__ l_state, uint8_t* __restrict__ l_planes, tafl_move* __restrict__ l_moves,
                  uint32_t* __restrict__ l_counts, const float* __restrict__ l_priors, const float* __restrict__ l_values,
                  uint32_t* __restrict__ any_overflow, int do_expand, int do_select) {
    constexpr int TPW = 32 / G;
    enum { M_NONE = 0, M_BACKUP, M_CREATE, M_LEAF };
    const int lane = threadIdx.x & 31, gl = lane % G, gshift = (lane / G) * G;
    const unsigned gmask = G == 32 ? 0xFFFFFFFFu : (((1u << (G & 31)) - 1u) << gshift);
    const uint32_t warp_tree0 = (blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5)) * TPW;
    // ---- phase 1: the leaves evaluated since the last round.  Normalising the priors needs the full warp per tree
    // (np.sum's association order), so the warp's trees take turns — with the evaluator outputs of all of them
    // requested up front — and each tree's own group then backs its value up, all groups at once.
    if (do_expand) {
        const uint32_t my_tree = warp_tree0 + (uint32_t)(lane / G);
        const uint32_t my_pending = my_tree < n_trees ? mv.t_pending[my_tree] : MCTS_NO_PARENT;
        const double my_v = my_pending != MCTS_NO_PARENT ? (double)l_values[my_tree] : 0.0;
        float pfv[TPW][16];
        uint32_t pend[TPW];
        const int blk = lane >> 3, acc = lane & 7;
#pragma unroll
        for (int t = 0; t < TPW; ++t) {
            const uint32_t tr = warp_tree0 + (uint32_t)t;
            pend[t] = tr < n_trees ? mv.t_pending[tr] : MCTS_NO_PARENT;   // warp-uniform
            if (pend[t] != MCTS_NO_PARENT) {
                const float* const pf = l_priors + (size_t)tr * TAFL_MAX_MOVES;
#pragma unroll
                for (int i = 0; i < 16; ++i) pfv[t][i] = pf[blk * 128 + acc + 8 * i];
            }
        }
#pragma unroll
        for (int t = 0; t < TPW; ++t)
            if (pend[t] != MCTS_NO_PARENT) warp_expand_tree(mv, warp_tree0 + (uint32_t)t, pend[t], pfv[t], l_priors, lane);
        __syncwarp();
        if (my_pending != MCTS_NO_PARENT)   // mcts.py:102: return -v
            grp_backup<false>(mv.n_hdr + (size_t)my_tree * mv.node_cap, mv.edges + (size_t)my_tree * mv.edge_cap, my_pending, -my_v,
                              gmask, gl);
        __syncwarp();
    }
    if (!do_select) return;
    // ---- phase 2: MCTS.search down to the next leaf (mcts.py:55-123), G lanes per tree
    const uint32_t tree_raw = warp_tree0 + (uint32_t)(lane / G);
    const bool live = tree_raw < n_trees;
    const uint32_t tree = live ? tree_raw : 0u;
    tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    const double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap;
    uint32_t overflow = 0, want = 0, pending = MCTS_NO_PARENT;
    const bool stop = !live || mv.t_overflow[tree] != 0;
    int mode = M_NONE;
    uint32_t node = 0, depth = 0, sel_a = 0, sel_base = 0;
    uint4 h = make_uint4(0u, 0u, 0u, 0u);
    double v = 0.0;
    unsigned long long c_sims = 0, c_nodes = 0, c_depth = 0, c_term = 0;
    if (!stop) {
        while (true) {
            h = n_hdr[node];
            const uint32_t es = hdr_es(h.w);
            if (es != ES_ONGOING) {   // mcts.py:77-81: no evaluation, the simulation completes here
                v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                c_term += 1;
                c_depth += depth;
                mode = M_BACKUP;
                break;
            }
            if (!hdr_expanded(h.w)) { mode = M_LEAF; break; }   // the root of a fresh tree
            const uint32_t L = hdr_n_edges(h.w);
            if (L == 0) { v = 0.0; mode = M_BACKUP; break; }
            if (gl < 3) {   // the node's state and piece table, wanted if the descent ends here with a new child (as k_mcts_sims_g)
                const void* pf = gl == 0 ? (const void*)(n_state + node) : (const void*)(n_ptab + (size_t)node * MCTS_PT_MAX + (gl - 1) * 32);
                asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
            }
            uint32_t a, bN, bchild;
            grp_puct_select_priors<G>(edges + h.x, prior + h.x, L, h.y, cpuct, gmask, gl, a, bN, bchild);
            if (bN == 0) { mode = M_CREATE; sel_a = a; sel_base = h.x; break; }
            node = bchild;
            depth += 1;
        }
    }
    __syncwarp();   // every tree of the warp opens its new node together (a CTA-wide barrier here, as in k_mcts_sims_g, changes nothing: 1.30 x 10^8 either way)
    if (mode == M_CREATE || mode == M_LEAF) {
        tafl_state s = load_state(n_state + node);
        uint32_t target = node, parent = h.z, pedge = hdr_pedge(h.w), es = ES_ONGOING;
        bool leaf = true;
        if (mode == M_CREATE) {   // getNextState (mcts.py:122-123)
            const uint32_t n_nodes = mv.t_nodes[tree];
            if (n_nodes >= mv.node_cap) {
                overflow = OVF_NODES;
                leaf = false;
                mode = M_NONE;
            } else {
                const tafl_move m = grp_decode_edge<G>(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                                       (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)sel_a, gmask, gshift, gl);
                do_valid_move(cx, s, m);
                target = n_nodes;
                c_nodes += 1;
                parent = node;
                pedge = sel_a;
                depth += 1;
                es = es_code(s);
                __syncwarp(gmask);   // every lane has read t_nodes
                if (gl == 0) {
                    mv.t_nodes[tree] = n_nodes + 1u;
                    store_state(n_state + target, s);
                    edges[sel_base + sel_a].child = target;
                }
                if (es != ES_ONGOING) {   // the new node is terminal (mcts.py:77-81)
                    if (gl == 0) n_hdr[target] = make_uint4(0u, 0u, parent, hdr_pack(0, pedge, 0, es));
                    v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                    c_term += 1;
                    c_depth += depth;
                    leaf = false;
                    mode = M_BACKUP;
                }
            }
        }
        if (leaf) {   // mcts.py:83-86: ask the evaluator
            const Pos<BB> p = load_pos(cx, s);
            const int def_side = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
            uint32_t* const ptab = n_ptab + (size_t)target * MCTS_PT_MAX;
            int n_pieces;
            const int L = grp_build_ptab<G>(cx, p, def_side, ptab, gmask, gl, &n_pieces);
            if ((uint32_t)L > mv.edge_cap - mv.t_edges[tree] || n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) {
                overflow = (n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) ? OVF_PIECES : OVF_EDGES;
                if (mode == M_CREATE && gl == 0) n_hdr[target] = make_uint4(0u, 0u, parent, hdr_pack(0, pedge, 0, es));
                mode = M_NONE;
            } else if (L == 0) {   // Ongoing position without plays: unreachable through do_play; expand empty, v = 0
                if (gl == 0) n_hdr[target] = make_uint4(mv.t_edges[tree], 0u, parent, hdr_pack(0u, pedge, 1u, es, (uint32_t)n_pieces));
                v = -0.0;
                mode = M_BACKUP;
            } else {
                // the leaf batch row: position, planes, canonical move list (action index -> move)
                if (gl == 0) n_hdr[target] = make_uint4(0u, 0u, parent, hdr_pack((uint32_t)L, pedge, 0u, es, (uint32_t)n_pieces));
                grp_emit_leaf_row<BB, SP, G>(cx, s, p, def_side, ptab, n_pieces, l_state + tree, l_planes + (size_t)tree * (size_t)(cx.n() * cx.n()),
                                             l_moves + (size_t)tree * TAFL_MAX_MOVES, gmask, gl);
                want = (uint32_t)L;
                pending = target;
                c_depth += depth;
                mode = M_NONE;
            }
            if (mode == M_BACKUP) node = target;
        } else if (mode == M_BACKUP) {
            node = target;
        }
        __syncwarp(gmask);
    }
    if (mode == M_BACKUP) {   // terminal node / empty expansion: the simulation completes without the evaluator
        grp_backup<false>(n_hdr, edges, node, v, gmask, gl);
        c_sims += 1;
    }
    if (gl == 0 && live) {
        if (!stop) {
            l_counts[tree] = want;
            mv.t_pending[tree] = pending;
        } else {
            l_counts[tree] = 0u;
            mv.t_pending[tree] = MCTS_NO_PARENT;
        }
        if (overflow) { mv.t_overflow[tree] = overflow; atomicOr(any_overflow, overflow); }
        if (c_sims) atomicAdd(&mv.stats[ST_SIMS], c_sims);
        if (c_nodes) atomicAdd(&mv.stats[ST_NODES], c_nodes);
        if (c_depth) atomicAdd(&mv.stats[ST_DEPTH], c_depth);
        if (c_term) atomicAdd(&mv.stats[ST_TERMINAL], c_term);
    }
}

// ---------------------------------------------------------------------------------------------
// Virtual loss through the evaluator hook: W leaves per tree and round, so that one evaluator call sees a batch of
// n_trees x W rows (row = tree * W + simulation).  The schedule is oracle/mcts_oracle.hpp::run_from_root_vl with an
// evaluator: phase 2 selects the W simulations of a batch in index order under virtual loss, creates the new
// children (one lane per child, as k_mcts_sims_vl) and emits a leaf row for every new, non-terminal leaf; after the
// evaluator phase 1 expands those leaves in order (mask / renormalise, mcts.py:87-101) and backs every simulation
// up in order, taking its virtual losses off; a collision takes the value of the simulation that opened the child.
// What crosses the launch boundary is the batch table (kind and leaf of each simulation) and the per-tree count of
// simulations done.  Every edge of an expanded node is initialised (no prefix rule with real priors).
// ---------------------------------------------------------------------------------------------
struct VlBatch {
    uint32_t* leaf;   // [n_trees][16]
    uint8_t* kind;    // [n_trees][16]
    uint32_t* w;      // [n_trees] simulations of the pending batch
    uint32_t* done;   // [n_trees] simulations finished since tafl_mcts_begin
};
enum { VK_NONE = 0, VK_TERM, VK_ROOT, VK_EMPTY, VK_NEW, VK_COLL };

template <int G>
__device__ __forceinline__ void grp_puct_select_priors_vl(const MctsEdge* __restrict__ ed, const double* __restrict__ prior, uint32_t L,
                                                          uint32_t ns_raw, double cpuct, unsigned gmask, int gl, uint32_t& out_a,
                                                          uint32_t& out_Nraw, uint32_t& out_child) {
    const uint32_t ns_eff = (ns_raw & MCTS_VL_MASK) + (ns_raw >> 24);
    const double sq_vis = sqrt((double)ns_eff), sq_new = sqrt((double)ns_eff + MCTS_EPS);
    double best_u = -INFINITY;
    uint32_t best_a = 0xFFFFFFFFu, best_N = 0, best_child = 0;
    for (uint32_t a0 = gl; a0 < L; a0 += 4 * G) {
        MctsEdge e4[4];
        double p4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) { e4[j] = ed[a]; p4[j] = prior[a]; }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t a = a0 + (uint32_t)(G * j);
            if (a < L) {
                const uint32_t N = e4[j].N & MCTS_VL_MASK, vl = e4[j].N >> 24, nn = N + vl;
                const double c = cpuct * p4[j];
                double u;
                if (nn > 0) {
                    const double qv = vl ? ((double)N * e4[j].Q - (double)vl) / (double)nn : e4[j].Q;
                    u = qv + (c * sq_vis) / (double)(1u + nn);
                } else {
                    u = c * sq_new;
                }
                if (u > best_u) { best_u = u; best_a = a; best_N = e4[j].N; best_child = e4[j].child; }
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        const double ou = __shfl_xor_sync(gmask, best_u, o, G);
        const uint32_t oa = __shfl_xor_sync(gmask, best_a, o, G);
        const uint32_t on = __shfl_xor_sync(gmask, best_N, o, G);
        const uint32_t oc = __shfl_xor_sync(gmask, best_child, o, G);
        if (ou > best_u || (ou == best_u && oa < best_a)) { best_u = ou; best_a = oa; best_N = on; best_child = oc; }
    }
    out_a = best_a; out_Nraw = best_N; out_child = best_child;
}

template <class BB, int SP>
__global__ void __launch_bounds__(MCTS_WARPS * 32, 4)
k_mcts_hook_vl(const __grid_constant__ Ctx<BB, SP> cx, MctsView mv, uint32_t n_trees, uint32_t sims, uint32_t W, double cpuct,
               double draw_value, tafl_state* __restrict__ l_state, uint8_t* __restrict__ l_planes, tafl_move* __restrict__ l_moves,
               uint32_t* __restrict__ l_counts, const float* __restrict__ l_priors, float* __restrict__ l_values, VlBatch vb,
               uint32_t* __restrict__ any_overflow, int do_expand, int do_select) {
    constexpr int G = 16, TPW = 32 / G;
    const int lane = threadIdx.x & 31, gl = lane % G, gshift = (lane / G) * G;
    const unsigned gmask = ((1u << G) - 1u) << gshift;
    const uint32_t warp_tree0 = (blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5)) * TPW;
    const uint32_t tree_raw = warp_tree0 + (uint32_t)(lane / G);
    const bool live = tree_raw < n_trees;
    const uint32_t tree = live ? tree_raw : 0u;
    uint4* const n_hdr = mv.n_hdr + (size_t)tree * mv.node_cap;
    MctsEdge* const edges = mv.edges + (size_t)tree * mv.edge_cap;
    unsigned long long c_sims = 0, c_nodes = 0, c_depth = 0, c_term = 0;
    // ---- phase 1: the batch evaluated since the last round
    if (do_expand) {
        // expansions, simulation order inside a tree; np.sum needs the full warp, so the warp's trees take turns
#pragma unroll 1
        for (int t = 0; t < TPW; ++t) {
            const uint32_t tr = warp_tree0 + (uint32_t)t;
            if (tr >= n_trees) continue;
            const uint32_t w = vb.w[tr];
            const int blk = lane >> 3, acc = lane & 7;
#pragma unroll 1
            for (uint32_t i = 0; i < w; ++i) {
                const int kind = (int)vb.kind[(size_t)tr * 16 + i];
                if (kind != VK_NEW && kind != VK_ROOT) continue;
                const uint32_t leaf = vb.leaf[(size_t)tr * 16 + i];
                const uint4 h = mv.n_hdr[(size_t)tr * mv.node_cap + leaf];
                if (hdr_es(h.w) != ES_ONGOING || hdr_expanded(h.w)) continue;   // terminal child / empty expansion: no evaluation
                const float* const pf = l_priors + ((size_t)tr * W + i) * TAFL_MAX_MOVES;
                float pfv[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) pfv[k] = pf[blk * 128 + acc + 8 * k];
                warp_expand_leaf(mv, tr, leaf, pfv, pf, lane, false);
                __syncwarp();
            }
        }
        __syncwarp();
        // backups, simulation order, every tree's own group
        const uint32_t w = live ? vb.w[tree] : 0u;
        const int my_kind = (live && (uint32_t)gl < w) ? (int)vb.kind[(size_t)tree * 16 + gl] : VK_NONE;
        const uint32_t my_leaf = (live && (uint32_t)gl < w) ? vb.leaf[(size_t)tree * 16 + gl] : 0u;
        for (uint32_t i = 0; i < w; ++i) {
            const int kind = __shfl_sync(gmask, my_kind, (int)i, G);
            const uint32_t leaf = __shfl_sync(gmask, my_leaf, (int)i, G);
            if (kind == VK_NONE) continue;
            double v = 0.0;
            if (kind != VK_EMPTY) {
                const uint32_t es = hdr_es(n_hdr[leaf].w);
                if (es != ES_ONGOING) {
                    v = -(es == ES_WIN ? 1.0 : (es == ES_LOSS ? -1.0 : draw_value));
                } else {
                    uint32_t src = i;   // a collision takes the value of the simulation that opened the child
                    if (kind == VK_COLL) {
                        const unsigned opener = __ballot_sync(gmask, my_kind == VK_NEW && my_leaf == leaf && (uint32_t)gl < i) >> gshift;
                        src = opener ? (uint32_t)(__ffs((int)opener) - 1) : i;
                    }
                    v = -(double)l_values[(size_t)tree * W + src];
                }
            }
            grp_backup_vl(n_hdr, edges, leaf, v, gmask, gl);
            c_sims += 1;
        }
        if (gl == 0 && live && w) {
            vb.done[tree] += w;
            vb.w[tree] = 0u;
        }
        __syncwarp();
    }
    if (do_select) {
        // ---- phase 2: the next batch (run_from_root_vl, select sweep), G lanes per tree
        tafl_state* const n_state = mv.n_state + (size_t)tree * mv.node_cap;
        uint32_t* const n_ptab = mv.n_ptab + (size_t)tree * mv.node_cap * MCTS_PT_MAX;
        const double* const prior = mv.e_prior + (size_t)tree * mv.edge_cap;
        uint32_t overflow = 0, n_nodes = mv.t_nodes[tree];
        if (W > (uint32_t)G) W = (uint32_t)G;
        uint32_t w = 0;
        if (live && mv.t_overflow[tree] == 0) {
            const uint32_t done = vb.done[tree];
            if (done < sims) {
                w = sims - done < W ? sims - done : W;
                const uint4 h0 = n_hdr[0];
                if (!hdr_expanded(h0.w) && hdr_es(h0.w) == ES_ONGOING) w = 1;   // one simulation per batch while the root is unexpanded
            }
        }
        int my_kind = VK_NONE;
        uint32_t my_leaf = 0, my_parent = 0, my_pedge = 0;
        tafl_move my_move = 0;
        for (uint32_t i = 0; i < w; ++i) {
            int kind = VK_NONE;
            uint32_t node = 0, depth = 0, leaf = 0, parent = 0, pedge = 0;
            tafl_move mvv = 0;
            while (true) {
                const uint4 h = n_hdr[node];
                if (hdr_es(h.w) != ES_ONGOING) { kind = VK_TERM; leaf = node; c_term += 1; break; }   // mcts.py:77-81
                if (!hdr_expanded(h.w)) { kind = VK_ROOT; leaf = node; break; }
                const uint32_t L = hdr_n_edges(h.w);
                if (L == 0) { kind = VK_EMPTY; leaf = node; break; }
                if (gl < 3) {   // the node's state and piece table, wanted if the descent ends here with a new child
                    const void* pf = gl == 0 ? (const void*)(n_state + node) : (const void*)(n_ptab + (size_t)node * MCTS_PT_MAX + (gl - 1) * 32);
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
                }
                uint32_t a, bNraw, bchild;
                grp_puct_select_priors_vl<G>(edges + h.x, prior + h.x, L, h.y, cpuct, gmask, gl, a, bNraw, bchild);
                depth += 1;
                const bool fresh = (bNraw & MCTS_VL_MASK) == 0;        // (s,a) not in Qsa
                const bool collide = fresh && (bNraw >> 24) != 0;      // ... and an earlier simulation of this batch is opening it
                uint32_t child = bchild;
                if (fresh && !collide) {
                    if (n_nodes >= mv.node_cap) { overflow = OVF_NODES; break; }
                    child = n_nodes++;
                    c_nodes += 1;
                    const tafl_state ps = load_state(n_state + node);
                    mvv = grp_decode_edge<G>(cx, n_ptab + (size_t)node * MCTS_PT_MAX, (int)hdr_n_pieces(h.w),
                                             (ps.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, (int)a, gmask, gshift, gl);
                }
                __syncwarp(gmask);
                if (gl == 0) {   // the virtual loss of this simulation on (s,a) and on s
                    MctsEdge* const ep = edges + h.x + a;
                    ep->N = bNraw + MCTS_VL_ONE;
                    if (fresh && !collide) ep->child = child;
                    reinterpret_cast<uint32_t*>(n_hdr + node)[1] = h.y + MCTS_VL_ONE;
                }
                __syncwarp(gmask);
                if (fresh) {
                    kind = collide ? VK_COLL : VK_NEW;
                    leaf = child; parent = node; pedge = a;
                    break;
                }
                node = child;
            }
            if (overflow) break;
            if (kind != VK_EMPTY) c_depth += depth;
            if (gl == (int)i) { my_kind = kind; my_leaf = leaf; my_parent = parent; my_pedge = pedge; my_move = mvv; }
        }
        if (overflow) { w = 0; my_kind = VK_NONE; }
        __syncwarp();   // ---- create: every new child of every tree of the warp, one lane each
        if (my_kind == VK_NEW) {
            tafl_state s = load_state(n_state + my_parent);
            do_valid_move(cx, s, my_move);
            store_state(n_state + my_leaf, s);
            n_hdr[my_leaf] = make_uint4(0u, 0u, my_parent, hdr_pack(0, my_pedge, 0, es_code(s)));
        }
        __syncwarp();
        // ---- leaf rows, one simulation after the other, group-cooperative
        uint32_t pend_edges = 0;
        for (uint32_t i = 0; i < W; ++i) {
            const int kind = __shfl_sync(gmask, my_kind, (int)i, G);
            const uint32_t leaf = __shfl_sync(gmask, my_leaf, (int)i, G);
            const size_t row = (size_t)tree * W + i;
            uint32_t want = 0;
            if (i < w && !overflow && (kind == VK_NEW || kind == VK_ROOT || kind == VK_COLL)) {
                const uint4 h = n_hdr[leaf];
                const uint32_t es = hdr_es(h.w);
                if (es != ES_ONGOING) {
                    c_term += 1;   // the new (or collided) child is terminal
                } else if (kind != VK_COLL) {
                    const tafl_state s = load_state(n_state + leaf);
                    const Pos<BB> p = load_pos(cx, s);
                    const int def_side = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
                    uint32_t* const ptab = n_ptab + (size_t)leaf * MCTS_PT_MAX;
                    int n_pieces;
                    const int L = grp_build_ptab<G>(cx, p, def_side, ptab, gmask, gl, &n_pieces);
                    if ((uint32_t)L > mv.edge_cap - mv.t_edges[tree] - pend_edges || n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) {
                        overflow = (n_pieces > MCTS_PT_MAX || L > TAFL_MAX_MOVES) ? OVF_PIECES : OVF_EDGES;
                    } else if (L == 0) {   // Ongoing position without plays: unreachable through do_play; expand empty, v = 0
                        if (gl == 0) {
                            n_hdr[leaf] = make_uint4(mv.t_edges[tree], 0u, h.z, hdr_pack(0u, hdr_pedge(h.w), 1u, es, (uint32_t)n_pieces));
                            l_values[row] = 0.0f;
                        }
                    } else {
                        if (gl == 0) n_hdr[leaf] = make_uint4(0u, 0u, h.z, hdr_pack((uint32_t)L, hdr_pedge(h.w), 0u, es, (uint32_t)n_pieces));
                        grp_emit_leaf_row<BB, SP, G>(cx, s, p, def_side, ptab, n_pieces, l_state + row, l_planes + row * (size_t)(cx.n() * cx.n()),
                                                     l_moves + row * TAFL_MAX_MOVES, gmask, gl);
                        want = (uint32_t)L;
                        pend_edges += (uint32_t)L;
                    }
                }
            }
            if (gl == 0 && live) l_counts[row] = want;
            __syncwarp(gmask);
        }
        if (live) {
            if ((uint32_t)gl < w && !overflow) {
                vb.kind[(size_t)tree * 16 + gl] = (uint8_t)my_kind;
                vb.leaf[(size_t)tree * 16 + gl] = my_leaf;
            }
            if (gl == 0) {
                vb.w[tree] = overflow ? 0u : w;
                mv.t_nodes[tree] = n_nodes;
                if (overflow) { mv.t_overflow[tree] = overflow; atomicOr(any_overflow, overflow); }
            }
        }
    }
    if (gl == 0 && live) {
        if (c_sims) atomicAdd(&mv.stats[ST_SIMS], c_sims);
        if (c_nodes) atomicAdd(&mv.stats[ST_NODES], c_nodes);
        if (c_depth) atomicAdd(&mv.stats[ST_DEPTH], c_depth);
        if (c_term) atomicAdd(&mv.stats[ST_TERMINAL], c_term);
    }
}

// built-in evaluators (device side of TAFL_EVAL_STUB / TAFL_EVAL_HASH), one warp per leaf
__global__ void __launch_bounds__(MCTS_WARPS * 32)
k_eval_builtin(int kind, uint64_t seed, const tafl_state* __restrict__ l_state, const uint32_t* __restrict__ l_counts, uint32_t n,
               float* __restrict__ priors, float* __restrict__ values) {
    const int lane = threadIdx.x & 31;
    const uint32_t g = blockIdx.x * MCTS_WARPS + (threadIdx.x >> 5);
    if (g >= n || l_counts[g] == 0) return;
    float* const P = priors + (size_t)g * TAFL_MAX_MOVES;
    if (kind == TAFL_EVAL_STUB) {   // uniform over the action space, value 0 (the north-star stub)
        for (int a = lane; a < TAFL_MAX_MOVES; a += 32) P[a] = 1.0f / (float)TAFL_MAX_MOVES;
        if (lane == 0) values[g] = 0.0f;
        return;
    }
    const tafl_state s = load_state(l_state + g);
    const uint64_t side = (s.flags & TAFL_FLAG_DEF_TO_PLAY) ? 1u : 0u;
    uint64_t h = splitmix64(s.def[1] ^ side);
    h = splitmix64(s.def[0] ^ h);
    h = splitmix64(s.att[1] ^ h);
    h = splitmix64(s.att[0] ^ h);
    h = splitmix64(seed ^ h);
    const bool masked = (h & 15u) == 0;
    for (int a = lane; a < TAFL_MAX_MOVES; a += 32) {
        const uint64_t x = splitmix64(h + (uint64_t)(a + 1) * 0x9E3779B97F4A7C15ull);
        P[a] = masked ? 0.0f : (float)((x >> 40) + 1) * 5.9604644775390625e-08f;
    }
    if (lane == 0) {
        const uint64_t y = splitmix64(h ^ 0xD1B54A32D192ED03ull);
        values[g] = ((float)(long long)(y >> 40) - 8388608.0f) * 5.9604644775390625e-08f;
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace tafl {

// W leaves per tree and round through the evaluator hook (1 = the sequential search of mcts.py)
static inline uint32_t hook_width(const tafl_mcts_config& c) { return c.vl_width > 1 ? c.vl_width : 1u; }
// rounds a hook search needs: a fresh tree spends one batch of one simulation on its root, then ceil((sims - 1) / W)
static inline uint32_t hook_rounds(const tafl_mcts_config& c) {
    const uint32_t W = hook_width(c);
    return W == 1 ? c.num_sims : (c.num_sims ? 1u + (c.num_sims - 1u + W - 1u) / W : 0u);
}
// the evaluator-hook buffers (sized by vl_width) and the graph that holds their addresses
static void eval_free(MctsPools* m) {
    if (m->round_graph) { cudaGraphExecDestroy(m->round_graph); m->round_graph = nullptr; }
    cudaFree(m->v.e_prior); cudaFree(m->v.t_pending);
    cudaFree(m->l_state); cudaFree(m->l_planes); cudaFree(m->l_moves); cudaFree(m->l_counts); cudaFree(m->l_priors); cudaFree(m->l_values);
    cudaFree(m->b_leaf); cudaFree(m->b_kind); cudaFree(m->b_w); cudaFree(m->b_done);
    m->v.e_prior = nullptr; m->v.t_pending = nullptr;
    m->l_state = nullptr; m->l_planes = nullptr; m->l_moves = nullptr; m->l_counts = nullptr; m->l_priors = nullptr; m->l_values = nullptr;
    m->b_leaf = nullptr; m->b_kind = nullptr; m->b_w = nullptr; m->b_done = nullptr;
}
void mcts_free(tafl_engine* e) {
    MctsPools* m = e->mcts;
    if (!m) return;
    cudaFree(m->v.n_state); cudaFree(m->v.n_hdr); cudaFree(m->v.n_ptab); cudaFree(m->v.edges); cudaFree(m->v.t_nodes); cudaFree(m->v.t_edges); cudaFree(m->v.t_overflow); cudaFree(m->v.stats);
    cudaFree(m->r_moves); cudaFree(m->r_visits); cudaFree(m->r_q); cudaFree(m->r_counts); cudaFree(m->r_best); cudaFree(m->r_probs); cudaFree(m->remap);
    cudaFree(m->d_any_overflow); cudaFree(m->row_of);
    eval_free(m);
    cudaFree(m->d_noise); cudaFree(m->t_noised);
    delete m;
    e->mcts = nullptr;
}

}  // namespace tafl

#define MCTS_WIDTH_DISPATCH(e, body)                                                        \
    do {                                                                                    \
        if ((e)->width == 64) { using BB = BB64; const Ctx<BB64>& cx = (e)->cx64; body; }   \
        else { using BB = BB128; const Ctx<BB128>& cx = (e)->cx128; body; }                 \
    } while (0)
#define MCTS_LAUNCHED(e)                         \
    do {                                         \
        (e)->launches += 1;                      \
        TAFL_CUDA_TRY((e), cudaGetLastError());  \
    } while (0)

// fresh trees, or (tree reuse) carried trees whose root still is the game's state
static void launch_tree_init(tafl_engine* e, int tree_mode) {
    MctsPools* m = e->mcts;
    // every tree is re-initialised (or re-validated) below: the engine-wide overflow word describes THIS search only
    cudaMemsetAsync(m->d_any_overflow, 0, sizeof(uint32_t), e->stream);
    if ((m->cfg.flags & TAFL_MCTS_TREE_REUSE) && m->tree_mode != tree_mode)   // a carried tree of the other layout is dropped
        cudaMemsetAsync(m->v.t_nodes, 0, (size_t)e->n_games * sizeof(uint32_t), e->stream);
    m->tree_mode = tree_mode;
    if (m->cfg.flags & TAFL_MCTS_TREE_REUSE)
        k_mcts_init_carry<<<grid_for(e->n_games, 128), 128, 0, e->stream>>>(e->d_states, e->n_games, m->v);
    else
        k_mcts_init<<<grid_for(e->n_games, 128), 128, 0, e->stream>>>(e->d_states, e->n_games, m->v);
}


// engine-wide overflow word -> error (the word is cleared at the start of every search)
static int overflow_error(tafl_engine* e, uint32_t of) {
    if (!of) return TAFL_OK;
    std::string msg = "MCTS search stopped early in at least one tree:";
    if (of & OVF_EDGES) msg += " edge pool exhausted (raise tafl_mcts_config.edge_capacity);";
    if (of & OVF_NODES) msg += " node pool exhausted (raise tafl_mcts_config.node_capacity; with tree reuse the carried subtree counts);";
    if (of & OVF_PIECES) msg += " a position has more pieces / legal moves than a node can hold;";
    e->err = msg;
    return TAFL_ERR_CAPACITY;
}

static int need_mcts(tafl_engine* e) {
    if (!e) return TAFL_ERR_INVALID_ARG;
    if (!e->mcts) { e->err = "tafl_mcts_configure has not been called"; return TAFL_ERR_INVALID_ARG; }
    return TAFL_OK;
}

extern "C" {

int tafl_mcts_configure(tafl_engine* e, const tafl_mcts_config* cfg) {
    if (!e || !cfg || cfg->num_sims == 0 || !(cfg->draw_value != 0.0)) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    tafl_mcts_config c = *cfg;
    const bool reuse = (c.flags & TAFL_MCTS_TREE_REUSE) != 0;
    if (c.node_capacity == 0) c.node_capacity = (c.num_sims + 1) * (reuse ? 4u : 1u);
    if (c.node_capacity < c.num_sims + 1) return TAFL_ERR_INVALID_ARG;
    if (c.edge_capacity == 0) c.edge_capacity = c.node_capacity * 160u;
    if (c.vl_width > 16) { e->err = "tafl_mcts_configure: vl_width is at most 16"; return TAFL_ERR_INVALID_ARG; }
    if (e->mcts && e->mcts->cfg.num_sims == c.num_sims && e->mcts->cfg.edge_capacity == c.edge_capacity &&
        e->mcts->cfg.node_capacity == c.node_capacity && e->mcts->cfg.flags == c.flags) {
        if (hook_width(c) != e->mcts->l_width && e->mcts->v.e_prior) {   // the leaf batch is sized by vl_width
            TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
            eval_free(e->mcts);
        }
        e->mcts->cfg = c;   // same pools, possibly new cpuct / draw value
        return TAFL_OK;
    }
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    mcts_free(e);
    MctsPools* m = new MctsPools();
    m->cfg = c;
    e->mcts = m;
    const size_t n = e->n_games, nc = (size_t)c.node_capacity, ec = c.edge_capacity;
    m->v.node_cap = (uint32_t)nc;
    m->v.edge_cap = (uint32_t)ec;
#define MCTS_ALLOC(ptr, bytes)                                          \
    do {                                                                \
        cudaError_t _ce = cudaMalloc((void**)&(ptr), (bytes));          \
        if (_ce != cudaSuccess) {                                       \
            int _rc = set_cuda_error(e, _ce, "cudaMalloc(" #ptr ")");   \
            mcts_free(e);                                               \
            return _rc;                                                 \
        }                                                               \
    } while (0)
    MCTS_ALLOC(m->v.n_state, n * nc * sizeof(tafl_state));
    MCTS_ALLOC(m->v.n_hdr, n * nc * sizeof(uint4));
    MCTS_ALLOC(m->v.n_ptab, n * nc * MCTS_PT_MAX * sizeof(uint32_t));
    MCTS_ALLOC(m->v.edges, n * ec * sizeof(MctsEdge));
    MCTS_ALLOC(m->v.t_nodes, n * sizeof(uint32_t));
    MCTS_ALLOC(m->v.t_edges, n * sizeof(uint32_t));
    MCTS_ALLOC(m->v.t_overflow, n * sizeof(uint32_t));
    MCTS_ALLOC(m->v.stats, ST_COUNT * sizeof(unsigned long long));
    MCTS_ALLOC(m->r_moves, n * TAFL_MAX_MOVES * sizeof(tafl_move));
    MCTS_ALLOC(m->r_visits, n * TAFL_MAX_MOVES * sizeof(uint32_t));
    MCTS_ALLOC(m->r_q, n * TAFL_MAX_MOVES * sizeof(double));
    MCTS_ALLOC(m->r_counts, n * sizeof(uint32_t));
    MCTS_ALLOC(m->r_best, n * sizeof(uint32_t));
    MCTS_ALLOC(m->d_any_overflow, sizeof(uint32_t));
    if (reuse) MCTS_ALLOC(m->remap, n * nc * sizeof(uint32_t));
#undef MCTS_ALLOC
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->v.t_nodes, 0, n * sizeof(uint32_t), e->stream));   // no tree is carried yet
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->v.stats, 0, ST_COUNT * sizeof(unsigned long long), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->d_any_overflow, 0, sizeof(uint32_t), e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return TAFL_OK;
}

int tafl_mcts_search_async(tafl_engine* e) {
    int rc = need_mcts(e);
    if (rc) return rc;
    cudaSetDevice(e->device);
    MctsPools* m = e->mcts;
    if (m->noise_eps != 0.0) {
        e->err = "tafl_mcts_search: root noise is set, but the fused stub search has implicit uniform priors (use the evaluator path)";
        return TAFL_ERR_INVALID_ARG;
    }
    const uint32_t n = e->n_games;
    launch_tree_init(e, (m->cfg.vl_width > 1 || e->mcts_group == 32 || e->mcts_group == 16 || e->mcts_group == 8) ? 1 : 0);
    MCTS_LAUNCHED(e);
    const unsigned grid = grid_for(n, MCTS_WARPS);
#define TAFL_MCTS_LAUNCH(MINB)                                                                                          \
    TAFL_SPEC_DISPATCH(e, (k_mcts_sims<BB, SP, MINB><<<grid, MCTS_WARPS * 32, 0, e->stream>>>(cx, m->v, n, m->cfg.num_sims,   \
                                                                                           m->cfg.cpuct, m->cfg.draw_value, \
                                                                                           m->d_any_overflow)))
#define TAFL_MCTS_LAUNCH_G(G, MINB)                                                                                         \
    TAFL_SPEC_DISPATCH(e, (k_mcts_sims_g<BB, SP, G, MINB><<<grid_for(n, SIMS_WARPS * (32 / G)), SIMS_WARPS * 32, 0, e->stream>>>( \
                              cx, m->v, n, m->cfg.num_sims, m->cfg.cpuct, m->cfg.draw_value, m->d_any_overflow)))
    if (m->cfg.vl_width > 1) {        // virtual loss: vl_width (<= 16) simulations in flight per tree, 16 lanes per tree
        TAFL_SPEC_DISPATCH(e, (k_mcts_sims_vl<BB, SP, 16><<<grid_for(n, MCTS_WARPS * 2), MCTS_WARPS * 32, 0, e->stream>>>(
                                  cx, m->v, n, m->cfg.num_sims, m->cfg.vl_width, m->cfg.cpuct, m->cfg.draw_value, m->d_any_overflow)));
    } else if (e->mcts_group == 16) {        // lanes per tree (TAFL_MCTS_G): 2 trees per warp
        TAFL_MCTS_LAUNCH_G(16, 4);
    } else if (e->mcts_group == 8) {  // 4 trees per warp
        TAFL_MCTS_LAUNCH_G(8, 4);
    } else if (e->mcts_group == 32) { // one tree per warp, converged create + expand
        TAFL_MCTS_LAUNCH_G(32, 4);
    } else {
        switch (e->mcts_minb) {   // the first-generation kernel: resident CTAs of 4 warps per SM (register cap)
            case 4: TAFL_MCTS_LAUNCH(4); break;
            case 6: TAFL_MCTS_LAUNCH(6); break;
            case 10: TAFL_MCTS_LAUNCH(10); break;
            default: TAFL_MCTS_LAUNCH(8); break;
        }
    }
#undef TAFL_MCTS_LAUNCH
#undef TAFL_MCTS_LAUNCH_G
    MCTS_LAUNCHED(e);
    TAFL_SPEC_DISPATCH(e, (k_mcts_results<BB, SP><<<grid, MCTS_WARPS * 32, 0, e->stream>>>(cx, m->v, n, m->r_moves, m->r_visits, m->r_q,
                                                                                        m->r_counts, m->r_best, m->tree_mode)));
    MCTS_LAUNCHED(e);
    m->searched = true;
    return TAFL_OK;
}

int tafl_mcts_search(tafl_engine* e) {
    int rc = tafl_mcts_search_async(e);
    if (rc) return rc;
    uint32_t of = 0;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&of, e->mcts->d_any_overflow, sizeof(of), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return overflow_error(e, of);
}

int tafl_mcts_results_device(tafl_engine* e, tafl_move** d_moves, uint32_t** d_visits, double** d_q, uint32_t** d_counts,
                             uint32_t** d_best) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (d_moves) *d_moves = m->r_moves;
    if (d_visits) *d_visits = m->r_visits;
    if (d_q) *d_q = m->r_q;
    if (d_counts) *d_counts = m->r_counts;
    if (d_best) *d_best = m->r_best;
    return TAFL_OK;
}

int tafl_mcts_results_host(tafl_engine* e, uint32_t first, uint32_t n, tafl_move* h_moves, uint32_t* h_visits, double* h_q,
                           uint32_t* h_counts, uint32_t* h_best) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if ((uint64_t)first + n > e->n_games) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    MctsPools* m = e->mcts;
    const size_t o = (size_t)first * TAFL_MAX_MOVES, k = (size_t)n * TAFL_MAX_MOVES;
    if (h_moves) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_moves, m->r_moves + o, k * sizeof(tafl_move), cudaMemcpyDeviceToHost, e->stream));
    if (h_visits) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_visits, m->r_visits + o, k * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    if (h_q) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_q, m->r_q + o, k * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    if (h_counts) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_counts, m->r_counts + first, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    if (h_best) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_best, m->r_best + first, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return TAFL_OK;
}

int tafl_mcts_play_best(tafl_engine* e, uint8_t* h_status) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!e->mcts->searched) { e->err = "tafl_mcts_search has not been run"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    MctsPools* m = e->mcts;
    const uint32_t n = e->n_games;
    TAFL_SPEC_DISPATCH(e, (k_mcts_play_best<BB, SP><<<grid_for(n, 64), 64, 0, e->stream>>>(cx, e->d_states, n, m->r_moves, m->r_counts,
                                                                                       m->r_best, e->d_status)));
    MCTS_LAUNCHED(e);
    hist_forget_all(e);   // not a Game::do_play: see tafl_engine.cu
    if (h_status) {
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_status, e->d_status, n, cudaMemcpyDeviceToHost, e->stream));
        TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    }
    return TAFL_OK;
}

int tafl_mcts_reroot(tafl_engine* e) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (!(m->cfg.flags & TAFL_MCTS_TREE_REUSE)) { e->err = "tafl_mcts_reroot: configure with TAFL_MCTS_TREE_REUSE"; return TAFL_ERR_INVALID_ARG; }
    if (!m->searched) { e->err = "tafl_mcts_reroot: no search results to re-root at"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    k_mcts_reroot<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, n, m->r_counts, m->r_best, m->remap, m->tree_mode);
    MCTS_LAUNCHED(e);
    m->searched = false;   // the root results describe the previous root
    return TAFL_OK;
}

int tafl_mcts_advance(tafl_engine* e, uint8_t* h_status) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!(e->mcts->cfg.flags & TAFL_MCTS_TREE_REUSE)) { e->err = "tafl_mcts_advance: configure with TAFL_MCTS_TREE_REUSE"; return TAFL_ERR_INVALID_ARG; }
    rc = tafl_mcts_play_best(e, nullptr);
    if (rc) return rc;
    rc = tafl_mcts_reroot(e);
    if (rc) return rc;
    if (h_status) {
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_status, e->d_status, e->n_games, cudaMemcpyDeviceToHost, e->stream));
        TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    }
    return TAFL_OK;
}

int tafl_mcts_policy(tafl_engine* e, double temp, int choose, uint64_t seed, double* h_probs) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (!m->searched) { e->err = "tafl_mcts_policy: no search results"; return TAFL_ERR_INVALID_ARG; }
    if (!(temp >= 0.0)) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    if (h_probs && !m->r_probs) {
        cudaError_t ce = cudaMalloc((void**)&m->r_probs, (size_t)n * TAFL_MAX_MOVES * sizeof(double));
        if (ce != cudaSuccess) return set_cuda_error(e, ce, "cudaMalloc(r_probs)");
    }
    k_mcts_policy<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, n, e->game_id_base, m->r_visits, m->r_counts, temp,
                                                                             seed, choose, h_probs ? m->r_probs : nullptr, m->r_best);
    MCTS_LAUNCHED(e);
    if (h_probs) {
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_probs, m->r_probs, (size_t)n * TAFL_MAX_MOVES * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
        TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    }
    return TAFL_OK;
}

int tafl_mcts_stats_get(tafl_engine* e, tafl_mcts_stats* out) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!out) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    unsigned long long h[ST_COUNT];
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h, e->mcts->v.stats, sizeof(h), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    out->sims = h[ST_SIMS]; out->nodes = h[ST_NODES]; out->depth_sum = h[ST_DEPTH]; out->edges = h[ST_EDGES];
    out->terminal_hits = h[ST_TERMINAL]; out->max_edges_used = h[ST_MAXEDGES];
    return TAFL_OK;
}

int tafl_pack_records(tafl_engine* e, void* d_out) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!d_out || !e->mcts->searched) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    MctsPools* m = e->mcts;
    const uint32_t n = e->n_games;
    k_pack_records<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, e->d_states, n, e->game_id_base, m->r_moves,
                                                                              m->r_visits, m->r_counts, m->r_best, e->d_episode, nullptr,
                                                                              (uint8_t*)d_out);
    MCTS_LAUNCHED(e);
    return TAFL_OK;
}

int tafl_pack_records_compact(tafl_engine* e, void* d_out, uint32_t* d_count, uint32_t* h_count) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!d_out || !e->mcts->searched) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    MctsPools* m = e->mcts;
    const uint32_t n = e->n_games;
    if (!m->row_of) TAFL_CUDA_TRY(e, cudaMalloc((void**)&m->row_of, ((size_t)n + 1) * sizeof(uint32_t)));
    uint32_t* h_alias = nullptr;   // pinned host word: written by the kernel itself, no copy-engine hand-off
    if (h_count) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, h_count) == cudaSuccess && a.type == cudaMemoryTypeHost) h_alias = (uint32_t*)a.devicePointer;
        else cudaGetLastError();
    }
    uint32_t* const d_cnt = d_count ? d_count : m->row_of + n;
    k_record_rows<<<1, 1024, 0, e->stream>>>(m->r_counts, n, m->row_of, d_cnt, h_alias, m->d_any_overflow);
    MCTS_LAUNCHED(e);
    k_pack_records<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, e->d_states, n, e->game_id_base, m->r_moves,
                                                                              m->r_visits, m->r_counts, m->r_best, e->d_episode, m->row_of,
                                                                              (uint8_t*)d_out);
    MCTS_LAUNCHED(e);
    if (h_count && !h_alias) {   // pageable destination: staged copies (the host still must not read it before the stream reaches here)
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_count, d_cnt, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_count + 1, m->d_any_overflow, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    }
    return TAFL_OK;
}

int tafl_mcts_overflow(tafl_engine* e, uint32_t* h_flags) {
    int rc = need_mcts(e);
    if (rc) return rc;
    if (!h_flags) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_flags, e->mcts->d_any_overflow, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return TAFL_OK;
}

}  // extern "C"

// ---- evaluator hook -------------------------------------------------------------------------
static int eval_alloc(tafl_engine* e) {
    MctsPools* m = e->mcts;
    if (m->v.e_prior) return TAFL_OK;
    const size_t games = e->n_games, cells = (size_t)e->N * e->N;
    m->l_width = hook_width(m->cfg);
    const size_t n = games * m->l_width;   // rows of the leaf batch
#define EVAL_ALLOC(ptr, bytes)                                                                   \
    do {                                                                                         \
        cudaError_t _ce = cudaMalloc((void**)&(ptr), (bytes));                                   \
        if (_ce != cudaSuccess) return set_cuda_error(e, _ce, "cudaMalloc(" #ptr ")");           \
    } while (0)
    EVAL_ALLOC(m->v.t_pending, games * sizeof(uint32_t));
    EVAL_ALLOC(m->b_leaf, games * 16 * sizeof(uint32_t));
    EVAL_ALLOC(m->b_kind, games * 16);
    EVAL_ALLOC(m->b_w, games * sizeof(uint32_t));
    EVAL_ALLOC(m->b_done, games * sizeof(uint32_t));
    EVAL_ALLOC(m->l_state, n * sizeof(tafl_state));
    EVAL_ALLOC(m->l_planes, n * cells);
    EVAL_ALLOC(m->l_moves, n * TAFL_MAX_MOVES * sizeof(tafl_move));
    EVAL_ALLOC(m->l_counts, n * sizeof(uint32_t));
    EVAL_ALLOC(m->l_priors, n * TAFL_MAX_MOVES * sizeof(float));
    EVAL_ALLOC(m->l_values, n * sizeof(float));
    EVAL_ALLOC(m->v.e_prior, games * (size_t)m->v.edge_cap * sizeof(double));   // last: marks the set complete
#undef EVAL_ALLOC
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->l_priors, 0, n * TAFL_MAX_MOVES * sizeof(float), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->l_values, 0, n * sizeof(float), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->l_moves, 0, n * TAFL_MAX_MOVES * sizeof(tafl_move), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->l_counts, 0, n * sizeof(uint32_t), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->b_w, 0, games * sizeof(uint32_t), e->stream));
    return TAFL_OK;
}

static void fill_batch(tafl_engine* e, tafl_leaf_batch* b) {
    MctsPools* m = e->mcts;
    b->d_states = m->l_state; b->d_planes = m->l_planes; b->d_moves = m->l_moves; b->d_counts = m->l_counts;
    b->d_priors = m->l_priors; b->d_values = m->l_values;
    b->n_games = e->n_games * m->l_width; b->side_len = (uint32_t)e->N;   // rows: game * W + simulation
    b->stream = (void*)e->stream;
}

// one launch of the second-generation hook kernel (phase 1 = expand + backup of the evaluated leaves, phase 2 = descent)
static int hook_round_launch(tafl_engine* e, int do_expand, int do_select) {
    MctsPools* m = e->mcts;
    const uint32_t n = e->n_games;
    if (m->l_width > 1) {   // W leaves per tree: k_mcts_hook_vl, 16 lanes per tree
        const VlBatch vb{m->b_leaf, m->b_kind, m->b_w, m->b_done};
        TAFL_SPEC_DISPATCH(e, (k_mcts_hook_vl<BB, SP><<<grid_for(n, MCTS_WARPS * 2), MCTS_WARPS * 32, 0, e->stream>>>(
                                  cx, m->v, n, m->cfg.num_sims, m->l_width, m->cfg.cpuct, m->cfg.draw_value, m->l_state, m->l_planes,
                                  m->l_moves, m->l_counts, m->l_priors, m->l_values, vb, m->d_any_overflow, do_expand, do_select)));
        MCTS_LAUNCHED(e);
        return TAFL_OK;
    }
    const int G = (e->mcts_group == 32 || e->mcts_group == 8) ? e->mcts_group : 16;
#define TAFL_HOOK_LAUNCH(GG)                                                                                                        \
    TAFL_SPEC_DISPATCH(e, (k_mcts_hook_round<BB, SP, GG><<<grid_for(n, MCTS_WARPS * (32 / GG)), MCTS_WARPS * 32, 0, e->stream>>>(     \
                              cx, m->v, n, m->cfg.cpuct, m->cfg.draw_value, m->l_state, m->l_planes, m->l_moves, m->l_counts,       \
                              m->l_priors, m->l_values, m->d_any_overflow, do_expand, do_select)))
    if (G == 32) TAFL_HOOK_LAUNCH(32);
    else if (G == 8) TAFL_HOOK_LAUNCH(8);
    else TAFL_HOOK_LAUNCH(16);
#undef TAFL_HOOK_LAUNCH
    MCTS_LAUNCHED(e);
    return TAFL_OK;
}

extern "C" {

int tafl_mcts_begin(tafl_engine* e) {
    int rc = need_mcts(e);
    if (rc) return rc;
    cudaSetDevice(e->device);
    rc = eval_alloc(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    launch_tree_init(e, 0);
    MCTS_LAUNCHED(e);
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->b_done, 0, (size_t)e->n_games * sizeof(uint32_t), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(m->b_w, 0, (size_t)e->n_games * sizeof(uint32_t), e->stream));
    if (m->noise_eps != 0.0) {   // carried (already expanded) roots take the noise now, fresh ones after round 1
        TAFL_CUDA_TRY(e, cudaMemsetAsync(m->t_noised, 0, e->n_games, e->stream));
        k_mcts_root_noise<<<grid_for(e->n_games, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, e->n_games, m->d_noise, m->noise_eps, m->t_noised);
        MCTS_LAUNCHED(e);
    }
    m->phase = 1;
    m->rounds = 0;
    m->searched = false;
    return TAFL_OK;
}

int tafl_mcts_root_noise(tafl_engine* e, const double* noise, double eps) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase == 2) { e->err = "tafl_mcts_root_noise: a leaf batch is pending"; return TAFL_ERR_INVALID_ARG; }
    if (!(eps >= 0.0 && eps <= 1.0)) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    if (!noise || eps == 0.0) { m->noise_eps = 0.0; return TAFL_OK; }
    const size_t bytes = (size_t)e->n_games * TAFL_MAX_MOVES * sizeof(double);
    if (!m->d_noise) {
        TAFL_CUDA_TRY(e, cudaMalloc((void**)&m->d_noise, bytes));
        TAFL_CUDA_TRY(e, cudaMalloc((void**)&m->t_noised, e->n_games));
    }
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(m->d_noise, noise, bytes, cudaMemcpyDefault, e->stream));   // host or device source
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    m->noise_eps = eps;
    return TAFL_OK;
}

int tafl_mcts_select_leaves(tafl_engine* e, tafl_leaf_batch* out) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase != 1) { e->err = "tafl_mcts_select_leaves: call tafl_mcts_begin / tafl_mcts_expand_backup first"; return TAFL_ERR_INVALID_ARG; }
    if (m->rounds >= hook_rounds(m->cfg)) { e->err = "tafl_mcts_select_leaves: num_sims simulations already run (node pool is sized for num_sims)"; return TAFL_ERR_CAPACITY; }
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    if (e->hook_gen >= 2 || m->l_width > 1) {
        rc = hook_round_launch(e, 0, 1);
        if (rc) return rc;
    } else {   // first generation: one warp per tree (TAFL_HOOK_GEN=1; kept as the comparison point)
        TAFL_SPEC_DISPATCH(e, (k_mcts_select_leaf<BB, SP><<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(
                                  cx, m->v, n, m->cfg.cpuct, m->cfg.draw_value, m->l_state, m->l_planes, m->l_moves, m->l_counts,
                                  m->d_any_overflow)));
        MCTS_LAUNCHED(e);
    }
    m->phase = 2;
    if (out) fill_batch(e, out);
    return TAFL_OK;
}

int tafl_mcts_expand_backup(tafl_engine* e) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase != 2) { e->err = "tafl_mcts_expand_backup: no leaf batch is pending"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    if (e->hook_gen >= 2 || m->l_width > 1) {
        rc = hook_round_launch(e, 1, 0);
        if (rc) return rc;
    } else {
        k_mcts_expand_backup<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, n, m->l_priors, m->l_values);
        MCTS_LAUNCHED(e);
    }
    if (m->noise_eps != 0.0 && m->rounds == 0) {   // the roots expanded by this first round
        k_mcts_root_noise<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(m->v, n, m->d_noise, m->noise_eps, m->t_noised);
        MCTS_LAUNCHED(e);
    }
    m->phase = 1;
    m->rounds += 1;
    return TAFL_OK;
}

int tafl_mcts_finish(tafl_engine* e) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase != 1) { e->err = "tafl_mcts_finish: a leaf batch is still pending (or no search was begun)"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    TAFL_SPEC_DISPATCH(e, (k_mcts_results<BB, SP><<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(
                              cx, m->v, n, m->r_moves, m->r_visits, m->r_q, m->r_counts, m->r_best, m->tree_mode)));
    MCTS_LAUNCHED(e);
    m->searched = true;
    m->phase = 0;
    uint32_t of = 0;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&of, m->d_any_overflow, sizeof(of), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return overflow_error(e, of);
}

int tafl_eval_builtin(tafl_engine* e, int evaluator, uint64_t seed) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (evaluator != TAFL_EVAL_STUB && evaluator != TAFL_EVAL_HASH) return TAFL_ERR_INVALID_ARG;
    if (!m->l_state) { e->err = "tafl_eval_builtin: no leaf batch (tafl_mcts_begin has not been called)"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games * m->l_width;
    k_eval_builtin<<<grid_for(n, MCTS_WARPS), MCTS_WARPS * 32, 0, e->stream>>>(evaluator, seed, m->l_state, m->l_counts, n, m->l_priors,
                                                                              m->l_values);
    MCTS_LAUNCHED(e);
    return TAFL_OK;
}

// host copies of the pending leaf batch / host-computed evaluator outputs (callers without a CUDA runtime)
int tafl_leaf_batch_host(tafl_engine* e, tafl_state* h_states, uint8_t* h_planes, tafl_move* h_moves, uint32_t* h_counts) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase != 2) { e->err = "tafl_leaf_batch_host: no leaf batch is pending"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const size_t n = (size_t)e->n_games * m->l_width, cells = (size_t)e->N * e->N;
    if (h_states) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_states, m->l_state, n * sizeof(tafl_state), cudaMemcpyDeviceToHost, e->stream));
    if (h_planes) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_planes, m->l_planes, n * cells, cudaMemcpyDeviceToHost, e->stream));
    if (h_moves) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_moves, m->l_moves, n * TAFL_MAX_MOVES * sizeof(tafl_move), cudaMemcpyDeviceToHost, e->stream));
    if (h_counts) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_counts, m->l_counts, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return TAFL_OK;
}
int tafl_leaf_eval_host(tafl_engine* e, const float* h_priors, const float* h_values) {
    int rc = need_mcts(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    if (m->phase != 2 || !h_priors || !h_values) { e->err = "tafl_leaf_eval_host: no leaf batch is pending"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(e->device);
    const size_t n = (size_t)e->n_games * m->l_width;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(m->l_priors, h_priors, n * TAFL_MAX_MOVES * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(m->l_values, h_values, n * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    return TAFL_OK;
}

int tafl_mcts_search_with(tafl_engine* e, tafl_evaluator_fn fn, void* user) {
    if (!fn) return TAFL_ERR_INVALID_ARG;
    int rc = tafl_mcts_begin(e);
    if (rc) return rc;
    tafl_leaf_batch b;
    for (uint32_t i = 0, rounds = hook_rounds(e->mcts->cfg); i < rounds; ++i) {
        rc = tafl_mcts_select_leaves(e, &b);
        if (rc) return rc;
        rc = fn(user, &b);
        if (rc) { e->err = "evaluator callback failed"; e->mcts->phase = 0; return rc < 0 ? rc : TAFL_ERR_INVALID_ARG; }
        rc = tafl_mcts_expand_backup(e);
        if (rc) return rc;
    }
    return tafl_mcts_finish(e);
}

struct BuiltinEval { tafl_engine* e; int kind; uint64_t seed; };
static int builtin_eval_cb(void* user, const tafl_leaf_batch*) {
    BuiltinEval* b = (BuiltinEval*)user;
    return tafl_eval_builtin(b->e, b->kind, b->seed);
}
int tafl_mcts_search_builtin(tafl_engine* e, int evaluator, uint64_t seed) {
    if (evaluator != TAFL_EVAL_STUB && evaluator != TAFL_EVAL_HASH) return TAFL_ERR_INVALID_ARG;
    int rc = tafl_mcts_begin(e);
    if (rc) return rc;
    MctsPools* m = e->mcts;
    // All num_sims rounds (select, evaluate, expand/backup) have static arguments: one CUDA graph,
    // captured once per (evaluator, seed, config), replaces 3 x num_sims launches.
    if (!m->round_graph || m->rg_kind != evaluator || m->rg_seed != seed || m->rg_sims != m->cfg.num_sims ||
        m->rg_cpuct != m->cfg.cpuct || m->rg_draw != m->cfg.draw_value || m->rg_spec != e->spec || m->rg_eps != m->noise_eps ||
        m->rg_gen != (int)(e->hook_gen * 2 + e->hook_fuse + 8 * m->l_width)) {
        if (m->round_graph) { cudaGraphExecDestroy(m->round_graph); m->round_graph = nullptr; }
        TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        const uint64_t before = e->launches;
        cudaGraph_t graph = nullptr;
        TAFL_CUDA_TRY(e, cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        const bool fuse = (e->hook_gen >= 2 || m->l_width > 1) && e->hook_fuse;
        for (uint32_t i = 0, rounds = hook_rounds(m->cfg); i < rounds && rc == TAFL_OK; ++i) {
            // round i: descent -> evaluator -> expand + backup.  Second generation: the expand + backup of round
            // i is fused into the launch that runs the descent of round i + 1 (two launches per simulation);
            // the root-noise kernel has to sit between the two after round 0, so that pair stays split.
            const bool fused_in = fuse && i > 0 && !(m->noise_eps != 0.0 && i == 1);
            if (fused_in) {
                rc = hook_round_launch(e, 1, 1);
                m->rounds += 1;
            } else {
                if (fuse && i > 0) rc = tafl_mcts_expand_backup(e);      // the split pair (root noise after round 0)
                if (!rc) rc = tafl_mcts_select_leaves(e, nullptr);
            }
            if (!rc) rc = tafl_eval_builtin(e, evaluator, seed);
            if (!rc && !fuse) rc = tafl_mcts_expand_backup(e);
            if (!rc && fuse) m->phase = 2;
        }
        if (!rc && fuse) rc = tafl_mcts_expand_backup(e);   // the last round's leaves
        cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
        m->rg_launches = e->launches - before;   // kernels in the graph
        e->launches = before;
        m->phase = 1;
        m->rounds = 0;
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        TAFL_CUDA_TRY(e, ce);
        ce = cudaGraphInstantiate(&m->round_graph, graph, 0);
        cudaGraphDestroy(graph);
        TAFL_CUDA_TRY(e, ce);
        m->rg_kind = evaluator; m->rg_seed = seed; m->rg_sims = m->cfg.num_sims; m->rg_cpuct = m->cfg.cpuct;
        m->rg_draw = m->cfg.draw_value; m->rg_spec = e->spec; m->rg_eps = m->noise_eps; m->rg_gen = (int)(e->hook_gen * 2 + e->hook_fuse + 8 * m->l_width);
    }
    TAFL_CUDA_TRY(e, cudaGraphLaunch(m->round_graph, e->stream));
    e->launches += m->rg_launches;
    m->rounds = hook_rounds(m->cfg);
    return tafl_mcts_finish(e);
}

}  // extern "C"
