// tafl_engine.hpp — internal definition of the opaque `tafl_engine` handle (include/tafl.h).
#pragma once
#include <cuda_runtime.h>

#include <cstring>
#include <string>

#include "tafl_ctx.hpp"

namespace tafl {

// device-side mirror of tafl_stats (without the checksum)
struct DevStats {
    unsigned long long steps, legal, caps, finished, plies;
    unsigned long long win[2][8];
    unsigned long long draw[2];
};

struct MctsPools;  // tafl_mcts.cu

// Game<T>::state_history / play_history (game/mod.rs:77-108) for every resident game: a ring of the
// last `cap` (state before the play, PlayRecord) pairs per game, ply-major so a push is coalesced.
struct HistView {
    tafl_state* st = nullptr;          // [cap][n_games]
    tafl_play_record* rec = nullptr;   // [cap][n_games]
    uint32_t* len = nullptr;           // [n_games] plays recorded so far (Vec::len of play_history)
    uint32_t* floor = nullptr;         // [n_games] oldest entry still held (len - floor <= cap)
    uint32_t cap = 0, n = 0;
};

// device aliases of the caller's pinned result buffers (tafl_step_status), null = not in use
struct HostMirror {
    uint8_t* status = nullptr;
    uint8_t* err = nullptr;
};

}  // namespace tafl

struct tafl_engine {
    int device = 0, width = 0, N = 0, sm_count = 0;
    uint32_t n_games = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_block = nullptr;    // blocking-sync event for host waits that may be long (tafl_gather_records_on)
    tafl_rules rules{};
    tafl_state start{};
    tafl::Ctx<tafl::BB64> cx64;
    tafl::Ctx<tafl::BB128> cx128;
    tafl_state* d_states = nullptr;
    tafl_move* d_moves = nullptr;      // [n_games][TAFL_MAX_MOVES]
    uint32_t* d_counts = nullptr;      // [n_games]
    uint8_t* d_status = nullptr;       // [n_games]
    uint32_t* d_episode = nullptr;     // [n_games] restarts of each game id (record header: (game_id, episode) names an episode)
    uint32_t* d_actions = nullptr;     // [n_games] tafl_action of the last random ply
    tafl::DevStats* d_stats = nullptr;
    unsigned long long* d_checksum = nullptr;
    uint64_t game_id_base = 0, ply = 0, launches = 0;
    int spec = 0;                      // kernel specialisation in use (tafl::SPEC_*), 0 = generic
    int spec_detected = 0;             // what the Ruleset + board matched at create time
    static constexpr int N_AUX = 4;    // streams of the pipelined host-buffer path (tafl_step_random_host)
    cudaStream_t aux[N_AUX] = {nullptr, nullptr, nullptr, nullptr};
    bool aux_ready = false;
    int host_chunks = 4;   // profiles/sweeps/r01r_e2e_modes.txt
    int host_io = 3;   // tafl_step_status with pinned caller buffers: bit 0 kernel reads the plays in place, bit 1 writes results in place
    int mcts_minb = 8;                 // k_mcts_sims variant: CTAs of 4 warps resident per SM
    int mcts_group = 16;               // TAFL_MCTS_G: lanes per tree of k_mcts_sims_g (32/16/8); 0 = first-generation k_mcts_sims
    int hook_gen = 2;                  // TAFL_HOOK_GEN: evaluator-hook kernels, 2 = k_mcts_hook_round (G lanes per tree), 1 = warp-per-tree pair
    int hook_fuse = 1;                 // TAFL_HOOK_FUSE: built-in searches fuse expand + next descent into one launch
    int lm_group = 0;                  // TAFL_LM_G: lanes per game of k_legal_moves_g (16/8; 0 = by batch size); 1 = thread-per-game k_legal_moves
    int step_minb = 8;                 // k_step_random variant: CTAs of 64 threads resident per SM (8: one wave for 65 536 games)
    // lockstep plies as ONE CUDA graph of n_steps kernel nodes joined by programmatic (PDL) edges
    int step_single = 1;               // TAFL_STEP_SINGLE=0: one-ply launches run the general multi-ply kernel
    int step_graph_on = 1;             // TAFL_STEP_GRAPH=0 falls back to plain launches
    int step_pdl_on = 1;               // TAFL_STEP_PDL=0: plain (full-serialisation) edges in the graph
    cudaGraphExec_t step_graph = nullptr;
    struct { uint64_t seed = 0, gid_base = 0; uint32_t steps = 0; int auto_reset = 0, minb = 0, spec = 0, chain = 0; } step_graph_key;
    unsigned long long* d_ply = nullptr;   // [2]: RNG ply of the first node of the graph, graph-launch sequence number (copied before each launch)
    int step_chain_on = 1;             // TAFL_STEP_CHAIN=0: every node waits for the whole previous grid (griddepcontrol.wait)
    uint32_t* d_chain = nullptr;       // [CTAs + 1] token per CTA of the chained step graph, then the time-out word
    uint32_t chain_ctas = 0;
    bool chain_dirty = false;
    unsigned long long chain_seq = 0;  // graph launches so far
    // scratch for the by-value batch API (grown on demand)
    void* d_scratch[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t scratch_bytes[7] = {0, 0, 0, 0, 0, 0, 0};
    tafl::MctsPools* mcts = nullptr;
    tafl::HistView hist;
    std::string err;
};

namespace tafl {

int set_cuda_error(tafl_engine* e, cudaError_t ce, const char* what);
int scratch(tafl_engine* e, int slot, size_t bytes, void** out);
void mcts_free(tafl_engine* e);
void hist_forget_all(tafl_engine* e);   // tafl_engine.cu

#define TAFL_CUDA_TRY(e, call)                                              \
    do {                                                                    \
        cudaError_t _ce = (call);                                           \
        if (_ce != cudaSuccess) return tafl::set_cuda_error((e), _ce, #call); \
    } while (0)

// run `body` with BB / SP / cx bound to the engine's bitfield width and kernel specialisation
#define TAFL_SPEC_DISPATCH(e, body)                                                                     \
    do {                                                                                                \
        if ((e)->width == 64) {                                                                         \
            using BB = tafl::BB64;                                                                      \
            if ((e)->spec == tafl::SPEC_BRANDUBH) {                                                     \
                constexpr int SP = tafl::SPEC_BRANDUBH;                                                 \
                const auto cx = tafl::spec_cast<SP>((e)->cx64);                                         \
                body;                                                                                   \
            } else {                                                                                    \
                constexpr int SP = tafl::SPEC_GENERIC;                                                  \
                const auto& cx = (e)->cx64;                                                             \
                body;                                                                                   \
            }                                                                                           \
        } else {                                                                                        \
            using BB = tafl::BB128;                                                                     \
            if ((e)->spec == tafl::SPEC_COPENHAGEN) {                                                   \
                constexpr int SP = tafl::SPEC_COPENHAGEN;                                               \
                const auto cx = tafl::spec_cast<SP>((e)->cx128);                                        \
                body;                                                                                   \
            } else if ((e)->spec == tafl::SPEC_TABLUT) {                                                \
                constexpr int SP = tafl::SPEC_TABLUT;                                                   \
                const auto cx = tafl::spec_cast<SP>((e)->cx128);                                        \
                body;                                                                                   \
            } else {                                                                                    \
                constexpr int SP = tafl::SPEC_GENERIC;                                                  \
                const auto& cx = (e)->cx128;                                                            \
                body;                                                                                   \
            }                                                                                           \
        }                                                                                               \
    } while (0)

inline unsigned grid_for(uint32_t n, unsigned block) { return (n + block - 1) / block; }

// 64-byte state <-> registers as four 16-byte vector accesses.  Fields are packed/unpacked
// explicitly (no union type punning: nvcc kept a stale att[0] through a punned store).
// start position passed to kernels as plain words (see unpack_state)
struct StateWords { uint4 v[4]; };
inline StateWords state_words(const tafl_state& s) {
    StateWords w;
    static_assert(sizeof(StateWords) == sizeof(tafl_state), "64-byte state");
    std::memcpy(&w, &s, sizeof(w));
    return w;
}
__device__ __forceinline__ tafl_state unpack_state(const uint4 a, const uint4 d, const uint4 r, const uint4 m) {
    tafl_state s;
    s.att[0] = (uint64_t)a.x | ((uint64_t)a.y << 32);
    s.att[1] = (uint64_t)a.z | ((uint64_t)a.w << 32);
    s.def[0] = (uint64_t)d.x | ((uint64_t)d.y << 32);
    s.def[1] = (uint64_t)d.z | ((uint64_t)d.w << 32);
    s.recent[0] = r.x; s.recent[1] = r.y; s.recent[2] = r.z; s.recent[3] = r.w;
    s.att_reps = (uint8_t)(m.x & 0xFF);
    s.def_reps = (uint8_t)((m.x >> 8) & 0xFF);
    s.flags = (uint8_t)((m.x >> 16) & 0xFF);
    s.status = (uint8_t)(m.x >> 24);
    s.turn = m.y;
    s.plays_since_capture = m.z;
    s.ring_head = (uint8_t)(m.w & 0xFF);
    s.side_len = (uint8_t)((m.w >> 8) & 0xFF);
    s._pad[0] = (uint8_t)((m.w >> 16) & 0xFF);
    s._pad[1] = (uint8_t)(m.w >> 24);
    return s;
}
__device__ __forceinline__ tafl_state load_state(const tafl_state* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    return unpack_state(q[0], q[1], q[2], q[3]);
}
// the same through L2 only (ld.global.cg): for states another CTA of a still-running grid has just published
__device__ __forceinline__ tafl_state load_state_cg(const tafl_state* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    return unpack_state(__ldcg(q), __ldcg(q + 1), __ldcg(q + 2), __ldcg(q + 3));
}
__device__ __forceinline__ void store_state(tafl_state* p, const tafl_state& s) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4((uint32_t)s.att[0], (uint32_t)(s.att[0] >> 32), (uint32_t)s.att[1], (uint32_t)(s.att[1] >> 32));
    q[1] = make_uint4((uint32_t)s.def[0], (uint32_t)(s.def[0] >> 32), (uint32_t)s.def[1], (uint32_t)(s.def[1] >> 32));
    q[2] = make_uint4(s.recent[0], s.recent[1], s.recent[2], s.recent[3]);
    q[3] = make_uint4((uint32_t)s.att_reps | ((uint32_t)s.def_reps << 8) | ((uint32_t)s.flags << 16) | ((uint32_t)s.status << 24),
                      s.turn, s.plays_since_capture,
                      (uint32_t)s.ring_head | ((uint32_t)s.side_len << 8) | ((uint32_t)s._pad[0] << 16) | ((uint32_t)s._pad[1] << 24));
}

}  // namespace tafl
