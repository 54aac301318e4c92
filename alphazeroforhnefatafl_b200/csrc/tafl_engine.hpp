// tafl_engine.hpp — internal definition of the opaque `tafl_engine` handle (include/tafl.h).
#pragma once
#include <cuda_runtime.h>

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

}  // namespace tafl

struct tafl_engine {
    int device = 0, width = 0, N = 0, sm_count = 0;
    uint32_t n_games = 0;
    cudaStream_t stream = nullptr;
    tafl_rules rules{};
    tafl_state start{};
    tafl::Ctx<tafl::BB64> cx64;
    tafl::Ctx<tafl::BB128> cx128;
    tafl_state* d_states = nullptr;
    tafl_move* d_moves = nullptr;      // [n_games][TAFL_MAX_MOVES]
    uint32_t* d_counts = nullptr;      // [n_games]
    uint8_t* d_status = nullptr;       // [n_games]
    tafl::DevStats* d_stats = nullptr;
    unsigned long long* d_checksum = nullptr;
    uint64_t game_id_base = 0, ply = 0, launches = 0;
    // scratch for the by-value batch API (grown on demand)
    void* d_scratch[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t scratch_bytes[6] = {0, 0, 0, 0, 0, 0};
    tafl::MctsPools* mcts = nullptr;
    std::string err;
};

namespace tafl {

int set_cuda_error(tafl_engine* e, cudaError_t ce, const char* what);
int scratch(tafl_engine* e, int slot, size_t bytes, void** out);
void mcts_free(tafl_engine* e);

#define TAFL_CUDA_TRY(e, call)                                              \
    do {                                                                    \
        cudaError_t _ce = (call);                                           \
        if (_ce != cudaSuccess) return tafl::set_cuda_error((e), _ce, #call); \
    } while (0)

inline unsigned grid_for(uint32_t n, unsigned block) { return (n + block - 1) / block; }

// 64-byte state <-> registers as four 16-byte vectors
__device__ __forceinline__ tafl_state load_state(const tafl_state* p) {
    union { tafl_state s; uint4 v[4]; } u;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    u.v[0] = q[0]; u.v[1] = q[1]; u.v[2] = q[2]; u.v[3] = q[3];
    return u.s;
}
__device__ __forceinline__ void store_state(tafl_state* p, const tafl_state& s) {
    union { tafl_state s; uint4 v[4]; } u;
    u.s = s;
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = u.v[0]; q[1] = u.v[1]; q[2] = u.v[2]; q[3] = u.v[3];
}

}  // namespace tafl
