// tafl_engine.cu — engine handle, lockstep kernels (one game per thread) and the C ABI
// (include/tafl.h) except the MCTS entry points (tafl_mcts.cu) and host text helpers
// (tafl_host.cpp).  sm_100a only; no CPU path.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "tafl_engine.hpp"
#include "tafl_mcts.cuh"
#include "tafl_select.cuh"

using namespace tafl;

// ---------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------
namespace {

#ifndef TAFL_STEP_BLOCK
#define TAFL_STEP_BLOCK 64
#endif
constexpr int STEP_BLOCK = TAFL_STEP_BLOCK;   // threads per CTA for the per-game kernels
constexpr int step_minb(int minb) { return minb * 64 / STEP_BLOCK > 0 ? minb * 64 / STEP_BLOCK : 1; }   // MINB counts 64-thread CTAs

// NOTE: the start position travels as StateWords (plain uint4 words) and is rebuilt with
// unpack_state: assigning a tafl_state kernel parameter to a local that is modified afterwards
// (`s = start` on auto-reset) made nvcc 12.9 reload att[0] from the parameter at the final store.
// `episode` (may be null): the restart counter of each game id, bumped whenever its game starts over
__global__ void k_fill_states(tafl_state* __restrict__ states, uint32_t n, const __grid_constant__ StateWords start,
                              uint32_t* __restrict__ episode) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n) {
        uint4* q = reinterpret_cast<uint4*>(states + g);
        q[0] = start.v[0]; q[1] = start.v[1]; q[2] = start.v[2]; q[3] = start.v[3];
        if (episode) episode[g] += 1u;
    }
}
__global__ void k_reset_ids(tafl_state* __restrict__ states, const uint32_t* __restrict__ ids, uint32_t n,
                            uint32_t n_games, const __grid_constant__ StateWords start, uint32_t* __restrict__ episode) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && ids[i] < n_games) {
        uint4* q = reinterpret_cast<uint4*>(states + ids[i]);
        q[0] = start.v[0]; q[1] = start.v[1]; q[2] = start.v[2]; q[3] = start.v[3];
        atomicAdd(&episode[ids[i]], 1u);   // an id may be listed twice
    }
}
__global__ void k_bump_episodes(uint32_t* __restrict__ episode, uint32_t first, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) episode[first + i] += 1u;
}
// restart every finished game (self-play sweep: all trees stay busy); counts the restarts
__global__ void k_reset_finished(tafl_state* __restrict__ states, uint32_t n, const __grid_constant__ StateWords start,
                                 uint32_t* __restrict__ episode, HistView hv, uint32_t* __restrict__ n_reset) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    bool hit = false;
    if (g < n) {
        const uint32_t meta = reinterpret_cast<const uint4*>(states + g)[3].x;
        if ((meta >> 24) != TAFL_STATUS_ONGOING) {
            hit = true;
            uint4* q = reinterpret_cast<uint4*>(states + g);
            q[0] = start.v[0]; q[1] = start.v[1]; q[2] = start.v[2]; q[3] = start.v[3];
            episode[g] += 1u;
            if (hv.st) { hv.len[g] = 0; hv.floor[g] = 0; }
        }
    }
    const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(n_reset, (uint32_t)__popc(m));
}

// get_all_possible_moves (game/main.rs:33-43), one game per thread: first-generation kernel, kept as the
// comparison point (TAFL_LM_G=1): scalar u16 stores, consecutive lanes write 1 KB apart.
template <class BB>
__global__ void __launch_bounds__(STEP_BLOCK)
k_legal_moves(const __grid_constant__ Ctx<BB> cx, const tafl_state* __restrict__ states, uint32_t n,
              tafl_move* __restrict__ moves, uint32_t* __restrict__ counts) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    tafl_state s = load_state(states + g);
    counts[g] = (uint32_t)gen_moves(cx, s, moves + (size_t)g * TAFL_MAX_MOVES, TAFL_MAX_MOVES);
}

// get_all_possible_moves (game/main.rs:33-43 over ValidPlayIterator, play.rs:186-226), warp-cooperative:
// a group of G lanes owns one game.  Lane i takes the i-th piece of the side to move (ascending bit
// index = canonical order) and gets its four ray lengths from piece_rays; a __shfl_up prefix sum over
// the per-piece totals gives every lane the list offset of its piece; each lane writes its piece's
// plays (tile arithmetic, no per-move decode) into the game's shared-memory row; the group then copies
// the row out with 16-byte vector stores, consecutive lanes writing consecutive addresses — the list
// leaves the SM as full 32-byte sectors instead of the 2-byte scatter of the thread-per-game kernel.
constexpr int LM_BLOCK = 128;
template <class BB, int SP, int G>
__global__ void __launch_bounds__(LM_BLOCK)
k_legal_moves_g(const __grid_constant__ Ctx<BB, SP> cx, const tafl_state* __restrict__ states, uint32_t n,
                tafl_move* __restrict__ moves, uint32_t* __restrict__ counts) {
    constexpr int GPB = LM_BLOCK / G;   // games per CTA
    constexpr int S = BB::STRIDE;
    __shared__ __align__(16) tafl_move rows[GPB][TAFL_MAX_MOVES];
    const int lane = threadIdx.x & 31, gl = lane % G, gshift = (lane / G) * G;
    const unsigned gmask = G == 32 ? 0xFFFFFFFFu : (((1u << (G & 31)) - 1u) << gshift);
    const int slot = threadIdx.x / G;
    const uint32_t g_raw = blockIdx.x * GPB + (uint32_t)slot;
    const bool live = g_raw < n;
    const uint32_t g = live ? g_raw : n - 1u;   // a dead group mirrors the last game and writes nothing
    tafl_move* const row = rows[slot];
    // the lanes of a group read the same three words: one transaction per word, broadcast by the LSU
    const uint4* const q = reinterpret_cast<const uint4*>(states + g);
    const uint4 a = q[0], d = q[1];
    const uint32_t meta = q[3].x;
    tafl_state st;
    st.att[0] = (uint64_t)a.x | ((uint64_t)a.y << 32); st.att[1] = (uint64_t)a.z | ((uint64_t)a.w << 32);
    st.def[0] = (uint64_t)d.x | ((uint64_t)d.y << 32); st.def[1] = (uint64_t)d.z | ((uint64_t)d.w << 32);
    const bool ongoing = (meta >> 24) == TAFL_STATUS_ONGOING;   // logic.rs:165-167: a finished game has no plays
    const int def_side = ((meta >> 16) & TAFL_FLAG_DEF_TO_PLAY) != 0;
    const Pos<BB> p = load_pos(cx, st);
    const BB own = p.side(def_side), occ = p.occ();
    const int n_pieces = ongoing ? own.popc() : 0;
    int before = 0;
    for (int base = 0; base < n_pieces; base += G) {
        const int idx = base + gl;
        int c = 0, tile = 0, row_i = 0, col_i = 0, kind = KIND_AS;
        Rays r;
        r.cnt[0] = r.cnt[1] = r.cnt[2] = r.cnt[3] = 0;
        if (idx < n_pieces) {
            tile = nth_set_bit(own, idx);
            row_i = tile / S;
            col_i = tile - row_i * S;
            kind = def_side ? ((BB::bit(tile) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
            r = piece_rays(cx, occ, kind, row_i, col_i);
            c = r.cnt[0] + r.cnt[1] + r.cnt[2] + r.cnt[3];
        }
        int incl = c;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) {
            const int t = __shfl_up_sync(gmask, incl, o, G);
            if (gl >= o) incl += t;
        }
        emit_piece_plays(cx, tile, row_i, col_i, kind, r, row, before + incl - c, TAFL_MAX_MOVES);
        before += __shfl_sync(gmask, incl, G - 1, G);
    }
    const int L = before;   // the same on every lane of the group
    const int Lc = L < TAFL_MAX_MOVES ? L : TAFL_MAX_MOVES;
    if (gl < 8 && Lc + gl < ((Lc + 7) & ~7)) row[Lc + gl] = 0;   // the unused tail of the last vector
    __syncwarp(gmask);
    if (live) {
        const uint4* const src = reinterpret_cast<const uint4*>(row);
        uint4* const dst = reinterpret_cast<uint4*>(moves + (size_t)g * TAFL_MAX_MOVES);
        const int nvec = (Lc + 7) >> 3;
        for (int v = gl; v < nvec; v += G) dst[v] = src[v];
        if (gl == 0) counts[g] = (uint32_t)L;
    }
}

// Game::do_play (game/mod.rs:95-101) with validation.  HostMirror: the caller's pinned result buffers,
// written from here over PCIe (plays may likewise be the caller's pinned buffer).
template <class BB, int SP>
__global__ void __launch_bounds__(STEP_BLOCK)
k_step_plays(const __grid_constant__ Ctx<BB, SP> cx, tafl_state* __restrict__ states, uint32_t n,
             const tafl_play* __restrict__ plays, tafl_play_effects* __restrict__ fx_out,
             uint8_t* __restrict__ status_out, uint8_t* __restrict__ err_out, HistView hv, HostMirror hm,
             const __grid_constant__ StateWords start_w, uint32_t* __restrict__ restart_episode) {
    const uint32_t g = blockIdx.x * STEP_BLOCK + threadIdx.x;
    if (g < n) {
        const tafl_play pl = plays[g];   // may be the caller's pinned buffer, read over PCIe: issue it first
        tafl_state s = load_state(states + g);
        if (restart_episode && s.status != TAFL_STATUS_ONGOING) {   // tafl_step_status_restart: Game::new for a game that had ended
            s = unpack_state(start_w.v[0], start_w.v[1], start_w.v[2], start_w.v[3]);
            restart_episode[g] += 1u;
            if (hv.st) { hv.len[g] = 0; hv.floor[g] = 0; }
        }
        const tafl_state before = s;
        tafl_play_effects fx;
        fx.captures[0] = fx.captures[1] = 0;
        fx.n_captures = 0;
        for (int i = 0; i < 5; ++i) fx._pad[i] = 0;
        const int err = validate_play(cx, s, pl, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
        fx.err = (uint8_t)err;
        if (!err) {
            BB caps = do_valid_play(cx, s, pl.from_row, pl.from_col, pl.axis == TAFL_HORIZONTAL, pl.disp);
            put_bb(fx.captures, caps);
            fx.n_captures = (uint8_t)caps.popc();
            store_state(states + g, s);
        }
        fx.status = s.status;
        if (!err && hv.st) {   // Game::do_play: state_history.push(old state), play_history.push(record)
            const uint32_t len = hv.len[g], slot = len % hv.cap;
            store_state(hv.st + (size_t)slot * hv.n + g, before);
            tafl_play_record r;
            r.play = pl;
            r.side = (before.flags & TAFL_FLAG_DEF_TO_PLAY) ? TAFL_DEFENDER : TAFL_ATTACKER;
            r._pad[0] = r._pad[1] = r._pad[2] = 0;
            r.effects = fx;
            hv.rec[(size_t)slot * hv.n + g] = r;
            hv.len[g] = len + 1;
            if (len + 1 - hv.floor[g] > hv.cap) hv.floor[g] = len + 1 - hv.cap;
        }
        if (fx_out) fx_out[g] = fx;
        if (status_out) status_out[g] = s.status;
        if (err_out) err_out[g] = (uint8_t)err;
        if (hm.status) hm.status[g] = s.status;
        if (hm.err) hm.err[g] = (uint8_t)err;
    }
}

// Game::undo_last_play (game/mod.rs:103-108): pop the newest (state, record) pair of each selected game
__global__ void k_undo(tafl_state* __restrict__ states, uint32_t n, const uint8_t* __restrict__ mask, HistView hv) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n || (mask && !mask[g])) return;
    const uint32_t len = hv.len[g];
    if (len <= hv.floor[g]) return;   // state_history.pop() == None: nothing to undo
    const uint32_t slot = (len - 1) % hv.cap;
    store_state(states + g, load_state(hv.st + (size_t)slot * hv.n + g));
    hv.len[g] = len - 1;
}
// forget the history of games [first, first + n) or of the listed games
__global__ void k_hist_clear(HistView hv, uint32_t first, uint32_t n, const uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = ids ? ids[i] : first + i;
    if (g >= hv.n) return;
    hv.len[g] = 0;
    hv.floor[g] = 0;
}

// do_valid_play (logic.rs:782-820) on one compact move per game
template <class BB, int SP>
__global__ void __launch_bounds__(STEP_BLOCK)
k_step_moves(const __grid_constant__ Ctx<BB, SP> cx, tafl_state* __restrict__ states, uint32_t n,
             const tafl_move* __restrict__ mv, uint32_t mv_stride, const uint32_t* __restrict__ pick,
             uint8_t* __restrict__ status_out) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    tafl_state s = load_state(states + g);
    if (s.status == TAFL_STATUS_ONGOING) {
        const tafl_move m = mv[(size_t)g * mv_stride + (pick ? pick[g] : 0u)];
        do_valid_move(cx, s, m);
        store_state(states + g, s);
    }
    if (status_out) status_out[g] = s.status;
}

// Chaining pays once a grid has more than four 64-thread CTAs per SM (measured: 40 960 games 10.0 -> 7.6 us per ply, 65 536
// 10.3 -> 8.95, 196 608 24.4 -> 22.7; below that the token hand-off costs more than the tail it hides: 32 768 games 6.8 -> 7.4)
// and stops paying when the grid is many waves deep (524 288 games: 59.3 -> 60.1)
constexpr uint32_t CHAIN_MIN_CTAS_PER_SM = 4, CHAIN_MAX_CTAS_PER_SM = 28;   // in 64-game units per SM
#if !defined(TAFL_CHAIN_BLOCK)
#define TAFL_CHAIN_BLOCK 256
#endif
constexpr int CHAIN_BLOCK = TAFL_CHAIN_BLOCK;
#if !defined(TAFL_CHAIN_SLEEP)
#define TAFL_CHAIN_SLEEP 40
#endif
constexpr int CHAIN_POS_BITS = 12;                    // nodes per graph that can be chained (token = launch sequence | node)
constexpr uint32_t CHAIN_SPIN_LIMIT = 1u << 22;       // ~ 0.3 s of polling: a lost token raises the time-out word instead of hanging

// The fused random-playout step (K4): count -> pick -> select -> apply -> outcome, `steps`
// plies per launch with the state held in registers in between.  BYVAL (tafl_step_random_host): the
// states come from the staged copy of the caller's array (`src`); a row whose side_len is not the
// engine's is rejected here (flag raised, the game keeps its resident state) instead of by a host
// loop over 4 MiB of input before anything is enqueued.
#if defined(TAFL_STEP_MAXNREG)   // A/B builds: an explicit register cap instead of the minimum-blocks hint
#define TAFL_STEP_BOUNDS(MINB) __maxnreg__(TAFL_STEP_MAXNREG)
#else
#define TAFL_STEP_BOUNDS(MINB) __launch_bounds__(BLOCK, (MINB) * 64 / BLOCK > 0 ? (MINB) * 64 / BLOCK : 1)
#endif
// BLOCK: threads per CTA.  The nodes of a chained graph run CHAIN_BLOCK-thread CTAs (one token per 256 games instead
// of per 64: 7.7 us per 65 536-game ply against 9.0 with 64, 8.55 with 128 and 9.2 with 512), everything else STEP_BLOCK.
template <class BB, int SP, int MINB, bool BYVAL = false, bool SINGLE = false, int BLOCK = STEP_BLOCK>
__global__ void TAFL_STEP_BOUNDS(MINB)
k_step_random(const __grid_constant__ Ctx<BB, SP> cx, tafl_state* __restrict__ states, uint32_t n,
              const __grid_constant__ StateWords start_w, uint64_t seed, uint64_t gid_base, uint64_t ply0,
              uint32_t steps, int auto_reset, DevStats* __restrict__ stats, uint32_t* __restrict__ actions,
              const unsigned long long* __restrict__ ply_base, const tafl_state* __restrict__ src = nullptr,
              uint32_t* __restrict__ rejected = nullptr, uint32_t* chain = nullptr) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t c_steps = 0, c_legal = 0, c_caps = 0;
    // CTA chaining (graph nodes only, `chain` = one word per CTA + a time-out word): game g at ply i+1 needs nothing
    // but game g at ply i, i.e. CTA b of this node needs CTA b of the previous node and no other.  A node of the
    // chain therefore lets its successor launch at once, waits for the token its own predecessor CTA publishes
    // after storing its states (release / acquire at GPU scope) instead of for the whole previous grid, and
    // publishes its own.  The time of a grid is the time of its slowest warp (a rare long path: a fill that runs to
    // the end, a finished game); chained, the next ply's CTAs run in the slots the early finishers leave.
    const bool chained = SINGLE && chain != nullptr;
    const uint32_t chain_pos = (uint32_t)ply0;          // node index inside the graph
    uint32_t chain_seq = 0;
    if (chained && chain_pos > 0) {
        asm volatile("griddepcontrol.launch_dependents;");
    } else {
        // Programmatic dependent launch (no-ops for a plain launch): wait for the previous ply's grid
        // before touching its states, then let the next ply's grid start filling free CTA slots so its
        // launch latency and ramp overlap this grid's tail.
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
    }
    if (ply_base) ply0 += *ply_base;
    if (chained) {
        chain_seq = (uint32_t)ply_base[1] << CHAIN_POS_BITS;
        if (chain_pos > 0) {
            if (threadIdx.x == 0) {
                const uint32_t want = chain_seq | chain_pos;
                uint32_t got, spins = 0;
                while (true) {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(got) : "l"(chain + blockIdx.x) : "memory");
                    if (got == want) break;
                    if (++spins > CHAIN_SPIN_LIMIT) { atomicOr(chain + gridDim.x, 1u); break; }   // reported by the host, never a hang
                    if ((spins & 4095u) == 0u && *(volatile uint32_t*)(chain + gridDim.x) != 0u) break;   // a token was already lost: do not wait again
                    if (TAFL_CHAIN_SLEEP) __nanosleep(TAFL_CHAIN_SLEEP);
                }
            }
            __syncthreads();
        }
    }
    bool live = g < n;
    if (BYVAL && live && src[g].side_len != (uint8_t)cx.n()) {
        atomicOr(rejected, 1u);
        live = false;
    }
    if (live) {
        tafl_state s = chained ? load_state_cg(states + g) : load_state((BYVAL ? src : states) + g);
        uint32_t act = 0;   // tafl_action of the last ply of this launch
        // one ply; false = this game stops stepping (finished, no auto-reset)
        auto ply = [&](uint32_t i) {
            act = (uint32_t)s.status << 16;
            if (s.status != TAFL_STATUS_ONGOING) {   // rare: only these lanes pay for the reset
                if (!auto_reset) return false;
                s = unpack_state(start_w.v[0], start_w.v[1], start_w.v[2], start_w.v[3]);
            }
            StepEvt ev = random_step_ongoing(cx, s, seed, gid_base + g, ply0 + i);
            c_steps += ev.stepped;
            c_legal += ev.legal;
            c_caps += ev.caps;
            act = (uint32_t)ev.move | ((uint32_t)s.status << 16) | (ev.stepped << 24);
            if (ev.status) {  // rare: a game ended
                atomicAdd(&stats->finished, 1ull);
                atomicAdd(&stats->plies, (unsigned long long)ev.turn);
                if (TAFL_STATUS_IS_DRAW(ev.status)) atomicAdd(&stats->draw[TAFL_STATUS_REASON(ev.status) & 1], 1ull);
                else atomicAdd(&stats->win[ev.status & 1][TAFL_STATUS_REASON(ev.status)], 1ull);
            }
            return true;
        };
        if constexpr (SINGLE) {   // the lockstep graph launches one ply per kernel: no loop-carried copies of the state
            ply(0);
        } else {
            for (uint32_t i = 0; i < steps; ++i)
                if (!ply(i)) break;
        }
        store_state(states + g, s);
        actions[g] = act;
    }
    if (chained) {
        __syncthreads();                               // every state of this CTA is stored
        if (threadIdx.x == 0) {
            uint32_t token = chain_seq | (chain_pos + 1u);   // release: ordered after the barrier, hence after every thread's stores
            if (blockIdx.x == 0 && chain[gridDim.x + 1] == chain_pos + 1u) token = ~token;   // test hook (TAFL_STEP_CHAIN_LOSE): this token is lost
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(chain + blockIdx.x), "r"(token) : "memory");
        }
        // completion order of the nodes stays the launch order: this grid does not end before its predecessor has
        if (chain_pos > 0) asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    // warp-level reduction of the dense counters (REDUX), one atomic per warp
    c_steps = __reduce_add_sync(0xFFFFFFFFu, c_steps);
    c_legal = __reduce_add_sync(0xFFFFFFFFu, c_legal);
    c_caps = __reduce_add_sync(0xFFFFFFFFu, c_caps);
    if ((threadIdx.x & 31) == 0 && c_steps) {
        atomicAdd(&stats->steps, (unsigned long long)c_steps);
        atomicAdd(&stats->legal, (unsigned long long)c_legal);
        atomicAdd(&stats->caps, (unsigned long long)c_caps);
    }
}

__global__ void k_checksum(const tafl_state* __restrict__ states, uint32_t n, unsigned long long* out) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long acc = 0;
    if (g < n) {
        const unsigned long long* w = reinterpret_cast<const unsigned long long*>(states + g);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc ^= splitmix64(w[j] ^ (((unsigned long long)g * 8ull + j) * 0x9E3779B97F4A7C15ull));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_down_sync(0xFFFFFFFFu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicXor(out, acc);
}

// ---- by-value batch kernels -------------------------------------------------------------------
template <class BB>
__global__ void k_bv_validate(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const tafl_play* pl, uint32_t n,
                              uint8_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    out[i] = (uint8_t)validate_play(cx, s, pl[i], (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
}
template <class BB>
__global__ void k_bv_do_plays(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const tafl_play* pl, uint32_t n,
                              tafl_state* out_st, tafl_play_effects* out_fx) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    const tafl_play p = pl[i];
    tafl_play_effects fx;
    fx.captures[0] = fx.captures[1] = 0;
    fx.n_captures = 0;
    for (int k = 0; k < 5; ++k) fx._pad[k] = 0;
    const int err = validate_play(cx, s, p, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
    fx.err = (uint8_t)err;
    if (!err) {
        BB caps = do_valid_play(cx, s, p.from_row, p.from_col, p.axis == TAFL_HORIZONTAL, p.disp);
        put_bb(fx.captures, caps);
        fx.n_captures = (uint8_t)caps.popc();
    }
    fx.status = s.status;
    out_st[i] = s;
    out_fx[i] = fx;
}
template <class BB>
__global__ void k_bv_piece_plays(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const uint8_t* rc, uint32_t n,
                                 tafl_move* moves, int32_t* counts) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    counts[i] = piece_plays(cx, s, rc[2 * i], rc[2 * i + 1], moves + (size_t)i * 32);
}
template <class BB>
__global__ void k_bv_captures(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const tafl_play* pl, uint32_t n,
                              tafl_play_effects* out, int shieldwall_only) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    out[i] = shieldwall_only ? query_shieldwall(cx, s, pl[i]) : query_captures(cx, s, pl[i]);
}
template <class BB>
__global__ void k_bv_enclosure(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const tafl_enclosure_query* q,
                               uint32_t n, tafl_enclosure* out, uint8_t* secure) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    uint8_t sec;
    tafl_enclosure e = query_enclosure(cx, s, q[i], &sec);
    if (out) out[i] = e;
    if (secure) secure[i] = sec;
}
template <class BB>
__global__ void k_bv_exit_fort(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, uint32_t n, uint8_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    out[i] = detect_exit_fort(cx, load_pos(cx, s)) ? 1 : 0;
}
template <class BB>
__global__ void k_bv_can_play(const __grid_constant__ Ctx<BB> cx, const tafl_state* st, const uint8_t* sides, uint32_t n,
                              uint8_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state s = st[i];
    out[i] = query_can_play(cx, s, sides[i] != 0);
}

std::string g_create_error;

}  // namespace

// ---------------------------------------------------------------------------------------------
// internal helpers
// ---------------------------------------------------------------------------------------------
namespace tafl {

int set_cuda_error(tafl_engine* e, cudaError_t ce, const char* what) {
    std::string msg = std::string("CUDA error: ") + cudaGetErrorString(ce) + " in " + what;
    if (e) e->err = msg; else g_create_error = msg;
    return TAFL_ERR_CUDA;
}
int scratch(tafl_engine* e, int slot, size_t bytes, void** out) {
    if (bytes == 0) bytes = 16;
    if (e->scratch_bytes[slot] < bytes) {
        if (e->d_scratch[slot]) cudaFree(e->d_scratch[slot]);
        e->d_scratch[slot] = nullptr;
        e->scratch_bytes[slot] = 0;
        size_t cap = bytes + bytes / 2;
        TAFL_CUDA_TRY(e, cudaMalloc(&e->d_scratch[slot], cap));
        e->scratch_bytes[slot] = cap;
    }
    *out = e->d_scratch[slot];
    return TAFL_OK;
}

}  // namespace tafl

#define TAFL_CHECK_E(e) \
    if (!(e)) return TAFL_ERR_INVALID_ARG
// run `body` with BB / cx bound to the engine's bitfield width
#define TAFL_WIDTH_DISPATCH(e, body)                                     \
    do {                                                                 \
        if ((e)->width == 64) { using BB = BB64; const Ctx<BB64>& cx = (e)->cx64; body; }       \
        else { using BB = BB128; const Ctx<BB128>& cx = (e)->cx128; body; }                     \
    } while (0)
#define TAFL_LAUNCHED(e)                                                 \
    do {                                                                 \
        (e)->launches += 1;                                              \
        TAFL_CUDA_TRY((e), cudaGetLastError());                          \
    } while (0)

// Game::do_play is the reference's only mutator, so its histories are always complete (game/mod.rs:95-108).
// The batch mutators that have no reference counterpart (random stepping, compact-move stepping, the MCTS
// move) do not record plays; they forget the history of the games they advance instead of leaving a
// history with holes behind (tafl_undo_last_play then has nothing to undo, as after tafl_set_states).
namespace tafl {
void hist_forget_all(tafl_engine* e) {
    if (e->hist.st) k_hist_clear<<<grid_for(e->n_games, 256), 256, 0, e->stream>>>(e->hist, 0, e->n_games, nullptr);
}
}  // namespace tafl

static int sync_engine(tafl_engine* e) {
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    if (e->d_chain && e->chain_dirty) {   // a chained step graph ran since the last look at its time-out word
        uint32_t lost = 0;
        TAFL_CUDA_TRY(e, cudaMemcpy(&lost, e->d_chain + e->chain_ctas, sizeof(lost), cudaMemcpyDeviceToHost));
        e->chain_dirty = false;
        if (lost) {
            cudaMemset(e->d_chain + e->chain_ctas, 0, sizeof(uint32_t));
            e->err = "chained step graph: a CTA never saw its predecessor's token (TAFL_STEP_CHAIN=0 disables chaining); states are not valid";
            return TAFL_ERR_CUDA;
        }
    }
    return TAFL_OK;
}
static int check_side_len(tafl_engine* e, const tafl_state* st, uint32_t n) {
    for (uint32_t i = 0; i < n; ++i)
        if (st[i].side_len != e->N) {
            e->err = "state side_len does not match the engine's board";
            return TAFL_ERR_INVALID_ARG;
        }
    return TAFL_OK;
}
template <class T>
static int upload(tafl_engine* e, int slot, const T* h, size_t n, T** d) {
    void* p;
    int rc = scratch(e, slot, n * sizeof(T), &p);
    if (rc) return rc;
    if (n) TAFL_CUDA_TRY(e, cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, e->stream));
    *d = (T*)p;
    return TAFL_OK;
}
template <class T>
static int device_out(tafl_engine* e, int slot, size_t n, T** d) {
    void* p;
    int rc = scratch(e, slot, n * sizeof(T), &p);
    if (rc) return rc;
    *d = (T*)p;
    return TAFL_OK;
}
template <class T>
static int download(tafl_engine* e, T* h, const T* d, size_t n) {
    if (n) TAFL_CUDA_TRY(e, cudaMemcpyAsync(h, d, n * sizeof(T), cudaMemcpyDeviceToHost, e->stream));
    return TAFL_OK;
}

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
extern "C" {

const char* tafl_last_error(const tafl_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }

int tafl_engine_create(const tafl_rules* rules, const char* start_fen, int width_bits, uint32_t n_games, int device,
                       tafl_engine** out) {
    if (!rules || !start_fen || !out || n_games == 0) { g_create_error = "invalid argument"; return TAFL_ERR_INVALID_ARG; }
    *out = nullptr;
    // parse once at 128 bits to learn the side length, then at the chosen width
    tafl_state probe;
    int rc = tafl_state_from_fen(start_fen, 128, rules->starting_side, &probe);
    if (rc) { g_create_error = "could not parse the starting FEN"; return rc; }
    const int N = probe.side_len;
    if (width_bits == 0) width_bits = N <= 7 ? 64 : 128;
    if ((width_bits != 64 && width_bits != 128) || N < 1 || N > (width_bits == 64 ? 7 : 11)) {
        g_create_error = "board does not fit the requested bitfield width (u64: <=7, u128: <=11)";
        return TAFL_ERR_UNSUPPORTED;
    }
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no usable CUDA device (") + cudaGetErrorString(ce) +
                         "); libtafl_b200 has no CPU fallback";
        return TAFL_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return TAFL_ERR_INVALID_ARG; }
    tafl_engine* e = new tafl_engine();
    e->device = device; e->width = width_bits; e->N = N; e->n_games = n_games; e->rules = *rules;
    rc = tafl_state_from_fen(start_fen, width_bits, rules->starting_side, &e->start);
    if (rc) { delete e; g_create_error = "could not parse the starting FEN"; return rc; }
    e->cx64 = make_ctx<BB64>(*rules, N <= 7 ? N : 7);
    e->cx128 = make_ctx<BB128>(*rules, N);
    if (width_bits == 64) {
        if (spec_matches<SPEC_BRANDUBH>(e->cx64)) e->spec_detected = SPEC_BRANDUBH;
    } else {
        if (spec_matches<SPEC_COPENHAGEN>(e->cx128)) e->spec_detected = SPEC_COPENHAGEN;
        else if (spec_matches<SPEC_TABLUT>(e->cx128)) e->spec_detected = SPEC_TABLUT;
    }
    e->spec = e->spec_detected;
#define TAFL_CREATE_TRY(call)                                                                \
    do {                                                                                     \
        cudaError_t _ce = (call);                                                            \
        if (_ce != cudaSuccess) {                                                            \
            g_create_error = std::string("CUDA error: ") + cudaGetErrorString(_ce) + " in " #call; \
            tafl_engine_destroy(e);                                                          \
            return TAFL_ERR_CUDA;                                                            \
        }                                                                                    \
    } while (0)
    TAFL_CREATE_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    TAFL_CREATE_TRY(cudaGetDeviceProperties(&prop, device));
    e->sm_count = prop.multiProcessorCount;
    if (const char* v = std::getenv("TAFL_STEP_MINB")) e->step_minb = std::atoi(v);   // tuning knobs (profiles/sweeps)
    if (const char* v = std::getenv("TAFL_STEP_GRAPH")) e->step_graph_on = std::atoi(v);
    if (const char* v = std::getenv("TAFL_STEP_SINGLE")) e->step_single = std::atoi(v);
    if (const char* v = std::getenv("TAFL_STEP_PDL")) e->step_pdl_on = std::atoi(v);
    if (const char* v = std::getenv("TAFL_STEP_CHAIN")) e->step_chain_on = std::atoi(v);
    if (const char* v = std::getenv("TAFL_HOST_CHUNKS")) e->host_chunks = std::atoi(v);
    if (const char* v = std::getenv("TAFL_HOST_IO")) e->host_io = std::atoi(v);
    if (const char* v = std::getenv("TAFL_MCTS_MINB")) e->mcts_minb = std::atoi(v);
    if (const char* v = std::getenv("TAFL_MCTS_G")) e->mcts_group = std::atoi(v);
    if (const char* v = std::getenv("TAFL_LM_G")) e->lm_group = std::atoi(v);
    if (const char* v = std::getenv("TAFL_HOOK_GEN")) e->hook_gen = std::atoi(v);
    if (const char* v = std::getenv("TAFL_HOOK_FUSE")) e->hook_fuse = std::atoi(v);
    TAFL_CREATE_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_states, (size_t)n_games * sizeof(tafl_state)));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_counts, (size_t)n_games * sizeof(uint32_t)));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_status, (size_t)n_games));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_actions, (size_t)n_games * sizeof(uint32_t)));
    TAFL_CREATE_TRY(cudaMemsetAsync(e->d_actions, 0, (size_t)n_games * sizeof(uint32_t), e->stream));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_stats, sizeof(DevStats)));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_checksum, sizeof(unsigned long long)));
    TAFL_CREATE_TRY(cudaMalloc(&e->d_episode, (size_t)n_games * sizeof(uint32_t)));
    TAFL_CREATE_TRY(cudaMemsetAsync(e->d_episode, 0, (size_t)n_games * sizeof(uint32_t), e->stream));
    TAFL_CREATE_TRY(cudaMemsetAsync(e->d_stats, 0, sizeof(DevStats), e->stream));
    k_fill_states<<<grid_for(n_games, 256), 256, 0, e->stream>>>(e->d_states, n_games, state_words(e->start), nullptr);
    e->launches += 1;
    TAFL_CREATE_TRY(cudaGetLastError());
    TAFL_CREATE_TRY(cudaStreamSynchronize(e->stream));
#undef TAFL_CREATE_TRY
    *out = e;
    return TAFL_OK;
}

int tafl_engine_destroy(tafl_engine* e) {
    if (!e) return TAFL_OK;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    mcts_free(e);
    cudaFree(e->d_states); cudaFree(e->d_moves); cudaFree(e->d_counts); cudaFree(e->d_status); cudaFree(e->d_actions);
    cudaFree(e->d_stats); cudaFree(e->d_checksum); cudaFree(e->d_ply); cudaFree(e->d_chain); cudaFree(e->d_episode);
    cudaFree(e->hist.st); cudaFree(e->hist.rec); cudaFree(e->hist.len); cudaFree(e->hist.floor);
    if (e->step_graph) cudaGraphExecDestroy(e->step_graph);
    if (e->ev_block) cudaEventDestroy(e->ev_block);
    for (auto& p : e->d_scratch) cudaFree(p);
    if (e->aux_ready) for (auto& st : e->aux) cudaStreamDestroy(st);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
    return TAFL_OK;
}

int tafl_engine_info(const tafl_engine* e, uint32_t* n_games, int* width_bits, int* side_len, int* device, int* sm_count) {
    TAFL_CHECK_E(e);
    if (n_games) *n_games = e->n_games;
    if (width_bits) *width_bits = e->width;
    if (side_len) *side_len = e->N;
    if (device) *device = e->device;
    if (sm_count) *sm_count = e->sm_count;
    return TAFL_OK;
}
void* tafl_engine_stream(const tafl_engine* e) { return e ? (void*)e->stream : nullptr; }
int tafl_engine_sync(tafl_engine* e) { TAFL_CHECK_E(e); cudaSetDevice(e->device); return sync_engine(e); }
uint64_t tafl_launch_count(const tafl_engine* e) { return e ? e->launches : 0; }
int tafl_engine_spec(const tafl_engine* e) { return e ? e->spec : TAFL_ERR_INVALID_ARG; }
int tafl_engine_force_generic(tafl_engine* e, int on) {
    TAFL_CHECK_E(e);
    e->spec = on ? (int)SPEC_GENERIC : e->spec_detected;
    return TAFL_OK;
}
tafl_state* tafl_states_device(tafl_engine* e) { return e ? e->d_states : nullptr; }

int tafl_reset(tafl_engine* e, const uint32_t* h_ids, uint32_t n) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    if (!h_ids) {
        k_fill_states<<<grid_for(e->n_games, 256), 256, 0, e->stream>>>(e->d_states, e->n_games, state_words(e->start), e->d_episode);
        TAFL_LAUNCHED(e);
        if (e->hist.st) k_hist_clear<<<grid_for(e->n_games, 256), 256, 0, e->stream>>>(e->hist, 0, e->n_games, nullptr);
    } else if (n) {
        uint32_t* d_ids;
        int rc = upload(e, 0, h_ids, n, &d_ids);
        if (rc) return rc;
        k_reset_ids<<<grid_for(n, 256), 256, 0, e->stream>>>(e->d_states, d_ids, n, e->n_games, state_words(e->start), e->d_episode);
        TAFL_LAUNCHED(e);
        if (e->hist.st) k_hist_clear<<<grid_for(n, 256), 256, 0, e->stream>>>(e->hist, 0, n, d_ids);
    }
    return sync_engine(e);
}
int tafl_set_states(tafl_engine* e, const tafl_state* h, uint32_t first, uint32_t n) {
    TAFL_CHECK_E(e);
    if (!h || (uint64_t)first + n > e->n_games) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    int rc = check_side_len(e, h, n);
    if (rc) return rc;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(e->d_states + first, h, (size_t)n * sizeof(tafl_state), cudaMemcpyHostToDevice, e->stream));
    if (n) k_bump_episodes<<<grid_for(n, 256), 256, 0, e->stream>>>(e->d_episode, first, n);   // replaced positions start new episodes
    if (e->hist.st && n) k_hist_clear<<<grid_for(n, 256), 256, 0, e->stream>>>(e->hist, first, n, nullptr);
    return sync_engine(e);
}
int tafl_reset_finished(tafl_engine* e, uint32_t* h_n) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    void* p;
    int rc = scratch(e, 4, sizeof(uint32_t), &p);
    if (rc) return rc;
    TAFL_CUDA_TRY(e, cudaMemsetAsync(p, 0, sizeof(uint32_t), e->stream));
    k_reset_finished<<<grid_for(e->n_games, 256), 256, 0, e->stream>>>(e->d_states, e->n_games, state_words(e->start), e->d_episode,
                                                                      e->hist, (uint32_t*)p);
    TAFL_LAUNCHED(e);
    if (h_n) {
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_n, p, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
        return sync_engine(e);
    }
    return TAFL_OK;
}
int tafl_get_states(tafl_engine* e, tafl_state* h, uint32_t first, uint32_t n) {
    TAFL_CHECK_E(e);
    if (!h || (uint64_t)first + n > e->n_games) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h, e->d_states + first, (size_t)n * sizeof(tafl_state), cudaMemcpyDeviceToHost, e->stream));
    return sync_engine(e);
}

static int ensure_moves(tafl_engine* e) {
    if (!e->d_moves) TAFL_CUDA_TRY(e, cudaMalloc(&e->d_moves, (size_t)e->n_games * TAFL_MAX_MOVES * sizeof(tafl_move)));
    return TAFL_OK;
}
int tafl_moves_device(tafl_engine* e, tafl_move** d_moves, uint32_t** d_counts) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    int rc = ensure_moves(e);
    if (rc) return rc;
    if (d_moves) *d_moves = e->d_moves;
    if (d_counts) *d_counts = e->d_counts;
    return TAFL_OK;
}
static int legal_moves_enqueue(tafl_engine* e, const tafl_state* d_states, uint32_t n, tafl_move* d_moves, uint32_t* d_counts) {
    if (n == 0) return TAFL_OK;
    // lanes per game (TAFL_LM_G; 0 = by batch size): 8 lanes keep more lanes busy when the batch fills the GPU (38.9 us
    // against 49.1 us for 65 536 games), 16 lanes halve the passes a game needs when a launch is latency-bound
    // (8.2 us against 9.2 us for 2048 games)
    const int lm_g = e->lm_group ? e->lm_group : (n <= 8192u ? 16 : 8);
    if (lm_g == 16) {
        TAFL_SPEC_DISPATCH(e, (k_legal_moves_g<BB, SP, 16><<<grid_for(n, LM_BLOCK / 16), LM_BLOCK, 0, e->stream>>>(cx, d_states, n, d_moves, d_counts)));
    } else if (lm_g == 8) {
        TAFL_SPEC_DISPATCH(e, (k_legal_moves_g<BB, SP, 8><<<grid_for(n, LM_BLOCK / 8), LM_BLOCK, 0, e->stream>>>(cx, d_states, n, d_moves, d_counts)));
    } else {                          // first generation: one game per thread
        TAFL_WIDTH_DISPATCH(e, (k_legal_moves<BB><<<grid_for(n, STEP_BLOCK), STEP_BLOCK, 0, e->stream>>>(cx, d_states, n, d_moves, d_counts)));
    }
    TAFL_LAUNCHED(e);
    return TAFL_OK;
}
int tafl_legal_moves(tafl_engine* e, tafl_move* d_moves, uint32_t* d_counts) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    if (!d_moves || !d_counts) {
        int rc = ensure_moves(e);
        if (rc) return rc;
        if (!d_moves) d_moves = e->d_moves;
        if (!d_counts) d_counts = e->d_counts;
    }
    int rc = legal_moves_enqueue(e, e->d_states, e->n_games, d_moves, d_counts);
    if (rc) return rc;
    return sync_engine(e);
}
int tafl_legal_moves_async(tafl_engine* e, tafl_move* d_moves, uint32_t* d_counts) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    if (!d_moves || !d_counts) {
        int rc = ensure_moves(e);
        if (rc) return rc;
        if (!d_moves) d_moves = e->d_moves;
        if (!d_counts) d_counts = e->d_counts;
    }
    return legal_moves_enqueue(e, e->d_states, e->n_games, d_moves, d_counts);
}
int tafl_legal_moves_host(tafl_engine* e, tafl_move* h_moves, uint32_t* h_counts, uint32_t first, uint32_t n) {
    TAFL_CHECK_E(e);
    if (!h_moves || !h_counts || (uint64_t)first + n > e->n_games) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    int rc = ensure_moves(e);
    if (rc) return rc;
    rc = legal_moves_enqueue(e, e->d_states, e->n_games, e->d_moves, e->d_counts);
    if (rc) return rc;
    rc = download(e, h_moves, e->d_moves + (size_t)first * TAFL_MAX_MOVES, (size_t)n * TAFL_MAX_MOVES);
    if (rc) return rc;
    rc = download(e, h_counts, e->d_counts + first, n);
    if (rc) return rc;
    return sync_engine(e);
}

int tafl_step(tafl_engine* e, const tafl_play* h_plays, tafl_play_effects* h_fx) {
    TAFL_CHECK_E(e);
    if (!h_plays) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    tafl_play* d_pl;
    tafl_play_effects* d_fx;
    int rc = upload(e, 0, h_plays, n, &d_pl);
    if (rc) return rc;
    rc = device_out(e, 1, n, &d_fx);
    if (rc) return rc;
    TAFL_SPEC_DISPATCH(e, (k_step_plays<BB, SP><<<grid_for(n, STEP_BLOCK), STEP_BLOCK, 0, e->stream>>>(cx, e->d_states, n, d_pl, d_fx, nullptr, nullptr, e->hist, HostMirror{}, state_words(e->start), nullptr)));
    TAFL_LAUNCHED(e);
    if (h_fx) { rc = download(e, h_fx, d_fx, n); if (rc) return rc; }
    return sync_engine(e);
}
// device alias of a page-locked, device-mapped host pointer (cudaHostAlloc / cudaHostRegister), else null
static void* mapped_alias(const void* h) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}
static int step_status_impl(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err, bool restart);
int tafl_step_status(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err) {
    return step_status_impl(e, h_plays, h_status, h_err, false);
}
int tafl_step_status_restart(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err) {
    return step_status_impl(e, h_plays, h_status, h_err, true);
}
static int step_status_impl(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err, bool restart) {
    TAFL_CHECK_E(e);
    if (!h_plays || !h_status) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    tafl_play* d_pl = nullptr;
    uint8_t* d_err;
    int rc;
    // One ply is ~8 us of kernel but four copy-engine / launch hand-offs; with pinned caller buffers the
    // kernel reads the plays and writes the result bytes through their device aliases instead
    // (host_io bit 0: inputs, bit 1: outputs), which leaves one launch + one sync per call:
    // 41.4 -> 26.5 us per 65 536-game ply (profiles/sweeps/r01r_e2e_modes.txt).  Pageable buffers take
    // the copy-engine path.  Tried and dropped: 128-byte result rows staged through shared memory (no
    // change) and a pinned completion word polled by the host instead of the stream sync (2 us slower).
    if (e->host_io & 1) d_pl = (tafl_play*)mapped_alias(h_plays);
    HostMirror hm{};
    if (e->host_io & 2) {
        hm.status = (uint8_t*)mapped_alias(h_status);
        hm.err = h_err ? (uint8_t*)mapped_alias(h_err) : nullptr;
        if (!hm.status || (h_err && !hm.err)) hm = HostMirror{};
    }
    if (!d_pl && (rc = upload(e, 0, h_plays, n, &d_pl))) return rc;
    rc = device_out(e, 1, n, &d_err);
    if (rc) return rc;
    TAFL_SPEC_DISPATCH(e, (k_step_plays<BB, SP><<<grid_for(n, STEP_BLOCK), STEP_BLOCK, 0, e->stream>>>(cx, e->d_states, n, d_pl, nullptr,
                                                                                                    e->d_status, d_err, e->hist, hm,
                                                                                                    state_words(e->start), restart ? e->d_episode : nullptr)));
    TAFL_LAUNCHED(e);
    if (!hm.status) {
        if ((rc = download(e, h_status, e->d_status, n))) return rc;
        if (h_err && (rc = download(e, h_err, d_err, n))) return rc;
    }
    return sync_engine(e);
}
// page-locked host memory for the HOST-buffer calls (see tafl_step_status)
int tafl_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return TAFL_ERR_INVALID_ARG;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) { cudaGetLastError(); return TAFL_ERR_CUDA; }
    return TAFL_OK;
}
int tafl_host_free(void* p) {
    if (!p) return TAFL_OK;
    if (cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return TAFL_ERR_INVALID_ARG; }
    return TAFL_OK;
}
int tafl_host_register(void* p, size_t bytes) {
    if (!p || bytes == 0) return TAFL_ERR_INVALID_ARG;
    if (cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) { cudaGetLastError(); return TAFL_ERR_CUDA; }
    return TAFL_OK;
}
int tafl_host_unregister(void* p) {
    if (!p) return TAFL_ERR_INVALID_ARG;
    if (cudaHostUnregister(p) != cudaSuccess) { cudaGetLastError(); return TAFL_ERR_INVALID_ARG; }
    return TAFL_OK;
}
int tafl_step_moves_device(tafl_engine* e, const tafl_move* d_moves, uint8_t* d_status) {
    TAFL_CHECK_E(e);
    if (!d_moves) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    TAFL_SPEC_DISPATCH(e, (k_step_moves<BB, SP><<<grid_for(n, STEP_BLOCK), STEP_BLOCK, 0, e->stream>>>(cx, e->d_states, n, d_moves, 1u, nullptr, d_status)));
    TAFL_LAUNCHED(e);
    hist_forget_all(e);   // only Game::do_play (tafl_step / tafl_step_status) records history
    return sync_engine(e);
}

// ---- Game history / undo (game/mod.rs:77-108) -------------------------------------------------
int tafl_history_enable(tafl_engine* e, uint32_t capacity) {
    TAFL_CHECK_E(e);
    if (capacity == 0) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    cudaFree(e->hist.st); cudaFree(e->hist.rec); cudaFree(e->hist.len); cudaFree(e->hist.floor);
    e->hist = HistView{};
    const size_t n = e->n_games;
    HistView hv;
    hv.cap = capacity;
    hv.n = e->n_games;
    TAFL_CUDA_TRY(e, cudaMalloc(&hv.len, n * sizeof(uint32_t)));
    TAFL_CUDA_TRY(e, cudaMalloc(&hv.floor, n * sizeof(uint32_t)));
    TAFL_CUDA_TRY(e, cudaMalloc(&hv.rec, n * capacity * sizeof(tafl_play_record)));
    TAFL_CUDA_TRY(e, cudaMalloc(&hv.st, n * capacity * sizeof(tafl_state)));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(hv.len, 0, n * sizeof(uint32_t), e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(hv.floor, 0, n * sizeof(uint32_t), e->stream));
    e->hist = hv;
    return sync_engine(e);
}
static int need_hist(tafl_engine* e) {
    if (!e->hist.st) { e->err = "tafl_history_enable has not been called"; return TAFL_ERR_INVALID_ARG; }
    return TAFL_OK;
}
int tafl_undo_last_play(tafl_engine* e, const uint8_t* h_mask) {
    TAFL_CHECK_E(e);
    int rc = need_hist(e);
    if (rc) return rc;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    uint8_t* d_mask = nullptr;
    if (h_mask && (rc = upload(e, 0, h_mask, n, &d_mask))) return rc;
    k_undo<<<grid_for(n, 256), 256, 0, e->stream>>>(e->d_states, n, d_mask, e->hist);
    TAFL_LAUNCHED(e);
    return sync_engine(e);
}
int tafl_history_len(tafl_engine* e, uint32_t* h_len) {
    TAFL_CHECK_E(e);
    int rc = need_hist(e);
    if (rc) return rc;
    if (!h_len) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_len, e->hist.len, (size_t)e->n_games * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    return sync_engine(e);
}
int tafl_play_history(tafl_engine* e, uint32_t game, tafl_play_record* h_out, uint32_t cap, uint32_t* n_out) {
    TAFL_CHECK_E(e);
    int rc = need_hist(e);
    if (rc) return rc;
    if (game >= e->n_games || !n_out || (cap && !h_out)) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    uint32_t lf[2];
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&lf[0], e->hist.len + game, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&lf[1], e->hist.floor + game, sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    uint32_t held = lf[0] - lf[1];
    if (held > cap) { lf[1] += held - cap; held = cap; }   // the newest `cap` records
    for (uint32_t k = 0; k < held; ++k) {
        const uint32_t slot = (lf[1] + k) % e->hist.cap;
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_out + k, e->hist.rec + (size_t)slot * e->hist.n + game, sizeof(tafl_play_record),
                                         cudaMemcpyDeviceToHost, e->stream));
    }
    *n_out = held;
    return sync_engine(e);
}

int tafl_set_game_id_base(tafl_engine* e, uint64_t v) { TAFL_CHECK_E(e); e->game_id_base = v; return TAFL_OK; }
int tafl_set_random_ply(tafl_engine* e, uint64_t v) { TAFL_CHECK_E(e); e->ply = v; return TAFL_OK; }
uint64_t tafl_random_ply(const tafl_engine* e) { return e ? e->ply : 0; }

// enqueue k_step_random over games [first, first + n) on `stream`: `steps` plies starting at RNG ply `ply0`
// (+ *d_ply_base when given); pdl = launch with the programmatic-stream-serialization attribute
static int launch_step_random(tafl_engine* e, cudaStream_t stream, uint32_t first, uint32_t n, uint64_t seed, uint64_t ply0,
                              uint32_t steps, int auto_reset, const unsigned long long* d_ply_base = nullptr, bool pdl = false,
                              uint32_t* d_chain = nullptr) {
    cudaLaunchConfig_t cfg{};
    const bool chained = d_chain != nullptr && steps == 1 && e->step_single && e->step_minb == 8;
    const unsigned block = chained ? CHAIN_BLOCK : STEP_BLOCK;
    cfg.gridDim = dim3(grid_for(n, block));
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t lerr = cudaSuccess;
#define TAFL_STEP_LAUNCH_(MINB, SINGLE)                                                                               \
    TAFL_SPEC_DISPATCH(e, (lerr = cudaLaunchKernelEx(&cfg, k_step_random<BB, SP, MINB, false, SINGLE>, cx, e->d_states + first, n, \
                                                     state_words(e->start), seed, e->game_id_base + first, ply0, steps, \
                                                     auto_reset, e->d_stats, e->d_actions + first, d_ply_base,         \
                                                     (const tafl_state*)nullptr, (uint32_t*)nullptr, d_chain)))
#define TAFL_STEP_LAUNCH(MINB)                                     \
    do {                                                           \
        if (chained)                                               \
            TAFL_SPEC_DISPATCH(e, (lerr = cudaLaunchKernelEx(&cfg, k_step_random<BB, SP, 8, false, true, CHAIN_BLOCK>, cx, e->d_states + first, n, \
                                                             state_words(e->start), seed, e->game_id_base + first, ply0, steps, \
                                                             auto_reset, e->d_stats, e->d_actions + first, d_ply_base,         \
                                                             (const tafl_state*)nullptr, (uint32_t*)nullptr, d_chain)));      \
        else if (steps == 1 && e->step_single && MINB == 8) TAFL_STEP_LAUNCH_(8, true); \
        else TAFL_STEP_LAUNCH_(MINB, false);                       \
    } while (0)
    switch (e->step_minb) {   // resident CTAs per SM the kernel is compiled for (register cap), see DESIGN.md
        case 4: TAFL_STEP_LAUNCH(4); break;
        case 6: TAFL_STEP_LAUNCH(6); break;
        case 10: TAFL_STEP_LAUNCH(10); break;
        case 12: TAFL_STEP_LAUNCH(12); break;
        default: TAFL_STEP_LAUNCH(8); break;
    }
#undef TAFL_STEP_LAUNCH
#undef TAFL_STEP_LAUNCH_
    e->launches += 1;
    TAFL_CUDA_TRY(e, lerr);
    return TAFL_OK;
}

// n_steps single-ply launches as one graph: node i steps RNG ply *d_ply + i; nodes are joined by
// programmatic edges (captured PDL launches).  Re-captured when (n_steps, seed, flags) change.
static int step_graph_launch(tafl_engine* e, uint64_t seed, uint32_t n_steps, int auto_reset) {
    auto& k = e->step_graph_key;
    // CTA chaining needs the single-ply kernel, programmatic edges and a node index that fits the token
    // ... and pays when the grid is about one wave of CTAs (TAFL_STEP_CHAIN: 0 off, 1 by batch size, 2 always)
    const uint32_t n_ctas = grid_for(e->n_games, 64);
    const bool chain_fits = e->step_chain_on == 2 || (n_ctas > CHAIN_MIN_CTAS_PER_SM * (uint32_t)e->sm_count && n_ctas <= CHAIN_MAX_CTAS_PER_SM * (uint32_t)e->sm_count);
    const int chain = e->step_chain_on && chain_fits && e->step_pdl_on && e->step_single && e->step_minb == 8 && n_steps < (1u << CHAIN_POS_BITS);
    if (!e->step_graph || k.seed != seed || k.steps != n_steps || k.auto_reset != auto_reset || k.minb != e->step_minb ||
        k.spec != e->spec || k.gid_base != e->game_id_base || k.chain != chain) {
        if (e->step_graph) { cudaGraphExecDestroy(e->step_graph); e->step_graph = nullptr; }
        if (!e->d_ply) TAFL_CUDA_TRY(e, cudaMalloc(&e->d_ply, 2 * sizeof(unsigned long long)));
        const uint32_t ctas = grid_for(e->n_games, CHAIN_BLOCK);
        if (chain && (!e->d_chain || e->chain_ctas != ctas)) {
            if (e->d_chain) cudaFree(e->d_chain);
            e->d_chain = nullptr;
            TAFL_CUDA_TRY(e, cudaMalloc(&e->d_chain, ((size_t)ctas + 2) * sizeof(uint32_t)));   // tokens, time-out word, test hook
            TAFL_CUDA_TRY(e, cudaMemsetAsync(e->d_chain, 0, ((size_t)ctas + 2) * sizeof(uint32_t), e->stream));
            if (const char* v = std::getenv("TAFL_STEP_CHAIN_LOSE")) {   // tests: node v (1-based) of every graph launch loses CTA 0's token
                const uint32_t lose = (uint32_t)std::atoi(v);
                TAFL_CUDA_TRY(e, cudaMemcpyAsync(e->d_chain + ctas + 1, &lose, sizeof(lose), cudaMemcpyHostToDevice, e->stream));
                TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
            }
            e->chain_ctas = ctas;
        }
        const uint64_t before = e->launches;
        cudaGraph_t graph = nullptr;
        TAFL_CUDA_TRY(e, cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        int rc = TAFL_OK;
        for (uint32_t i = 0; i < n_steps && rc == TAFL_OK; ++i)
            rc = launch_step_random(e, e->stream, 0, e->n_games, seed, i, 1, auto_reset, e->d_ply, i > 0 && e->step_pdl_on,
                                    chain ? e->d_chain : nullptr);
        cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
        e->launches = before;   // capturing launched nothing
        if (rc != TAFL_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        TAFL_CUDA_TRY(e, ce);
        ce = cudaGraphInstantiate(&e->step_graph, graph, 0);
        cudaGraphDestroy(graph);
        TAFL_CUDA_TRY(e, ce);
        k.seed = seed; k.steps = n_steps; k.auto_reset = auto_reset; k.minb = e->step_minb; k.spec = e->spec; k.gid_base = e->game_id_base;
        k.chain = chain;
    }
    const unsigned long long ply[2] = {e->ply, ++e->chain_seq};   // the sequence number makes every token of every launch distinct
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(e->d_ply, ply, sizeof(ply), cudaMemcpyHostToDevice, e->stream));   // pageable: staged at call time
    TAFL_CUDA_TRY(e, cudaGraphLaunch(e->step_graph, e->stream));
    e->chain_dirty = chain != 0;
    e->launches += n_steps;
    e->ply += n_steps;
    return TAFL_OK;
}

int tafl_step_random_async(tafl_engine* e, uint64_t seed, uint32_t n_steps, int auto_reset, int spl) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    if (spl <= 0) spl = 1;
    if (n_steps) hist_forget_all(e);
    if (spl == 1 && n_steps >= 4 && e->step_graph_on) return step_graph_launch(e, seed, n_steps, auto_reset);
    uint32_t done = 0;
    while (done < n_steps) {
        const uint32_t chunk = (n_steps - done) < (uint32_t)spl ? (n_steps - done) : (uint32_t)spl;
        int rc = launch_step_random(e, e->stream, 0, e->n_games, seed, e->ply, chunk, auto_reset);
        if (rc) return rc;
        e->ply += chunk;
        done += chunk;
    }
    return TAFL_OK;
}

// By-value random stepping with HOST buffers: == tafl_set_states(h_in) + tafl_step_random(n_steps) +
// tafl_get_states(h_out), but the games are cut into chunks that flow through a few streams, so the
// H2D copy of chunk i+1, the kernel of chunk i and the D2H copy of chunk i-1 overlap (PCIe is full
// duplex).  Games are independent, so chunking cannot change any result.
int tafl_step_random_host(tafl_engine* e, const tafl_state* h_in, tafl_state* h_out, uint64_t seed, uint32_t n_steps,
                          int auto_reset) {
    TAFL_CHECK_E(e);
    if (!h_in || !h_out) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    const uint32_t n = e->n_games;
    int rc;
    if (!e->aux_ready) {
        for (auto& st : e->aux) TAFL_CUDA_TRY(e, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        e->aux_ready = true;
    }
    if (n_steps == 0) {   // set + get: validate on the host, nothing to overlap
        if ((rc = tafl_set_states(e, h_in, 0, n))) return rc;
        return tafl_get_states(e, h_out, 0, n);
    }
    hist_forget_all(e);
    // rows are staged next to the resident states and validated (side_len) by the step kernel itself
    void* p;
    if ((rc = scratch(e, 6, (size_t)n * sizeof(tafl_state) + 16, &p))) return rc;
    tafl_state* const d_in = (tafl_state*)p;
    uint32_t* const d_rej = (uint32_t*)(d_in + n);
    TAFL_CUDA_TRY(e, cudaMemsetAsync(d_rej, 0, sizeof(uint32_t), e->stream));
    TAFL_CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    uint32_t chunks = (uint32_t)e->host_chunks;
    if (chunks < 1) chunks = 1;
    uint32_t per = (n + chunks - 1) / chunks;
    per = (per + STEP_BLOCK - 1) / STEP_BLOCK * STEP_BLOCK;   // whole CTAs per chunk
    uint32_t c = 0;
    for (uint32_t first = 0; first < n; first += per, ++c) {
        const uint32_t cnt = (n - first) < per ? (n - first) : per;
        cudaStream_t st = e->aux[c % tafl_engine::N_AUX];
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(d_in + first, h_in + first, (size_t)cnt * sizeof(tafl_state), cudaMemcpyHostToDevice, st));
        cudaError_t lerr = cudaSuccess;
        TAFL_SPEC_DISPATCH(e, ((k_step_random<BB, SP, 8, true><<<grid_for(cnt, STEP_BLOCK), STEP_BLOCK, 0, st>>>(
                                   cx, e->d_states + first, cnt, state_words(e->start), seed, e->game_id_base + first, e->ply, n_steps,
                                   auto_reset, e->d_stats, e->d_actions + first, nullptr, d_in + first, d_rej)),
                               lerr = cudaGetLastError()));
        TAFL_CUDA_TRY(e, lerr);
        e->launches += 1;
        TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_out + first, e->d_states + first, (size_t)cnt * sizeof(tafl_state), cudaMemcpyDeviceToHost, st));
    }
    e->ply += n_steps;
    for (auto& st : e->aux) TAFL_CUDA_TRY(e, cudaStreamSynchronize(st));
    uint32_t rej = 0;
    TAFL_CUDA_TRY(e, cudaMemcpy(&rej, d_rej, sizeof(rej), cudaMemcpyDeviceToHost));
    if (rej) {
        e->err = "state side_len does not match the engine's board (those games kept their resident state)";
        return TAFL_ERR_INVALID_ARG;
    }
    return TAFL_OK;
}
int tafl_step_random(tafl_engine* e, uint64_t seed, uint32_t n_steps, int auto_reset, int spl) {
    int rc = tafl_step_random_async(e, seed, n_steps, auto_reset, spl);
    if (rc) return rc;
    return sync_engine(e);
}

uint32_t* tafl_actions_device(tafl_engine* e) { return e ? e->d_actions : nullptr; }
int tafl_actions_host(tafl_engine* e, uint32_t* h_actions, uint32_t first, uint32_t n) {
    TAFL_CHECK_E(e);
    if (!h_actions || (uint64_t)first + n > e->n_games) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_actions, e->d_actions + first, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    return sync_engine(e);
}

int tafl_stats_reset(tafl_engine* e) {
    TAFL_CHECK_E(e);
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemsetAsync(e->d_stats, 0, sizeof(DevStats), e->stream));
    return sync_engine(e);
}
int tafl_stats_get(tafl_engine* e, tafl_stats* out) {
    TAFL_CHECK_E(e);
    if (!out) return TAFL_ERR_INVALID_ARG;
    cudaSetDevice(e->device);
    TAFL_CUDA_TRY(e, cudaMemsetAsync(e->d_checksum, 0, sizeof(unsigned long long), e->stream));
    k_checksum<<<grid_for(e->n_games, 256), 256, 0, e->stream>>>(e->d_states, e->n_games, e->d_checksum);
    TAFL_LAUNCHED(e);
    DevStats ds;
    unsigned long long cs;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&ds, e->d_stats, sizeof(ds), cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(&cs, e->d_checksum, sizeof(cs), cudaMemcpyDeviceToHost, e->stream));
    int rc = sync_engine(e);
    if (rc) return rc;
    std::memset(out, 0, sizeof(*out));
    out->steps = ds.steps; out->legal_moves = ds.legal; out->captures = ds.caps;
    out->games_finished = ds.finished; out->finished_plies = ds.plies;
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 8; ++j) out->win_reason[i][j] = ds.win[i][j];
    for (int j = 0; j < 2; ++j) out->draw_reason[j] = ds.draw[j];
    out->state_checksum = cs;
    return TAFL_OK;
}

// ---- functional batch API -----------------------------------------------------------------------
#define TAFL_BV_PROLOGUE(e, st, n)                        \
    TAFL_CHECK_E(e);                                      \
    if (!(st) && (n)) return TAFL_ERR_INVALID_ARG;        \
    cudaSetDevice((e)->device);                           \
    {                                                     \
        int _rc = check_side_len((e), (st), (n));         \
        if (_rc) return _rc;                              \
    }

int tafl_validate_plays(tafl_engine* e, const tafl_state* h_st, const tafl_play* h_pl, uint32_t n, uint8_t* h_out) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; tafl_play* d_pl; uint8_t* d_out;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_pl, n, &d_pl)) || (rc = device_out(e, 2, n, &d_out))) return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_validate<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, d_pl, n, d_out)));
    TAFL_LAUNCHED(e);
    if ((rc = download(e, h_out, d_out, n))) return rc;
    return sync_engine(e);
}
int tafl_do_plays(tafl_engine* e, const tafl_state* h_st, const tafl_play* h_pl, uint32_t n, tafl_state* h_out_st,
                  tafl_play_effects* h_out_fx) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state *d_st, *d_ost; tafl_play* d_pl; tafl_play_effects* d_fx;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_pl, n, &d_pl)) || (rc = device_out(e, 2, n, &d_ost)) ||
        (rc = device_out(e, 3, n, &d_fx)))
        return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_do_plays<BB><<<grid_for(n, STEP_BLOCK), STEP_BLOCK, 0, e->stream>>>(cx, d_st, d_pl, n, d_ost, d_fx)));
    TAFL_LAUNCHED(e);
    if (h_out_st && (rc = download(e, h_out_st, d_ost, n))) return rc;
    if (h_out_fx && (rc = download(e, h_out_fx, d_fx, n))) return rc;
    return sync_engine(e);
}
int tafl_legal_moves_of(tafl_engine* e, const tafl_state* h_st, uint32_t n, tafl_move* h_moves, uint32_t* h_counts) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; tafl_move* d_mv; uint32_t* d_cnt;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = device_out(e, 1, (size_t)n * TAFL_MAX_MOVES, &d_mv)) ||
        (rc = device_out(e, 2, n, &d_cnt)))
        return rc;
    if ((rc = legal_moves_enqueue(e, d_st, n, d_mv, d_cnt))) return rc;
    if ((rc = download(e, h_moves, d_mv, (size_t)n * TAFL_MAX_MOVES)) || (rc = download(e, h_counts, d_cnt, n))) return rc;
    return sync_engine(e);
}
int tafl_piece_plays(tafl_engine* e, const tafl_state* h_st, const uint8_t* h_rc, uint32_t n, tafl_move* h_moves, int32_t* h_counts) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; uint8_t* d_rc; tafl_move* d_mv; int32_t* d_cnt;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_rc, (size_t)2 * n, &d_rc)) ||
        (rc = device_out(e, 2, (size_t)n * 32, &d_mv)) || (rc = device_out(e, 3, n, &d_cnt)))
        return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_piece_plays<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, d_rc, n, d_mv, d_cnt)));
    TAFL_LAUNCHED(e);
    if ((rc = download(e, h_moves, d_mv, (size_t)n * 32)) || (rc = download(e, h_counts, d_cnt, n))) return rc;
    return sync_engine(e);
}
static int captures_common(tafl_engine* e, const tafl_state* h_st, const tafl_play* h_pl, uint32_t n, tafl_play_effects* h_out, int sw) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; tafl_play* d_pl; tafl_play_effects* d_fx;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_pl, n, &d_pl)) || (rc = device_out(e, 2, n, &d_fx))) return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_captures<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, d_pl, n, d_fx, sw)));
    TAFL_LAUNCHED(e);
    if ((rc = download(e, h_out, d_fx, n))) return rc;
    return sync_engine(e);
}
int tafl_get_captures(tafl_engine* e, const tafl_state* s, const tafl_play* p, uint32_t n, tafl_play_effects* o) { return captures_common(e, s, p, n, o, 0); }
int tafl_detect_shieldwall(tafl_engine* e, const tafl_state* s, const tafl_play* p, uint32_t n, tafl_play_effects* o) { return captures_common(e, s, p, n, o, 1); }

static int enclosure_common(tafl_engine* e, const tafl_state* h_st, const tafl_enclosure_query* h_q, uint32_t n, tafl_enclosure* h_enc, uint8_t* h_sec) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; tafl_enclosure_query* d_q; tafl_enclosure* d_enc; uint8_t* d_sec;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_q, n, &d_q)) || (rc = device_out(e, 2, n, &d_enc)) ||
        (rc = device_out(e, 3, n, &d_sec)))
        return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_enclosure<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, d_q, n, d_enc, d_sec)));
    TAFL_LAUNCHED(e);
    if (h_enc && (rc = download(e, h_enc, d_enc, n))) return rc;
    if (h_sec && (rc = download(e, h_sec, d_sec, n))) return rc;
    return sync_engine(e);
}
int tafl_find_enclosure(tafl_engine* e, const tafl_state* s, const tafl_enclosure_query* q, uint32_t n, tafl_enclosure* o) { return enclosure_common(e, s, q, n, o, nullptr); }
int tafl_enclosure_secure(tafl_engine* e, const tafl_state* s, const tafl_enclosure_query* q, uint32_t n, uint8_t* o) { return enclosure_common(e, s, q, n, nullptr, o); }

int tafl_detect_exit_fort(tafl_engine* e, const tafl_state* h_st, uint32_t n, uint8_t* h_out) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; uint8_t* d_out;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = device_out(e, 1, n, &d_out))) return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_exit_fort<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, n, d_out)));
    TAFL_LAUNCHED(e);
    if ((rc = download(e, h_out, d_out, n))) return rc;
    return sync_engine(e);
}
int tafl_side_can_play(tafl_engine* e, const tafl_state* h_st, const uint8_t* h_sides, uint32_t n, uint8_t* h_out) {
    TAFL_BV_PROLOGUE(e, h_st, n);
    if (!n) return TAFL_OK;
    tafl_state* d_st; uint8_t *d_sides, *d_out;
    int rc;
    if ((rc = upload(e, 0, h_st, n, &d_st)) || (rc = upload(e, 1, h_sides, n, &d_sides)) || (rc = device_out(e, 2, n, &d_out))) return rc;
    TAFL_WIDTH_DISPATCH(e, (k_bv_can_play<BB><<<grid_for(n, 128), 128, 0, e->stream>>>(cx, d_st, d_sides, n, d_out)));
    TAFL_LAUNCHED(e);
    if ((rc = download(e, h_out, d_out, n))) return rc;
    return sync_engine(e);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// tafl_legal_moves_of_static: tafl_legal_moves_of without any device allocation
// ---------------------------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (code under graph capture, tight serving loops, harnesses that forbid it).  The device
// staging that tafl_legal_moves_of keeps in the engine's grown-on-demand scratch slots (states in, move
// rows and counts out) lives in ONE file-scope array instead: every GPU that loads the library holds
// TAFL_STATIC_SCRATCH_BYTES of it until the process exits.  All calls on a device share the array and it keeps
// the previous call's contents, so the move rows are cleared on the stream before every launch, and two calls
// that overlap on the GPU (two engines / two streams on one device) race: callers serialise them.
namespace {

constexpr size_t SS_STATES_OFF = 0;
constexpr size_t SS_MOVES_OFF = SS_STATES_OFF + (size_t)TAFL_STATIC_MAX_GAMES * sizeof(tafl_state);
constexpr size_t SS_COUNTS_OFF = SS_MOVES_OFF + (size_t)TAFL_STATIC_MAX_GAMES * TAFL_MAX_MOVES * sizeof(tafl_move);
constexpr size_t SS_BYTES = SS_COUNTS_OFF + (size_t)TAFL_STATIC_MAX_GAMES * sizeof(uint32_t);
static_assert(SS_BYTES == TAFL_STATIC_SCRATCH_BYTES, "include/tafl.h states the size of the static scratch");
static_assert(SS_MOVES_OFF % 256 == 0 && SS_COUNTS_OFF % 256 == 0, "sections stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[SS_BYTES];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address a kernel
// may use comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int SS_MAX_DEVICES = 64;
std::mutex g_ss_mutex;
unsigned char* g_ss_ptr[SS_MAX_DEVICES] = {};

int static_scratch_for(tafl_engine* e, unsigned char** out) {
    if (e->device < 0 || e->device >= SS_MAX_DEVICES) { e->err = "device index beyond the static scratch table"; return TAFL_ERR_UNSUPPORTED; }
    std::lock_guard<std::mutex> lk(g_ss_mutex);
    if (!g_ss_ptr[e->device]) {
        TAFL_CUDA_TRY(e, cudaSetDevice(e->device));
        void* p = nullptr;
        TAFL_CUDA_TRY(e, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { e->err = "cudaGetSymbolAddress returned a null address for the static scratch"; return TAFL_ERR_CUDA; }
        g_ss_ptr[e->device] = (unsigned char*)p;
    }
    *out = g_ss_ptr[e->device];
    return TAFL_OK;
}

}  // namespace

extern "C" {

uint32_t tafl_static_max_games(void) { return TAFL_STATIC_MAX_GAMES; }

int tafl_legal_moves_of_static(tafl_engine* e, const tafl_state* h_st, uint32_t n, tafl_move* h_moves, uint32_t* h_counts) {
    TAFL_CHECK_E(e);
    if (n > TAFL_STATIC_MAX_GAMES) {
        e->err = "tafl_legal_moves_of_static: n = " + std::to_string(n) + " states exceed the static scratch (TAFL_STATIC_MAX_GAMES = " +
                 std::to_string((unsigned)TAFL_STATIC_MAX_GAMES) + "); split the batch or use tafl_legal_moves_of";
        return TAFL_ERR_CAPACITY;
    }
    if (n && (!h_st || !h_moves || !h_counts)) return TAFL_ERR_INVALID_ARG;
    TAFL_CUDA_TRY(e, cudaSetDevice(e->device));
    int rc = check_side_len(e, h_st, n);
    if (rc) return rc;
    if (!n) return TAFL_OK;
    unsigned char* base;
    if ((rc = static_scratch_for(e, &base))) return rc;
    tafl_state* d_st = (tafl_state*)(base + SS_STATES_OFF);
    tafl_move* d_mv = (tafl_move*)(base + SS_MOVES_OFF);
    uint32_t* d_cnt = (uint32_t*)(base + SS_COUNTS_OFF);
    const size_t mv_bytes = (size_t)n * TAFL_MAX_MOVES * sizeof(tafl_move);
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(d_st, h_st, (size_t)n * sizeof(tafl_state), cudaMemcpyHostToDevice, e->stream));
    TAFL_CUDA_TRY(e, cudaMemsetAsync(d_mv, 0, mv_bytes, e->stream));          // rows past a game's count read as zero
    TAFL_CUDA_TRY(e, cudaMemsetAsync(d_cnt, 0, (size_t)n * sizeof(uint32_t), e->stream));
    if ((rc = legal_moves_enqueue(e, d_st, n, d_mv, d_cnt))) return rc;
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_moves, d_mv, mv_bytes, cudaMemcpyDeviceToHost, e->stream));
    TAFL_CUDA_TRY(e, cudaMemcpyAsync(h_counts, d_cnt, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    return sync_engine(e);
}

}  // extern "C"
