// tafl_core.cuh — the rules engine as branch-light bitboard code, one game per thread.
//
// Everything here is __host__ __device__: the CUDA kernels (tafl_kernels.cu) inline it, and
// tests/hostsim compiles the very same functions with g++ so the kernel math can be checked
// against the CPU oracle on millions of positions without a GPU.  (The host build is a test
// harness only; the product library has no CPU path.)
//
// Reference behaviour reproduced (paths relative to the reference tree, `game/`):
//   board/state.rs   bit layout, king nibbles, get/set/move/clear, count_pieces
//   game/logic.rs    validate_play_for_side, hostility, get_captures, detect_shieldwall,
//                    find_enclosure, enclosure_secure, detect_exit_fort, get_game_outcome,
//                    do_valid_play, side_can_play
//   play.rs          ValidPlayIterator (ray order V+,V-,H+,H-, distance ascending)
//   game/state.rs    RepetitionTracker
//   game/main.rs     get_all_possible_moves (canonical move order)
#pragma once
#include "../../include/tafl.h"
#include <string.h>

#include "tafl_bb.cuh"

namespace tafl {

enum { KIND_AS = 0, KIND_DS = 1, KIND_K = 2, KIND_NONE = 3 };
enum { DIR_VP = 0, DIR_VM = 1, DIR_HP = 2, DIR_HM = 3 };  // play.rs:157,166-183

// ---------------------------------------------------------------------------------------------
// Ctx: `GameLogic` (logic.rs:61-72) = rules + board geometry, pre-digested into masks on the
// host (tafl_ctx.hpp) and handed to every kernel as a __grid_constant__ parameter.
// ---------------------------------------------------------------------------------------------
// Kernel specialisations: for the reference's presets (preset.rs:12-134) every rule flag below is
// a compile-time constant, so the rule branches fold and the hot kernels shrink.  SPEC_GENERIC
// reads the same flags from the Ctx at run time and serves every other Ruleset / board size.  The
// engine only picks a specialisation after spec_matches() has compared each constant with the
// run-time Ctx built from the caller's Ruleset (tafl_ctx.hpp), so a table typo can cost speed but
// never correctness.
enum { SPEC_GENERIC = 0, SPEC_COPENHAGEN = 1, SPEC_TABLUT = 2, SPEC_BRANDUBH = 3, SPEC_COUNT = 4 };

template <int SP> struct SpecTraits;
// rules::COPENHAGEN + boards::COPENHAGEN (preset.rs:12-34,127-128)
template <> struct SpecTraits<SPEC_COPENHAGEN> {
    static constexpr int N = 11;
    TAFL_HD static constexpr bool noentry(int k) { return k != KIND_K; }      // ThroneRule::KingEntry
    TAFL_HD static constexpr bool pass_block(int) { return false; }
    TAFL_HD static constexpr bool may_corner(int k) { return k == KIND_K; }   // may_enter_corners: King
    TAFL_HD static constexpr bool slow(int) { return false; }
    TAFL_HD static constexpr bool edge_hostile(int) { return false; }
    TAFL_HD static constexpr bool sw_caps(int k) { return k != KIND_K; }      // shieldwall captures Soldier
    static constexpr bool king_hostile = true, king_captures = true;          // KingAttack::Armed
    static constexpr bool sbt_special = false;                                // KingStrength::Strong
    static constexpr int king_strength = TAFL_KING_STRONG;
    static constexpr bool shieldwall = true, sw_cmc = true, linnaean = false, exit_fort = true, edge_escape = false;
    static constexpr int enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
    static constexpr bool rep_enabled = true, rep_is_loss = true, draw_no_plays = false;
    static constexpr uint32_t n_reps = 3;
};
// rules::TABLUT (Linnaeus) + boards::TABLUT (preset.rs:80-100,134)
template <> struct SpecTraits<SPEC_TABLUT> {
    static constexpr int N = 9;
    TAFL_HD static constexpr bool noentry(int) { return true; }               // ThroneRule::NoEntry
    TAFL_HD static constexpr bool pass_block(int) { return false; }
    TAFL_HD static constexpr bool may_corner(int) { return true; }
    TAFL_HD static constexpr bool slow(int) { return false; }
    TAFL_HD static constexpr bool edge_hostile(int) { return false; }
    TAFL_HD static constexpr bool sw_caps(int) { return false; }
    static constexpr bool king_hostile = true, king_captures = true;
    static constexpr bool sbt_special = true;                                 // StrongByThrone + NoEntry
    static constexpr int king_strength = TAFL_KING_STRONG_BY_THRONE;
    static constexpr bool shieldwall = false, sw_cmc = false, linnaean = true, exit_fort = false, edge_escape = true;
    static constexpr int enclosure_win = TAFL_ENCL_NONE;
    static constexpr bool rep_enabled = true, rep_is_loss = false, draw_no_plays = true;
    static constexpr uint32_t n_reps = 3;
};
// rules::BRANDUBH (Federation) + boards::BRANDUBH (preset.rs:36-56,130)
template <> struct SpecTraits<SPEC_BRANDUBH> {
    static constexpr int N = 7;
    TAFL_HD static constexpr bool noentry(int k) { return k != KIND_K; }      // ThroneRule::KingEntry
    TAFL_HD static constexpr bool pass_block(int) { return false; }
    TAFL_HD static constexpr bool may_corner(int k) { return k == KIND_K; }
    TAFL_HD static constexpr bool slow(int) { return false; }
    TAFL_HD static constexpr bool edge_hostile(int) { return false; }
    TAFL_HD static constexpr bool sw_caps(int) { return false; }
    static constexpr bool king_hostile = true, king_captures = true;
    static constexpr bool sbt_special = true;                                 // StrongByThrone + KingEntry
    static constexpr int king_strength = TAFL_KING_STRONG_BY_THRONE;
    static constexpr bool shieldwall = false, sw_cmc = false, linnaean = false, exit_fort = false, edge_escape = false;
    static constexpr int enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
    static constexpr bool rep_enabled = true, rep_is_loss = true, draw_no_plays = false;
    static constexpr uint32_t n_reps = 3;
};

template <class BB, int SP = SPEC_GENERIC>
struct Ctx {
    BB board;        // every in-bounds tile
    BB edge;         // geometry.rs:111-116
    BB corners;      // geometry.rs:13-24
    BB throne;       // centre tile
    BB throne_adj;   // in-bounds neighbours of the throne
    BB not_col0, not_colL;   // board minus first / last column (guards for +-1 shifts)
    BB row0, rowL, col0, colL;
    BB sh[3];          // special_tile_hostile (logic.rs:76-82), in-bounds part, per victim kind
    BB dest_block[3];  // tiles a piece of this kind can never stop on / occupy
                       //   (logic.rs:190-193,201-206 == coords_occupiable logic.rs:250-266)
    int32_t N, mid;
    uint8_t edge_hostile[3];   // hostility.edge contains kind
    uint8_t pass_block[3];     // ray may not continue past the throne (NoPass / KingPass&&!king)
    uint8_t noentry[3];        // may not stop on the throne but passes it (NoEntry / KingEntry&&!king)
    uint8_t may_corner[3];
    uint8_t slow[3];
    uint8_t sw_caps[3];        // shieldwall.captures contains kind
    uint8_t king_hostile;      // king_attack in {Armed, Anvil}   (logic.rs:89-93)
    uint8_t king_captures;     // king_attack in {Armed, Hammer}  (logic.rs:609-611)
    uint8_t sbt_special;       // StrongByThrone && throne rule in {NoEntry, KingEntry} (logic.rs:621-625)
    uint8_t _pad[5];
    tafl_rules rules;

    // rule flags: run-time fields for SPEC_GENERIC, compile-time constants otherwise
#define TAFL_FLAG_K(name)                                                   \
    TAFL_HD bool f_##name(int k) const {                                    \
        if constexpr (SP == SPEC_GENERIC) return name[k] != 0;              \
        else return SpecTraits<SP>::name(k);                                \
    }
    TAFL_FLAG_K(noentry) TAFL_FLAG_K(pass_block) TAFL_FLAG_K(may_corner) TAFL_FLAG_K(slow) TAFL_FLAG_K(edge_hostile)
    TAFL_FLAG_K(sw_caps)
#undef TAFL_FLAG_K
#define TAFL_FLAG_0(acc, field, trait)                                      \
    TAFL_HD auto acc() const {                                              \
        if constexpr (SP == SPEC_GENERIC) return field;                     \
        else return SpecTraits<SP>::trait;                                  \
    }
    TAFL_FLAG_0(f_king_hostile, king_hostile != 0, king_hostile)
    TAFL_FLAG_0(f_king_captures, king_captures != 0, king_captures)
    TAFL_FLAG_0(f_sbt_special, sbt_special != 0, sbt_special)
    TAFL_FLAG_0(r_king_strength, (int)rules.king_strength, king_strength)
    TAFL_FLAG_0(r_shieldwall_enabled, rules.shieldwall_enabled != 0, shieldwall)
    TAFL_FLAG_0(r_sw_corners_may_close, rules.sw_corners_may_close != 0, sw_cmc)
    TAFL_FLAG_0(r_linnaean_capture, rules.linnaean_capture != 0, linnaean)
    TAFL_FLAG_0(r_exit_fort, rules.exit_fort != 0, exit_fort)
    TAFL_FLAG_0(r_edge_escape, rules.edge_escape != 0, edge_escape)
    TAFL_FLAG_0(r_enclosure_win, (int)rules.enclosure_win, enclosure_win)
    TAFL_FLAG_0(r_repetition_enabled, rules.repetition_enabled != 0, rep_enabled)
    TAFL_FLAG_0(r_repetition_is_loss, rules.repetition_is_loss != 0, rep_is_loss)
    TAFL_FLAG_0(r_draw_on_no_plays, rules.draw_on_no_plays != 0, draw_no_plays)
    TAFL_FLAG_0(r_n_repetitions, (uint32_t)rules.n_repetitions, n_reps)
    TAFL_FLAG_0(n, (int)N, N)
#undef TAFL_FLAG_0
    TAFL_HD int m() const { return n() / 2; }
};

// does the run-time Ctx (built from the caller's Ruleset and board size) equal the constants of SP?
template <int SP, class BB>
inline bool spec_matches(const Ctx<BB, SPEC_GENERIC>& c) {
    using T = SpecTraits<SP>;
    bool ok = c.N == T::N && c.mid == T::N / 2;
    for (int k = 0; k < 3; ++k)
        ok = ok && (c.noentry[k] != 0) == T::noentry(k) && (c.pass_block[k] != 0) == T::pass_block(k) &&
             (c.may_corner[k] != 0) == T::may_corner(k) && (c.slow[k] != 0) == T::slow(k) &&
             (c.edge_hostile[k] != 0) == T::edge_hostile(k) &&
             // shieldwall.captures only matters when the rule is on
             (!T::shieldwall || (c.sw_caps[k] != 0) == T::sw_caps(k));
    const tafl_rules& r = c.rules;
    ok = ok && (c.king_hostile != 0) == T::king_hostile && (c.king_captures != 0) == T::king_captures &&
         (c.sbt_special != 0) == T::sbt_special && (int)r.king_strength == T::king_strength &&
         (r.shieldwall_enabled != 0) == T::shieldwall && (!T::shieldwall || (r.sw_corners_may_close != 0) == T::sw_cmc) &&
         (r.linnaean_capture != 0) == T::linnaean && (r.exit_fort != 0) == T::exit_fort &&
         (r.edge_escape != 0) == T::edge_escape && (int)r.enclosure_win == T::enclosure_win &&
         (r.repetition_enabled != 0) == T::rep_enabled &&
         (!T::rep_enabled || ((r.repetition_is_loss != 0) == T::rep_is_loss && r.n_repetitions == T::n_reps)) &&
         (r.draw_on_no_plays != 0) == T::draw_no_plays;
    return ok;
}
// same bytes, different compile-time view
template <int SP, class BB>
inline Ctx<BB, SP> spec_cast(const Ctx<BB, SPEC_GENERIC>& c) {
    static_assert(sizeof(Ctx<BB, SP>) == sizeof(Ctx<BB, SPEC_GENERIC>), "Ctx layout is independent of SP");
    Ctx<BB, SP> out;
    memcpy(&out, &c, sizeof(out));
    return out;
}

template <class BB>
struct Pos {
    BB att, def;   // occupancy without the king nibbles
    int kr, kc;    // raw king nibbles (board/state.rs:127-131); survive the king's capture
    BB kbit;       // bit of the nibble tile when it is in bounds (else empty) — NOT occupancy
    TAFL_HD BB occ() const { return att | def; }
    TAFL_HD BB king() const { return def & kbit; }     // get_piece == King (board/state.rs:173-180)
    TAFL_HD BB dsold() const { return def & ~kbit; }
    TAFL_HD BB kind(int k) const { return k == KIND_AS ? att : (k == KIND_DS ? dsold() : king()); }
    TAFL_HD BB side(int def_side) const { return def_side ? def : att; }
};

template <class BB> TAFL_HD int idx_of(int r, int c) { return r * BB::STRIDE + c; }

template <class BB, int SP>
TAFL_HD BB king_bit(const Ctx<BB, SP>& cx, int kr, int kc) {
    return (kr < cx.n() && kc < cx.n()) ? BB::bit(idx_of<BB>(kr, kc)) : BB::zero();
}

template <class BB> struct Nib;
template <> struct Nib<BB64> {
    using State = tafl_state;
    TAFL_HD static void load(const tafl_state& s, Pos<BB64>& p) {
        p.att = BB64{s.att[0] & 0x0FFFFFFFFFFFFFFFull};
        p.def = BB64{s.def[0] & 0x0FFFFFFFFFFFFFFFull};
        p.kr = (int)(s.def[0] >> 60);
        p.kc = (int)(s.att[0] >> 60);
    }
    TAFL_HD static void store(const Pos<BB64>& p, tafl_state& s) {
        s.att[0] = p.att.v | ((uint64_t)p.kc << 60);
        s.def[0] = p.def.v | ((uint64_t)p.kr << 60);
        s.att[1] = 0;
        s.def[1] = 0;
    }
    // raw get_piece bits incl. nibble region (validate_play probes `from` before the bounds check)
    TAFL_HD static bool raw_att(const tafl_state& s, int i) { return i < 64 && ((s.att[0] >> i) & 1); }
    TAFL_HD static bool raw_def(const tafl_state& s, int i) { return i < 64 && ((s.def[0] >> i) & 1); }
};
template <> struct Nib<BB128> {
    using State = tafl_state;
    TAFL_HD static void load(const tafl_state& s, Pos<BB128>& p) {
        p.att = BB128{s.att[0], s.att[1] & 0x0FFFFFFFFFFFFFFFull};
        p.def = BB128{s.def[0], s.def[1] & 0x0FFFFFFFFFFFFFFFull};
        p.kr = (int)(s.def[1] >> 60);
        p.kc = (int)(s.att[1] >> 60);
    }
    TAFL_HD static void store(const Pos<BB128>& p, tafl_state& s) {
        s.att[0] = p.att.lo;
        s.att[1] = p.att.hi | ((uint64_t)p.kc << 60);
        s.def[0] = p.def.lo;
        s.def[1] = p.def.hi | ((uint64_t)p.kr << 60);
    }
    TAFL_HD static bool raw_att(const tafl_state& s, int i) {
        return i < 128 && (((i < 64 ? s.att[0] >> i : s.att[1] >> (i - 64))) & 1);
    }
    TAFL_HD static bool raw_def(const tafl_state& s, int i) {
        return i < 128 && (((i < 64 ? s.def[0] >> i : s.def[1] >> (i - 64))) & 1);
    }
};

// BitfieldBoardState<U256> / <U512> in tafl_state_w: NW words per side, king nibbles in the top 4 bits of the last one
template <int NW, int S> struct Nib<BBW<NW, S>> {
    using State = tafl_state_w;
    using BB = BBW<NW, S>;
    TAFL_HD static void load(const tafl_state_w& s, Pos<BB>& p) {
#pragma unroll
        for (int i = 0; i < NW; ++i) { p.att.w[i] = s.att[i]; p.def.w[i] = s.def[i]; }
        p.att.w[NW - 1] &= 0x0FFFFFFFFFFFFFFFull;
        p.def.w[NW - 1] &= 0x0FFFFFFFFFFFFFFFull;
        p.kr = (int)(s.def[NW - 1] >> 60);
        p.kc = (int)(s.att[NW - 1] >> 60);
    }
    TAFL_HD static void store(const Pos<BB>& p, tafl_state_w& s) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { s.att[i] = i < NW ? p.att.w[i < NW ? i : 0] : 0ull; s.def[i] = i < NW ? p.def.w[i < NW ? i : 0] : 0ull; }
        s.att[NW - 1] |= (uint64_t)p.kc << 60;
        s.def[NW - 1] |= (uint64_t)p.kr << 60;
    }
    TAFL_HD static bool raw_att(const tafl_state_w& s, int i) { return i >= 0 && i < BB::BITS && ((s.att[i >> 6] >> (i & 63)) & 1); }
    TAFL_HD static bool raw_def(const tafl_state_w& s, int i) { return i >= 0 && i < BB::BITS && ((s.def[i >> 6] >> (i & 63)) & 1); }
};

template <class BB, int SP, class ST>
TAFL_HD Pos<BB> load_pos(const Ctx<BB, SP>& cx, const ST& s) {
    Pos<BB> p;
    Nib<BB>::load(s, p);
    p.kbit = king_bit(cx, p.kr, p.kc);
    return p;
}
template <class BB, class ST>
TAFL_HD void store_pos(const Pos<BB>& p, ST& s) { Nib<BB>::store(p, s); }

// one-tile steps of a tile set, results clipped to the board
template <class BB, int SP> TAFL_HD BB step_up(const Ctx<BB, SP>&, BB x) { return x.template shr<BB::STRIDE>(); }              // row-1
template <class BB, int SP> TAFL_HD BB step_down(const Ctx<BB, SP>& cx, BB x) { return x.template shl<BB::STRIDE>() & cx.board; }  // row+1
template <class BB, int SP> TAFL_HD BB step_left(const Ctx<BB, SP>& cx, BB x) { return x.template shr<1>() & cx.not_colL; }     // col-1
template <class BB, int SP> TAFL_HD BB step_right(const Ctx<BB, SP>& cx, BB x) { return x.template shl<1>() & cx.not_col0; }    // col+1
template <class BB, int SP> TAFL_HD BB dilate4(const Ctx<BB, SP>& cx, BB x) {
    return step_up(cx, x) | step_down(cx, x) | step_left(cx, x) | step_right(cx, x);
}
template <class BB, int SP> TAFL_HD BB step_dir(const Ctx<BB, SP>& cx, BB x, int dir) {
    switch (dir) {
        case DIR_VP: return step_down(cx, x);
        case DIR_VM: return step_up(cx, x);
        case DIR_HP: return step_right(cx, x);
        default: return step_left(cx, x);
    }
}

// kind of the piece on tile bit `b` (board/state.rs:173-187)
template <class BB>
TAFL_HD int kind_at(const Pos<BB>& p, BB b) {
    if ((p.def & b).any()) return (b == p.kbit) ? KIND_K : KIND_DS;
    if ((p.att & b).any()) return KIND_AS;
    return KIND_NONE;
}

// tile_hostile (logic.rs:85-99) as a tile set, for a victim of kind vk
template <class BB, int SP>
TAFL_HD BB hostile_set(const Ctx<BB, SP>& cx, const Pos<BB>& p, int vk) {
    BB empty = cx.board & ~p.occ();
    BB pieces = (vk == KIND_AS) ? (cx.f_king_hostile() ? p.def : p.dsold()) : p.att;
    return pieces | (empty & cx.sh[vk]);
}

// ---------------------------------------------------------------------------------------------
// Rays of one piece, in "line space" (bit i of a line word = i-th tile of the row / column).
// ---------------------------------------------------------------------------------------------
template <class BB, int SP>
TAFL_HD uint32_t extract_row(const Ctx<BB, SP>& cx, BB x, int row) {
    return x.extract32(row * BB::STRIDE) & ((1u << cx.n()) - 1u);
}
// Column gather by multiplication: after the shift the column's bits sit at multiples of STRIDE; the
// multiplier sum_i 2^(T - (STRIDE-1) i) moves bit STRIDE*i to T + i, and no two partial products land
// on the same bit (STRIDE (i-i') = (STRIDE-1)(j-j') has no solution inside one word), so there are no
// carries.  The 64-bit multiplies run on the FMA pipe, which this integer code leaves idle.
TAFL_HD uint32_t gather_col(BB64 t) {
    constexpr uint64_t M = (1ull << 0) | (1ull << 7) | (1ull << 14) | (1ull << 21) | (1ull << 28) | (1ull << 35) | (1ull << 42);
    constexpr uint64_t K = (1ull << 57) | (1ull << 51) | (1ull << 45) | (1ull << 39) | (1ull << 33) | (1ull << 27) | (1ull << 21);
    return (uint32_t)(((t.v & M) * K) >> 57);
}
TAFL_HD uint32_t gather_col(BB128 t) {
    constexpr uint64_t M6 = (1ull << 0) | (1ull << 11) | (1ull << 22) | (1ull << 33) | (1ull << 44) | (1ull << 55);
    constexpr uint64_t M5 = (1ull << 0) | (1ull << 11) | (1ull << 22) | (1ull << 33) | (1ull << 44);
    constexpr uint64_t K = (1ull << 58) | (1ull << 48) | (1ull << 38) | (1ull << 28) | (1ull << 18) | (1ull << 8);
    const uint32_t lo = (uint32_t)(((t.lo & M6) * K) >> 58);            // rows 0..5: bits 0, 11, .., 55
    const uint32_t hi = (uint32_t)((((t.hi >> 2) & M5) * K) >> 58);     // rows 6..10: bits 66, 77, .., 110
    return lo | (hi << 6);
}
template <int NW, int S>
TAFL_HD uint32_t gather_col(BBW<NW, S> t) {   // wide boards: the rules kernels are not the hot path, a plain bit loop
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 21 && i * S < 64 * NW; ++i) r |= (uint32_t)((t.w[(i * S) >> 6] >> ((i * S) & 63)) & 1ull) << i;
    return r;
}
template <class BB, int SP>
TAFL_HD uint32_t extract_col(const Ctx<BB, SP>& cx, BB x, int col) {
    return gather_col(x.shrv(col)) & ((1u << cx.n()) - 1u);
}

struct Rays {
    uint32_t m[4];  // destination positions along the line per direction (DIR_*), bit = row (V) / col (H)
    int cnt[4];
};

// ValidPlayIterator (play.rs:189-225) + validate_play_for_side (logic.rs:159-214) for the piece
// of kind `k` on (row, col): the set of tiles it may move to, per direction.
template <class BB, int SP>
TAFL_HD Rays piece_rays(const Ctx<BB, SP>& cx, BB occ, int k, int row, int col) {
    Rays r;
    const int N = cx.n(), mid = cx.m();
    const uint32_t full = (1u << N) - 1u;
#pragma unroll
    for (int ax = 0; ax < 2; ++ax) {
        // ax 0: vertical (line = column `col`, position = row); ax 1: horizontal
        const int line = ax == 0 ? col : row, p = ax == 0 ? row : col;
        uint32_t lw = ax == 0 ? extract_col(cx, occ, col) : extract_row(cx, occ, row);
        // plus direction: free run above p
        uint32_t x = (lw >> (p + 1)) | (1u << (N - 1 - p));
        int lenp = ctz32(x);
        uint32_t mp = (((1u << lenp) - 1u) << (p + 1)) & full;
        // minus direction
        uint32_t y = lw & ((1u << p) - 1u);
        int top = 32 - clz32(y);
        uint32_t mm = ((1u << p) - 1u) & ~((1u << top) - 1u);
        const bool has_throne = (line == mid);
        if (has_throne && cx.f_pass_block(k)) {  // logic.rs:194-200: may reach, not cross, the throne
            if (p < mid) mp &= (2u << mid) - 1u;
            if (p > mid) mm &= ~((1u << mid) - 1u);
        }
        uint32_t blocked = 0;
        if (has_throne && cx.f_noentry(k)) blocked |= 1u << mid;                                   // logic.rs:201-206
        if ((line == 0 || line == N - 1) && !cx.f_may_corner(k)) blocked |= 1u | (1u << (N - 1));   // logic.rs:190-193
        mp &= ~blocked;
        mm &= ~blocked;
        if (cx.f_slow(k)) {  // logic.rs:207-210 (a skipped throne does not extend a slow piece's reach)
            mp &= 1u << (p + 1);
            mm &= (p > 0) ? (1u << (p - 1)) : 0u;
        }
        r.m[ax * 2 + 0] = mp;
        r.m[ax * 2 + 1] = mm;
        r.cnt[ax * 2 + 0] = popc32(mp);
        r.cnt[ax * 2 + 1] = popc32(mm);
    }
    return r;
}

// position of the j-th (0-based, ascending) set bit of m; caller guarantees j < popc(m).
// Branch-free 5-level search on masked popcounts (no data-dependent loop: lanes stay converged).
TAFL_HD int nth_bit32(uint32_t m, int j) {
    int pos = 0;
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const int c = popc32((m >> pos) & ((1u << half) - 1u));
        const bool up = j >= c;
        j -= up ? c : 0;
        pos += up ? half : 0;
    }
    return pos;
}
// j-th destination (distance ascending) of direction `dir` -> position along the line
TAFL_HD int ray_nth(uint32_t m, int dir, int j) {
    // plus directions enumerate ascending positions, minus directions descending ones
    return nth_bit32(m, (dir & 1) == 0 ? j : popc32(m) - 1 - j);
}

// Full legal-move list of the side to move, canonical order (game/main.rs:33-43).
// Simple per-piece generator; the kernels use it for list materialisation.
template <class BB, int SP>
TAFL_HD int gen_moves_side(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side, tafl_move* out, int cap) {
    BB own = p.side(def_side);
    BB occ = p.occ();
    int n = 0;
    while (own.any()) {
        int s = own.ctz();
        own = own & ~BB::bit(s);
        int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
        int k = def_side ? ((BB::bit(s) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
        Rays r = piece_rays(cx, occ, k, row, col);
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            uint32_t m = r.m[d];
            while (m) {
                int pos;
                if ((d & 1) == 0) { pos = ctz32(m); m &= m - 1; }
                else { pos = 31 - clz32(m); m &= ~(1u << pos); }
                int to = (d < 2) ? idx_of<BB>(pos, col) : idx_of<BB>(row, pos);
                if (n < cap) out[n] = (tafl_move)(s | (to << 7));
                ++n;
            }
        }
    }
    return n;
}
template <class BB, int SP, class ST>
TAFL_HD int gen_moves(const Ctx<BB, SP>& cx, const ST& st, tafl_move* out, int cap) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;  // logic.rs:165-167
    Pos<BB> p = load_pos(cx, st);
    return gen_moves_side(cx, p, (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, out, cap);
}

// side_can_play (logic.rs:837-846): does any piece of the side have a legal play?
// Timing-only builds (scripts/gpu_ablate.sh, build.py -DTAFL_ABL_<BLOCK>): one block of a ply switched off, to read what it
// costs in TIME — the instruction counts do not say (profiles/sweeps/r02n_step_ablation.txt).  Results are wrong by construction.
#define TAFL_ABL_ENCL_ON 1
#define TAFL_ABL_EXITFORT_ON 1
#define TAFL_ABL_CANPLAY_ON 1
#define TAFL_ABL_SW_ON 1
#define TAFL_ABL_REP_ON 1
#if defined(TAFL_ABL_ENCL)
#undef TAFL_ABL_ENCL_ON
#define TAFL_ABL_ENCL_ON 0
#endif
#if defined(TAFL_ABL_EXITFORT)
#undef TAFL_ABL_EXITFORT_ON
#define TAFL_ABL_EXITFORT_ON 0
#endif
#if defined(TAFL_ABL_CANPLAY)
#undef TAFL_ABL_CANPLAY_ON
#define TAFL_ABL_CANPLAY_ON 0
#endif
#if defined(TAFL_ABL_SW)
#undef TAFL_ABL_SW_ON
#define TAFL_ABL_SW_ON 0
#endif
#if defined(TAFL_ABL_REP)
#undef TAFL_ABL_REP_ON
#define TAFL_ABL_REP_ON 0
#endif
#define TAFL_ABL_KING_ON 1
#define TAFL_ABL_CAPS_ON 1
#if defined(TAFL_ABL_KING)
#undef TAFL_ABL_KING_ON
#define TAFL_ABL_KING_ON 0
#endif
#if defined(TAFL_ABL_CAPS)
#undef TAFL_ABL_CAPS_ON
#define TAFL_ABL_CAPS_ON 0
#endif
#define TAFL_ABL_KEEP(B) (TAFL_ABL_##B##_ON != 0)

template <class BB, int SP>
TAFL_HD bool side_can_play(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side) {
    BB empty = cx.board & ~p.occ();
    // Fast path (almost always decides): some piece has an empty neighbour it may stop on.  The
    // one-step test below does not depend on the direction, so the four directions fold into one
    // dilation per piece class; only the hop over the empty throne needs the per-direction walk.
    {
        const int ks = def_side ? KIND_DS : KIND_AS;
        const BB srcS = def_side ? p.dsold() : p.att;
        const BB free = empty & ~cx.dest_block[ks];
#if !defined(TAFL_CANPLAY_NO_SHORTCUT)
        if ((step_right(cx, srcS) & free).any()) return true;   // some soldier can step one tile to the right: nearly always
#endif
        if ((dilate4(cx, srcS) & free).any()) return true;
        if (def_side && (dilate4(cx, p.king()) & empty & ~cx.dest_block[KIND_K]).any()) return true;
    }
    bool any = false;
#pragma unroll 1
    for (int kk = 0; kk < 2; ++kk) {
        int k = def_side ? (kk == 0 ? KIND_DS : KIND_K) : KIND_AS;
        if (!def_side && kk == 1) break;
        BB src = p.kind(k);
        BB ok = ~cx.dest_block[k];
#pragma unroll 1
        for (int d = 0; d < 4; ++d) {
            BB s1 = step_dir(cx, src, d) & empty;
            any |= (s1 & ok).any();
            if (cx.f_noentry(k) && !cx.f_slow(k)) {  // hop over the empty throne (play.rs:208-211)
                BB s2 = step_dir(cx, s1 & cx.throne, d) & empty & ok;
                any |= s2.any();
            }
        }
    }
    return any;
}

// ---------------------------------------------------------------------------------------------
// Captures
// the tile-by-tile walk of dir_sw_search (logic.rs:471-530), reached only through the pre-filter
template <class BB, int SP>
TAFL_COLD bool shieldwall_walk(const Ctx<BB, SP>& cx, const Pos<BB>& p, int mover_def, int to_r, int to_c, int horizontal, int away,
                               BB* caps) {
    const int N = cx.n();
    const BB own = p.side(mover_def), opp = p.side(!mover_def), occ = p.occ();
    const bool cmc = cx.r_sw_corners_may_close() != 0;
    BB wall = BB::zero();
    bool found = false;
    for (int dd = 0; dd < 2 && !found; ++dd) {
        const int dir = dd == 0 ? -1 : 1;
        wall = BB::zero();
        int cnt = 0, r = to_r, c = to_c;
        for (int it = 0; it <= N; ++it) {   // at most N - 1 tiles along an edge
            if (horizontal) c += dir; else r += dir;
            if (r < 0 || c < 0 || r >= N || c >= N) break;  // ran off the board: None
            BB tb = BB::bit(idx_of<BB>(r, c));
            bool is_occ = (occ & tb).any(), is_corner = (cx.corners & tb).any();
            if (!(is_occ || (cmc && is_corner))) break;
            if (!is_occ) { found = cnt >= 2; break; }  // empty closing corner
            if ((opp & tb).any()) {
                int pr = horizontal ? r + away : r, pc = horizontal ? c : c + away;
                BB pin = BB::bit(idx_of<BB>(pr, pc));
                if ((own & pin).any()) { wall = wall | tb; ++cnt; }
                else break;  // not pinned by a piece of the mover
            }
            if ((own & tb).any() || (is_corner && cmc)) { found = cnt >= 2; break; }
        }
    }
    if (!found) return false;
    BB keep = BB::zero();
    const int osk = mover_def ? KIND_AS : KIND_DS;
    if (cx.f_sw_caps(osk)) keep = keep | ~p.kbit;
    if (cx.f_sw_caps(KIND_K)) keep = keep | p.kbit;
    *caps = wall & keep;
    return true;
}

// ---------------------------------------------------------------------------------------------
// detect_shieldwall (logic.rs:535-569) + dir_sw_search (logic.rs:471-530).
// `mover_def` is state.side_to_play.  Returns true for Some(..) and the captured tiles in *caps.
template <class BB, int SP>
TAFL_HD bool detect_shieldwall(const Ctx<BB, SP>& cx, const Pos<BB>& p, int mover_def, int to_r, int to_c, BB* caps) {
    *caps = BB::zero();
    if (!TAFL_ABL_KEEP(SW) || !cx.r_shieldwall_enabled()) return false;
    const int N = cx.n();
    if (to_r < 0 || to_c < 0 || to_r >= N || to_c >= N) return false;
    const BB own = p.side(mover_def), opp = p.side(!mover_def);
#if defined(TAFL_SW_PREFILTER_TILES)   // measured r02b: 11.63 us per ply against 11.40 us for the index form below; kept for reference
    {   // Exact pre-filter, as tile sets, so that every lane of a warp runs the same few dozen instructions
        // whether or not its move landed on an edge.  The search counts enemy pieces pinned by a piece of the
        // mover, walking away from `to` along the edge, and needs cnt >= 2 (logic.rs:485-522); any other tile
        // ends the run.  So a wall exists only if the first TWO tiles on one side of `to` are both pinned
        // enemies.  Ph / Pv = pinned enemies on the horizontal / vertical edges (the pinning piece sits one
        // step inward); they are confined to the two edge rows / columns, so the shifted tests cannot wrap.
        constexpr int S = BB::STRIDE;
        const BB tob = BB::bit(idx_of<BB>(to_r, to_c));
        const BB hrows = cx.row0 | cx.rowL;
        const BB Ph = opp & ((cx.row0 & own.template shr<S>()) | (cx.rowL & own.template shl<S>()));
        const BB Pv = opp & ((cx.col0 & own.template shr<1>()) | (cx.colL & own.template shl<1>()));
        const BB hit_h = hrows & ((Ph.template shr<1>() & Ph.template shr<2>()) | (Ph.template shl<1>() & Ph.template shl<2>()));
        const BB hit_v = (cx.col0 | cx.colL) & ~hrows &
                         ((Pv.template shr<S>() & Pv.template shr<2 * S>()) | (Pv.template shl<S>() & Pv.template shl<2 * S>()));
        if (!(tob & (hit_h | hit_v)).any()) return false;
    }
    int horizontal, away;  // away: +1 / -1 along the other axis (rows first: a corner tile searches along its row)
    if (to_r == 0) { horizontal = 1; away = 1; }
    else if (to_r == N - 1) { horizontal = 1; away = -1; }
    else if (to_c == 0) { horizontal = 0; away = 1; }
    else { horizontal = 0; away = -1; }
#else
    int horizontal, away;  // away: +1 / -1 along the other axis
    if (to_r == 0) { horizontal = 1; away = 1; }
    else if (to_r == N - 1) { horizontal = 1; away = -1; }
    else if (to_c == 0) { horizontal = 0; away = 1; }
    else if (to_c == N - 1) { horizontal = 0; away = -1; }
    else return false;
    {   // Exact pre-filter.  The search counts enemy pieces pinned by a piece of the mover, walking
        // away from `to` along the edge, and needs cnt >= 2 (logic.rs:485-522); any other tile ends
        // the run.  So a wall exists only if the first TWO tiles on one side of `to` are both pinned
        // enemies: eight single-tile tests in index arithmetic; the tile-by-tile walk runs for real candidates only.
        const int t = idx_of<BB>(to_r, to_c);
        const int de = horizontal ? 1 : BB::STRIDE;               // along the edge
        const int in = (horizontal ? BB::STRIDE : 1) * away;      // from the edge tile to its inward neighbour
        const int x = horizontal ? to_c : to_r;                   // position along the edge
        auto pinned = [&](int i) { return opp.test(i) & own.test(i + in); };
        const bool minus = (x >= 2) & pinned(t - de) & pinned(t - 2 * de);
        const bool plus = (x + 2 <= N - 1) & pinned(t + de) & pinned(t + 2 * de);
        if (!(minus | plus)) return false;
    }
#endif
    return shieldwall_walk(cx, p, mover_def, to_r, to_c, horizontal, away, caps);
}

// get_captures (logic.rs:604-699) on the post-move board.  piece_def / piece_kind describe the
// moving piece, mover_def is state.side_to_play (used by shield wall and Linnaean capture).
template <class BB, int SP>
TAFL_HD BB get_captures(const Ctx<BB, SP>& cx, const Pos<BB>& p, int piece_def, int piece_kind, int mover_def,
                        int to_r, int to_c) {
    BB caps = BB::zero();
    const int N = cx.n();
    const bool to_in = to_r >= 0 && to_c >= 0 && to_r < N && to_c < N;
    if (TAFL_ABL_KEEP(CAPS) && to_in && (piece_kind != KIND_K || cx.f_king_captures())) {
        const BB tob = BB::bit(idx_of<BB>(to_r, to_c));
        const BB opp = p.side(!piece_def);
        const BB kingb = p.king();
        // the four neighbours of `to`, per direction (n = to + d)
        const BB nVP = step_down(cx, tob), nVM = step_up(cx, tob), nHP = step_right(cx, tob), nHM = step_left(cx, tob);
        // ---- soldier victims (logic.rs:634-646,675): custodial against the far tile, all four
        // directions as masks.  X aligned to n: bit n is set iff the tile n + d is hostile to the
        // victim; an off-board far tile is hostile iff hostility.edge says so (logic.rs:103-114).
        {
            const int vs = piece_def ? KIND_AS : KIND_DS;
            const BB oppS = opp & ~kingb;
            const BB host = hostile_set(cx, p, vs);
            const bool eh = cx.f_edge_hostile(vs) != 0;
            BB hit = (nVP & (step_up(cx, host) | (eh ? cx.rowL : BB::zero()))) |
                     (nVM & (step_down(cx, host) | (eh ? cx.row0 : BB::zero()))) |
                     (nHP & (step_left(cx, host) | (eh ? cx.colL : BB::zero()))) |
                     (nHM & (step_right(cx, host) | (eh ? cx.col0 : BB::zero())));
            caps = hit & oppS;
        }
        // ---- Linnaean capture (logic.rs:676-685, 859-879): the far tile is the throne holding the
        // king and exactly 3 of the throne's neighbours are hostile to the king
        if (cx.r_linnaean_capture() && !mover_def && p.kbit == cx.throne) {
            const BB hostK = hostile_set(cx, p, KIND_K);
            if ((cx.throne_adj & hostK).popc() == 3) {
                const BB t = cx.throne;
                caps = caps | (opp & ((nVP & step_up(cx, t)) | (nVM & step_down(cx, t)) | (nHP & step_left(cx, t)) |
                                      (nHM & step_right(cx, t))));
            }
        }
        // ---- king victim (rare: the king stands beside `to`)
        if (((nVP | nVM | nHP | nHM) & kingb & opp).any()) {
            // king strength (logic.rs:225-245), from the nibble tile
            const bool beside = (p.kbit & cx.throne_adj).any();
            const bool strong = cx.r_king_strength() == TAFL_KING_STRONG ||
                                (cx.r_king_strength() == TAFL_KING_STRONG_BY_THRONE &&
                                 (beside || (p.kbit & cx.throne).any()));
            const BB hostK = hostile_set(cx, p, KIND_K);
            const BB n = kingb;
            const int d = (nVP & kingb).any() ? DIR_VP : (nVM & kingb).any() ? DIR_VM : (nHP & kingb).any() ? DIR_HP : DIR_HM;
            bool taken = false;
            if (beside && cx.f_sbt_special()) {  // logic.rs:621-632
                BB nb = dilate4(cx, n);
                taken = !(nb & ~cx.throne & ~hostK).any();
            }
            if (!taken) {
                BB far = step_dir(cx, n, d);
                const bool far_h = far.any() ? (far & hostK).any() : (cx.f_edge_hostile(KIND_K) != 0);  // logic.rs:103-114
                if (far_h) {
                    taken = true;
                    if (strong) {  // logic.rs:647-674
                        BB q1 = step_dir(cx, n, d < 2 ? DIR_HP : DIR_VP), q2 = step_dir(cx, n, d < 2 ? DIR_HM : DIR_VM);
                        bool h1 = q1.any() ? (q1 & hostK).any() : (cx.f_edge_hostile(KIND_K) != 0);
                        bool h2 = q2.any() ? (q2 & hostK).any() : (cx.f_edge_hostile(KIND_K) != 0);
                        taken = h1 && h2;
                    }
                }
            }
            if (taken) caps = caps | n;
        }
    }
    // A shield wall starts with an enemy piece on the edge tile next to `to` (the first tile dir_sw_search looks at must be
    // a pinned enemy, logic.rs:485-522): one uniform test for every lane, so that only the few with such a neighbour enter
    // the edge case analysis below (a quarter of all moves land on an edge)
    BB walled;
#if !defined(TAFL_SW_NO_QUICK_REJECT)
    if (to_in && cx.r_shieldwall_enabled()) {
        const BB tob = BB::bit(idx_of<BB>(to_r, to_c));
        if (!(tob & cx.edge).any() || !(dilate4(cx, tob) & cx.edge & p.side(!mover_def)).any()) return caps;
    }
#endif
    if (detect_shieldwall(cx, p, mover_def, to_r, to_c, &walled)) caps = caps | walled;  // logic.rs:691-696
    return caps;
}

// ---------------------------------------------------------------------------------------------
// Enclosures
// ---------------------------------------------------------------------------------------------
template <class BB>
struct Encl {
    BB region;    // occupied | unoccupied (logic.rs:25-38)
    BB boundary;
    bool found;
};
TAFL_HD bool pieceset_has(uint16_t set, int kind) {
    uint16_t m = kind == KIND_AS ? TAFL_PIECESET_PIECE(TAFL_SOLDIER, TAFL_ATTACKER)
               : kind == KIND_DS ? TAFL_PIECESET_PIECE(TAFL_SOLDIER, TAFL_DEFENDER)
                                 : TAFL_PIECESET_PIECE(TAFL_KING, TAFL_DEFENDER);
    return (set & m) != 0;
}
template <class BB>
TAFL_HD BB pieceset_bb(const Pos<BB>& p, uint16_t set) {
    BB r = BB::zero();
    if (pieceset_has(set, KIND_AS)) r = r | p.att;
    if (pieceset_has(set, KIND_DS)) r = r | p.dsold();
    if (pieceset_has(set, KIND_K)) r = r | p.king();
    return r;
}
// find_enclosure (logic.rs:309-401).  The reference's span fill visits exactly the 4-connected
// component of fillable tiles (empty or `enclosed` pieces) around the start and probes exactly
// that component's 4-neighbourhood, so the result is computed by mask dilation.
template <class BB, int SP>
TAFL_HD Encl<BB> find_enclosure(const Ctx<BB, SP>& cx, const Pos<BB>& p, int r, int c, BB enclosed, BB enclosing,
                                bool abort_on_edge, bool abort_on_corner) {
    Encl<BB> e;
    e.region = BB::zero();
    e.boundary = BB::zero();
    e.found = false;
    if (r < 0 || c < 0 || r >= cx.n() || c >= cx.n()) return e;
    const BB occ = p.occ();
    const BB fillable = (cx.board & ~occ) | enclosed;
    const BB neither = occ & ~enclosed & ~enclosing;
    const BB stop = (abort_on_edge ? cx.edge : BB::zero()) | (abort_on_corner ? cx.corners : BB::zero());
    BB reg = BB::bit(idx_of<BB>(r, c)) & fillable;
    if (!reg.any()) return e;
    for (int it = 0; it < BB::BITS; ++it) {
        if ((reg & stop).any()) return e;
        const BB around = dilate4(cx, reg);
        // a piece of neither set beside the region means None whatever the rest of the fill finds (the region only
        // grows): leave at the first contact instead of flooding the open board first (the exit-fort test of a king
        // on the edge in an ordinary position: an attacker is met within a step or two)
        if ((around & neither).any()) return e;
        BB grown = reg | (around & fillable);
        if (grown == reg) break;
        reg = grown;
    }
    if ((reg & stop).any()) return e;
    BB nb = dilate4(cx, reg) & ~reg;
    if ((nb & neither).any()) return e;
    e.region = reg;
    e.boundary = nb & enclosing;
    e.found = true;
    return e;
}
// Exact pre-filter for the enclosure WIN test of get_game_outcome (logic.rs:720-734).  The win needs
// the king's region (fillable = empty or defender) to hold EVERY defender and to avoid the abort
// tiles (corners, and the edge under WithoutEdgeAccess).  If that holds, the region equals the union
// of the components of all defenders — so when a fill started from all defenders at once reaches an
// abort tile, either the king's region reaches it too (find_enclosure returns None) or some
// defender lies outside it (the occupied count differs): no win either way.  The multi-source fill
// escapes in a step or two in ordinary positions, where the single-source fill from the king needs
// many; only a genuinely sealed position pays for the full test afterwards.
template <class BB, int SP>
TAFL_HD bool enclosure_win_possible(const Ctx<BB, SP>& cx, const Pos<BB>& p, bool abort_on_edge) {
    const BB fillable = (cx.board & ~p.occ()) | p.def;
    const BB stop = (abort_on_edge ? cx.edge : BB::zero()) | cx.corners;
#if !defined(TAFL_ENCL_TWO_SIDED)
    BB reg = p.def;
    bool escaped = false;   // single exit: the lanes of a warp leave the loop together
    for (int it = 0; it < BB::BITS; ++it) {
        escaped = (reg & stop).any();
        if (escaped) break;
        const BB grown = reg | (dilate4(cx, reg) & fillable);
        if (grown == reg) break;
        reg = grown;
    }
    return !escaped;
#else
    // Tried (r02n): reachability is symmetric, so the fill can run from both ends — from the defenders and from the abort
    // tiles — until the fronts meet: half the trips for the warp's slowest lane, two independent dilations per trip.
    // Measured 10.96 us per ply against 10.79 us for the one-sided fill above; kept for reference only.
    BB a = p.def, b = stop & fillable;
    bool escaped = false;   // single exit: the lanes of a warp leave the loop together
    for (int it = 0; it < BB::BITS; ++it) {
        escaped = (a & b).any();
        if (escaped) break;
        const BB ga = a | (dilate4(cx, a) & fillable), gb = b | (dilate4(cx, b) & fillable);
        escaped = (ga & b).any();   // fronts one step apart: no second trip for that
        if (escaped) break;
        if (ga == a || gb == b) break;   // one side is complete without having met the other
        a = ga;
        b = gb;
    }
    return !escaped;
#endif
}
// enclosure_secure (logic.rs:408-463)
template <class BB, int SP>
TAFL_HD bool enclosure_secure(const Ctx<BB, SP>& cx, const Pos<BB>& p, const Encl<BB>& e, bool inside_safe,
                              bool outside_safe) {
    if (inside_safe && outside_safe) return true;
    const BB occ = p.occ();
    const BB empty = cx.board & ~occ;
    const BB safe_zone = (inside_safe ? e.region : BB::zero()) | (outside_safe ? (cx.board & ~e.region) : BB::zero());
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        BB bk = e.boundary & p.kind(k);
        if (!bk.any()) continue;
        const int ek = (k == KIND_AS) ? KIND_DS : KIND_AS;  // hostile soldier (logic.rs:426)
        const BB hostile = hostile_set(cx, p, k);
        const BB prot = (safe_zone & ~cx.sh[k]) | (cx.board & ~hostile & (occ | cx.dest_block[ek]));
        const bool oob_ok = !cx.f_edge_hostile(k);
        // value of `prot` at the four neighbours, aligned to the boundary tile
        BB up = prot.template shl<BB::STRIDE>() | (oob_ok ? cx.row0 : BB::zero());           // neighbour row-1
        BB down = prot.template shr<BB::STRIDE>() | (oob_ok ? cx.rowL : BB::zero());         // neighbour row+1
        BB left = (prot.template shl<1>() & cx.not_col0) | (oob_ok ? cx.col0 : BB::zero());  // neighbour col-1
        BB right = (prot.template shr<1>() & cx.not_colL) | (oob_ok ? cx.colL : BB::zero()); // neighbour col+1
        (void)empty;
        if ((bk & ~((up | down) & (left | right))).any()) ok = false;
    }
    return ok;
}
// The full tests behind the cheap pre-filters are rare (a sealed position, a king on the edge inside
// a defender fort): out of line, so that the per-ply instruction stream stays short and i-cache resident.
template <class BB, int SP>
TAFL_COLD bool enclosure_win_full(const Ctx<BB, SP>& cx, const Pos<BB>& p) {
    const bool woe = cx.r_enclosure_win() == TAFL_ENCL_WITHOUT_EDGE_ACCESS;
    Encl<BB> e = find_enclosure(cx, p, p.kr, p.kc, p.def, p.att, woe, true);
    return e.found && (e.region & p.occ()).popc() == p.def.popc() && enclosure_secure(cx, p, e, false, true);
}
template <class BB, int SP>
TAFL_COLD bool exit_fort_full(const Ctx<BB, SP>& cx, const Pos<BB>& p) {
    Encl<BB> e = find_enclosure(cx, p, p.kr, p.kc, p.king(), p.dsold(), false, true);
    if (!e.found) return false;
    return enclosure_secure(cx, p, e, true, false);
}

// Exact pre-filter for the exit-fort test.  find_enclosure(king; enclosed = King, enclosing = defenders, abort on corner) is
// None when the king's region — the empty tiles connected to the king's tile — holds a corner or touches an attacker
// (a piece of neither set), i.e. when it holds a VOID tile: an empty corner or an empty tile beside an attacker.  On an
// ordinary board attackers are everywhere, so a fill from the void tiles meets the fill from the king within a step
// or two, where the fill from the king alone needs as many steps as the open area is wide.  Both fills run through
// empty tiles only; when either is complete without having met the other they never meet.
template <class BB, int SP>
TAFL_HD bool exit_fort_possible(const Ctx<BB, SP>& cx, const Pos<BB>& p) {
    const BB empty = cx.board & ~p.occ();
    BB k = p.kbit;
    BB v = (dilate4(cx, p.att) | cx.corners) & (empty | p.kbit);
    bool voided = false;   // single exit: the lanes of a warp leave the loop together
    for (int it = 0; it < BB::BITS; ++it) {
        voided = (k & v).any();
        if (voided) break;
        const BB gk = k | (dilate4(cx, k) & empty), gv = v | (dilate4(cx, v) & empty);
        if (gk == k || gv == v) break;
        k = gk;
        v = gv;
    }
    return !voided;
}

// detect_exit_fort (logic.rs:572-601)
template <class BB, int SP>
TAFL_HD bool detect_exit_fort(const Ctx<BB, SP>& cx, const Pos<BB>& p) {
    const int N = cx.n();
    if (!(p.kr == 0 || p.kr == N - 1 || p.kc == 0 || p.kc == N - 1)) return false;
    if (!p.kbit.any()) return false;
    // enclosed = PieceSet::from(King), enclosing = PieceSet::from(Defender); attackers are "neither"
    // find_enclosure -> king has an empty neighbour -> enclosure_secure (logic.rs:580-600); the empty-
    // neighbour test has no side effects, so it is taken first and the fill runs out of line
    const BB beside = dilate4(cx, p.kbit);
    if (!(beside & ~p.occ()).any()) return false;
    if ((beside & p.att).any()) return false;   // an attacker beside the king is beside the region: find_enclosure is None
#if !defined(TAFL_EXITFORT_NO_PREFILTER)
    if (!exit_fort_possible(cx, p)) return false;
#endif
    return exit_fort_full(cx, p);
}

// ---------------------------------------------------------------------------------------------
// RepetitionTracker::track_play (game/state.rs:71-113) on the packed state
// ---------------------------------------------------------------------------------------------
TAFL_HD uint32_t pack_rec(int from_r, int from_c, int horizontal, int disp, int def_side, bool captures) {
    uint32_t fl = TAFL_REC_SOME | (captures ? TAFL_REC_CAPTURES : 0u) | (def_side ? TAFL_REC_DEFENDER : 0u) |
                  (horizontal ? TAFL_REC_HORIZONTAL : 0u);
    return (uint32_t)(from_r & 0xFF) | ((uint32_t)(from_c & 0xFF) << 8) | ((uint32_t)(disp & 0xFF) << 16) | (fl << 24);
}
template <class ST>
TAFL_HD void track_play(ST& s, uint32_t rec, int def_side, bool captures) {
    const uint32_t head = s.ring_head & 3u;
    uint32_t first = s.recent[0];
    first = head == 1 ? s.recent[1] : first;
    first = head == 2 ? s.recent[2] : first;
    first = head == 3 ? s.recent[3] : first;
    const uint8_t mid_bit = def_side ? TAFL_FLAG_DEF_MID : TAFL_FLAG_ATT_MID;
    uint8_t reps = def_side ? s.def_reps : s.att_reps;
    if (!captures && first == rec) {
        bool is_rep = !(s.flags & mid_bit);
        s.flags ^= mid_bit;
        if (is_rep && reps < 255) ++reps;
    } else {
        reps = 0;
        s.flags &= (uint8_t)~mid_bit;
    }
    if (def_side) s.def_reps = reps; else s.att_reps = reps;
    if (head == 0) s.recent[0] = rec;
    if (head == 1) s.recent[1] = rec;
    if (head == 2) s.recent[2] = rec;
    if (head == 3) s.recent[3] = rec;
    s.ring_head = (uint8_t)((head + 1u) & 3u);
}

// ---------------------------------------------------------------------------------------------
// get_game_outcome (logic.rs:702-771) — returns the packed status (0 = still Ongoing)
// ---------------------------------------------------------------------------------------------
template <class BB, int SP, class ST>
TAFL_HD uint8_t game_outcome(const Ctx<BB, SP>& cx, const Pos<BB>& p, const ST& s, int mover_def, int piece_kind,
                             int to_r, int to_c, BB caps) {
    // Single exit on purpose: the side-specific tests diverge inside a warp (mixed sides to move,
    // data-dependent fill lengths) and the common tail — side_can_play — must be issued once.
    TAFL_LANES_HERE(lanes);
    uint8_t st = TAFL_STATUS_ONGOING;
    bool done = false;
    const int N = cx.n();
    if (!p.side(!mover_def).any()) {
        st = TAFL_STATUS_WIN(TAFL_WIN_ALL_CAPTURED, mover_def);
        done = true;
    } else if (!mover_def) {
        if ((caps & p.kbit).any()) {
            st = TAFL_STATUS_WIN(TAFL_WIN_KING_CAPTURED, 0);
            done = true;
        } else if (TAFL_ABL_KEEP(ENCL) && cx.r_enclosure_win() != TAFL_ENCL_NONE &&
                   enclosure_win_possible(cx, p, cx.r_enclosure_win() == TAFL_ENCL_WITHOUT_EDGE_ACCESS)) {
            if (enclosure_win_full(cx, p)) {
                st = TAFL_STATUS_WIN(TAFL_WIN_ENCLOSED, 0);
                done = true;
            }
        }
    } else {
        bool escaped = false;
        if (piece_kind == KIND_K) {
            const bool at_edge = to_r == 0 || to_r == N - 1 || to_c == 0 || to_c == N - 1;
            const bool at_corner = (to_r == 0 || to_r == N - 1) && (to_c == 0 || to_c == N - 1);
            escaped = cx.r_edge_escape() ? at_edge : at_corner;
        }
        if (escaped) {
            st = TAFL_STATUS_WIN(TAFL_WIN_KING_ESCAPED, 1);
            done = true;
        } else if (TAFL_ABL_KEEP(EXITFORT) && cx.r_exit_fort() && detect_exit_fort(cx, p)) {
            st = TAFL_STATUS_WIN(TAFL_WIN_EXIT_FORT, 1);
            done = true;
        }
    }
    if (TAFL_ABL_KEEP(REP) && !done && cx.r_repetition_enabled()) {
        uint32_t reps = mover_def ? s.def_reps : s.att_reps;
        if (reps >= cx.r_n_repetitions()) {
            st = cx.r_repetition_is_loss() ? TAFL_STATUS_WIN(TAFL_WIN_REPETITION, !mover_def)
                                           : TAFL_STATUS_DRAW(TAFL_DRAW_REPETITION);
            done = true;
        }
    }
    TAFL_RECONVERGE(lanes);
    if (TAFL_ABL_KEEP(CANPLAY) && !done && !side_can_play(cx, p, !mover_def))
        st = cx.r_draw_on_no_plays() ? TAFL_STATUS_DRAW(TAFL_DRAW_NO_PLAYS) : TAFL_STATUS_WIN(TAFL_WIN_NO_PLAYS, mover_def);
    return st;
}

// ---------------------------------------------------------------------------------------------
// do_valid_play (logic.rs:782-820).  (from, axis, disp) is the Play; the caller guarantees a
// piece on `from` and an in-bounds destination.  Returns the captured tiles.
// ---------------------------------------------------------------------------------------------
template <class BB, int SP, class ST>
TAFL_HD BB do_valid_play(const Ctx<BB, SP>& cx, ST& s, int from_r, int from_c, int horizontal, int disp) {
    Pos<BB> p = load_pos(cx, s);
    const int mover_def = (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    const int to_r = horizontal ? from_r : from_r + disp, to_c = horizontal ? from_c + disp : from_c;
    const BB fb = BB::bit(idx_of<BB>(from_r, from_c)), tb = BB::bit(idx_of<BB>(to_r, to_c));
    // move_piece (board/state.rs:218-223)
    const int pk = kind_at(p, fb);
    const int piece_def = pk != KIND_AS;
    if (piece_def) { p.def = p.def | tb; p.att = p.att & ~tb; }
    else { p.att = p.att | tb; p.def = p.def & ~tb; }
    if (pk == KIND_K) { p.kr = to_r; p.kc = to_c; p.kbit = tb; }
    p.att = p.att & ~fb;
    p.def = p.def & ~fb;
    const BB caps = get_captures(cx, p, piece_def, pk, mover_def, to_r, to_c);
    p.att = p.att & ~caps;
    p.def = p.def & ~caps;
    const bool captured = caps.any();
    track_play(s, pack_rec(from_r, from_c, horizontal, disp, mover_def, captured), mover_def, captured);
    if (!captured) s.plays_since_capture += 1;
    const uint8_t status = game_outcome(cx, p, s, mover_def, pk, to_r, to_c, caps);
    s.turn += 1;
    s.flags ^= TAFL_FLAG_DEF_TO_PLAY;
    s.status = status;
    store_pos(p, s);
    return caps;
}
template <class BB, int SP>
TAFL_HD BB do_valid_move(const Ctx<BB, SP>& cx, tafl_state& s, tafl_move m) {
    const int f = TAFL_MOVE_FROM(m), t = TAFL_MOVE_TO(m);
    const int fr = f / BB::STRIDE, fc = f - fr * BB::STRIDE, tr = t / BB::STRIDE, tc = t - tr * BB::STRIDE;
    const int horizontal = fr == tr;  // Play::from_tiles (play.rs:36-49)
    return do_valid_play(cx, s, fr, fc, horizontal, horizontal ? tc - fc : tr - fr);
}

// validate_play_for_side (logic.rs:159-214)
template <class BB, int SP, class ST>
TAFL_HD int validate_play(const Ctx<BB, SP>& cx, const ST& s, const tafl_play& pl, int side_def) {
    if (s.status != TAFL_STATUS_ONGOING) return TAFL_GAME_OVER;
    const int N = cx.n();
    const int fr = pl.from_row, fc = pl.from_col;
    const int horizontal = pl.axis == TAFL_HORIZONTAL;
    // Coords arithmetic in i8, destination cast to u8 (play.rs:59-67)
    const int tr = horizontal ? fr : (int)(uint8_t)(int8_t)((int8_t)fr + pl.disp);
    const int tc = horizontal ? (int)(uint8_t)(int8_t)((int8_t)fc + pl.disp) : fc;
    const int fi = fr * BB::STRIDE + fc;
    int pk;
    {
        Pos<BB> q = load_pos(cx, s);
        if (Nib<BB>::raw_def(s, fi)) pk = (fr == q.kr && fc == q.kc) ? KIND_K : KIND_DS;
        else if (Nib<BB>::raw_att(s, fi)) pk = KIND_AS;
        else return TAFL_NO_PIECE;
    }
    if ((pk != KIND_AS) != (side_def != 0)) return TAFL_WRONG_PLAYER;
    if (!(fr < N && fc < N && tr < N && tc < N)) return TAFL_OUT_OF_BOUNDS;
    if (fr != tr && fc != tc) return TAFL_NO_COMMON_AXIS;
    Pos<BB> p = load_pos(cx, s);
    const BB occ = p.occ();
    const BB tb = BB::bit(idx_of<BB>(tr, tc));
    if ((occ & tb).any()) return TAFL_BLOCKED_BY_PIECE;
    BB between = BB::zero();  // geometry.rs:85-108
    if (fr == tr) { int lo = fc < tc ? fc : tc, hi = fc < tc ? tc : fc; for (int c = lo + 1; c < hi; ++c) between = between | BB::bit(idx_of<BB>(fr, c)); }
    else { int lo = fr < tr ? fr : tr, hi = fr < tr ? tr : fr; for (int r = lo + 1; r < hi; ++r) between = between | BB::bit(idx_of<BB>(r, fc)); }
    if ((between & occ).any()) return TAFL_BLOCKED_BY_PIECE;
    if (!cx.f_may_corner(pk) && (cx.corners & tb).any()) return TAFL_MOVE_ONTO_BLOCKED_TILE;
    if (cx.f_pass_block(pk) && (between & cx.throne).any()) return TAFL_MOVE_THROUGH_BLOCKED_TILE;
    if (cx.f_noentry(pk) && tb == cx.throne) return TAFL_MOVE_ONTO_BLOCKED_TILE;
    int dist = pl.disp < 0 ? -(int)pl.disp : (int)pl.disp;
    if (cx.f_slow(pk) && dist > 1) return TAFL_TOO_FAR;
    return TAFL_PLAY_OK;
}

// Counter-based RNG of the random-playout workload, identical to the oracle's.
TAFL_HD uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
TAFL_HD uint32_t pick_index(uint64_t seed, uint64_t game_id, uint64_t ply, uint32_t n) {
    uint64_t x = splitmix64(seed ^ (game_id * 0x9E3779B97F4A7C15ull) ^ (ply * 0xBF58476D1CE4E5B9ull));
    return (uint32_t)mulhi64(x, (uint64_t)n);
}

}  // namespace tafl
