// tafl_select.cuh — branch-free legal-move COUNT and k-th-move SELECT for one game per thread,
// plus the fused random-playout step and the by-value query helpers.
//
// The random-play step never materialises the move list.  Instead (all 128-bit bit-parallel):
//   1. for each of the 4 ray directions and each distance d, E_d = set of own soldiers that can
//      legally move exactly d tiles that way (source-aligned ray fill; play.rs:189-225 semantics);
//   2. the <= 40 E_d boards are summed per source tile with a carry-save-adder tree into 5
//      bit-sliced planes (plane b = bit b of "number of legal plays of the piece on this tile");
//      the king's count (different throne/corner rights) is computed in line space and OR-ed in;
//   3. total L = sum_b 2^b popc(plane_b); the k-th play in canonical order (pieces ascending
//      bit index, game/main.rs:33-43) is found by a word-level then 5-level binary search on
//      masked plane popcounts, and resolved to (direction, distance) with piece_rays().
#pragma once
#include "tafl_core.cuh"

namespace tafl {

template <class BB>
struct MoveSel {
    BB plane[5];   // bit b of the per-tile count (before the correction below)
    BB minus;      // tiles whose count is one less than the planes say: a soldier whose run crosses the empty
                   // throne cannot stop there (play.rs:208-211); kept apart so that no bit-sliced decrement is needed
    int total;
    int def_side;
};

template <class BB>
struct CsaAcc {
    BB ones, twos, fours, eights, sixteens, p1, p2, p4;
    TAFL_HD void init() { ones = twos = fours = eights = sixteens = p1 = p2 = p4 = BB::zero(); }
    TAFL_HD static void csa(BB& h, BB& l, BB a, BB b, BB c) {
        BB u = a ^ b;
        h = (a & b) | (u & c);
        l = u ^ c;
    }
    // i = 0,1,2,... must be a compile-time constant after unrolling
    TAFL_HD void add(BB x, int i) {
        if ((i & 1) == 0) { p1 = x; return; }
        BB t, f, e;
        csa(t, ones, ones, p1, x);
        if ((i & 2) == 0) { p2 = t; return; }
        csa(f, twos, twos, p2, t);
        if ((i & 4) == 0) { p4 = f; return; }
        csa(e, fours, fours, p4, f);
        BB carry = eights & e;
        eights = eights ^ e;
        sixteens = sixteens ^ carry;
    }
};

template <class BB> struct MaxD;
template <> struct MaxD<BB64> { static constexpr int v = 6; };
template <> struct MaxD<BB128> { static constexpr int v = 10; };

// longest ray the count loop has to walk: N-1 tiles when the board size is a compile-time constant
template <class BB, int SP> struct RayLen { static constexpr int v = SpecTraits<SP>::N - 1; };
template <class BB> struct RayLen<BB, SPEC_GENERIC> { static constexpr int v = MaxD<BB>::v; };

// [s + d*step == throne] as a source-aligned tile set; folds to an immediate when N, dir and d are
// compile-time constants (preset specialisations)
template <class BB>
TAFL_HD BB throne_aligned(int N, int dir, int d) {
    int r = N / 2, c = N / 2;
    if (dir == DIR_VP) r -= d;
    else if (dir == DIR_VM) r += d;
    else if (dir == DIR_HP) c -= d;
    else c += d;
    return (r >= 0 && r < N && c >= 0 && c < N) ? BB::bit(r * BB::STRIDE + c) : BB::zero();
}

// board columns lo..hi of every row as a tile set; folds to an immediate when its arguments are constants
template <class BB>
TAFL_HD BB cols_between(int N, int lo, int hi) {
    BB r = BB::zero();
    for (int row = 0; row < N; ++row)
        for (int c = lo; c <= hi; ++c) r = r | BB::bit(row * BB::STRIDE + c);
    return r;
}
// bit-sliced x + y of two 4-plane numbers -> 5 planes
template <class BB>
TAFL_HD void add4(const BB (&x)[4], const BB (&y)[4], BB (&out)[5]) {
    BB c = x[0] & y[0];
    out[0] = x[0] ^ y[0];
#pragma unroll
    for (int i = 1; i < 4; ++i) {
        const BB u = x[i] ^ y[i];
        out[i] = u ^ c;
        c = (x[i] & y[i]) | (c & u);
    }
    out[4] = c;
}

// count legal plays of `def_side`'s pieces on position p (status must be Ongoing)
template <class BB, int SP>
TAFL_HD MoveSel<BB> count_moves(const Ctx<BB, SP>& cx, const Pos<BB>& p, int def_side) {
    constexpr int S = BB::STRIDE;
    constexpr int MD = RayLen<BB, SP>::v;
    static_assert((4 * MD) % 8 == 0, "CsaAcc folds its inputs in groups of 8");
    MoveSel<BB> ms;
    ms.def_side = def_side;
    const int k = def_side ? KIND_DS : KIND_AS;
    const BB occ = p.occ();
    // a blocked corner always ends a ray, so it can be treated as not walkable (logic.rs:190-193)
    const BB walk = cx.board & ~occ & ~(cx.f_may_corner(k) ? BB::zero() : cx.corners);
    const BB src = p.kind(k);
    const BB thr = cx.throne;
    const bool noentry = cx.f_noentry(k) != 0, passblk = cx.f_pass_block(k) != 0, slow = cx.f_slow(k) != 0;
    ms.minus = BB::zero();
#if !defined(TAFL_COUNT_NESTED)
    if constexpr (SP != SPEC_GENERIC) {
        // Preset rules have no pass-block and no slow pieces, so a piece's number of plays along one ray is the
        // LENGTH n of its run of walkable tiles (the empty throne counts as walkable here; see `minus`).  The binary
        // digits of n come from run-length DOUBLING: R1, R2, R4, R8 = "the next 1, 2, 4, 8 tiles are walkable"
        // (R2 = R1 & R1>>1 step, R4 = R2 & R2>>2 steps, ...), then b3 = R8, b2 = R4 & ~R8, and the lower digits are
        // read off the same masks shifted past the part already accounted for (b1 from R2 moved 0 / 4 / 8 steps,
        // b0 from R1 moved 0 / 2 / ... / 8 steps).  10 shifts and 11 logic ops per word and direction instead of the
        // 10 shift+AND steps, the throne fix-up and the digit formulas of the distance-by-distance chain.
        // Sideways no column guards are needed: the walk mask of a rightward ray drops column 0 (a run that wraps
        // into the next row would have to include its column-0 tile), of a leftward ray the last column.
        static_assert(MD <= 10, "digit formulas cover runs of up to 10 tiles");
        BB dig[4][4];
#pragma unroll
        for (int dir = 0; dir < 4; ++dir) {
            const BB Wd = dir == DIR_HP ? (walk & cx.not_col0) : dir == DIR_HM ? (walk & cx.not_colL) : walk;
            // value of X at the tile k steps ahead, aligned to the source tile
            auto T1 = [&](BB X) { return dir == DIR_VP ? X.template shr<S>() : dir == DIR_VM ? X.template shl<S>() : dir == DIR_HP ? X.template shr<1>() : X.template shl<1>(); };
            auto T2 = [&](BB X) { return dir == DIR_VP ? X.template shr<2 * S>() : dir == DIR_VM ? X.template shl<2 * S>() : dir == DIR_HP ? X.template shr<2>() : X.template shl<2>(); };
            auto T4 = [&](BB X) { return dir == DIR_VP ? X.template shr<4 * S>() : dir == DIR_VM ? X.template shl<4 * S>() : dir == DIR_HP ? X.template shr<4>() : X.template shl<4>(); };
            auto T6 = [&](BB X) { return dir == DIR_VP ? X.template shr<6 * S>() : dir == DIR_VM ? X.template shl<6 * S>() : dir == DIR_HP ? X.template shr<6>() : X.template shl<6>(); };
            auto T8 = [&](BB X) { return dir == DIR_VP ? X.template shr<8 * S>() : dir == DIR_VM ? X.template shl<8 * S>() : dir == DIR_HP ? X.template shr<8>() : X.template shl<8>(); };
            const BB R1 = T1(Wd);
            const BB R2 = R1 & T1(R1);
            const BB R4 = R2 & T2(R2);
            const BB R8 = MD >= 8 ? (R4 & T4(R4)) : BB::zero();
            const BB b3 = R8;
            const BB b2 = R4 & ~R8;
            // b1: two more walkable tiles behind the 0 / 4 / 8 already counted
            const BB hi2 = MD >= 10 ? (R8 & T8(R2)) : BB::zero();                 // n >= 10
            const BB mid2 = MD >= 6 ? (b2 & T4(R2)) : BB::zero();                 // n in {6, 7}
            const BB b1 = (R2 & ~R4) | mid2 | hi2;
            // b0: one more walkable tile behind the even part 8 b3 + 4 b2 + 2 b1
            const BB x0 = (R1 & ~b1) | (T2(R1) & b1);                             // even part 0 / 2
            const BB x1 = MD >= 5 ? ((T4(R1) & ~b1) | ((MD >= 7 ? T6(R1) : BB::zero()) & b1)) : BB::zero();   // 4 / 6
            const BB x2 = MD >= 9 ? (T8(R1) & ~b1) : BB::zero();                  // 8 (10 is the longest run: no tile behind it)
            const BB b0 = (~R4 & x0) | (b2 & x1) | (b3 & x2);
            dig[dir][0] = b0; dig[dir][1] = b1; dig[dir][2] = b2; dig[dir][3] = b3;
            if (noentry) {
                // sources on the throne's line whose run reaches the throne: d tiles before it with n >= d
                const BB ge3 = R2 & T2(R1), ge5 = R4 & T4(R1);
                BB reach = (R1 & throne_aligned<BB>(cx.n(), dir, 1)) | (R2 & throne_aligned<BB>(cx.n(), dir, 2)) |
                           (ge3 & throne_aligned<BB>(cx.n(), dir, 3));
                if (cx.m() >= 4) reach = reach | (R4 & throne_aligned<BB>(cx.n(), dir, 4));
                if (cx.m() >= 5) reach = reach | (ge5 & throne_aligned<BB>(cx.n(), dir, 5));
                static_assert(SpecTraits<SP>::N / 2 <= 5, "reach covers thrones up to 5 tiles from the edge");
                ms.minus = ms.minus | reach;
            }
        }
        // (VP + VM) + (HP + HM), ripple-carry over bit planes; a piece has at most 2(N-1) <= 20 plays
        BB a[5], b[5];
        add4(dig[0], dig[1], a);
        add4(dig[2], dig[3], b);
        BB c = a[0] & b[0];
        ms.plane[0] = (a[0] ^ b[0]) & src;
#pragma unroll
        for (int i = 1; i < 5; ++i) {
            ms.plane[i] = (a[i] ^ b[i] ^ c) & src;
            c = (a[i] & b[i]) | (c & (a[i] ^ b[i]));
        }
        ms.minus = ms.minus & src;
    } else {
#else
    if constexpr (SP != SPEC_GENERIC) {
        // Preset rules have no pass-block and no slow pieces, so along one ray the sets E_1 >= E_2 >= ... are
        // nested and a piece's number of plays that way is the LENGTH of its run: its binary digits come
        // straight from the nested sets (no adder per distance), and the four directions are then added
        // bit-sliced.  A non-king piece passes an empty throne without stopping on it (noentry): the run that
        // crosses the throne is one shorter, which keeps the sets nested if its tail is moved up by one.
        static_assert(MD <= 10, "digit formulas cover runs of up to 10 tiles");
        BB dig[4][4];
#pragma unroll
        for (int dir = 0; dir < 4; ++dir) {
            BB X = walk, E = src;
            BB m[MD + 2];
#pragma unroll
            for (int d = 1; d <= MD; ++d) {
                // X(s) = walkable(s + d*step); sideways the source column must leave room for d steps
                if (dir == DIR_VP) { X = X.template shr<S>(); E = E & X; }
                else if (dir == DIR_VM) { X = X.template shl<S>(); E = E & X; }
                else if (dir == DIR_HP) { X = X.template shr<1>(); E = E & X & cols_between<BB>(cx.n(), 0, cx.n() - 1 - d); }
                else { X = X.template shl<1>(); E = E & X & cols_between<BB>(cx.n(), d, cx.n() - 1); }
                m[d] = E;
            }
            m[MD + 1] = BB::zero();
            if (noentry) {
                BB U = BB::zero();   // sources whose ray has reached the throne by distance d
#pragma unroll
                for (int d = 1; d <= MD; ++d) {
                    U = U | throne_aligned<BB>(cx.n(), dir, d);
                    m[d] = (m[d] & ~U) | (m[d + 1] & U);
                }
            }
            auto M = [&](int d) { return d <= MD ? m[d] : BB::zero(); };
            dig[dir][0] = (M(1) & ~M(2)) | (M(3) & ~M(4)) | (M(5) & ~M(6)) | (M(7) & ~M(8)) | (M(9) & ~M(10));
            dig[dir][1] = (M(2) & ~M(4)) | (M(6) & ~M(8)) | M(10);
            dig[dir][2] = M(4) & ~M(8);
            dig[dir][3] = M(8);
        }
        // (VP + VM) + (HP + HM), ripple-carry over bit planes; a piece has at most 2(N-1) <= 20 plays
        BB a[5], b[5];
        add4(dig[0], dig[1], a);
        add4(dig[2], dig[3], b);
        BB c = a[0] & b[0];
        ms.plane[0] = a[0] ^ b[0];
#pragma unroll
        for (int i = 1; i < 5; ++i) {
            ms.plane[i] = a[i] ^ b[i] ^ c;
            c = (a[i] & b[i]) | (c & (a[i] ^ b[i]));
        }
    } else {
#endif
        CsaAcc<BB> acc;
        acc.init();
        int n = 0;
#pragma unroll
        for (int dir = 0; dir < 4; ++dir) {
            BB X = walk, E = src, T = thr, Tprev = BB::zero();
#pragma unroll
            for (int d = 1; d <= MD; ++d) {
                // X(s) = walkable(s + d*step); T(s) = [s + d*step == throne]
                if (dir == DIR_VP) X = X.template shr<S>();
                else if (dir == DIR_VM) X = X.template shl<S>();
                else if (dir == DIR_HP) X = X.template shr<1>() & cx.not_colL;
                else X = X.template shl<1>() & cx.not_col0;
                if constexpr (SP == SPEC_GENERIC) {
                    if (dir == DIR_VP) T = T.template shr<S>();
                    else if (dir == DIR_VM) T = T.template shl<S>();
                    else if (dir == DIR_HP) T = T.template shr<1>();
                    else T = T.template shl<1>();
                } else {
                    T = throne_aligned<BB>(cx.n(), dir, d);
                }
                E = E & X;
                if (passblk) E = E & ~Tprev;                  // logic.rs:194-200
                BB emit = noentry ? (E & ~T) : E;             // logic.rs:201-206 + play.rs:208-211
                if (slow && d > 1) emit = BB::zero();         // logic.rs:207-210
                acc.add(emit, n);
                ++n;
                Tprev = T;
            }
        }
        // flush pending partial sums when the number of inputs is not a multiple of 8 (BB64: 24 = 3*8 ok)
        ms.plane[0] = acc.ones; ms.plane[1] = acc.twos; ms.plane[2] = acc.fours; ms.plane[3] = acc.eights;
        ms.plane[4] = acc.sixteens;
    }
    if (TAFL_ABL_KEEP(KING) && def_side && (p.def & p.kbit).any()) {
        Rays r = piece_rays(cx, occ, KIND_K, p.kr, p.kc);
        int ck = r.cnt[0] + r.cnt[1] + r.cnt[2] + r.cnt[3];
#pragma unroll
        for (int b = 0; b < 5; ++b)
            if ((ck >> b) & 1) ms.plane[b] = ms.plane[b] | p.kbit;
    }
    int total = -ms.minus.popc();
#pragma unroll
    for (int b = 0; b < 5; ++b) total += ms.plane[b].popc() << b;
    ms.total = total;
    return ms;
}

// k-th legal play (0 <= k < ms.total) in canonical order
template <class BB, int SP>
TAFL_HD tafl_move select_move(const Ctx<BB, SP>& cx, const Pos<BB>& p, const MoveSel<BB>& ms, uint32_t k) {
    // word-level prefix
    uint32_t pw[5] = {0, 0, 0, 0, 0}, pm = 0;
    int wsel = 0;
    bool done = false;
#pragma unroll
    for (int j = 0; j < BB::WORDS; ++j) {
        uint32_t w = 0u - (uint32_t)popc32(ms.minus.word(j));
#pragma unroll
        for (int b = 0; b < 5; ++b) w += (uint32_t)popc32(ms.plane[b].word(j)) << b;
        const bool here = !done && k < w;
        if (here) {
            wsel = j;
#pragma unroll
            for (int b = 0; b < 5; ++b) pw[b] = ms.plane[b].word(j);
            pm = ms.minus.word(j);
            done = true;
        }
        if (!done) k -= w;
    }
    // binary search for the source bit inside the word
    int lo = 0;
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const uint32_t m = ((1u << half) - 1u) << lo;
        uint32_t w = 0u - (uint32_t)popc32(pm & m);
#pragma unroll
        for (int b = 0; b < 5; ++b) w += (uint32_t)popc32(pw[b] & m) << b;
        if (k >= w) { k -= w; lo += half; }
    }
    const int s = wsel * 32 + lo;
    const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
    const int kind = ms.def_side ? ((BB::bit(s) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
    Rays r = piece_rays(cx, p.occ(), kind, row, col);
    int dir = 0;
    uint32_t kk = k;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        if (dir == d && kk >= (uint32_t)r.cnt[d]) { kk -= r.cnt[d]; dir = d + 1; }
    }
    // the destinations of a ray are consecutive tiles from the piece outward, except for the one hole
    // at the throne a non-king slides across (play.rs:208-211): position by arithmetic, as pt_decode
    const bool vertical = dir < 2;
    const int pp = vertical ? row : col, line = vertical ? col : row, mid = cx.m();
    const bool hole = cx.f_noentry(kind) && line == mid;
    int pos;
    if ((dir & 1) == 0) {
        pos = pp + 1 + (int)kk;
        if (hole && pp < mid && pos >= mid) ++pos;
    } else {
        pos = pp - 1 - (int)kk;
        if (hole && pp > mid && pos <= mid) --pos;
    }
    const int to = vertical ? idx_of<BB>(pos, col) : idx_of<BB>(row, pos);
    return (tafl_move)(s | (to << 7));
}

// ---------------------------------------------------------------------------------------------
// fused random-playout step (BASELINE.json configs 1-2)
// ---------------------------------------------------------------------------------------------
struct StepEvt {
    tafl_move move;   // the play that was applied (valid iff stepped)
    uint32_t stepped, legal, caps;
    uint8_t status;   // != 0: the game ended on this step
    uint32_t turn;
};
// one random ply of a game that is Ongoing (the caller has already reset a finished game)
template <class BB, int SP>
TAFL_HD StepEvt random_step_ongoing(const Ctx<BB, SP>& cx, tafl_state& s, uint64_t seed, uint64_t gid, uint64_t ply) {
    StepEvt ev;
    ev.stepped = ev.legal = ev.caps = ev.turn = 0;
    ev.status = 0;
    ev.move = 0;
    const Pos<BB> p = load_pos(cx, s);
    const MoveSel<BB> ms = count_moves(cx, p, (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
    if (ms.total == 0) return ev;
    const uint32_t k = pick_index(seed, gid, ply, (uint32_t)ms.total);
    const tafl_move m = select_move(cx, p, ms, k);
    const BB caps = do_valid_move(cx, s, m);
    ev.stepped = 1;
    ev.move = m;
    ev.legal = (uint32_t)ms.total;
    ev.caps = (uint32_t)caps.popc();
    ev.status = s.status;
    ev.turn = s.turn;
    return ev;
}
template <class BB, int SP>
TAFL_HD StepEvt random_step(const Ctx<BB, SP>& cx, tafl_state& s, const tafl_state& start, uint64_t seed, uint64_t gid,
                            uint64_t ply, bool auto_reset) {
    if (s.status != TAFL_STATUS_ONGOING) {
        if (!auto_reset) {
            StepEvt ev;
            ev.stepped = ev.legal = ev.caps = ev.turn = 0;
            ev.status = 0;
            ev.move = 0;
            return ev;
        }
        s = start;
    }
    return random_step_ongoing(cx, s, seed, gid, ply);
}

struct StepStats {
    uint64_t steps, legal, caps, finished, plies, win[2][8], draw[2];
};
TAFL_HD void add_evt(StepStats& a, const StepEvt& e) {
    a.steps += e.stepped;
    a.legal += e.legal;
    a.caps += e.caps;
    if (e.status) {
        a.finished += 1;
        a.plies += e.turn;
        if (TAFL_STATUS_IS_DRAW(e.status)) a.draw[TAFL_STATUS_REASON(e.status) & 1] += 1;
        else a.win[e.status & 1][TAFL_STATUS_REASON(e.status)] += 1;
    }
}
template <class BB, int SP>
TAFL_HD void random_step(const Ctx<BB, SP>& cx, tafl_state& s, const tafl_state& start, uint64_t seed, uint64_t gid,
                         uint64_t ply, bool auto_reset, StepStats& acc) {
    add_evt(acc, random_step(cx, s, start, seed, gid, ply, auto_reset));
}
inline void fold_stats(const StepStats& a, tafl_stats* o) {
    o->steps += a.steps; o->legal_moves += a.legal; o->captures += a.caps;
    o->games_finished += a.finished; o->finished_plies += a.plies;
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 8; ++j) o->win_reason[i][j] += a.win[i][j];
    for (int j = 0; j < 2; ++j) o->draw_reason[j] += a.draw[j];
}

// ---------------------------------------------------------------------------------------------
// by-value query helpers (one item each; used by the functional batch kernels)
// ---------------------------------------------------------------------------------------------
template <class BB> TAFL_HD void put_bb(uint64_t o[2], BB b) { o[0] = b.lo64(); o[1] = b.hi64(); }

// GameLogic::iter_plays(tile) (logic.rs:850-856): -1 == Err(NoPiece)
template <class BB, int SP>
TAFL_HD int piece_plays(const Ctx<BB, SP>& cx, const tafl_state& st, int row, int col, tafl_move* out) {
    Pos<BB> p = load_pos(cx, st);
    const int s = idx_of<BB>(row, col);
    int kind;
    if (Nib<BB>::raw_def(st, s)) kind = (row == p.kr && col == p.kc) ? KIND_K : KIND_DS;
    else if (Nib<BB>::raw_att(st, s)) kind = KIND_AS;
    else return -1;
    if (st.status != TAFL_STATUS_ONGOING) return 0;
    if (row >= cx.n() || col >= cx.n()) return 0;
    Rays r = piece_rays(cx, p.occ(), kind, row, col);
    int n = 0;
    for (int d = 0; d < 4; ++d)
        for (int j = 0; j < r.cnt[d]; ++j) {
            int pos = ray_nth(r.m[d], d, j);
            int to = d < 2 ? idx_of<BB>(pos, col) : idx_of<BB>(row, pos);
            if (n < 32) out[n] = (tafl_move)(s | (to << 7));
            ++n;
        }
    return n;
}
// GameLogic::get_captures (logic.rs:604-699) after moving the piece as do_valid_play does
template <class BB, int SP>
TAFL_HD tafl_play_effects query_captures(const Ctx<BB, SP>& cx, const tafl_state& st, const tafl_play& pl) {
    tafl_play_effects fx;
    fx.captures[0] = fx.captures[1] = 0;
    fx.status = 0; fx.err = 0; fx.n_captures = 0;   // status is not part of a capture query
    for (int i = 0; i < 5; ++i) fx._pad[i] = 0;
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
    BB caps = get_captures(cx, p, piece_def, pk, (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, tr, tc);
    put_bb(fx.captures, caps);
    fx.n_captures = (uint8_t)caps.popc();
    return fx;
}
// GameLogic::detect_shieldwall (logic.rs:535-569) on the board as given
template <class BB, int SP>
TAFL_HD tafl_play_effects query_shieldwall(const Ctx<BB, SP>& cx, const tafl_state& st, const tafl_play& pl) {
    tafl_play_effects fx;
    fx.captures[0] = fx.captures[1] = 0;
    fx.status = 0; fx.err = 0; fx.n_captures = 0xFF;
    for (int i = 0; i < 5; ++i) fx._pad[i] = 0;
    Pos<BB> p = load_pos(cx, st);
    const int fr = pl.from_row, fc = pl.from_col, hz = pl.axis == TAFL_HORIZONTAL;
    const int tr = hz ? fr : fr + pl.disp, tc = hz ? fc + pl.disp : fc;
    BB caps;
    if (detect_shieldwall(cx, p, (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0, tr, tc, &caps)) {
        put_bb(fx.captures, caps);
        fx.n_captures = (uint8_t)caps.popc();
    }
    return fx;
}
// GameLogic::find_enclosure (logic.rs:309-401) [+ enclosure_secure logic.rs:408-463 -> *secure]
template <class BB, int SP>
TAFL_HD tafl_enclosure query_enclosure(const Ctx<BB, SP>& cx, const tafl_state& st, const tafl_enclosure_query& q,
                                       uint8_t* secure) {
    tafl_enclosure o;
    o.occupied[0] = o.occupied[1] = o.unoccupied[0] = o.unoccupied[1] = o.boundary[0] = o.boundary[1] = 0;
    o.found = 0;
    for (int i = 0; i < 7; ++i) o._pad[i] = 0;
    Pos<BB> p = load_pos(cx, st);
    BB enclosed = pieceset_bb(p, q.enclosed_pieces);
    BB enclosing = pieceset_bb(p, q.enclosing_pieces) & ~enclosed;
    Encl<BB> e = find_enclosure(cx, p, q.row, q.col, enclosed, enclosing, q.abort_on_edge != 0, q.abort_on_corner != 0);
    *secure = 0xFF;
    if (!e.found) return o;
    o.found = 1;
    put_bb(o.occupied, e.region & p.occ());
    put_bb(o.unoccupied, e.region & ~p.occ());
    put_bb(o.boundary, e.boundary);
    *secure = enclosure_secure(cx, p, e, q.inside_safe != 0, q.outside_safe != 0) ? 1 : 0;
    return o;
}
// GameLogic::side_can_play (logic.rs:837-846)
template <class BB, int SP>
TAFL_HD uint8_t query_can_play(const Ctx<BB, SP>& cx, const tafl_state& st, int def_side) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;  // every play validates as GameOver (logic.rs:165-167)
    return side_can_play(cx, load_pos(cx, st), def_side) ? 1 : 0;
}


// ---------------------------------------------------------------------------------------------
// Piece table: the legal-move list of a position in compressed form (used by the MCTS node pools).
// One 32-bit entry per piece of the side to move, pieces in ascending bit order:
//   bits 0..6  tile index, bit 7 = the piece is the king, bits 8..23 = number of plays in each of the
//   four directions (4 bits each, DIR_* order), bits 24..29 = their sum.
// The plays of one direction are the tiles p+1, p+2, ... along the ray (ValidPlayIterator,
// play.rs:189-225); the only tile a ray can pass without being allowed to stop on it is the empty
// throne (play.rs:208-211), so the j-th play is recovered from the count alone: step j+1 tiles,
// one more if the piece may not enter the throne and the walk reaches or crosses it.
// The canonical move index a (game/main.rs:33-43) lives in the piece whose running total covers it.
// ---------------------------------------------------------------------------------------------
TAFL_HD uint32_t pt_pack(int tile, bool is_king, const Rays& r) {
    const uint32_t total = (uint32_t)(r.cnt[0] + r.cnt[1] + r.cnt[2] + r.cnt[3]);
    return (uint32_t)tile | (is_king ? 0x80u : 0u) | ((uint32_t)r.cnt[0] << 8) | ((uint32_t)r.cnt[1] << 12) |
           ((uint32_t)r.cnt[2] << 16) | ((uint32_t)r.cnt[3] << 20) | (total << 24);
}
TAFL_HD int pt_total(uint32_t e) { return (int)((e >> 24) & 63u); }
// j-th play (0 <= j < pt_total(e)) of the piece described by entry e
template <class BB, int SP>
TAFL_HD tafl_move pt_decode(const Ctx<BB, SP>& cx, uint32_t e, int def_side, int j) {
    const int tile = (int)(e & 127u);
    const int row = tile / BB::STRIDE, col = tile - row * BB::STRIDE;
    const int k = (e & 0x80u) ? KIND_K : (def_side ? KIND_DS : KIND_AS);
    const int c0 = (int)((e >> 8) & 15u), c1 = (int)((e >> 12) & 15u), c2 = (int)((e >> 16) & 15u);
    int dir = 0;
    if (j >= c0) { j -= c0; dir = 1; if (j >= c1) { j -= c1; dir = 2; if (j >= c2) { j -= c2; dir = 3; } } }
    const int mid = cx.m();
    const bool vertical = dir < 2;
    const int p = vertical ? row : col, line = vertical ? col : row;
    const bool hole = cx.f_noentry(k) && line == mid;   // the ray may cross the throne but not stop on it
    int q;
    if ((dir & 1) == 0) {
        q = p + 1 + j;
        if (hole && p < mid && q >= mid) ++q;
    } else {
        q = p - 1 - j;
        if (hole && p > mid && q <= mid) --q;
    }
    const int to = vertical ? idx_of<BB>(q, col) : idx_of<BB>(row, q);
    return (tafl_move)(tile | (to << 7));
}
// The plays of ONE piece in ValidPlayIterator order (V+, V-, H+, H-, distance ascending; play.rs:189-225),
// written to out[off ...] by tile arithmetic alone: the destinations of a ray are consecutive tiles from the
// piece outward, and the one tile a ray can cross without stopping is the empty throne (play.rs:208-211,
// noentry) — one extra step when the walk reaches it.  Returns the offset behind the last play; writes
// behind `cap` are dropped.  Used per lane by k_legal_moves_g (out = the game's shared-memory row).
template <class BB, int SP>
TAFL_HD int emit_piece_plays(const Ctx<BB, SP>& cx, int tile, int row, int col, int kind, const Rays& r, tafl_move* out, int off,
                             int cap) {
    constexpr int S = BB::STRIDE;
    constexpr int MD = RayLen<BB, SP>::v;   // a ray holds at most N-1 plays
    const int mid = cx.m();
    const bool ne = cx.f_noentry(kind);
    const int total = r.cnt[0] + r.cnt[1] + r.cnt[2] + r.cnt[3];
    if (off + total > cap) return off + total;   // cannot happen on boards up to 11x11 (L <= 4 (N^2 - 1) = 480 < 512)
#pragma unroll
    for (int dir = 0; dir < 4; ++dir) {
        const int cnt = r.cnt[dir];
        const int step = (dir == DIR_VP ? S : dir == DIR_VM ? -S : dir == DIR_HP ? 1 : -1) * 128;   // on the `to` field
        const int pp = dir < 2 ? row : col, line = dir < 2 ? col : row;
        int hole_at = 99;   // play index whose destination lies just behind the throne
        if (ne && line == mid) {
            if ((dir & 1) == 0) { if (pp < mid) hole_at = mid - pp - 1; }
            else if (pp > mid) hole_at = pp - mid - 1;
        }
        int m = tile * 129;   // from | to << 7 with to == from
        // fully unrolled, stores predicated on j < cnt: no data-dependent trip count, so the lanes of a warp
        // (one piece each, rays of very different lengths) stay converged
#pragma unroll
        for (int j = 0; j < MD; ++j) {
            m += step;
            if (j == hole_at) m += step;
            if (j < cnt) out[off + j] = (tafl_move)m;
        }
        off += cnt;
    }
    return off;
}
// host restatement of k_legal_moves_g's per-piece path (tests/hostsim): pieces ascending, emit_piece_plays each
template <class BB, int SP>
TAFL_HD int emit_gen_moves(const Ctx<BB, SP>& cx, const tafl_state& st, tafl_move* out, int cap) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;
    const Pos<BB> p = load_pos(cx, st);
    const int def_side = (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    BB own = p.side(def_side);
    const BB occ = p.occ();
    int off = 0;
    while (own.any()) {
        const int s = own.ctz();
        own = own & ~BB::bit(s);
        const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
        const int kind = def_side ? ((BB::bit(s) == p.kbit) ? KIND_K : KIND_DS) : KIND_AS;
        off = emit_piece_plays(cx, s, row, col, kind, piece_rays(cx, occ, kind, row, col), out, off, cap);
    }
    return off;
}

// whole move list through the piece table (host check of the encoding; the kernels build the table
// one piece per lane, tafl_mcts.cuh)
template <class BB, int SP>
TAFL_HD int pt_gen_moves(const Ctx<BB, SP>& cx, const tafl_state& st, tafl_move* out, int cap) {
    if (st.status != TAFL_STATUS_ONGOING) return 0;
    const Pos<BB> p = load_pos(cx, st);
    const int def_side = (st.flags & TAFL_FLAG_DEF_TO_PLAY) != 0;
    BB own = p.side(def_side);
    const BB occ = p.occ();
    int n = 0;
    while (own.any()) {
        const int s = own.ctz();
        own = own & ~BB::bit(s);
        const int row = s / BB::STRIDE, col = s - row * BB::STRIDE;
        const bool is_king = def_side && BB::bit(s) == p.kbit;
        const uint32_t e = pt_pack(s, is_king, piece_rays(cx, occ, is_king ? KIND_K : (def_side ? KIND_DS : KIND_AS), row, col));
        for (int j = 0; j < pt_total(e); ++j) {
            if (n < cap) out[n] = pt_decode(cx, e, def_side, j);
            ++n;
        }
    }
    return n;
}

}  // namespace tafl
