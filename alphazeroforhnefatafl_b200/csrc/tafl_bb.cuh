// tafl_bb.cuh — 64/128-bit tile bitboards for the tafl kernels.
//
// A board of BitfieldBoardState<u64> (row stride 7) lives in one BB64, a board of
// BitfieldBoardState<u128> (row stride 11) in one BB128 = two 64-bit halves (the SM has no
// 128-bit ALU; every op below lowers to 32-bit LOP3/SHF pairs).  Bit index = row*STRIDE + col,
// the reference layout (bitfield.rs:72-81), so converting to/from the HBM state is a mask.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TAFL_HD __host__ __device__ __forceinline__
// rare paths: out of line where the translation unit asks for it (tafl_mcts.cu: shorter hot stream,
// +4.5 % sims/s), inline elsewhere (the call's stack frame costs the one-thread-per-game step kernel 2 %)
#if defined(TAFL_OUTLINE_RARE)
#define TAFL_COLD __host__ __device__ __noinline__
#else
#define TAFL_COLD __host__ __device__ __forceinline__
#endif
#else
#define TAFL_HD inline
#define TAFL_COLD inline
#endif

// Re-convergence points for the one-thread-per-game kernels: lanes that entered a region together
// (TAFL_LANES_HERE) wait for each other again (TAFL_RECONVERGE) before a long common tail, so the
// tail is issued once per warp instead of once per divergent group.  No-ops in the host build.
#if defined(__CUDA_ARCH__)
#define TAFL_LANES_HERE(m) const unsigned m = __activemask()
#define TAFL_RECONVERGE(m) __syncwarp(m)
#else
#define TAFL_LANES_HERE(m) const unsigned m = 0u
#define TAFL_RECONVERGE(m) (void)(m)
#endif

namespace tafl {

TAFL_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
TAFL_HD int popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
// count trailing zeros; 32 / 64 for x == 0
TAFL_HD int ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz(__brev(x));
#else
    return x ? __builtin_ctz(x) : 32;
#endif
}
TAFL_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz(x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
TAFL_HD int ctz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll(__brevll(x));
#else
    return x ? __builtin_ctzll(x) : 64;
#endif
}
TAFL_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}

struct BB64 {
    uint64_t v;
    static constexpr int BITS = 64;
    static constexpr int STRIDE = 7;   // bitfield.rs:178
    static constexpr int WORDS = 2;    // 32-bit words
    TAFL_HD static BB64 zero() { return BB64{0}; }
    TAFL_HD static BB64 bit(int i) { return BB64{(i >= 0 && i < 64) ? (1ull << i) : 0ull}; }
    TAFL_HD static BB64 from_words(uint64_t lo, uint64_t) { return BB64{lo}; }
    TAFL_HD uint64_t lo64() const { return v; }
    TAFL_HD uint64_t hi64() const { return 0; }
    TAFL_HD uint32_t word(int i) const { return (uint32_t)(v >> (32 * i)); }
    TAFL_HD void set_word(int i, uint32_t w) { v = (v & ~(0xFFFFFFFFull << (32 * i))) | ((uint64_t)w << (32 * i)); }
    TAFL_HD bool any() const { return v != 0; }
    TAFL_HD bool test(int i) const { return (i >= 0 && i < 64) ? ((v >> i) & 1) != 0 : false; }
    TAFL_HD int popc() const { return popc64(v); }
    TAFL_HD int ctz() const { return ctz64(v); }
    template <int K> TAFL_HD BB64 shl() const { return BB64{K >= 64 ? 0 : (v << (K & 63))}; }
    template <int K> TAFL_HD BB64 shr() const { return BB64{K >= 64 ? 0 : (v >> (K & 63))}; }
    TAFL_HD BB64 shlv(int k) const { return BB64{k >= 64 ? 0 : (v << k)}; }
    TAFL_HD BB64 shrv(int k) const { return BB64{k >= 64 ? 0 : (v >> k)}; }
    // low 32 bits of (*this >> k)
    TAFL_HD uint32_t extract32(int k) const { return k >= 64 ? 0u : (uint32_t)(v >> k); }
};
TAFL_HD BB64 operator&(BB64 a, BB64 b) { return BB64{a.v & b.v}; }
TAFL_HD BB64 operator|(BB64 a, BB64 b) { return BB64{a.v | b.v}; }
TAFL_HD BB64 operator^(BB64 a, BB64 b) { return BB64{a.v ^ b.v}; }
TAFL_HD BB64 operator~(BB64 a) { return BB64{~a.v}; }
TAFL_HD bool operator==(BB64 a, BB64 b) { return a.v == b.v; }
TAFL_HD bool operator!=(BB64 a, BB64 b) { return a.v != b.v; }

struct BB128 {
    uint64_t lo, hi;
    static constexpr int BITS = 128;
    static constexpr int STRIDE = 11;  // bitfield.rs:179
    static constexpr int WORDS = 4;
    TAFL_HD static BB128 zero() { return BB128{0, 0}; }
    TAFL_HD static BB128 bit(int i) {
        return BB128{(i >= 0 && i < 64) ? (1ull << i) : 0ull, (i >= 64 && i < 128) ? (1ull << (i - 64)) : 0ull};
    }
    TAFL_HD static BB128 from_words(uint64_t l, uint64_t h) { return BB128{l, h}; }
    TAFL_HD uint64_t lo64() const { return lo; }
    TAFL_HD uint64_t hi64() const { return hi; }
    TAFL_HD uint32_t word(int i) const { return (uint32_t)((i < 2 ? lo : hi) >> (32 * (i & 1))); }
    TAFL_HD void set_word(int i, uint32_t w) {
        uint64_t& h = (i < 2 ? lo : hi);
        h = (h & ~(0xFFFFFFFFull << (32 * (i & 1)))) | ((uint64_t)w << (32 * (i & 1)));
    }
    TAFL_HD bool any() const { return (lo | hi) != 0; }
    TAFL_HD bool test(int i) const {
        if (i < 0 || i >= 128) return false;
        return (((i < 64) ? lo >> i : hi >> (i - 64)) & 1) != 0;
    }
    TAFL_HD int popc() const { return popc64(lo) + popc64(hi); }
    TAFL_HD int ctz() const { return lo ? ctz64(lo) : 64 + ctz64(hi); }
    template <int K> TAFL_HD BB128 shl() const {
        if (K == 0) return *this;
        if (K >= 128) return zero();
        if (K >= 64) return BB128{0, lo << ((K - 64) & 63)};
        return BB128{lo << (K & 63), (hi << (K & 63)) | (lo >> ((64 - K) & 63))};
    }
    template <int K> TAFL_HD BB128 shr() const {
        if (K == 0) return *this;
        if (K >= 128) return zero();
        if (K >= 64) return BB128{hi >> ((K - 64) & 63), 0};
        return BB128{(lo >> (K & 63)) | (hi << ((64 - K) & 63)), hi >> (K & 63)};
    }
    TAFL_HD BB128 shlv(int k) const {
        if (k <= 0) return *this;
        if (k >= 128) return zero();
        if (k >= 64) return BB128{0, lo << (k - 64)};
        return BB128{lo << k, (hi << k) | (lo >> (64 - k))};
    }
    TAFL_HD BB128 shrv(int k) const {
        if (k <= 0) return *this;
        if (k >= 128) return zero();
        if (k >= 64) return BB128{hi >> (k - 64), 0};
        return BB128{(lo >> k) | (hi << (64 - k)), hi >> k};
    }
    TAFL_HD uint32_t extract32(int k) const { return (uint32_t)shrv(k).lo; }
};
TAFL_HD BB128 operator&(BB128 a, BB128 b) { return BB128{a.lo & b.lo, a.hi & b.hi}; }
TAFL_HD BB128 operator|(BB128 a, BB128 b) { return BB128{a.lo | b.lo, a.hi | b.hi}; }
TAFL_HD BB128 operator^(BB128 a, BB128 b) { return BB128{a.lo ^ b.lo, a.hi ^ b.hi}; }
TAFL_HD BB128 operator~(BB128 a) { return BB128{~a.lo, ~a.hi}; }
TAFL_HD bool operator==(BB128 a, BB128 b) { return a.lo == b.lo && a.hi == b.hi; }
TAFL_HD bool operator!=(BB128 a, BB128 b) { return !(a == b); }

// ---------------------------------------------------------------------------------------------
// BitfieldBoardState<U256> (row stride 15) and <U512> (row stride 21), bitfield.rs:180-181: boards up to
// 15x15 / 21x21 as NW 64-bit words.  Used by the by-value rules kernels (tafl_wide.cu); the word loops have
// compile-time bounds and unroll.  The same operations as BB64 / BB128, nothing board-size specific.
// ---------------------------------------------------------------------------------------------
template <int NW, int STRIDE_>
struct BBW {
    uint64_t w[NW];
    static constexpr int BITS = 64 * NW;
    static constexpr int STRIDE = STRIDE_;
    static constexpr int WORDS = 2 * NW;
    TAFL_HD static BBW zero() {
        BBW r;
#pragma unroll
        for (int i = 0; i < NW; ++i) r.w[i] = 0;
        return r;
    }
    TAFL_HD static BBW bit(int i) {
        BBW r = zero();
#pragma unroll
        for (int k = 0; k < NW; ++k)
            if (i >= 64 * k && i < 64 * (k + 1)) r.w[k] = 1ull << (i - 64 * k);
        return r;
    }
    TAFL_HD uint32_t word(int i) const { return (uint32_t)(w[i >> 1] >> (32 * (i & 1))); }
    TAFL_HD bool any() const {
        uint64_t a = 0;
#pragma unroll
        for (int i = 0; i < NW; ++i) a |= w[i];
        return a != 0;
    }
    TAFL_HD bool test(int i) const {
        if (i < 0 || i >= BITS) return false;
        uint64_t v = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k)
            if ((i >> 6) == k) v = w[k];
        return ((v >> (i & 63)) & 1) != 0;
    }
    TAFL_HD int popc() const {
        int c = 0;
#pragma unroll
        for (int i = 0; i < NW; ++i) c += popc64(w[i]);
        return c;
    }
    TAFL_HD int ctz() const {
        int r = BITS;
#pragma unroll
        for (int i = NW - 1; i >= 0; --i)
            if (w[i]) r = 64 * i + ctz64(w[i]);
        return r;
    }
    TAFL_HD BBW shlv(int k) const {
        if (k <= 0) return *this;
        BBW r = zero();
        if (k >= BITS) return r;
        const int ws = k >> 6, bs = k & 63;
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            uint64_t lo = 0, hi = 0;   // r.w[i] = w[i - ws] << bs | w[i - ws - 1] >> (64 - bs)
#pragma unroll
            for (int j = 0; j < NW; ++j) {
                if (j == i - ws) lo = w[j];
                if (j == i - ws - 1) hi = w[j];
            }
            r.w[i] = (lo << bs) | (bs ? (hi >> (64 - bs)) : 0ull);
        }
        return r;
    }
    TAFL_HD BBW shrv(int k) const {
        if (k <= 0) return *this;
        BBW r = zero();
        if (k >= BITS) return r;
        const int ws = k >> 6, bs = k & 63;
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            uint64_t lo = 0, hi = 0;   // r.w[i] = w[i + ws] >> bs | w[i + ws + 1] << (64 - bs)
#pragma unroll
            for (int j = 0; j < NW; ++j) {
                if (j == i + ws) lo = w[j];
                if (j == i + ws + 1) hi = w[j];
            }
            r.w[i] = (lo >> bs) | (bs ? (hi << (64 - bs)) : 0ull);
        }
        return r;
    }
    template <int K> TAFL_HD BBW shl() const {
        if (K == 0) return *this;
        BBW r = zero();
        if (K >= BITS) return r;
        constexpr int ws = (K >> 6), bs = (K & 63);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            const uint64_t lo = (i - ws >= 0 && i - ws < NW) ? w[(i - ws >= 0 && i - ws < NW) ? i - ws : 0] : 0ull;
            const uint64_t hi = (i - ws - 1 >= 0 && i - ws - 1 < NW) ? w[(i - ws - 1 >= 0 && i - ws - 1 < NW) ? i - ws - 1 : 0] : 0ull;
            r.w[i] = (lo << bs) | (bs ? (hi >> ((64 - bs) & 63)) : 0ull);
        }
        return r;
    }
    template <int K> TAFL_HD BBW shr() const {
        if (K == 0) return *this;
        BBW r = zero();
        if (K >= BITS) return r;
        constexpr int ws = (K >> 6), bs = (K & 63);
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            const uint64_t lo = (i + ws < NW) ? w[(i + ws < NW) ? i + ws : 0] : 0ull;
            const uint64_t hi = (i + ws + 1 < NW) ? w[(i + ws + 1 < NW) ? i + ws + 1 : 0] : 0ull;
            r.w[i] = (lo >> bs) | (bs ? (hi << ((64 - bs) & 63)) : 0ull);
        }
        return r;
    }
    TAFL_HD uint32_t extract32(int k) const { return (uint32_t)shrv(k).w[0]; }
};
template <int NW, int S> TAFL_HD BBW<NW, S> operator&(BBW<NW, S> a, BBW<NW, S> b) {
#pragma unroll
    for (int i = 0; i < NW; ++i) a.w[i] &= b.w[i];
    return a;
}
template <int NW, int S> TAFL_HD BBW<NW, S> operator|(BBW<NW, S> a, BBW<NW, S> b) {
#pragma unroll
    for (int i = 0; i < NW; ++i) a.w[i] |= b.w[i];
    return a;
}
template <int NW, int S> TAFL_HD BBW<NW, S> operator^(BBW<NW, S> a, BBW<NW, S> b) {
#pragma unroll
    for (int i = 0; i < NW; ++i) a.w[i] ^= b.w[i];
    return a;
}
template <int NW, int S> TAFL_HD BBW<NW, S> operator~(BBW<NW, S> a) {
#pragma unroll
    for (int i = 0; i < NW; ++i) a.w[i] = ~a.w[i];
    return a;
}
template <int NW, int S> TAFL_HD bool operator==(BBW<NW, S> a, BBW<NW, S> b) {
    uint64_t d = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) d |= a.w[i] ^ b.w[i];
    return d == 0;
}
template <int NW, int S> TAFL_HD bool operator!=(BBW<NW, S> a, BBW<NW, S> b) { return !(a == b); }
using BB256 = BBW<4, 15>;   // bitfield.rs:180
using BB512 = BBW<8, 21>;   // bitfield.rs:181

template <int WBITS> struct BBSel;
template <> struct BBSel<64> { using type = BB64; };
template <> struct BBSel<128> { using type = BB128; };
template <> struct BBSel<256> { using type = BB256; };
template <> struct BBSel<512> { using type = BB512; };

}  // namespace tafl
