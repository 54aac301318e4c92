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

template <int WBITS> struct BBSel;
template <> struct BBSel<64> { using type = BB64; };
template <> struct BBSel<128> { using type = BB128; };

}  // namespace tafl
