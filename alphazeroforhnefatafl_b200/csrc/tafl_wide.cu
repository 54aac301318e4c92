// tafl_wide.cu — large boards: BitfieldBoardState<U256> / <U512> (15x15 / 21x21, game/board/state.rs:337-340,
// game/bitfield.rs:180-181) behind the C ABI's `tafl_w_*` calls (include/tafl.h, "Large boards").
//
// The reference is generic over the bitfield width and instantiates its logic tests at all four widths
// (game/game/logic.rs:1015-1021,1081-1087); no preset, benchmark configuration or MCTS use touches the two
// large ones, so they get the by-value GameLogic surface — one kernel per call over the caller's batch — and
// nothing resident.  The rules are the width-generic device functions of tafl_core.cuh; sm_100a only, no CPU path.
#include <cuda_runtime.h>

#include <cstring>
#include <string>

#include "tafl_ctx.hpp"
#include "tafl_wide.cuh"

using namespace tafl;

struct tafl_wide {
    int device = 0, width = 0, N = 0;
    tafl_rules rules{};
    Ctx<BB256> cx256;
    Ctx<BB512> cx512;
    cudaStream_t stream = nullptr;
    void* d_buf[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t bytes[4] = {0, 0, 0, 0};
    std::string err;
};

namespace {

std::string g_wide_create_error;

// ---- host helpers on one wide state (board/state.rs) -------------------------------------------------------
struct WGeo { int bits, stride, top; };   // top = index of the word that holds the king nibble
inline bool wgeo_for(int width_bits, WGeo* g) {
    if (width_bits == 256) { *g = WGeo{256, 15, 3}; return true; }   // bitfield.rs:180
    if (width_bits == 512) { *g = WGeo{512, 21, 7}; return true; }   // bitfield.rs:181
    return false;
}
inline bool wget(const uint64_t* w, unsigned i, const WGeo& g) { return i < (unsigned)g.bits && ((w[i >> 6] >> (i & 63)) & 1); }
inline void wput(uint64_t* w, unsigned i, bool v, const WGeo& g) {
    if (i >= (unsigned)g.bits) return;
    if (v) w[i >> 6] |= 1ull << (i & 63);
    else w[i >> 6] &= ~(1ull << (i & 63));
}
inline void wking_set(tafl_state_w* s, const WGeo& g, int r, int c) {   // board/state.rs:136-147: four bits each
    s->def[g.top] = (s->def[g.top] & 0x0FFFFFFFFFFFFFFFull) | ((uint64_t)(r & 0xF) << 60);
    s->att[g.top] = (s->att[g.top] & 0x0FFFFFFFFFFFFFFFull) | ((uint64_t)(c & 0xF) << 60);
}
inline void wpiece_set(tafl_state_w* s, const WGeo& g, int r, int c, bool king, bool defender) {   // board/state.rs:149-165
    const unsigned i = (unsigned)r * g.stride + c;
    if (!defender) { wput(s->att, i, true, g); wput(s->def, i, false, g); }
    else { wput(s->def, i, true, g); wput(s->att, i, false, g); }
    if (king) wking_set(s, g, r, c);
}
inline int wpiece_of_char(char ch, bool* king, bool* defender) {   // pieces.rs:119-143
    const bool up = ch >= 'A' && ch <= 'Z', low = ch >= 'a' && ch <= 'z';
    if (!up && !low) return TAFL_ERR_PARSE_BAD_CHAR;
    *defender = up;
    switch (up ? ch - 'A' + 'a' : ch) {
        case 'k': *king = true; return 0;
        case 't': case 'n': case 'c': case 'g': case 'm': *king = false; return 0;
    }
    return TAFL_ERR_PARSE_BAD_CHAR;
}

// ---- kernels: one query per thread -----------------------------------------------------------------------------
template <class BB>
__global__ void k_w_validate(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const tafl_play* pl, uint32_t n, uint8_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    out[i] = (uint8_t)validate_play(cx, s, pl[i], (s.flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
}
template <class BB>
__global__ void k_w_do_plays(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const tafl_play* pl, uint32_t n,
                             tafl_state_w* out_st, tafl_play_effects_w* out_fx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    tafl_state_w s = st[i];
    const tafl_play_effects_w fx = do_play_w(cx, s, pl[i]);
    out_st[i] = s;
    out_fx[i] = fx;
}
template <class BB>
__global__ void k_w_legal(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, uint32_t n, tafl_play* plays, uint32_t* counts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    counts[i] = (uint32_t)gen_plays(cx, s, plays + (size_t)i * TAFL_W_MAX_PLAYS, TAFL_W_MAX_PLAYS);
}
template <class BB>
__global__ void k_w_piece_plays(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const uint8_t* rc, uint32_t n,
                                tafl_play* plays, int32_t* counts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    counts[i] = piece_plays_w(cx, s, rc[2 * i], rc[2 * i + 1], plays + (size_t)i * 64, 64);
}
template <class BB>
__global__ void k_w_captures(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const tafl_play* pl, uint32_t n,
                             tafl_play_effects_w* out, int shieldwall_only) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    out[i] = shieldwall_only ? query_shieldwall_w(cx, s, pl[i]) : query_captures_w(cx, s, pl[i]);
}
template <class BB>
__global__ void k_w_enclosure(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const tafl_enclosure_query* q, uint32_t n,
                              tafl_enclosure_w* out, uint8_t* secure) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    uint8_t sec;
    const tafl_enclosure_w e = query_enclosure_w(cx, s, q[i], &sec);
    if (out) out[i] = e;
    if (secure) secure[i] = sec;
}
template <class BB>
__global__ void k_w_exit_fort(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, uint32_t n, uint8_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    out[i] = detect_exit_fort(cx, load_pos(cx, s)) ? 1 : 0;
}
template <class BB>
__global__ void k_w_can_play(const __grid_constant__ Ctx<BB> cx, const tafl_state_w* st, const uint8_t* sides, uint32_t n, uint8_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const tafl_state_w s = st[i];
    out[i] = query_can_play_w(cx, s, sides[i] != 0);
}

int w_cuda_error(tafl_wide* w, cudaError_t ce, const char* what) {
    w->err = std::string("CUDA error: ") + cudaGetErrorString(ce) + " in " + what;
    return TAFL_ERR_CUDA;
}
#define W_TRY(w, call)                                                  \
    do {                                                                \
        cudaError_t _ce = (call);                                       \
        if (_ce != cudaSuccess) return w_cuda_error((w), _ce, #call);   \
    } while (0)

int w_buf(tafl_wide* w, int slot, size_t bytes, void** out) {
    if (bytes == 0) bytes = 16;
    if (w->bytes[slot] < bytes) {
        if (w->d_buf[slot]) cudaFree(w->d_buf[slot]);
        w->d_buf[slot] = nullptr;
        w->bytes[slot] = 0;
        W_TRY(w, cudaMalloc(&w->d_buf[slot], bytes + bytes / 2));
        w->bytes[slot] = bytes + bytes / 2;
    }
    *out = w->d_buf[slot];
    return TAFL_OK;
}
template <class T>
int w_up(tafl_wide* w, int slot, const T* h, size_t n, T** d) {
    void* p;
    int rc = w_buf(w, slot, n * sizeof(T), &p);
    if (rc) return rc;
    W_TRY(w, cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, w->stream));
    *d = (T*)p;
    return TAFL_OK;
}
template <class T>
int w_out(tafl_wide* w, int slot, size_t n, T** d) {
    void* p;
    int rc = w_buf(w, slot, n * sizeof(T), &p);
    if (rc) return rc;
    *d = (T*)p;
    return TAFL_OK;
}
template <class T>
int w_down(tafl_wide* w, T* h, const T* d, size_t n) {
    W_TRY(w, cudaMemcpyAsync(h, d, n * sizeof(T), cudaMemcpyDeviceToHost, w->stream));
    return TAFL_OK;
}
int w_check(tafl_wide* w, const tafl_state_w* st, uint32_t n) {
    if (!w) return TAFL_ERR_INVALID_ARG;
    if (!st && n) return TAFL_ERR_INVALID_ARG;
    for (uint32_t i = 0; i < n; ++i)
        if (st[i].side_len != w->N) { w->err = "state side_len does not match the board this tafl_wide was created for"; return TAFL_ERR_INVALID_ARG; }
    cudaSetDevice(w->device);
    return TAFL_OK;
}
#define W_DISPATCH(w, body)                                                             \
    do {                                                                                \
        if ((w)->width == 256) { using BB = BB256; const Ctx<BB256>& cx = (w)->cx256; body; }   \
        else { using BB = BB512; const Ctx<BB512>& cx = (w)->cx512; body; }                     \
    } while (0)
#define W_FINISH(w)                                      \
    do {                                                 \
        W_TRY((w), cudaGetLastError());                  \
        W_TRY((w), cudaStreamSynchronize((w)->stream));  \
        return TAFL_OK;                                  \
    } while (0)
inline unsigned wgrid(uint32_t n) { return (n + 63u) / 64u; }

}  // namespace

extern "C" {

// board/state.rs:225-250 + game/state.rs:136-145
int tafl_w_state_from_fen(const char* fen, int width_bits, int side, tafl_state_w* out) {
    WGeo g;
    if (!fen || !out || !wgeo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    tafl_state_w s;
    std::memset(&s, 0, sizeof(s));
    if (side == TAFL_DEFENDER) s.flags |= TAFL_FLAG_DEF_TO_PLAY;
    int row = 0;
    unsigned n_empty = 0, col = 0;
    for (const char* p = fen;; ++p) {
        const char ch = *p;
        if (ch == '/' || ch == '\0') {
            col = (col + n_empty) & 0xFF;
            if (s.side_len == 0) s.side_len = (uint8_t)col;
            else if (s.side_len != col) return TAFL_ERR_PARSE_BAD_LINE_LEN;
            if (ch == '\0') break;
            ++row; col = 0; n_empty = 0;
        } else if (ch >= '0' && ch <= '9') {
            n_empty = (n_empty * 10 + (unsigned)(ch - '0')) & 0xFF;
        } else {
            col = (col + n_empty) & 0xFF;
            n_empty = 0;
            bool king, defender;
            const int rc = wpiece_of_char(ch, &king, &defender);
            if (rc) return rc;
            wpiece_set(&s, g, row, (int)col, king, defender);
            col = (col + 1) & 0xFF;
        }
    }
    *out = s;
    return TAFL_OK;
}
// board/state.rs:271-301
int tafl_w_state_to_fen(const tafl_state_w* st, int width_bits, char* buf, size_t buf_len) {
    WGeo g;
    if (!st || !buf || !wgeo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    const int kr = (int)(st->def[g.top] >> 60), kc = (int)(st->att[g.top] >> 60);
    std::string s;
    for (int r = 0; r < st->side_len; ++r) {
        int run = 0;
        for (int c = 0; c < st->side_len; ++c) {
            const unsigned i = (unsigned)r * g.stride + c;
            char ch = 0;
            if (wget(st->def, i, g)) ch = (kr == r && kc == c) ? 'K' : 'T';
            else if (wget(st->att, i, g)) ch = 't';
            if (!ch) { ++run; continue; }
            if (run) { s += std::to_string(run); run = 0; }
            s += ch;
        }
        if (run) s += std::to_string(run);
        if (r + 1 < st->side_len) s += '/';
    }
    if (s.size() + 1 > buf_len) return TAFL_ERR_INVALID_ARG;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return TAFL_OK;
}
// board/state.rs:218-223
int tafl_w_board_move_piece(tafl_state_w* st, int width_bits, uint8_t fr, uint8_t fc, uint8_t tr, uint8_t tc) {
    WGeo g;
    if (!st || !wgeo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    const unsigned f = (unsigned)fr * g.stride + fc;
    const int kr = (int)(st->def[g.top] >> 60), kc = (int)(st->att[g.top] >> 60);
    bool king = false, defender = false;
    if (wget(st->def, f, g)) { defender = true; king = (kr == fr && kc == fc); }
    else if (!wget(st->att, f, g)) return TAFL_ERR_NO_PIECE;
    wpiece_set(st, g, tr, tc, king, defender);
    wput(st->att, f, false, g);
    wput(st->def, f, false, g);
    return TAFL_OK;
}

int tafl_wide_create(const tafl_rules* rules, int side_len, int width_bits, int device, tafl_wide** out) {
    if (!rules || !out) { g_wide_create_error = "invalid argument"; return TAFL_ERR_INVALID_ARG; }
    *out = nullptr;
    if ((width_bits != 256 && width_bits != 512) || side_len < 1 || side_len > (width_bits == 256 ? 15 : 21)) {
        g_wide_create_error = "tafl_wide_create: width 256 holds boards up to 15x15, width 512 up to 21x21";
        return TAFL_ERR_UNSUPPORTED;
    }
    int ndev = 0;
    const cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0) {
        g_wide_create_error = std::string("no usable CUDA device (") + cudaGetErrorString(ce) + "); libtafl_b200 has no CPU fallback";
        return TAFL_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { g_wide_create_error = "device index out of range"; return TAFL_ERR_INVALID_ARG; }
    tafl_wide* w = new tafl_wide();
    w->device = device; w->width = width_bits; w->N = side_len; w->rules = *rules;
    w->cx256 = make_ctx<BB256>(*rules, side_len <= 15 ? side_len : 15);
    w->cx512 = make_ctx<BB512>(*rules, side_len);
    cudaSetDevice(device);
    if (cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking) != cudaSuccess) {
        g_wide_create_error = "cudaStreamCreate failed";
        delete w;
        return TAFL_ERR_CUDA;
    }
    *out = w;
    return TAFL_OK;
}
int tafl_wide_destroy(tafl_wide* w) {
    if (!w) return TAFL_OK;
    cudaSetDevice(w->device);
    if (w->stream) { cudaStreamSynchronize(w->stream); cudaStreamDestroy(w->stream); }
    for (auto& p : w->d_buf) cudaFree(p);
    delete w;
    return TAFL_OK;
}
const char* tafl_wide_last_error(const tafl_wide* w) { return w ? w->err.c_str() : g_wide_create_error.c_str(); }

int tafl_w_validate_plays(tafl_wide* w, const tafl_state_w* h_st, const tafl_play* h_pl, uint32_t n, uint8_t* h_out) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; tafl_play* d_pl; uint8_t* d_out;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_pl, n, &d_pl)) || (rc = w_out(w, 2, n, &d_out))) return rc;
    W_DISPATCH(w, (k_w_validate<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_pl, n, d_out)));
    if ((rc = w_down(w, h_out, d_out, n))) return rc;
    W_FINISH(w);
}
int tafl_w_do_plays(tafl_wide* w, const tafl_state_w* h_st, const tafl_play* h_pl, uint32_t n, tafl_state_w* h_out_st,
                    tafl_play_effects_w* h_out_fx) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w *d_st, *d_ost; tafl_play* d_pl; tafl_play_effects_w* d_fx;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_pl, n, &d_pl)) || (rc = w_out(w, 2, n, &d_ost)) || (rc = w_out(w, 3, n, &d_fx)))
        return rc;
    W_DISPATCH(w, (k_w_do_plays<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_pl, n, d_ost, d_fx)));
    if (h_out_st && (rc = w_down(w, h_out_st, d_ost, n))) return rc;
    if (h_out_fx && (rc = w_down(w, h_out_fx, d_fx, n))) return rc;
    W_FINISH(w);
}
int tafl_w_legal_plays_of(tafl_wide* w, const tafl_state_w* h_st, uint32_t n, tafl_play* h_plays, uint32_t* h_counts) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    if (!h_plays || !h_counts) return TAFL_ERR_INVALID_ARG;
    tafl_state_w* d_st; tafl_play* d_pl; uint32_t* d_cnt;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_out(w, 1, (size_t)n * TAFL_W_MAX_PLAYS, &d_pl)) || (rc = w_out(w, 2, n, &d_cnt))) return rc;
    W_DISPATCH(w, (k_w_legal<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, n, d_pl, d_cnt)));
    if ((rc = w_down(w, h_plays, d_pl, (size_t)n * TAFL_W_MAX_PLAYS)) || (rc = w_down(w, h_counts, d_cnt, n))) return rc;
    W_FINISH(w);
}
int tafl_w_piece_plays(tafl_wide* w, const tafl_state_w* h_st, const uint8_t* h_rc, uint32_t n, tafl_play* h_plays, int32_t* h_counts) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; uint8_t* d_rc; tafl_play* d_pl; int32_t* d_cnt;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_rc, (size_t)2 * n, &d_rc)) || (rc = w_out(w, 2, (size_t)n * 64, &d_pl)) ||
        (rc = w_out(w, 3, n, &d_cnt)))
        return rc;
    W_DISPATCH(w, (k_w_piece_plays<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_rc, n, d_pl, d_cnt)));
    if ((rc = w_down(w, h_plays, d_pl, (size_t)n * 64)) || (rc = w_down(w, h_counts, d_cnt, n))) return rc;
    W_FINISH(w);
}
static int w_captures_common(tafl_wide* w, const tafl_state_w* h_st, const tafl_play* h_pl, uint32_t n, tafl_play_effects_w* h_out, int sw) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; tafl_play* d_pl; tafl_play_effects_w* d_fx;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_pl, n, &d_pl)) || (rc = w_out(w, 2, n, &d_fx))) return rc;
    W_DISPATCH(w, (k_w_captures<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_pl, n, d_fx, sw)));
    if ((rc = w_down(w, h_out, d_fx, n))) return rc;
    W_FINISH(w);
}
int tafl_w_get_captures(tafl_wide* w, const tafl_state_w* s, const tafl_play* p, uint32_t n, tafl_play_effects_w* o) { return w_captures_common(w, s, p, n, o, 0); }
int tafl_w_detect_shieldwall(tafl_wide* w, const tafl_state_w* s, const tafl_play* p, uint32_t n, tafl_play_effects_w* o) { return w_captures_common(w, s, p, n, o, 1); }
static int w_enclosure_common(tafl_wide* w, const tafl_state_w* h_st, const tafl_enclosure_query* h_q, uint32_t n, tafl_enclosure_w* h_enc, uint8_t* h_sec) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; tafl_enclosure_query* d_q; tafl_enclosure_w* d_enc; uint8_t* d_sec;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_q, n, &d_q)) || (rc = w_out(w, 2, n, &d_enc)) || (rc = w_out(w, 3, n, &d_sec))) return rc;
    W_DISPATCH(w, (k_w_enclosure<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_q, n, d_enc, d_sec)));
    if (h_enc && (rc = w_down(w, h_enc, d_enc, n))) return rc;
    if (h_sec && (rc = w_down(w, h_sec, d_sec, n))) return rc;
    W_FINISH(w);
}
int tafl_w_find_enclosure(tafl_wide* w, const tafl_state_w* s, const tafl_enclosure_query* q, uint32_t n, tafl_enclosure_w* o) { return w_enclosure_common(w, s, q, n, o, nullptr); }
int tafl_w_enclosure_secure(tafl_wide* w, const tafl_state_w* s, const tafl_enclosure_query* q, uint32_t n, uint8_t* o) { return w_enclosure_common(w, s, q, n, nullptr, o); }
int tafl_w_detect_exit_fort(tafl_wide* w, const tafl_state_w* h_st, uint32_t n, uint8_t* h_out) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; uint8_t* d_out;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_out(w, 1, n, &d_out))) return rc;
    W_DISPATCH(w, (k_w_exit_fort<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, n, d_out)));
    if ((rc = w_down(w, h_out, d_out, n))) return rc;
    W_FINISH(w);
}
int tafl_w_side_can_play(tafl_wide* w, const tafl_state_w* h_st, const uint8_t* h_sides, uint32_t n, uint8_t* h_out) {
    int rc = w_check(w, h_st, n);
    if (rc || !n) return rc;
    tafl_state_w* d_st; uint8_t *d_sides, *d_out;
    if ((rc = w_up(w, 0, h_st, n, &d_st)) || (rc = w_up(w, 1, h_sides, n, &d_sides)) || (rc = w_out(w, 2, n, &d_out))) return rc;
    W_DISPATCH(w, (k_w_can_play<BB><<<wgrid(n), 64, 0, w->stream>>>(cx, d_st, d_sides, n, d_out)));
    if ((rc = w_down(w, h_out, d_out, n))) return rc;
    W_FINISH(w);
}

}  // extern "C"
