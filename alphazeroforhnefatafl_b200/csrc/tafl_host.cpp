// tafl_host.cpp — host-only entry points of libtafl_b200.so: presets, FEN / notation text I/O
// and the `trait BoardState` accessors on one packed tafl_state.  None of this is on the hot
// path; it is the boundary the reference's callers use to build and inspect positions
// (reference tree: game/preset.rs, game/board/state.rs, game/tiles.rs, game/play.rs).
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/tafl.h"

namespace {

struct Geo {
    int bits, stride;
};
inline bool geo_for(int width_bits, Geo* g) {
    if (width_bits == 64) { *g = Geo{64, 7}; return true; }     // bitfield.rs:178
    if (width_bits == 128) { *g = Geo{128, 11}; return true; }  // bitfield.rs:179
    return false;
}
// the word holding the top nibble
inline uint64_t& top(uint64_t w[2], const Geo& g) { return g.bits == 64 ? w[0] : w[1]; }
inline uint64_t top(const uint64_t w[2], const Geo& g) { return g.bits == 64 ? w[0] : w[1]; }
inline bool get_bit(const uint64_t w[2], unsigned i, const Geo& g) {
    if (i >= (unsigned)g.bits) return false;
    return (w[i >> 6] >> (i & 63)) & 1;
}
inline void put_bit(uint64_t w[2], unsigned i, bool v, const Geo& g) {
    if (i >= (unsigned)g.bits) return;
    if (v) w[i >> 6] |= 1ull << (i & 63);
    else w[i >> 6] &= ~(1ull << (i & 63));
}
inline unsigned tile_index(int row, int col, const Geo& g) { return (unsigned)row * g.stride + col; }

// board/state.rs:127-147
inline void king_of(const tafl_state* s, const Geo& g, int* r, int* c) {
    *r = (int)(top(s->def, g) >> 60);
    *c = (int)(top(s->att, g) >> 60);
}
inline void king_set(tafl_state* s, const Geo& g, int r, int c) {
    uint64_t& d = top(s->def, g);
    uint64_t& a = top(s->att, g);
    d = (d & 0x0FFFFFFFFFFFFFFFull) | ((uint64_t)(r & 0xF) << 60);
    a = (a & 0x0FFFFFFFFFFFFFFFull) | ((uint64_t)(c & 0xF) << 60);
}
// board/state.rs:149-165
inline void piece_set(tafl_state* s, const Geo& g, int r, int c, int type, int side) {
    unsigned i = tile_index(r, c, g);
    if (side == TAFL_ATTACKER) { put_bit(s->att, i, true, g); put_bit(s->def, i, false, g); }
    else { put_bit(s->def, i, true, g); put_bit(s->att, i, false, g); }
    if (type == TAFL_KING) king_set(s, g, r, c);
}
// board/state.rs:173-187
inline bool piece_get(const tafl_state* s, const Geo& g, int r, int c, int* type, int* side) {
    unsigned i = tile_index(r, c, g);
    if (get_bit(s->def, i, g)) {
        int kr, kc;
        king_of(s, g, &kr, &kc);
        *type = (kr == r && kc == c) ? TAFL_KING : TAFL_SOLDIER;
        *side = TAFL_DEFENDER;
        return true;
    }
    if (get_bit(s->att, i, g)) { *type = TAFL_SOLDIER; *side = TAFL_ATTACKER; return true; }
    *type = 0; *side = 0;
    return false;
}
inline void tile_clear(tafl_state* s, const Geo& g, int r, int c) {
    unsigned i = tile_index(r, c, g);
    put_bit(s->att, i, false, g);
    put_bit(s->def, i, false, g);
}
// pieces.rs:119-143
inline int piece_of_char(char ch, int* type, int* side) {
    bool up = ch >= 'A' && ch <= 'Z', low = ch >= 'a' && ch <= 'z';
    if (!up && !low) return TAFL_ERR_PARSE_BAD_CHAR;
    *side = up ? TAFL_DEFENDER : TAFL_ATTACKER;
    switch (up ? ch - 'A' + 'a' : ch) {
        case 't': *type = TAFL_SOLDIER; return 0;
        case 'k': *type = TAFL_KING; return 0;
        case 'n': *type = TAFL_KNIGHT; return 0;
        case 'c': *type = TAFL_COMMANDER; return 0;
        case 'g': *type = TAFL_GUARD; return 0;
        case 'm': *type = TAFL_MERCENARY; return 0;
    }
    return TAFL_ERR_PARSE_BAD_CHAR;
}
inline char char_of_piece(int type, int side) {  // pieces.rs:100-117 (only k / t are ever stored)
    char c = type == TAFL_KING ? 'k' : 't';
    return side == TAFL_DEFENDER ? (char)(c - 'a' + 'A') : c;
}
inline void fresh_state(tafl_state* s, int side) {
    std::memset(s, 0, sizeof(*s));
    if (side == TAFL_DEFENDER) s->flags |= TAFL_FLAG_DEF_TO_PLAY;
}
inline int copy_out(const std::string& s, char* buf, size_t len) {
    if (!buf || s.size() + 1 > len) return TAFL_ERR_INVALID_ARG;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return TAFL_OK;
}
// tiles.rs:137-157
int parse_tile(const std::string& s, int* row, int* col) {
    if (s.empty()) return TAFL_ERR_PARSE_EMPTY;
    unsigned char b = (unsigned char)s[0];
    if (b < 97 || b > 122) return TAFL_ERR_PARSE_BAD_CHAR;
    size_t i = 1;
    if (i < s.size() && s[i] == '+') ++i;
    if (i >= s.size()) return TAFL_ERR_PARSE_BAD_INT;
    unsigned v = 0;
    for (; i < s.size(); ++i) {
        if (s[i] < '0' || s[i] > '9') return TAFL_ERR_PARSE_BAD_INT;
        v = v * 10 + (unsigned)(s[i] - '0');
        if (v > 255) return TAFL_ERR_PARSE_BAD_INT;
    }
    if (v == 0) return TAFL_ERR_PARSE_BAD_INT;
    *row = (int)v - 1;
    *col = b - 97;
    return TAFL_OK;
}
std::string show_tile(int row, int col) { return std::string(1, (char)(col + 97)) + std::to_string(row + 1); }

}  // namespace

extern "C" {

const char* tafl_version(void) { return "tafl_b200 0.1.0 (sm_100a)"; }

// preset.rs:12-124
int tafl_preset_rules(int id, tafl_rules* out) {
    if (!out) return TAFL_ERR_INVALID_ARG;
    tafl_rules r;
    std::memset(&r, 0, sizeof(r));
    const uint16_t K = TAFL_PIECESET_TYPE(TAFL_KING), S = TAFL_PIECESET_TYPE(TAFL_SOLDIER);
    r.king_attack = TAFL_KING_ARMED;
    r.starting_side = TAFL_ATTACKER;
    r.throne_movement = TAFL_THRONE_KINGENTRY;
    r.may_enter_corners = K;
    r.hostility_edge = TAFL_PIECESET_NONE;
    r.slow_pieces = TAFL_PIECESET_NONE;
    auto rep3 = [&](bool loss) { r.repetition_enabled = 1; r.n_repetitions = 3; r.repetition_is_loss = loss; };
    switch (id) {
        case TAFL_PRESET_COPENHAGEN:
            r.king_strength = TAFL_KING_STRONG;
            r.shieldwall_enabled = 1; r.sw_corners_may_close = 1; r.sw_captures = S;
            r.exit_fort = 1;
            r.hostility_throne = TAFL_PIECESET_ALL; r.hostility_corners = S;
            r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            rep3(true);
            break;
        case TAFL_PRESET_BRANDUBH:
            r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.hostility_throne = S; r.hostility_corners = TAFL_PIECESET_ALL;
            r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            rep3(true);
            break;
        case TAFL_PRESET_MAGPIE:
            r.king_strength = TAFL_KING_STRONG;
            r.hostility_throne = TAFL_PIECESET_ALL; r.hostility_corners = TAFL_PIECESET_ALL;
            r.slow_pieces = K;
            r.enclosure_win = TAFL_ENCL_NONE;
            break;
        case TAFL_PRESET_TABLUT:
            r.edge_escape = 1;
            r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.throne_movement = TAFL_THRONE_NOENTRY;
            r.may_enter_corners = TAFL_PIECESET_ALL;
            r.hostility_throne = TAFL_PIECESET_ALL; r.hostility_corners = TAFL_PIECESET_NONE;
            r.enclosure_win = TAFL_ENCL_NONE;
            rep3(false);
            r.draw_on_no_plays = 1;
            r.linnaean_capture = 1;
            break;
        case TAFL_PRESET_KOCH:
            r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.hostility_throne = TAFL_PIECESET_ALL; r.hostility_corners = S;
            r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            rep3(true);
            break;
        default:
            return TAFL_ERR_INVALID_ARG;
    }
    *out = r;
    return TAFL_OK;
}

// preset.rs:127-134
int tafl_preset_board_fen(int id, const char** out) {
    if (!out) return TAFL_ERR_INVALID_ARG;
    switch (id) {
        case TAFL_PRESET_COPENHAGEN:
            *out = "3ttttt3/5t5/11/t4T4t/t3TTT3t/tt1TTKTT1tt/t3TTT3t/t4T4t/11/5t5/3ttttt3";
            return TAFL_OK;
        case TAFL_PRESET_BRANDUBH:
        case TAFL_PRESET_KOCH:
            *out = "3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3";
            return TAFL_OK;
        case TAFL_PRESET_MAGPIE:
            *out = "3t3/1t3t1/3T3/t1TKT1t/3T3/1t3t1/3t3";
            return TAFL_OK;
        case TAFL_PRESET_TABLUT:
            *out = "3ttt3/4t4/4T4/t3T3t/ttTTKTTtt/t3T3t/4T4/4t4/3ttt3";
            return TAFL_OK;
    }
    return TAFL_ERR_INVALID_ARG;
}

// board/state.rs:225-250 + game/state.rs:136-145
int tafl_state_from_fen(const char* fen, int width_bits, int side, tafl_state* out) {
    Geo g;
    if (!fen || !out || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    tafl_state s;
    fresh_state(&s, side);
    int row = 0;
    unsigned n_empty = 0, col = 0;
    for (const char* p = fen;; ++p) {
        char ch = *p;
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
            int type, sd;
            int rc = piece_of_char(ch, &type, &sd);
            if (rc) return rc;
            piece_set(&s, g, row, (int)col, type, sd);
            col = (col + 1) & 0xFF;
        }
    }
    *out = s;
    return TAFL_OK;
}

// board/state.rs:252-269
int tafl_state_from_display_str(const char* str, int width_bits, int side, tafl_state* out) {
    Geo g;
    if (!str || !out || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    std::string s(str);
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    s = (a == std::string::npos) ? std::string() : s.substr(a, b - a + 1);
    tafl_state st;
    fresh_state(&st, side);
    int row = 0;
    size_t pos = 0;
    while (!s.empty()) {
        size_t nl = s.find('\n', pos);
        std::string line = s.substr(pos, nl == std::string::npos ? std::string::npos : nl - pos);
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (st.side_len == 0) st.side_len = (uint8_t)line.size();
        else if (line.size() != st.side_len) return TAFL_ERR_PARSE_BAD_LINE_LEN;
        for (size_t c = 0; c < line.size(); ++c) {
            if (line[c] == '.') continue;
            int type, sd;
            int rc = piece_of_char(line[c], &type, &sd);
            if (rc) return rc;
            piece_set(&st, g, row, (int)c, type, sd);
        }
        if (nl == std::string::npos) break;
        pos = nl + 1;
        ++row;
    }
    *out = st;
    return TAFL_OK;
}

// board/state.rs:271-295
int tafl_state_to_fen(const tafl_state* st, int width_bits, char* buf, size_t len) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    std::string s;
    const int n = st->side_len;
    for (int r = 0; r < n; ++r) {
        int gap = 0;
        for (int c = 0; c < n; ++c) {
            int type, side;
            if (piece_get(st, g, r, c, &type, &side)) {
                if (gap) { s += std::to_string(gap); gap = 0; }
                s.push_back(char_of_piece(type, side));
            } else ++gap;
        }
        if (gap) s += std::to_string(gap);
        if (r < n - 1) s.push_back('/');
    }
    return copy_out(s, buf, len);
}

// board/state.rs:297-311
int tafl_state_to_display_str(const tafl_state* st, int width_bits, char* buf, size_t len) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    std::string s;
    const int n = st->side_len;
    for (int r = 0; r < n; ++r) {
        for (int c = 0; c < n; ++c) {
            int type, side;
            s.push_back(piece_get(st, g, r, c, &type, &side) ? char_of_piece(type, side) : '.');
        }
        s.push_back('\n');
    }
    return copy_out(s, buf, len);
}

int tafl_board_get_king(const tafl_state* st, int width_bits, uint8_t* row, uint8_t* col) {
    Geo g;
    if (!st || !row || !col || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    int r, c;
    king_of(st, g, &r, &c);
    *row = (uint8_t)r; *col = (uint8_t)c;
    return TAFL_OK;
}
int tafl_board_set_king(tafl_state* st, int width_bits, uint8_t row, uint8_t col) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    king_set(st, g, row, col);
    return TAFL_OK;
}
int tafl_board_is_king(const tafl_state* st, int width_bits, uint8_t row, uint8_t col) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    int r, c;
    king_of(st, g, &r, &c);
    return (r == row && c == col) ? 1 : 0;
}
int tafl_board_set_piece(tafl_state* st, int width_bits, uint8_t row, uint8_t col, int type, int side) {
    Geo g;
    if (!st || !geo_for(width_bits, &g) || (side != TAFL_ATTACKER && side != TAFL_DEFENDER)) return TAFL_ERR_INVALID_ARG;
    piece_set(st, g, row, col, type, side);
    return TAFL_OK;
}
int tafl_board_clear_tile(tafl_state* st, int width_bits, uint8_t row, uint8_t col) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    tile_clear(st, g, row, col);
    return TAFL_OK;
}
int tafl_board_get_piece(const tafl_state* st, int width_bits, uint8_t row, uint8_t col, int* type, int* side) {
    Geo g;
    if (!st || !type || !side || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    piece_get(st, g, row, col, type, side);
    return TAFL_OK;
}
int tafl_board_tile_occupied(const tafl_state* st, int width_bits, uint8_t row, uint8_t col) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    unsigned i = tile_index(row, col, g);
    return (get_bit(st->att, i, g) || get_bit(st->def, i, g)) ? 1 : 0;
}
// board/state.rs:195-200: popcount of the word shifted left by 4 (drops the nibble)
int tafl_board_count_pieces(const tafl_state* st, int width_bits, int side) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    const uint64_t* w = side == TAFL_DEFENDER ? st->def : st->att;
    if (g.bits == 64) return __builtin_popcountll(w[0] << 4);
    return __builtin_popcountll(w[0]) + __builtin_popcountll(w[1] << 4);
}
// board/state.rs:202-216 + :91-102
int tafl_board_iter_occupied(const tafl_state* st, int width_bits, int side, uint8_t* out_rc) {
    Geo g;
    if (!st || !out_rc || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    const uint64_t* w = side == TAFL_DEFENDER ? st->def : st->att;
    int n = 0;
    for (unsigned i = 0; i < (unsigned)g.bits - 4; ++i)
        if (get_bit(w, i, g)) {
            out_rc[2 * n] = (uint8_t)(i / g.stride);
            out_rc[2 * n + 1] = (uint8_t)(i % g.stride);
            ++n;
        }
    return n;
}
// board/state.rs:218-223
int tafl_board_move_piece(tafl_state* st, int width_bits, uint8_t fr, uint8_t fc, uint8_t tr, uint8_t tc) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    int type, side;
    if (!piece_get(st, g, fr, fc, &type, &side)) return TAFL_ERR_NO_PIECE;
    piece_set(st, g, tr, tc, type, side);
    tile_clear(st, g, fr, fc);
    return TAFL_OK;
}
// board/state.rs:69-79
int tafl_board_swap_pieces(tafl_state* st, int width_bits, uint8_t r1, uint8_t c1, uint8_t r2, uint8_t c2) {
    Geo g;
    if (!st || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    int t1, s1, t2, s2;
    bool h1 = piece_get(st, g, r1, c1, &t1, &s1), h2 = piece_get(st, g, r2, c2, &t2, &s2);
    if (h2) piece_set(st, g, r1, c1, t2, s2); else tile_clear(st, g, r1, c1);
    if (h1) piece_set(st, g, r2, c2, t1, s1); else tile_clear(st, g, r2, c2);
    return TAFL_OK;
}

// play.rs:36-49
int tafl_play_from_tiles(uint8_t fr, uint8_t fc, uint8_t tr, uint8_t tc, tafl_play* out) {
    if (!out) return TAFL_ERR_INVALID_ARG;
    tafl_play p;
    p.from_row = fr; p.from_col = fc;
    if (fr == tr) { p.axis = TAFL_HORIZONTAL; p.disp = (int8_t)((int8_t)tc - (int8_t)fc); }
    else if (fc == tc) { p.axis = TAFL_VERTICAL; p.disp = (int8_t)((int8_t)tr - (int8_t)fr); }
    else return TAFL_ERR_PARSE_BAD_PLAY;  // PlayError::DisjointTiles
    *out = p;
    return TAFL_OK;
}
int tafl_tile_from_str(const char* s, uint8_t* row, uint8_t* col) {
    if (!s || !row || !col) return TAFL_ERR_INVALID_ARG;
    int r, c;
    int rc = parse_tile(s, &r, &c);
    if (rc) return rc;
    *row = (uint8_t)r; *col = (uint8_t)c;
    return TAFL_OK;
}
int tafl_tile_to_str(uint8_t row, uint8_t col, char* buf, size_t len) { return copy_out(show_tile(row, col), buf, len); }
// play.rs:70-86
int tafl_play_from_str(const char* s, tafl_play* out) {
    if (!s || !out) return TAFL_ERR_INVALID_ARG;
    std::string str(s);
    size_t d = str.find('-');
    if (d == std::string::npos || str.find('-', d + 1) != std::string::npos) return TAFL_ERR_PARSE_BAD_STRING;
    int r1, c1, r2, c2;
    int rc = parse_tile(str.substr(0, d), &r1, &c1);
    if (rc) return rc;
    rc = parse_tile(str.substr(d + 1), &r2, &c2);
    if (rc) return rc;
    return tafl_play_from_tiles((uint8_t)r1, (uint8_t)c1, (uint8_t)r2, (uint8_t)c2, out);
}
// play.rs:88-92
int tafl_play_to_str(const tafl_play* p, char* buf, size_t len) {
    if (!p) return TAFL_ERR_INVALID_ARG;
    int tr = p->from_row, tc = p->from_col;
    if (p->axis == TAFL_HORIZONTAL) tc = (uint8_t)(int8_t)((int8_t)tc + p->disp);
    else tr = (uint8_t)(int8_t)((int8_t)tr + p->disp);
    return copy_out(show_tile(p->from_row, p->from_col) + "-" + show_tile(tr, tc), buf, len);
}
int tafl_move_to_play(tafl_move m, int width_bits, tafl_play* out) {
    Geo g;
    if (!out || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    unsigned f = TAFL_MOVE_FROM(m), t = TAFL_MOVE_TO(m);
    return tafl_play_from_tiles((uint8_t)(f / g.stride), (uint8_t)(f % g.stride), (uint8_t)(t / g.stride),
                                (uint8_t)(t % g.stride), out);
}

// game/state.rs:71-113 (same arithmetic as the device-side track_play in tafl_core.cuh)
int tafl_track_play(tafl_state* s, int side, const tafl_play* p, int captures) {
    if (!s || !p) return TAFL_ERR_INVALID_ARG;
    const bool def_side = side == TAFL_DEFENDER;
    uint32_t fl = TAFL_REC_SOME | (captures ? TAFL_REC_CAPTURES : 0u) | (def_side ? TAFL_REC_DEFENDER : 0u) |
                  (p->axis == TAFL_HORIZONTAL ? TAFL_REC_HORIZONTAL : 0u);
    uint32_t rec = (uint32_t)p->from_row | ((uint32_t)p->from_col << 8) | ((uint32_t)(uint8_t)p->disp << 16) | (fl << 24);
    unsigned head = s->ring_head & 3u;
    const uint8_t mid_bit = def_side ? TAFL_FLAG_DEF_MID : TAFL_FLAG_ATT_MID;
    uint8_t& reps = def_side ? s->def_reps : s->att_reps;
    if (!captures && s->recent[head] == rec) {
        bool is_rep = !(s->flags & mid_bit);
        s->flags ^= mid_bit;
        if (is_rep && reps < 255) ++reps;
    } else {
        reps = 0;
        s->flags &= (uint8_t)~mid_bit;
    }
    s->recent[head] = rec;
    s->ring_head = (uint8_t)((head + 1u) & 3u);
    return TAFL_OK;
}

// game/main.rs:55-83
int tafl_board_to_matrix(const tafl_state* st, int width_bits, uint8_t* out, size_t out_len) {
    Geo g;
    if (!st || !out || !geo_for(width_bits, &g)) return TAFL_ERR_INVALID_ARG;
    const int n = st->side_len;
    if (n == 0 || out_len < (size_t)n * n) return TAFL_ERR_INVALID_ARG;
    std::memset(out, 0, (size_t)n * n);
    out[0] = 20; out[n - 1] = 20; out[(size_t)(n - 1) * n] = 20; out[(size_t)(n - 1) * n + n - 1] = 20;
    out[(size_t)(n / 2) * n + n / 2] = 30;
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c) {
            int type, side;
            if (piece_get(st, g, r, c, &type, &side)) out[(size_t)r * n + c] += (type == TAFL_KING ? 5 : 1);
        }
    return TAFL_OK;
}

}  // extern "C"
