// tafl_oracle.hpp — CPU oracle: a tile-level C++17 restatement of the reference `hnefatafl`
// crate (reference tree: game/*.rs).
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may build, load or call anything in oracle/.
// The product (libtafl_b200.so) never links or calls it.
//
// It deliberately keeps the reference's data structures and control flow (per-tile loops,
// per-call containers, the same order of checks, the same quirks) rather than being a fast
// bitboard engine: it is both the parity oracle and the timed CPU baseline ("port").
// Parity pin: the reference's own 25 unit tests, carried over verbatim in oracle/kat_main.cpp
// and tests/golden/ref_kats.json (the Rust crate itself cannot be built here: no cargo/rustc).
//
// Every function cites the reference file:line it follows.
#pragma once
#include <array>
#include <cstdint>
#include <cstdlib>
#include <optional>
#include <set>
#include <string>
#include <tuple>
#include <vector>

#include "../include/tafl.h"

namespace orc {

using u128 = unsigned __int128;

// ---------------------------------------------------------------------------------------------
// tiles.rs
// ---------------------------------------------------------------------------------------------
struct Tile {  // tiles.rs:108-112
    uint8_t row = 0, col = 0;
    bool operator==(const Tile& o) const { return row == o.row && col == o.col; }
    bool operator!=(const Tile& o) const { return !(*this == o); }
    bool operator<(const Tile& o) const { return row != o.row ? row < o.row : col < o.col; }
};
struct Coords {  // tiles.rs:53-57
    int8_t row, col;
};
enum class Axis : uint8_t { Vertical = 0, Horizontal = 0x80 };  // tiles.rs:167-170
inline Axis other(Axis a) { return a == Axis::Vertical ? Axis::Horizontal : Axis::Vertical; }

// ---------------------------------------------------------------------------------------------
// pieces.rs
// ---------------------------------------------------------------------------------------------
enum class Side : uint8_t { Attacker = 0, Defender = 8 };  // pieces.rs:13-17
inline Side other(Side s) { return s == Side::Attacker ? Side::Defender : Side::Attacker; }
enum class PieceType : uint8_t {  // pieces.rs:31-38
    King = 1, Soldier = 2, Knight = 4, Commander = 8, Guard = 16, Mercenary = 32
};
struct Piece {  // pieces.rs:64-67
    PieceType type;
    Side side;
    bool operator==(const Piece& o) const { return type == o.type && side == o.side; }
};
static const Piece KING{PieceType::King, Side::Defender};  // pieces.rs:9

struct PieceSet {  // pieces.rs:157
    uint16_t v = 0;
    // pieces.rs:269-271 via get_mask :240-246
    bool contains(Piece p) const {
        return (v & (uint16_t)((uint16_t)p.type << (uint16_t)p.side)) > 0;
    }
    static PieceSet of_side(Side s) { return PieceSet{(uint16_t)(0x00FFu << (uint16_t)s)}; }  // :204-208
    static PieceSet of_type(PieceType t) {  // :226-228
        return PieceSet{(uint16_t)((uint16_t)t | ((uint16_t)t << 8))};
    }
    static PieceSet of_piece(Piece p) { return PieceSet{(uint16_t)((uint16_t)p.type << (uint16_t)p.side)}; }
};

struct PlacedPiece {  // pieces.rs:145-148
    Tile tile;
    Piece piece;
    bool operator<(const PlacedPiece& o) const { return tile < o.tile; }
    bool operator==(const PlacedPiece& o) const { return tile == o.tile && piece == o.piece; }
};

// error.rs:50-71 (codes: include/tafl.h)
using PlayInvalid = int;

// ---------------------------------------------------------------------------------------------
// play.rs
// ---------------------------------------------------------------------------------------------
struct Play {  // play.rs:23-27
    Tile from;
    Axis axis = Axis::Vertical;
    int8_t disp = 0;
    bool operator==(const Play& o) const { return from == o.from && axis == o.axis && disp == o.disp; }
    Coords to_coords() const {  // play.rs:65-67 + tiles.rs:92-99
        Coords c{(int8_t)from.row, (int8_t)from.col};
        if (axis == Axis::Vertical) c.row = (int8_t)(c.row + disp);
        else c.col = (int8_t)(c.col + disp);
        return c;
    }
    Tile to() const {  // play.rs:59-62 (unchecked i8 -> u8 cast)
        Coords c = to_coords();
        return Tile{(uint8_t)c.row, (uint8_t)c.col};
    }
    uint8_t distance() const { return (uint8_t)std::abs((int)disp); }  // play.rs:53-55
    static std::optional<Play> from_tiles(Tile s, Tile d) {  // play.rs:36-49
        Play p;
        p.from = s;
        if (s.row == d.row) { p.axis = Axis::Horizontal; p.disp = (int8_t)((int8_t)d.col - (int8_t)s.col); }
        else if (s.col == d.col) { p.axis = Axis::Vertical; p.disp = (int8_t)((int8_t)d.row - (int8_t)s.row); }
        else return std::nullopt;
        return p;
    }
};

// ---------------------------------------------------------------------------------------------
// bitfield.rs + board/state.rs
// ---------------------------------------------------------------------------------------------
// primitive_types::U256 / U512 (a dependency of the reference that is not in its tree; only the operations the
// board code uses: shifts, bit logic, comparison with zero, truncation): NW little-endian 64-bit words.
template <int NW>
struct BigU {
    uint64_t w[NW];
    BigU() { for (auto& x : w) x = 0; }
    BigU(uint64_t v) { for (auto& x : w) x = 0; w[0] = v; }   // NOLINT: the board code writes `W x = 0`, `(W)1 << sh`
    explicit operator uint64_t() const { return w[0]; }
    explicit operator uint8_t() const { return (uint8_t)w[0]; }
    explicit operator bool() const { for (auto x : w) if (x) return true; return false; }
    BigU operator<<(unsigned k) const {
        BigU r;
        if (k >= 64u * NW) return r;
        const unsigned ws = k / 64, bs = k % 64;
        for (int i = NW - 1; i >= (int)ws; --i) {
            r.w[i] = w[i - ws] << bs;
            if (bs && i - (int)ws - 1 >= 0) r.w[i] |= w[i - ws - 1] >> (64 - bs);
        }
        return r;
    }
    BigU operator>>(unsigned k) const {
        BigU r;
        if (k >= 64u * NW) return r;
        const unsigned ws = k / 64, bs = k % 64;
        for (int i = 0; i + (int)ws < NW; ++i) {
            r.w[i] = w[i + ws] >> bs;
            if (bs && i + (int)ws + 1 < NW) r.w[i] |= w[i + ws + 1] << (64 - bs);
        }
        return r;
    }
    BigU operator&(const BigU& o) const { BigU r; for (int i = 0; i < NW; ++i) r.w[i] = w[i] & o.w[i]; return r; }
    BigU operator|(const BigU& o) const { BigU r; for (int i = 0; i < NW; ++i) r.w[i] = w[i] | o.w[i]; return r; }
    BigU operator~() const { BigU r; for (int i = 0; i < NW; ++i) r.w[i] = ~w[i]; return r; }
    BigU& operator&=(const BigU& o) { for (int i = 0; i < NW; ++i) w[i] &= o.w[i]; return *this; }
    BigU& operator|=(const BigU& o) { for (int i = 0; i < NW; ++i) w[i] |= o.w[i]; return *this; }
    bool operator==(const BigU& o) const { for (int i = 0; i < NW; ++i) if (w[i] != o.w[i]) return false; return true; }
    bool operator!=(const BigU& o) const { return !(*this == o); }
    bool operator>(int z) const { (void)z; return (bool)*this; }   // only ever `> 0`
};
using U256 = BigU<4>;
using U512 = BigU<8>;

template <class W> struct BitTraits;
template <> struct BitTraits<uint64_t> { static constexpr int ROW_WIDTH = 7, BITS = 64; };   // bitfield.rs:178
template <> struct BitTraits<u128> { static constexpr int ROW_WIDTH = 11, BITS = 128; };     // bitfield.rs:179
template <> struct BitTraits<U256> { static constexpr int ROW_WIDTH = 15, BITS = 256; };     // bitfield.rs:180
template <> struct BitTraits<U512> { static constexpr int ROW_WIDTH = 21, BITS = 512; };     // bitfield.rs:181
inline int popcount_w(uint64_t w) { return __builtin_popcountll(w); }
inline int popcount_w(u128 w) { return __builtin_popcountll((uint64_t)w) + __builtin_popcountll((uint64_t)(w >> 64)); }
template <int NW> inline int popcount_w(const BigU<NW>& w) { int c = 0; for (auto x : w.w) c += __builtin_popcountll(x); return c; }

template <class W>
struct Board {  // board/state.rs:117-121
    W attackers = 0, defenders = 0;
    uint8_t side_len = 0;
    static constexpr int RW = BitTraits<W>::ROW_WIDTH;
    static constexpr int BITS = BitTraits<W>::BITS;

    bool operator==(const Board& o) const {
        return attackers == o.attackers && defenders == o.defenders && side_len == o.side_len;
    }
    static W tile_mask(Tile t) {  // bitfield.rs:72-74 (shift >= width: empty mask instead of a panic)
        unsigned sh = (unsigned)t.row * RW + t.col;
        if (sh >= (unsigned)BITS) return 0;
        return (W)1 << sh;
    }
    static Tile bit_to_tile(uint32_t bit) {  // bitfield.rs:77-81
        uint32_t row = bit / RW;
        return Tile{(uint8_t)row, (uint8_t)(bit - row * RW)};
    }
    Tile get_king() const {  // board/state.rs:127-131
        uint8_t row = (uint8_t)((defenders >> (BITS - 4)) & 0xF);
        uint8_t col = (uint8_t)((attackers >> (BITS - 4)) & 0xF);
        return Tile{row, col};
    }
    void set_king(Tile t) {  // board/state.rs:136-147 (u8 `<< 4` drops the high bits)
        W nib = ((W)0xF) << (BITS - 4);
        defenders = (defenders & ~nib) | ((W)((uint8_t)(t.row << 4) >> 4) << (BITS - 4));
        attackers = (attackers & ~nib) | ((W)((uint8_t)(t.col << 4) >> 4) << (BITS - 4));
    }
    bool is_king(Tile t) const { return get_king() == t; }  // board/state.rs:24-26
    void set_piece(Tile t, Piece p) {  // board/state.rs:149-165
        W m = tile_mask(t);
        if (p.side == Side::Attacker) { attackers |= m; defenders &= ~m; }
        else { defenders |= m; attackers &= ~m; }
        if (p.type == PieceType::King) set_king(t);
    }
    void clear_tile(Tile t) {  // board/state.rs:167-171
        W m = ~tile_mask(t);
        attackers &= m;
        defenders &= m;
    }
    std::optional<Piece> get_piece(Tile t) const {  // board/state.rs:173-187
        W m = tile_mask(t);
        if ((defenders & m) > 0) {
            if (is_king(t)) return KING;
            return Piece{PieceType::Soldier, Side::Defender};
        } else if ((attackers & m) > 0) {
            return Piece{PieceType::Soldier, Side::Attacker};
        }
        return std::nullopt;
    }
    bool tile_occupied(Tile t) const {  // board/state.rs:189-193
        return ((defenders | attackers) & tile_mask(t)) > 0;
    }
    static int popcount(W w) { return popcount_w(w); }
    uint8_t count_pieces(Side s) const {  // board/state.rs:195-200
        W w = (s == Side::Attacker ? attackers : defenders);
        return (uint8_t)popcount((W)(w << 4));
    }
    std::vector<Tile> iter_occupied(Side s) const {  // board/state.rs:202-216 + :91-102 (ascending bit)
        W w = (s == Side::Attacker ? attackers : defenders);
        w &= ~(((W)0xF) << (BITS - 4));
        std::vector<Tile> out;
        for (uint32_t b = 0; b < (uint32_t)BITS; ++b)
            if ((w >> b) & 1) out.push_back(bit_to_tile(b));
        return out;
    }
    Piece move_piece(Tile from, Tile to) {  // board/state.rs:218-223 (caller guarantees a piece)
        Piece p = *get_piece(from);
        set_piece(to, p);
        clear_tile(from);
        return p;
    }
    // pieces.rs:119-143  TryFrom<char> for Piece
    static int piece_from_char(char ch, Piece* out) {
        bool alpha = (ch >= 'a' && ch <= 'z') || (ch >= 'A' && ch <= 'Z');
        if (!alpha) return TAFL_ERR_PARSE_BAD_CHAR;
        Side side = Side::Attacker;
        if (ch >= 'A' && ch <= 'Z') { ch = (char)(ch - 'A' + 'a'); side = Side::Defender; }
        PieceType t;
        switch (ch) {
            case 't': t = PieceType::Soldier; break;
            case 'k': t = PieceType::King; break;
            case 'n': t = PieceType::Knight; break;
            case 'c': t = PieceType::Commander; break;
            case 'g': t = PieceType::Guard; break;
            case 'm': t = PieceType::Mercenary; break;
            default: return TAFL_ERR_PARSE_BAD_CHAR;
        }
        *out = Piece{t, side};
        return 0;
    }
    // board/state.rs:225-250
    static int from_fen(const std::string& fen, Board* out) {
        Board st;
        size_t pos = 0;
        unsigned r = 0;
        while (true) {
            size_t slash = fen.find('/', pos);
            std::string line = fen.substr(pos, slash == std::string::npos ? std::string::npos : slash - pos);
            uint8_t n_empty = 0, c = 0;
            for (char ch : line) {
                if (ch >= '0' && ch <= '9') {
                    n_empty = (uint8_t)(n_empty * 10 + (ch - '0'));
                } else {
                    c = (uint8_t)(c + n_empty);
                    n_empty = 0;
                    Piece p{PieceType::Soldier, Side::Attacker};
                    int rc = piece_from_char(ch, &p);
                    if (rc) return rc;
                    st.set_piece(Tile{(uint8_t)r, c}, p);
                    c = (uint8_t)(c + 1);
                }
            }
            if (n_empty > 0) c = (uint8_t)(c + n_empty);
            if (st.side_len == 0) st.side_len = c;
            else if (st.side_len != c) return TAFL_ERR_PARSE_BAD_LINE_LEN;
            if (slash == std::string::npos) break;
            pos = slash + 1;
            ++r;
        }
        *out = st;
        return 0;
    }
    // board/state.rs:252-269
    static int from_display_str(const std::string& s_in, Board* out) {
        size_t a = s_in.find_first_not_of(" \t\r\n"), b = s_in.find_last_not_of(" \t\r\n");
        std::string s = (a == std::string::npos) ? "" : s_in.substr(a, b - a + 1);
        Board st;
        size_t pos = 0;
        unsigned r = 0;
        while (pos <= s.size()) {
            size_t nl = s.find('\n', pos);
            std::string line = s.substr(pos, nl == std::string::npos ? std::string::npos : nl - pos);
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (s.empty()) break;
            uint8_t len = (uint8_t)line.size();
            if (st.side_len == 0) st.side_len = len;
            else if (len != st.side_len) return TAFL_ERR_PARSE_BAD_LINE_LEN;
            for (size_t c = 0; c < line.size(); ++c) {
                if (line[c] != '.') {
                    Piece p{PieceType::Soldier, Side::Attacker};
                    int rc = piece_from_char(line[c], &p);
                    if (rc) return rc;
                    st.set_piece(Tile{(uint8_t)r, (uint8_t)c}, p);
                }
            }
            if (nl == std::string::npos) break;
            pos = nl + 1;
            ++r;
        }
        *out = st;
        return 0;
    }
    static char piece_char(Piece p) {  // pieces.rs:100-117
        char c = 't';
        switch (p.type) {
            case PieceType::Soldier: c = 't'; break;
            case PieceType::King: c = 'k'; break;
            case PieceType::Knight: c = 'n'; break;
            case PieceType::Commander: c = 'c'; break;
            case PieceType::Guard: c = 'g'; break;
            case PieceType::Mercenary: c = 'm'; break;
        }
        return p.side == Side::Attacker ? c : (char)(c - 'a' + 'A');
    }
    std::string to_fen() const {  // board/state.rs:271-295
        std::string s;
        for (uint8_t row = 0; row < side_len; ++row) {
            int n_empty = 0;
            for (uint8_t col = 0; col < side_len; ++col) {
                auto p = get_piece(Tile{row, col});
                if (p) {
                    if (n_empty > 0) { s += std::to_string(n_empty); n_empty = 0; }
                    s.push_back(piece_char(*p));
                } else {
                    ++n_empty;
                }
            }
            if (n_empty > 0) s += std::to_string(n_empty);
            if (row < side_len - 1) s.push_back('/');
        }
        return s;
    }
    std::string to_display_str() const {  // board/state.rs:297-311
        std::string s;
        for (uint8_t r = 0; r < side_len; ++r) {
            for (uint8_t c = 0; c < side_len; ++c) {
                auto p = get_piece(Tile{r, c});
                s.push_back(p ? piece_char(*p) : '.');
            }
            s.push_back('\n');
        }
        return s;
    }
    void swap_pieces(Tile t1, Tile t2) {  // board/state.rs:69-79
        auto p1 = get_piece(t1);
        auto p2 = get_piece(t2);
        if (p2) set_piece(t1, *p2); else clear_tile(t1);
        if (p1) set_piece(t2, *p1); else clear_tile(t2);
    }
};

// ---------------------------------------------------------------------------------------------
// board/geometry.rs
// ---------------------------------------------------------------------------------------------
struct BoardGeometry {  // geometry.rs:30-33
    uint8_t side_len = 0;
    Tile throne;
    std::array<Tile, 4> corners;
    explicit BoardGeometry(uint8_t n = 0) : side_len(n) {  // geometry.rs:13-24
        corners = {Tile{0, 0}, Tile{0, (uint8_t)(n - 1)}, Tile{(uint8_t)(n - 1), (uint8_t)(n - 1)},
                   Tile{(uint8_t)(n - 1), 0}};
        throne = Tile{(uint8_t)(n / 2), (uint8_t)(n / 2)};
    }
    bool is_corner(Tile t) const {
        for (auto& c : corners) if (c == t) return true;
        return false;
    }
    bool tile_in_bounds(Tile t) const { return t.row < side_len && t.col < side_len; }  // :45-48
    bool coords_in_bounds(Coords c) const {  // :62-65
        return c.row >= 0 && c.row < (int8_t)side_len && c.col >= 0 && c.col < (int8_t)side_len;
    }
    std::optional<Tile> coords_to_tile(Coords c) const {  // :53-59
        if (coords_in_bounds(c)) return Tile{(uint8_t)c.row, (uint8_t)c.col};
        return std::nullopt;
    }
    std::vector<Tile> neighbors(Tile t) const {  // :5, :68-81  order: up, down, left, right
        static const int8_t OFF[4][2] = {{-1, 0}, {1, 0}, {0, -1}, {0, 1}};
        std::vector<Tile> out;
        for (auto& o : OFF) {
            Coords c{(int8_t)((int8_t)t.row + o[0]), (int8_t)((int8_t)t.col + o[1])};
            if (auto tt = coords_to_tile(c)) out.push_back(*tt);
        }
        return out;
    }
    std::vector<Tile> tiles_between(Tile t1, Tile t2) const {  // :85-108
        std::vector<Tile> out;
        if (t1.row == t2.row) {
            uint8_t lo = t1.col > t2.col ? t2.col : t1.col, hi = t1.col > t2.col ? t1.col : t2.col;
            for (unsigned c = (unsigned)lo + 1; c < hi; ++c) out.push_back(Tile{t1.row, (uint8_t)c});
        } else if (t1.col == t2.col) {
            uint8_t lo = t1.row > t2.row ? t2.row : t1.row, hi = t1.row > t2.row ? t1.row : t2.row;
            for (unsigned r = (unsigned)lo + 1; r < hi; ++r) out.push_back(Tile{(uint8_t)r, t1.col});
        }
        return out;
    }
    bool tile_at_edge(Tile t) const {  // :111-116
        return t.row == 0 || t.row == side_len - 1 || t.col == 0 || t.col == side_len - 1;
    }
};

// ---------------------------------------------------------------------------------------------
// rules.rs (POD = tafl_rules from include/tafl.h) — typed accessors
// ---------------------------------------------------------------------------------------------
struct Ruleset {
    tafl_rules r;
    PieceSet may_enter_corners() const { return PieceSet{r.may_enter_corners}; }
    PieceSet host_throne() const { return PieceSet{r.hostility_throne}; }
    PieceSet host_corners() const { return PieceSet{r.hostility_corners}; }
    PieceSet host_edge() const { return PieceSet{r.hostility_edge}; }
    PieceSet slow_pieces() const { return PieceSet{r.slow_pieces}; }
    PieceSet sw_captures() const { return PieceSet{r.sw_captures}; }
    Side starting_side() const { return r.starting_side ? Side::Defender : Side::Attacker; }
};

// ---------------------------------------------------------------------------------------------
// game/mod.rs:17-70 — status values use the packed byte of include/tafl.h
// ---------------------------------------------------------------------------------------------
using GameStatus = uint8_t;
using GameOutcome = std::optional<uint8_t>;  // packed "Over" byte
inline uint8_t Win(int reason, Side s) { return TAFL_STATUS_WIN(reason, s == Side::Defender); }
inline uint8_t Draw(int reason) { return TAFL_STATUS_DRAW(reason); }

// ---------------------------------------------------------------------------------------------
// game/state.rs
// ---------------------------------------------------------------------------------------------
struct ShortPlayRecord {  // game/state.rs:15-19
    Side side;
    Play play;
    bool captures;
    bool operator==(const ShortPlayRecord& o) const {
        return side == o.side && play == o.play && captures == o.captures;
    }
};

struct RepetitionTracker {  // game/state.rs:44-50
    size_t attacker_reps = 0, defender_reps = 0;
    bool attacker_mid_pair = false, defender_mid_pair = false;
    // utils.rs:30-81  FixedSizeQueue<Option<ShortPlayRecord>, 4>
    std::array<std::optional<ShortPlayRecord>, 4> queue{};
    size_t first_i = 0;

    bool is_mid_pair(Side s) const { return s == Side::Attacker ? attacker_mid_pair : defender_mid_pair; }
    void toggle_mid_pair(Side s) {
        if (s == Side::Attacker) attacker_mid_pair = !attacker_mid_pair;
        else defender_mid_pair = !defender_mid_pair;
    }
    std::pair<bool, bool> check_repetition(const ShortPlayRecord& rec) {  // :71-80
        const auto& first = queue[first_i];  // utils.rs:66-68
        if (!rec.captures && first.has_value() && *first == rec) {
            bool is_rep = !is_mid_pair(rec.side);
            toggle_mid_pair(rec.side);
            return {is_rep, false};
        }
        return {false, true};
    }
    size_t get_repetitions(Side s) const { return s == Side::Attacker ? attacker_reps : defender_reps; }  // :83-88
    void track_play(Side side, Play play, bool captures) {  // :92-113
        ShortPlayRecord rec{side, play, captures};
        auto [incr, reset] = check_repetition(rec);
        if (incr) {
            if (side == Side::Attacker) attacker_reps += 1; else defender_reps += 1;
        } else if (reset) {
            if (side == Side::Attacker) { attacker_reps = 0; attacker_mid_pair = false; }
            else { defender_reps = 0; defender_mid_pair = false; }
        }
        queue[first_i] = rec;  // utils.rs:57-64
        first_i = (first_i == 3) ? 0 : first_i + 1;
    }
};

template <class W>
struct GameState {  // game/state.rs:120-133
    Board<W> board;
    Side side_to_play = Side::Attacker;
    RepetitionTracker repetitions;
    size_t plays_since_capture = 0;
    GameStatus status = TAFL_STATUS_ONGOING;
    size_t turn = 0;
    static int create(const std::string& fen, Side side, GameState* out) {  // :136-145
        GameState s;
        int rc = Board<W>::from_fen(fen, &s.board);
        if (rc) return rc;
        s.side_to_play = side;
        *out = s;
        return 0;
    }
};

struct Enclosure {  // logic.rs:25-38
    std::set<Tile> occupied, unoccupied, boundary;
    bool contains(Tile t) const { return occupied.count(t) || unoccupied.count(t); }
};

struct PlayEffects {  // game/mod.rs:56-61
    std::set<PlacedPiece> captures;
    GameOutcome game_outcome;
};
struct PlayRecord {  // play.rs:106-113
    Side side;
    Play play;
    PlayEffects effects;
};

// ---------------------------------------------------------------------------------------------
// game/logic.rs
// ---------------------------------------------------------------------------------------------
struct GameLogic {  // logic.rs:61-72
    Ruleset rules;
    BoardGeometry geo;
    GameLogic() = default;
    GameLogic(const tafl_rules& r, uint8_t board_length) : rules{r}, geo(board_length) {}

    bool special_tile_hostile(Tile tile, Piece piece) const {  // logic.rs:76-82
        return (rules.host_throne().contains(piece) && tile == geo.throne) ||
               (rules.host_corners().contains(piece) && geo.is_corner(tile)) ||
               (rules.host_edge().contains(piece) && !geo.tile_in_bounds(tile));
    }
    template <class W>
    bool tile_hostile(Tile tile, Piece piece, const Board<W>& board) const {  // logic.rs:85-99
        if (auto other_piece = board.get_piece(tile)) {
            return (other_piece->side != piece.side) &&
                   (other_piece->type != PieceType::King || rules.r.king_attack == TAFL_KING_ARMED ||
                    rules.r.king_attack == TAFL_KING_ANVIL);
        }
        return special_tile_hostile(tile, piece);
    }
    template <class W>
    bool coords_hostile(Coords c, Piece piece, const Board<W>& board) const {  // logic.rs:103-114
        if (geo.coords_in_bounds(c)) return tile_hostile(Tile{(uint8_t)c.row, (uint8_t)c.col}, piece, board);
        return rules.host_edge().contains(piece);
    }

    // logic.rs:159-214 — returns 0 (Ok) or a PlayInvalid code
    template <class W>
    PlayInvalid validate_play_for_side(Play play, Side side, const GameState<W>& state) const {
        if (state.status != TAFL_STATUS_ONGOING) return TAFL_GAME_OVER;
        Tile from = play.from, to = play.to();
        auto maybe_piece = state.board.get_piece(from);
        if (!maybe_piece) return TAFL_NO_PIECE;
        Piece piece = *maybe_piece;
        if (piece.side != side) return TAFL_WRONG_PLAYER;
        if (!(geo.tile_in_bounds(from) && geo.tile_in_bounds(to))) return TAFL_OUT_OF_BOUNDS;
        if (from.row != to.row && from.col != to.col) return TAFL_NO_COMMON_AXIS;
        if (state.board.tile_occupied(to)) return TAFL_BLOCKED_BY_PIECE;
        auto between = geo.tiles_between(from, to);
        for (auto& t : between) if (state.board.tile_occupied(t)) return TAFL_BLOCKED_BY_PIECE;
        if (!rules.may_enter_corners().contains(piece) && geo.is_corner(to)) return TAFL_MOVE_ONTO_BLOCKED_TILE;
        const int tm = rules.r.throne_movement;
        bool between_has_throne = false;
        for (auto& t : between) if (t == geo.throne) between_has_throne = true;
        if ((tm == TAFL_THRONE_NOPASS || (tm == TAFL_THRONE_KINGPASS && piece.type != PieceType::King)) &&
            between_has_throne)
            return TAFL_MOVE_THROUGH_BLOCKED_TILE;
        if ((tm == TAFL_THRONE_NOENTRY || (tm == TAFL_THRONE_KINGENTRY && piece.type != PieceType::King)) &&
            to == geo.throne)
            return TAFL_MOVE_ONTO_BLOCKED_TILE;
        if (rules.slow_pieces().contains(piece) && play.distance() > 1) return TAFL_TOO_FAR;
        return TAFL_PLAY_OK;
    }
    template <class W>
    PlayInvalid validate_play(Play play, const GameState<W>& state) const {  // logic.rs:219-222
        return validate_play_for_side(play, state.side_to_play, state);
    }
    // logic.rs:119-154
    template <class W>
    std::pair<bool, bool> can_occupy_or_pass(Play play, Piece piece, const GameState<W>& state) const {
        PlayInvalid v = validate_play_for_side(play, piece.side, state);
        bool can_occupy = (v == TAFL_PLAY_OK);
        bool can_pass;
        if (v == TAFL_PLAY_OK) can_pass = true;
        else if (v == TAFL_MOVE_ONTO_BLOCKED_TILE) {
            if (play.to() == geo.throne) {
                switch (rules.r.throne_movement) {
                    case TAFL_THRONE_NOTHRONE: can_pass = true; break;
                    case TAFL_THRONE_NOPASS: can_pass = false; break;
                    case TAFL_THRONE_NOENTRY: can_pass = true; break;
                    case TAFL_THRONE_KINGPASS: can_pass = (piece.type == PieceType::King); break;
                    default: can_pass = true; break;  // KingEntry
                }
            } else can_pass = false;
        } else can_pass = false;
        return {can_occupy, can_pass};
    }

    template <class W> bool king_beside_throne(const Board<W>& b) const {  // logic.rs:225-227
        Tile k = b.get_king();
        for (auto& t : geo.neighbors(geo.throne)) if (t == k) return true;
        return false;
    }
    template <class W> bool king_on_throne(const Board<W>& b) const { return b.get_king() == geo.throne; }  // :230-232
    template <class W> bool king_is_strong(const Board<W>& b) const {  // logic.rs:237-245
        switch (rules.r.king_strength) {
            case TAFL_KING_STRONG: return true;
            case TAFL_KING_WEAK: return false;
            default: return king_beside_throne(b) || king_on_throne(b);
        }
    }
    bool coords_occupiable(Coords c, Piece piece) const {  // logic.rs:250-266
        if (!geo.coords_in_bounds(c)) return false;
        Tile t{(uint8_t)c.row, (uint8_t)c.col};
        const int tm = rules.r.throne_movement;
        if (geo.throne == t &&
            (tm == TAFL_THRONE_NOENTRY || (tm == TAFL_THRONE_KINGENTRY && piece.type != PieceType::King)))
            return false;
        if (!rules.may_enter_corners().contains(piece) && geo.is_corner(t)) return false;
        return true;
    }

    // logic.rs:270-301 — nullopt == `None` (abort the whole search)
    template <class W>
    std::optional<bool> row_col_enclosed(int8_t row, int8_t col, PieceSet enclosed, PieceSet enclosing,
                                         Enclosure& e, const Board<W>& board) const {
        if (auto tile = geo.coords_to_tile(Coords{row, col})) {
            if (auto p = board.get_piece(*tile)) {
                if (enclosed.contains(*p)) { e.occupied.insert(*tile); return true; }
                else if (enclosing.contains(*p)) { e.boundary.insert(*tile); return false; }
                return std::nullopt;
            }
            e.unoccupied.insert(*tile);
            return true;
        }
        return false;
    }

    // logic.rs:309-401 — literal transcription of the span fill incl. the UniqueStack (utils.rs:9-25)
    template <class W>
    std::optional<Enclosure> find_enclosure(Tile tile, PieceSet enclosed, PieceSet enclosing, bool abort_on_edge,
                                            bool abort_on_corner, const Board<W>& board) const {
        int8_t row = (int8_t)tile.row, col = (int8_t)tile.col;
        Enclosure enc;
#define ORC_RCE(R, C, DST)                                                     \
    {                                                                          \
        auto _v = row_col_enclosed((int8_t)(R), (int8_t)(C), enclosed, enclosing, enc, board); \
        if (!_v) return std::nullopt;                                          \
        DST = *_v;                                                             \
    }
        bool ok;
        ORC_RCE(row, col, ok);
        if (!ok) return std::nullopt;
        using Item = std::tuple<int8_t, int8_t, int8_t, int8_t>;
        std::vector<Item> stack;
        std::set<Item> added;
        auto push = [&](Item it) {
            if (!added.count(it)) { stack.push_back(it); added.insert(it); }
        };
        push(Item{col, col, row, 1});
        push(Item{col, col, (int8_t)(row - 1), -1});
        while (!stack.empty()) {
            auto [c1, c2, r, dr] = stack.back();
            stack.pop_back();
            int8_t c = c1;
            bool in;
            ORC_RCE(r, c, in);
            if (in) {
                while (true) {
                    bool left;
                    ORC_RCE(r, c - 1, left);
                    if (!left) break;
                    Tile t{(uint8_t)r, (uint8_t)(c - 1)};
                    if ((abort_on_edge && geo.tile_at_edge(t)) || (abort_on_corner && geo.is_corner(t)))
                        return std::nullopt;
                    c = (int8_t)(c - 1);
                }
                if (c < c1) push(Item{c, (int8_t)(c1 - 1), (int8_t)(r - dr), (int8_t)(-dr)});
            }
            while (c1 <= c2) {
                while (true) {
                    bool here;
                    ORC_RCE(r, c1, here);
                    if (!here) break;
                    Tile t{(uint8_t)r, (uint8_t)c1};
                    if (abort_on_edge && geo.tile_at_edge(t)) return std::nullopt;
                    if (abort_on_corner && geo.is_corner(t)) return std::nullopt;
                    c1 = (int8_t)(c1 + 1);
                }
                if (c1 > c) push(Item{c, (int8_t)(c1 - 1), (int8_t)(r + dr), dr});
                if (c1 - 1 > c2) push(Item{(int8_t)(c2 + 1), (int8_t)(c1 - 1), (int8_t)(r - dr), (int8_t)(-dr)});
                c1 = (int8_t)(c1 + 1);
                while (c1 < c2) {
                    bool here;
                    ORC_RCE(r, c1, here);
                    if (here) break;
                    c1 = (int8_t)(c1 + 1);
                }
                c = c1;
            }
        }
#undef ORC_RCE
        return enc;
    }

    // logic.rs:408-463
    template <class W>
    bool enclosure_secure(const Enclosure& encl, bool inside_safe, bool outside_safe, const Board<W>& board) const {
        if (inside_safe && outside_safe) return true;
        for (const Tile& t : encl.boundary) {
            Piece piece = *board.get_piece(t);
            Piece hostile_soldier{PieceType::Soldier, other(piece.side)};
            for (Axis axis : {Axis::Vertical, Axis::Horizontal}) {
                bool axis_ok = false;
                for (int d : {-1, 1}) {
                    Play step{t, axis, (int8_t)d};
                    Coords nc = step.to_coords();
                    if (auto nt = geo.coords_to_tile(nc)) {
                        bool is_inside = encl.contains(*nt);
                        if ((inside_safe && is_inside) || (outside_safe && !is_inside)) {
                            if (!special_tile_hostile(*nt, piece)) { axis_ok = true; break; }
                        }
                        if (!tile_hostile(*nt, piece, board) &&
                            (board.tile_occupied(*nt) || !coords_occupiable(nc, hostile_soldier))) {
                            axis_ok = true;
                            break;
                        }
                    } else {
                        if (!rules.host_edge().contains(piece)) { axis_ok = true; break; }
                    }
                }
                if (!axis_ok) return false;
            }
        }
        return true;
    }

    // logic.rs:471-530
    template <class W>
    std::optional<std::set<Tile>> dir_sw_search(Play play, Axis axis, int8_t away_from_edge, int8_t dir,
                                                const GameState<W>& state) const {
        Tile t = play.to();
        std::set<Tile> wall;
        const bool cmc = rules.r.sw_corners_may_close != 0;
        while (true) {
            Play step{t, axis, dir};
            t = step.to();
            if (!geo.tile_in_bounds(t)) return std::nullopt;
            if (!(state.board.tile_occupied(t) || (cmc && geo.is_corner(t)))) return std::nullopt;
            auto piece_opt = state.board.get_piece(t);
            if (!piece_opt) {
                if (wall.size() < 2) return std::nullopt;
                return wall;
            }
            Piece piece = *piece_opt;
            if (piece.side == other(state.side_to_play)) {
                Tile pin = Play{t, other(axis), away_from_edge}.to();
                if (auto p = state.board.get_piece(pin)) {
                    if (p->side == state.side_to_play) wall.insert(t);
                    else return std::nullopt;
                } else return std::nullopt;
            }
            if (piece.side == state.side_to_play || (geo.is_corner(t) && cmc)) {
                if (wall.size() < 2) return std::nullopt;
                return wall;
            }
        }
    }
    // logic.rs:535-569
    template <class W>
    std::optional<std::set<Tile>> detect_shieldwall(Play play, const GameState<W>& state) const {
        if (!rules.r.shieldwall_enabled) return std::nullopt;
        Tile to = play.to();
        Axis axis;
        int8_t away;
        if (to.row == 0) { axis = Axis::Horizontal; away = 1; }
        else if (to.row == geo.side_len - 1) { axis = Axis::Horizontal; away = -1; }
        else if (to.col == 0) { axis = Axis::Vertical; away = 1; }
        else if (to.col == geo.side_len - 1) { axis = Axis::Vertical; away = -1; }
        else return std::nullopt;
        auto wall = dir_sw_search(play, axis, away, -1, state);
        if (!wall) wall = dir_sw_search(play, axis, away, 1, state);
        if (!wall) return std::nullopt;
        if (wall->size() < 2) return std::nullopt;
        std::set<Tile> out;
        for (auto& t : *wall)
            if (rules.sw_captures().contains(*state.board.get_piece(t))) out.insert(t);
        return out;
    }

    // logic.rs:572-601
    template <class W>
    bool detect_exit_fort(const Board<W>& board) const {
        Tile king_tile = board.get_king();
        if (!geo.tile_at_edge(king_tile)) return false;
        auto encl = find_enclosure(king_tile, PieceSet::of_type(PieceType::King), PieceSet::of_side(Side::Defender),
                                   false, true, board);
        if (!encl) return false;
        bool any_free = false;
        for (auto& t : geo.neighbors(king_tile)) if (!board.tile_occupied(t)) any_free = true;
        if (!any_free) return false;
        if (!enclosure_secure(*encl, true, false, board)) return false;
        return true;
    }

    // logic.rs:859-879
    template <class W>
    std::optional<PlacedPiece> detect_linnaean_capture(Tile tile, Piece other_piece, Coords far,
                                                       const GameState<W>& state) const {
        if (auto far_tile = geo.coords_to_tile(far)) {
            if (*far_tile == geo.throne && state.board.is_king(*far_tile)) {
                int n = 0;
                for (auto& t : geo.neighbors(*far_tile)) if (tile_hostile(t, KING, state.board)) ++n;
                if (n == 3) return PlacedPiece{tile, other_piece};
            }
        }
        return std::nullopt;
    }

    // logic.rs:604-699
    template <class W>
    std::set<PlacedPiece> get_captures(Play play, Piece moving_piece, const GameState<W>& state) const {
        std::set<PlacedPiece> captures;
        Tile to = play.to();
        if (moving_piece.type != PieceType::King || rules.r.king_attack == TAFL_KING_ARMED ||
            rules.r.king_attack == TAFL_KING_HAMMER) {
            for (auto& n : geo.neighbors(to)) {
                auto op = state.board.get_piece(n);
                if (!op) continue;
                Piece other_piece = *op;
                if (other_piece.side == moving_piece.side) continue;
                if (other_piece.type == PieceType::King && king_beside_throne(state.board) &&
                    rules.r.king_strength == TAFL_KING_STRONG_BY_THRONE &&
                    (rules.r.throne_movement == TAFL_THRONE_NOENTRY || rules.r.throne_movement == TAFL_THRONE_KINGENTRY)) {
                    bool all = true;
                    for (auto& t : geo.neighbors(n))
                        if (!(t == geo.throne || tile_hostile(t, other_piece, state.board))) all = false;
                    if (all) {
                        captures.insert(PlacedPiece{n, other_piece});
                        continue;
                    }
                }
                int8_t tr = (int8_t)to.row, tc = (int8_t)to.col, nr = (int8_t)n.row, nc = (int8_t)n.col;
                Coords far{(int8_t)(tr + (nr - tr) * 2), (int8_t)(tc + (nc - tc) * 2)};
                if (coords_hostile(far, other_piece, state.board)) {
                    if (other_piece.type == PieceType::King && king_is_strong(state.board)) {
                        bool perp;
                        if (to.row == n.row) {
                            perp = coords_hostile(Coords{(int8_t)(nr + 1), nc}, other_piece, state.board) &&
                                   coords_hostile(Coords{(int8_t)(nr - 1), nc}, other_piece, state.board);
                        } else {
                            perp = coords_hostile(Coords{nr, (int8_t)(nc + 1)}, other_piece, state.board) &&
                                   coords_hostile(Coords{nr, (int8_t)(nc - 1)}, other_piece, state.board);
                        }
                        if (!perp) continue;
                    }
                    captures.insert(PlacedPiece{n, other_piece});
                } else if (rules.r.linnaean_capture && state.side_to_play == Side::Attacker) {
                    if (auto pp = detect_linnaean_capture(n, other_piece, far, state)) captures.insert(*pp);
                }
            }
        }
        if (auto walled = detect_shieldwall(play, state)) {
            for (auto& t : *walled) captures.insert(PlacedPiece{t, *state.board.get_piece(t)});
        }
        return captures;
    }

    // play.rs:139-226 — the per-piece legal play iterator, drained into a vector
    template <class W>
    bool iter_plays(Tile tile, const GameState<W>& state, std::vector<Play>* out, bool first_only = false) const {
        auto piece_opt = state.board.get_piece(tile);
        if (!piece_opt) return false;  // Err(BoardError::NoPiece)
        Piece piece = *piece_opt;
        Axis axis = Axis::Vertical;
        int8_t disp = 1;
        auto next_direction = [&]() -> bool {  // play.rs:166-183
            if (axis == Axis::Vertical) {
                if (disp > 0) { disp = -1; } else { axis = Axis::Horizontal; disp = 1; }
                return true;
            }
            if (disp > 0) { disp = -1; return true; }
            return false;
        };
        while (true) {  // play.rs:189-225
            Play play{tile, axis, disp};
            if (auto dest = geo.coords_to_tile(play.to_coords())) {
                disp = (int8_t)(disp + (disp > 0 ? 1 : -1));
                auto [can_occupy, can_pass] = can_occupy_or_pass(play, piece, state);
                if (can_occupy) {
                    out->push_back(*Play::from_tiles(tile, *dest));
                    if (first_only) return true;
                    continue;
                } else if (can_pass) {
                    continue;
                } else {
                    if (!next_direction()) return true;
                    continue;
                }
            } else {
                if (!next_direction()) return true;
                continue;
            }
        }
    }
    // logic.rs:837-846
    template <class W>
    bool side_can_play(Side side, const GameState<W>& state) const {
        for (auto& tile : state.board.iter_occupied(side)) {
            std::vector<Play> tmp;
            iter_plays(tile, state, &tmp, true);
            if (!tmp.empty()) return true;
        }
        return false;
    }
    // game/main.rs:33-43 — the canonical full move list
    template <class W>
    std::vector<Play> get_all_possible_moves(const GameState<W>& state) const {
        std::vector<Play> out;
        for (auto& tile : state.board.iter_occupied(state.side_to_play)) iter_plays(tile, state, &out);
        return out;
    }

    // logic.rs:702-771
    template <class W>
    GameOutcome get_game_outcome(Play play, Piece moving_piece, const std::set<PlacedPiece>& caps,
                                 const GameState<W>& state) const {
        if (state.board.count_pieces(other(state.side_to_play)) == 0)
            return Win(TAFL_WIN_ALL_CAPTURED, state.side_to_play);
        if (state.side_to_play == Side::Attacker) {
            for (auto& c : caps)
                if (state.board.is_king(c.tile)) return Win(TAFL_WIN_KING_CAPTURED, Side::Attacker);
            if (rules.r.enclosure_win != TAFL_ENCL_NONE) {
                auto encl = find_enclosure(state.board.get_king(), PieceSet::of_side(Side::Defender),
                                           PieceSet::of_side(Side::Attacker),
                                           rules.r.enclosure_win == TAFL_ENCL_WITHOUT_EDGE_ACCESS, true, state.board);
                if (encl) {
                    if (encl->occupied.size() == (size_t)state.board.count_pieces(Side::Defender) &&
                        enclosure_secure(*encl, false, true, state.board))
                        return Win(TAFL_WIN_ENCLOSED, Side::Attacker);
                }
            }
        } else {
            if (moving_piece.type == PieceType::King &&
                ((rules.r.edge_escape && geo.tile_at_edge(play.to())) ||
                 (!rules.r.edge_escape && geo.is_corner(play.to()))))
                return Win(TAFL_WIN_KING_ESCAPED, Side::Defender);
            if (rules.r.exit_fort && detect_exit_fort(state.board)) return Win(TAFL_WIN_EXIT_FORT, Side::Defender);
        }
        if (rules.r.repetition_enabled) {
            if (state.repetitions.get_repetitions(state.side_to_play) >= rules.r.n_repetitions) {
                if (rules.r.repetition_is_loss) return Win(TAFL_WIN_REPETITION, other(state.side_to_play));
                return Draw(TAFL_DRAW_REPETITION);
            }
        }
        if (!side_can_play(other(state.side_to_play), state)) {
            if (rules.r.draw_on_no_plays) return Draw(TAFL_DRAW_NO_PLAYS);
            return Win(TAFL_WIN_NO_PLAYS, state.side_to_play);
        }
        return std::nullopt;
    }

    // logic.rs:782-820
    template <class W>
    std::pair<GameState<W>, PlayRecord> do_valid_play(Play play, GameState<W> state) const {
        Piece moving_piece = state.board.move_piece(play.from, play.to());
        auto captures = get_captures(play, moving_piece, state);
        for (auto& c : captures) state.board.clear_tile(c.tile);
        state.repetitions.track_play(state.side_to_play, play, !captures.empty());
        if (captures.empty()) state.plays_since_capture += 1;
        GameOutcome outcome = get_game_outcome(play, moving_piece, captures, state);
        state.turn += 1;
        GameStatus status = outcome ? *outcome : (GameStatus)TAFL_STATUS_ONGOING;
        PlayRecord rec{state.side_to_play, play, PlayEffects{captures, outcome}};
        state.side_to_play = other(state.side_to_play);
        state.status = status;
        return {state, rec};
    }
    // logic.rs:827-834 — err != 0: state unchanged
    template <class W>
    PlayInvalid do_play(Play play, const GameState<W>& state, GameState<W>* new_state, PlayRecord* rec) const {
        PlayInvalid v = validate_play(play, state);
        if (v != TAFL_PLAY_OK) return v;
        auto res = do_valid_play(play, state);
        *new_state = res.first;
        if (rec) *rec = res.second;
        return TAFL_PLAY_OK;
    }
};

// ---------------------------------------------------------------------------------------------
// preset.rs
// ---------------------------------------------------------------------------------------------
inline int preset_rules(int id, tafl_rules* out) {
    tafl_rules r{};
    const uint16_t K = TAFL_PIECESET_TYPE(TAFL_KING), S = TAFL_PIECESET_TYPE(TAFL_SOLDIER);
    const uint16_t ALL = TAFL_PIECESET_ALL, NONE = TAFL_PIECESET_NONE;
    r.king_attack = TAFL_KING_ARMED;
    r.starting_side = TAFL_ATTACKER;
    switch (id) {
        case TAFL_PRESET_COPENHAGEN:  // preset.rs:12-34
            r.edge_escape = 0; r.king_strength = TAFL_KING_STRONG;
            r.shieldwall_enabled = 1; r.sw_corners_may_close = 1; r.sw_captures = S;
            r.exit_fort = 1; r.throne_movement = TAFL_THRONE_KINGENTRY; r.may_enter_corners = K;
            r.hostility_throne = ALL; r.hostility_corners = S; r.hostility_edge = NONE;
            r.slow_pieces = NONE; r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            r.repetition_enabled = 1; r.n_repetitions = 3; r.repetition_is_loss = 1;
            r.draw_on_no_plays = 0; r.linnaean_capture = 0;
            break;
        case TAFL_PRESET_BRANDUBH:  // preset.rs:37-56
            r.edge_escape = 0; r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.shieldwall_enabled = 0; r.exit_fort = 0; r.throne_movement = TAFL_THRONE_KINGENTRY;
            r.may_enter_corners = K; r.hostility_throne = S; r.hostility_corners = ALL; r.hostility_edge = NONE;
            r.slow_pieces = NONE; r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            r.repetition_enabled = 1; r.n_repetitions = 3; r.repetition_is_loss = 1;
            r.draw_on_no_plays = 0; r.linnaean_capture = 0;
            break;
        case TAFL_PRESET_MAGPIE:  // preset.rs:59-78
            r.edge_escape = 0; r.king_strength = TAFL_KING_STRONG;
            r.shieldwall_enabled = 0; r.exit_fort = 0; r.throne_movement = TAFL_THRONE_KINGENTRY;
            r.may_enter_corners = K; r.hostility_throne = ALL; r.hostility_corners = ALL; r.hostility_edge = NONE;
            r.slow_pieces = K; r.enclosure_win = TAFL_ENCL_NONE;
            r.repetition_enabled = 0; r.draw_on_no_plays = 0; r.linnaean_capture = 0;
            break;
        case TAFL_PRESET_TABLUT:  // preset.rs:81-100
            r.edge_escape = 1; r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.shieldwall_enabled = 0; r.exit_fort = 0; r.throne_movement = TAFL_THRONE_NOENTRY;
            r.may_enter_corners = ALL; r.hostility_throne = ALL; r.hostility_corners = NONE; r.hostility_edge = NONE;
            r.slow_pieces = NONE; r.enclosure_win = TAFL_ENCL_NONE;
            r.repetition_enabled = 1; r.n_repetitions = 3; r.repetition_is_loss = 0;
            r.draw_on_no_plays = 1; r.linnaean_capture = 1;
            break;
        case TAFL_PRESET_KOCH:  // preset.rs:105-124
            r.edge_escape = 0; r.king_strength = TAFL_KING_STRONG_BY_THRONE;
            r.shieldwall_enabled = 0; r.exit_fort = 0; r.throne_movement = TAFL_THRONE_KINGENTRY;
            r.may_enter_corners = K; r.hostility_throne = ALL; r.hostility_corners = S; r.hostility_edge = NONE;
            r.slow_pieces = NONE; r.enclosure_win = TAFL_ENCL_WITHOUT_EDGE_ACCESS;
            r.repetition_enabled = 1; r.n_repetitions = 3; r.repetition_is_loss = 1;
            r.draw_on_no_plays = 0; r.linnaean_capture = 0;
            break;
        default: return TAFL_ERR_INVALID_ARG;
    }
    *out = r;
    return 0;
}
inline const char* preset_board(int id) {  // preset.rs:127-134
    switch (id) {
        case TAFL_PRESET_COPENHAGEN: return "3ttttt3/5t5/11/t4T4t/t3TTT3t/tt1TTKTT1tt/t3TTT3t/t4T4t/11/5t5/3ttttt3";
        case TAFL_PRESET_BRANDUBH: return "3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3";
        case TAFL_PRESET_MAGPIE: return "3t3/1t3t1/3T3/t1TKT1t/3T3/1t3t1/3t3";
        case TAFL_PRESET_TABLUT: return "3ttt3/4t4/4T4/t3T3t/ttTTKTTtt/t3T3t/4T4/4t4/3ttt3";
        case TAFL_PRESET_KOCH: return "3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3";  // game/main.rs:137-140
        default: return nullptr;
    }
}

// ---------------------------------------------------------------------------------------------
// tiles.rs:131-157, play.rs:70-92 — algebraic notation
// ---------------------------------------------------------------------------------------------
inline int tile_from_str(const std::string& s, Tile* out) {
    if (s.empty()) return TAFL_ERR_PARSE_EMPTY;
    unsigned char b = (unsigned char)s[0];
    if (!(b >= 97 && b <= 122)) return TAFL_ERR_PARSE_BAD_CHAR;
    std::string num = s.substr(1);
    if (num.empty()) return TAFL_ERR_PARSE_BAD_INT;
    size_t i = 0;
    if (num[0] == '+') i = 1;  // Rust's u8 parser accepts a leading '+'
    if (i >= num.size()) return TAFL_ERR_PARSE_BAD_INT;
    unsigned v = 0;
    for (; i < num.size(); ++i) {
        if (num[i] < '0' || num[i] > '9') return TAFL_ERR_PARSE_BAD_INT;
        v = v * 10 + (unsigned)(num[i] - '0');
        if (v > 255) return TAFL_ERR_PARSE_BAD_INT;
    }
    if (v == 0) return TAFL_ERR_PARSE_BAD_INT;  // `parse::<u8>()? - 1` underflows upstream (panic in debug)
    *out = Tile{(uint8_t)(v - 1), (uint8_t)(b - 97)};
    return 0;
}
inline std::string tile_to_str(Tile t) {
    return std::string(1, (char)(t.col + 97)) + std::to_string((int)t.row + 1);
}
inline int play_from_str(const std::string& s, Play* out) {
    std::vector<std::string> tok;
    size_t pos = 0;
    while (true) {
        size_t d = s.find('-', pos);
        tok.push_back(s.substr(pos, d == std::string::npos ? std::string::npos : d - pos));
        if (d == std::string::npos) break;
        pos = d + 1;
    }
    if (tok.size() != 2) return TAFL_ERR_PARSE_BAD_STRING;
    Tile a, b;
    int rc = tile_from_str(tok[0], &a);
    if (rc) return rc;
    rc = tile_from_str(tok[1], &b);
    if (rc) return rc;
    auto p = Play::from_tiles(a, b);
    if (!p) return TAFL_ERR_PARSE_BAD_PLAY;
    *out = *p;
    return 0;
}
inline std::string play_to_str(const Play& p) { return tile_to_str(p.from) + "-" + tile_to_str(p.to()); }

// ---------------------------------------------------------------------------------------------
// packed tafl_state <-> GameState<W>
// ---------------------------------------------------------------------------------------------
inline uint32_t pack_record(const std::optional<ShortPlayRecord>& r) {
    if (!r) return 0;
    uint32_t fl = TAFL_REC_SOME | (r->captures ? TAFL_REC_CAPTURES : 0) |
                  (r->side == Side::Defender ? TAFL_REC_DEFENDER : 0) |
                  (r->play.axis == Axis::Horizontal ? TAFL_REC_HORIZONTAL : 0);
    return (uint32_t)r->play.from.row | ((uint32_t)r->play.from.col << 8) |
           ((uint32_t)(uint8_t)r->play.disp << 16) | (fl << 24);
}
inline std::optional<ShortPlayRecord> unpack_record(uint32_t v) {
    uint32_t fl = v >> 24;
    if (!(fl & TAFL_REC_SOME)) return std::nullopt;
    ShortPlayRecord r;
    r.side = (fl & TAFL_REC_DEFENDER) ? Side::Defender : Side::Attacker;
    r.captures = (fl & TAFL_REC_CAPTURES) != 0;
    r.play.from = Tile{(uint8_t)(v & 0xFF), (uint8_t)((v >> 8) & 0xFF)};
    r.play.axis = (fl & TAFL_REC_HORIZONTAL) ? Axis::Horizontal : Axis::Vertical;
    r.play.disp = (int8_t)((v >> 16) & 0xFF);
    return r;
}
template <class W>
inline void pack_state(const GameState<W>& s, tafl_state* o) {
    *o = tafl_state{};
    if constexpr (sizeof(W) == 8) {
        o->att[0] = s.board.attackers; o->def[0] = s.board.defenders;
    } else {
        o->att[0] = (uint64_t)s.board.attackers; o->att[1] = (uint64_t)(s.board.attackers >> 64);
        o->def[0] = (uint64_t)s.board.defenders; o->def[1] = (uint64_t)(s.board.defenders >> 64);
    }
    for (int i = 0; i < 4; ++i) o->recent[i] = pack_record(s.repetitions.queue[i]);
    o->att_reps = (uint8_t)(s.repetitions.attacker_reps > 255 ? 255 : s.repetitions.attacker_reps);
    o->def_reps = (uint8_t)(s.repetitions.defender_reps > 255 ? 255 : s.repetitions.defender_reps);
    o->flags = (s.repetitions.attacker_mid_pair ? TAFL_FLAG_ATT_MID : 0) |
               (s.repetitions.defender_mid_pair ? TAFL_FLAG_DEF_MID : 0) |
               (s.side_to_play == Side::Defender ? TAFL_FLAG_DEF_TO_PLAY : 0);
    o->status = s.status;
    o->turn = (uint32_t)s.turn;
    o->plays_since_capture = (uint32_t)s.plays_since_capture;
    o->ring_head = (uint8_t)s.repetitions.first_i;
    o->side_len = s.board.side_len;
}
template <class W>
inline void unpack_state(const tafl_state& i, GameState<W>* s) {
    *s = GameState<W>{};
    if constexpr (sizeof(W) == 8) {
        s->board.attackers = i.att[0]; s->board.defenders = i.def[0];
    } else {
        s->board.attackers = ((u128)i.att[1] << 64) | i.att[0];
        s->board.defenders = ((u128)i.def[1] << 64) | i.def[0];
    }
    s->board.side_len = i.side_len;
    for (int k = 0; k < 4; ++k) s->repetitions.queue[k] = unpack_record(i.recent[k]);
    s->repetitions.attacker_reps = i.att_reps;
    s->repetitions.defender_reps = i.def_reps;
    s->repetitions.attacker_mid_pair = (i.flags & TAFL_FLAG_ATT_MID) != 0;
    s->repetitions.defender_mid_pair = (i.flags & TAFL_FLAG_DEF_MID) != 0;
    s->repetitions.first_i = i.ring_head & 3;
    s->side_to_play = (i.flags & TAFL_FLAG_DEF_TO_PLAY) ? Side::Defender : Side::Attacker;
    s->status = i.status;
    s->turn = i.turn;
    s->plays_since_capture = i.plays_since_capture;
}
// tafl_state_w <-> GameState<U256 / U512>
template <int NW>
inline void pack_state_w(const GameState<BigU<NW>>& s, tafl_state_w* o) {
    *o = tafl_state_w{};
    for (int i = 0; i < NW; ++i) { o->att[i] = s.board.attackers.w[i]; o->def[i] = s.board.defenders.w[i]; }
    for (int i = 0; i < 4; ++i) o->recent[i] = pack_record(s.repetitions.queue[i]);
    o->att_reps = (uint8_t)(s.repetitions.attacker_reps > 255 ? 255 : s.repetitions.attacker_reps);
    o->def_reps = (uint8_t)(s.repetitions.defender_reps > 255 ? 255 : s.repetitions.defender_reps);
    o->flags = (s.repetitions.attacker_mid_pair ? TAFL_FLAG_ATT_MID : 0) |
               (s.repetitions.defender_mid_pair ? TAFL_FLAG_DEF_MID : 0) |
               (s.side_to_play == Side::Defender ? TAFL_FLAG_DEF_TO_PLAY : 0);
    o->status = s.status;
    o->turn = (uint32_t)s.turn;
    o->plays_since_capture = (uint32_t)s.plays_since_capture;
    o->ring_head = (uint8_t)s.repetitions.first_i;
    o->side_len = s.board.side_len;
}
template <int NW>
inline void unpack_state_w(const tafl_state_w& i, GameState<BigU<NW>>* s) {
    *s = GameState<BigU<NW>>{};
    for (int k = 0; k < NW; ++k) { s->board.attackers.w[k] = i.att[k]; s->board.defenders.w[k] = i.def[k]; }
    s->board.side_len = i.side_len;
    for (int k = 0; k < 4; ++k) s->repetitions.queue[k] = unpack_record(i.recent[k]);
    s->repetitions.attacker_reps = i.att_reps;
    s->repetitions.defender_reps = i.def_reps;
    s->repetitions.attacker_mid_pair = (i.flags & TAFL_FLAG_ATT_MID) != 0;
    s->repetitions.defender_mid_pair = (i.flags & TAFL_FLAG_DEF_MID) != 0;
    s->repetitions.first_i = i.ring_head & 3;
    s->side_to_play = (i.flags & TAFL_FLAG_DEF_TO_PLAY) ? Side::Defender : Side::Attacker;
    s->status = i.status;
    s->turn = i.turn;
    s->plays_since_capture = i.plays_since_capture;
}
template <class W>
inline tafl_move play_to_move(const Play& p) {
    Tile to = p.to();
    unsigned f = (unsigned)p.from.row * Board<W>::RW + p.from.col;
    unsigned t = (unsigned)to.row * Board<W>::RW + to.col;
    return (tafl_move)(f | (t << 7));
}
template <class W>
inline Play move_to_play(tafl_move m) {
    Tile f = Board<W>::bit_to_tile(TAFL_MOVE_FROM(m)), t = Board<W>::bit_to_tile(TAFL_MOVE_TO(m));
    return *Play::from_tiles(f, t);
}
inline tafl_play play_to_pod(const Play& p) {
    return tafl_play{p.from.row, p.from.col, (uint8_t)p.axis, p.disp};
}
inline Play play_from_pod(const tafl_play& p) {
    Play q;
    q.from = Tile{p.from_row, p.from_col};
    q.axis = p.axis == TAFL_HORIZONTAL ? Axis::Horizontal : Axis::Vertical;
    q.disp = p.disp;
    return q;
}

// ---------------------------------------------------------------------------------------------
// Counter-based RNG for random playouts (build-defined) — identical on CPU and GPU
// ---------------------------------------------------------------------------------------------
inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
inline uint32_t pick_index(uint64_t seed, uint64_t game_id, uint64_t ply, uint32_t n) {
    uint64_t x = splitmix64(seed ^ (game_id * 0x9E3779B97F4A7C15ull) ^ (ply * 0xBF58476D1CE4E5B9ull));
    return (uint32_t)(((u128)x * (u128)n) >> 64);
}

}  // namespace orc
