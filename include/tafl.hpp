// tafl.hpp — C++17 host mirror of the reference's public Rust API (the `hnefatafl` lib crate,
// game/lib.rs:66-95) above the C ABI of tafl.h.  Header-only; link with libtafl_b200.so.
//
// The reference is a Rust crate and no Rust toolchain exists in this build's image, so the compiled-
// language host side is written here in C++ with the reference's names, argument meaning and error
// behaviour: `Result<T, E>` values instead of exceptions for PlayInvalid / ParseError / BoardError
// (error.rs), a panic (std::logic_error) only where the reference panics (board/state.rs:219).
// Every method is a thin call into the C ABI — all game arithmetic runs in the CUDA kernels; there
// is no CPU implementation of the rules in this file.
//
//   reference                                   here
//   Tile, Coords            tiles.rs:11-28      hnefatafl::Tile
//   Axis, AxisOffset        tiles.rs:167-179    hnefatafl::Axis, Play::movement (axis + displacement)
//   Play, ValidPlay         play.rs:23-68       hnefatafl::Play
//   PlayInvalid, ParseError error.rs:6-71       hnefatafl::PlayInvalid, hnefatafl::ParseError
//   Side, PieceType, Piece  pieces.rs:13-98     hnefatafl::Side, PieceType, Piece
//   Ruleset, preset::*      rules.rs, preset.rs hnefatafl::Ruleset, preset::rules::*, preset::boards::*
//   BitfieldBoardState      board/state.rs      hnefatafl::BoardState (inside GameState)
//   GameState               game/state.rs       hnefatafl::GameState
//   GameStatus, GameOutcome game/mod.rs:17-70   hnefatafl::GameStatus
//   GameLogic               game/logic.rs       hnefatafl::GameLogic   (by value, batched underneath)
//   Game                    game/mod.rs:76-116  hnefatafl::Game (one game) / hnefatafl::Games (lockstep batch)
#ifndef TAFL_HPP_
#define TAFL_HPP_

#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <variant>
#include <vector>

#include "tafl.h"

namespace hnefatafl {

// ---- Result<T, E> (Rust's Result; std::expected is C++23) ------------------------------------------
template <class T, class E>
class Result {
    std::variant<T, E> v_;

public:
    Result(T ok) : v_(std::in_place_index<0>, std::move(ok)) {}
    static Result Err(E e) { Result r; r.v_.template emplace<1>(std::move(e)); return r; }
    bool is_ok() const { return v_.index() == 0; }
    bool is_err() const { return v_.index() == 1; }
    const T& unwrap() const {
        if (!is_ok()) throw std::logic_error("called `Result::unwrap()` on an `Err` value");
        return std::get<0>(v_);
    }
    const E& unwrap_err() const {
        if (!is_err()) throw std::logic_error("called `Result::unwrap_err()` on an `Ok` value");
        return std::get<1>(v_);
    }
    T take() {   // unwrap() by value (moves the Ok value out)
        if (!is_ok()) throw std::logic_error("called `Result::unwrap()` on an `Err` value");
        return std::move(std::get<0>(v_));
    }

private:
    Result() : v_(std::in_place_index<1>, E{}) {}
};

// ---- error.rs ---------------------------------------------------------------------------------------
enum class PlayInvalid : uint8_t {   // error.rs:50-71, declaration order; 0 is "Ok" on the C ABI
    WrongPlayer = TAFL_WRONG_PLAYER, NoPiece = TAFL_NO_PIECE, OutOfBounds = TAFL_OUT_OF_BOUNDS,
    NoCommonAxis = TAFL_NO_COMMON_AXIS, BlockedByPiece = TAFL_BLOCKED_BY_PIECE,
    MoveThroughBlockedTile = TAFL_MOVE_THROUGH_BLOCKED_TILE, MoveOntoBlockedTile = TAFL_MOVE_ONTO_BLOCKED_TILE,
    TooFar = TAFL_TOO_FAR, GameOver = TAFL_GAME_OVER
};
inline const char* to_string(PlayInvalid e) {
    static const char* n[] = {"Ok", "WrongPlayer", "NoPiece", "OutOfBounds", "NoCommonAxis", "BlockedByPiece",
                              "MoveThroughBlockedTile", "MoveOntoBlockedTile", "TooFar", "GameOver"};
    return (unsigned)e < 10 ? n[(unsigned)e] : "?";
}
enum class ParseError : int {   // error.rs:6-25
    BadLineLen = TAFL_ERR_PARSE_BAD_LINE_LEN, BadChar = TAFL_ERR_PARSE_BAD_CHAR, EmptyString = TAFL_ERR_PARSE_EMPTY,
    BadInt = TAFL_ERR_PARSE_BAD_INT, BadPlay = TAFL_ERR_PARSE_BAD_PLAY, BadString = TAFL_ERR_PARSE_BAD_STRING,
    Unsupported = TAFL_ERR_UNSUPPORTED
};
enum class BoardError : int { NoPiece = TAFL_ERR_NO_PIECE };   // error.rs:44

// CUDA / library failures are not part of the reference's error types: they throw.
struct EngineError : std::runtime_error {
    int code;
    EngineError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

// ---- pieces.rs --------------------------------------------------------------------------------------
enum class Side : uint8_t { Attacker = TAFL_ATTACKER, Defender = TAFL_DEFENDER };
inline Side other(Side s) { return s == Side::Attacker ? Side::Defender : Side::Attacker; }   // pieces.rs:20-26
enum class PieceType : uint8_t { King = TAFL_KING, Soldier = TAFL_SOLDIER };
struct Piece {
    PieceType piece_type;
    Side side;
    bool operator==(const Piece& o) const { return piece_type == o.piece_type && side == o.side; }
};
using PieceSet = uint16_t;   // pieces.rs:157: bit = piece_type << side

// ---- tiles.rs / play.rs -----------------------------------------------------------------------------
struct Tile {
    uint8_t row = 0, col = 0;
    bool operator==(const Tile& o) const { return row == o.row && col == o.col; }
    bool operator<(const Tile& o) const { return row != o.row ? row < o.row : col < o.col; }
    static Result<Tile, ParseError> from_str(const std::string& s) {   // tiles.rs:137-157 ("a1" = row 0, col 0)
        Tile t;
        int rc = tafl_tile_from_str(s.c_str(), &t.row, &t.col);
        if (rc) return Result<Tile, ParseError>::Err((ParseError)rc);
        return t;
    }
    std::string to_string() const {
        char buf[16];
        tafl_tile_to_str(row, col, buf, sizeof buf);
        return buf;
    }
};
enum class Axis : uint8_t { Vertical = TAFL_VERTICAL, Horizontal = TAFL_HORIZONTAL };

struct Play {   // play.rs:23-27
    Tile from;
    Axis axis = Axis::Vertical;
    int8_t displacement = 0;

    static Result<Play, PlayInvalid> from_tiles(Tile from, Tile to) {   // play.rs:36-49
        tafl_play p;
        int rc = tafl_play_from_tiles(from.row, from.col, to.row, to.col, &p);
        if (rc) return Result<Play, PlayInvalid>::Err(PlayInvalid::NoCommonAxis);
        return from_c(p);
    }
    static Result<Play, ParseError> from_str(const std::string& s) {   // play.rs:70-92 ("a4-a7")
        tafl_play p;
        int rc = tafl_play_from_str(s.c_str(), &p);
        if (rc) return Result<Play, ParseError>::Err((ParseError)rc);
        return from_c(p);
    }
    Tile to() const {   // play.rs:59-62: the i8 -> u8 cast is unchecked upstream as well
        Tile t = from;
        if (axis == Axis::Vertical) t.row = (uint8_t)(t.row + displacement);
        else t.col = (uint8_t)(t.col + displacement);
        return t;
    }
    std::string to_string() const {
        char buf[32];
        tafl_play c = to_c();
        tafl_play_to_str(&c, buf, sizeof buf);
        return buf;
    }
    tafl_play to_c() const { return tafl_play{from.row, from.col, (uint8_t)axis, displacement}; }
    static Play from_c(const tafl_play& p) { return Play{Tile{p.from_row, p.from_col}, (Axis)p.axis, p.disp}; }
    bool operator==(const Play& o) const { return from == o.from && axis == o.axis && displacement == o.displacement; }
};

// ---- game/mod.rs:17-70 ------------------------------------------------------------------------------
enum class WinReason : uint8_t { KingEscaped, ExitFort, KingCaptured, AllCaptured, Enclosed, NoPlays, Repetition };
enum class DrawReason : uint8_t { Repetition, NoPlays };
struct GameStatus {   // Ongoing | Over(Win(reason, side)) | Over(Draw(reason))
    uint8_t raw = TAFL_STATUS_ONGOING;
    bool is_ongoing() const { return raw == TAFL_STATUS_ONGOING; }
    bool is_draw() const { return TAFL_STATUS_IS_DRAW(raw); }
    bool is_win() const { return TAFL_STATUS_IS_OVER(raw) && !is_draw(); }
    WinReason win_reason() const { return (WinReason)TAFL_STATUS_REASON(raw); }
    DrawReason draw_reason() const { return (DrawReason)TAFL_STATUS_REASON(raw); }
    Side winner() const { return TAFL_STATUS_WINNER_IS_DEFENDER(raw) ? Side::Defender : Side::Attacker; }
    static GameStatus Ongoing() { return {}; }
    static GameStatus Win(WinReason r, Side s) { return {TAFL_STATUS_WIN((unsigned)r, s == Side::Defender)}; }
    static GameStatus Draw(DrawReason r) { return {TAFL_STATUS_DRAW((unsigned)r)}; }
    bool operator==(const GameStatus& o) const { return raw == o.raw; }
};

// ---- rules.rs / preset.rs ---------------------------------------------------------------------------
using Ruleset = tafl_rules;
namespace preset {
namespace rules {
inline Ruleset get(int id) {
    Ruleset r;
    if (tafl_preset_rules(id, &r)) throw std::logic_error("unknown preset");
    return r;
}
inline Ruleset COPENHAGEN() { return get(TAFL_PRESET_COPENHAGEN); }
inline Ruleset BRANDUBH() { return get(TAFL_PRESET_BRANDUBH); }
inline Ruleset MAGPIE() { return get(TAFL_PRESET_MAGPIE); }
inline Ruleset TABLUT() { return get(TAFL_PRESET_TABLUT); }
inline Ruleset KOCH() { return get(TAFL_PRESET_KOCH); }
}  // namespace rules
namespace boards {
inline std::string get(int id) {
    const char* s = nullptr;
    if (tafl_preset_board_fen(id, &s)) throw std::logic_error("unknown preset");
    return s;
}
inline std::string COPENHAGEN() { return get(TAFL_PRESET_COPENHAGEN); }
inline std::string BRANDUBH() { return get(TAFL_PRESET_BRANDUBH); }
inline std::string MAGPIE() { return get(TAFL_PRESET_MAGPIE); }
inline std::string TABLUT() { return get(TAFL_PRESET_TABLUT); }
}  // namespace boards
}  // namespace preset

// ---- board/state.rs + game/state.rs: one game by value ------------------------------------------------
// GameState<BitfieldBoardState<T>> in the 64-byte packed form of tafl.h; `width` (64 | 128) is the T.
struct GameState {
    tafl_state raw{};
    int width = 128;

    // GameState::new(fen, side_to_play) (game/state.rs:136-145) over BoardState::from_fen (board/state.rs:225-250)
    static Result<GameState, ParseError> create(const std::string& fen, Side side_to_play, int width_bits) {
        GameState g;
        g.width = width_bits;
        int rc = tafl_state_from_fen(fen.c_str(), width_bits, (int)side_to_play, &g.raw);
        if (rc) return Result<GameState, ParseError>::Err((ParseError)rc);
        return g;
    }
    std::string to_fen() const {   // board/state.rs:271-295
        char buf[512];
        tafl_state_to_fen(&raw, width, buf, sizeof buf);
        return buf;
    }
    std::string to_display_str() const {   // board/state.rs:297-311
        char buf[1024];
        tafl_state_to_display_str(&raw, width, buf, sizeof buf);
        return buf;
    }
    uint8_t side_len() const { return raw.side_len; }
    Side side_to_play() const { return (raw.flags & TAFL_FLAG_DEF_TO_PLAY) ? Side::Defender : Side::Attacker; }
    void set_side_to_play(Side s) {
        raw.flags = (uint8_t)(s == Side::Defender ? raw.flags | TAFL_FLAG_DEF_TO_PLAY : raw.flags & ~TAFL_FLAG_DEF_TO_PLAY);
    }
    GameStatus status() const { return GameStatus{raw.status}; }
    uint32_t turn() const { return raw.turn; }
    // trait BoardState (board/state.rs:13-81)
    Tile get_king() const { Tile t; tafl_board_get_king(&raw, width, &t.row, &t.col); return t; }
    void set_king(Tile t) { tafl_board_set_king(&raw, width, t.row, t.col); }
    bool is_king(Tile t) const { return tafl_board_is_king(&raw, width, t.row, t.col) == 1; }
    void set_piece(Tile t, Piece p) { tafl_board_set_piece(&raw, width, t.row, t.col, (int)p.piece_type, (int)p.side); }
    void clear_tile(Tile t) { tafl_board_clear_tile(&raw, width, t.row, t.col); }
    std::optional<Piece> get_piece(Tile t) const {
        int pt = 0, sd = 0;
        tafl_board_get_piece(&raw, width, t.row, t.col, &pt, &sd);
        if (!pt) return std::nullopt;
        return Piece{(PieceType)pt, (Side)sd};
    }
    bool tile_occupied(Tile t) const { return tafl_board_tile_occupied(&raw, width, t.row, t.col) == 1; }
    int count_pieces(Side s) const { return tafl_board_count_pieces(&raw, width, (int)s); }
    std::vector<Tile> iter_occupied(Side s) const {   // ascending bit order (board/state.rs:202-216)
        uint8_t rc[2 * 128];
        int n = tafl_board_iter_occupied(&raw, width, (int)s, rc);
        std::vector<Tile> out;
        for (int i = 0; i < n; ++i) out.push_back(Tile{rc[2 * i], rc[2 * i + 1]});
        return out;
    }
    void move_piece(Tile from, Tile to) {   // panics "No piece to move." upstream (board/state.rs:219)
        if (tafl_board_move_piece(&raw, width, from.row, from.col, to.row, to.col)) throw std::logic_error("No piece to move.");
    }
    bool operator==(const GameState& o) const { return std::memcmp(&raw, &o.raw, sizeof raw) == 0; }
};

// captured tiles of a play as Tiles (PlayEffects.captures, game/mod.rs:54-61)
inline std::vector<Tile> tiles_of(const uint64_t bb[2], int width) {
    const int stride = width == 64 ? 7 : 11;
    std::vector<Tile> out;
    for (int w = 0; w < 2; ++w)
        for (int b = 0; b < 64; ++b)
            if ((bb[w] >> b) & 1ull) {
                const int i = 64 * w + b;
                if (width == 64 ? i < 49 : i < 121) out.push_back(Tile{(uint8_t)(i / stride), (uint8_t)(i % stride)});
            }
    return out;
}

struct PlayEffects {   // game/mod.rs:54-61
    std::vector<Tile> captures;
    GameStatus game_outcome;
};
struct DoPlayResult {   // logic.rs:41-52
    GameState new_state;
    Play play;
    PlayEffects effects;
};
struct Enclosure {   // logic.rs:25-32
    std::vector<Tile> occupied, unoccupied, boundary;
};

namespace detail {
struct EngineDeleter {
    void operator()(tafl_engine* e) const { if (e) tafl_engine_destroy(e); }
};
inline void check(int rc, tafl_engine* e, const char* what) {
    if (rc) throw EngineError(rc, std::string(what) + ": " + tafl_last_error(e));
}
}  // namespace detail

// ---- game/logic.rs: GameLogic (immutable, by value).  One device engine behind it; each call is a batch
// of one through the functional API of tafl.h (tafl_do_plays, tafl_validate_plays, ...). --------------------
class GameLogic {
    std::shared_ptr<tafl_engine> e_;
    Ruleset rules_;
    int width_;
    uint8_t board_len_;

public:
    // GameLogic::new(rules, board_length) (logic.rs:70).  `width_bits` is the bitfield type (u64 | u128).
    GameLogic(const Ruleset& rules, uint8_t board_length, int width_bits, int device = 0)
        : rules_(rules), width_(width_bits), board_len_(board_length) {
        std::string fen;   // an empty board of that size: the engine only needs the geometry
        for (int r = 0; r < board_length; ++r) fen += (r ? "/" : "") + std::to_string((int)board_length);
        tafl_engine* e = nullptr;
        int rc = tafl_engine_create(&rules, fen.c_str(), width_bits, 1, device, &e);
        if (rc) throw EngineError(rc, std::string("tafl_engine_create: ") + tafl_last_error(nullptr));
        e_.reset(e, detail::EngineDeleter());
    }
    const Ruleset& rules() const { return rules_; }
    int width() const { return width_; }

    Result<Play, PlayInvalid> validate_play(Play play, const GameState& st) const {   // logic.rs:219-222
        tafl_play p = play.to_c();
        uint8_t err = 0;
        detail::check(tafl_validate_plays(e_.get(), &st.raw, &p, 1, &err), e_.get(), "validate_play");
        if (err) return Result<Play, PlayInvalid>::Err((PlayInvalid)err);
        return play;
    }
    Result<DoPlayResult, PlayInvalid> do_play(Play play, const GameState& st) const {   // logic.rs:827-834
        tafl_play p = play.to_c();
        DoPlayResult r;
        r.new_state.width = st.width;
        tafl_play_effects fx;
        detail::check(tafl_do_plays(e_.get(), &st.raw, &p, 1, &r.new_state.raw, &fx), e_.get(), "do_play");
        if (fx.err) return Result<DoPlayResult, PlayInvalid>::Err((PlayInvalid)fx.err);
        r.play = play;
        r.effects.captures = tiles_of(fx.captures, st.width);
        r.effects.game_outcome = GameStatus{fx.status};
        return r;
    }
    // iter_plays(tile, state) (logic.rs:850-856, play.rs:139-226): order V+1.., V-1.., H+1.., H-1.., distance ascending
    Result<std::vector<Play>, BoardError> iter_plays(Tile t, const GameState& st) const {
        uint8_t rc2[2] = {t.row, t.col};
        tafl_move mv[32];
        int32_t n = 0;
        detail::check(tafl_piece_plays(e_.get(), &st.raw, rc2, 1, mv, &n), e_.get(), "iter_plays");
        if (n < 0) return Result<std::vector<Play>, BoardError>::Err(BoardError::NoPiece);
        std::vector<Play> out;
        for (int i = 0; i < n; ++i) {
            tafl_play p;
            tafl_move_to_play(mv[i], st.width, &p);
            out.push_back(Play::from_c(p));
        }
        return out;
    }
    // get_all_possible_moves (game/main.rs:33-43): pieces of the side to move in ascending bit order x iter_plays
    std::vector<Play> get_all_possible_moves(const GameState& st) const {
        std::vector<tafl_move> mv(TAFL_MAX_MOVES);
        uint32_t n = 0;
        detail::check(tafl_legal_moves_of(e_.get(), &st.raw, 1, mv.data(), &n), e_.get(), "get_all_possible_moves");
        std::vector<Play> out;
        for (uint32_t i = 0; i < n; ++i) {
            tafl_play p;
            tafl_move_to_play(mv[i], st.width, &p);
            out.push_back(Play::from_c(p));
        }
        return out;
    }
    std::vector<Tile> get_captures(Play play, const GameState& st) const {   // logic.rs:604-699
        tafl_play p = play.to_c();
        tafl_play_effects fx;
        detail::check(tafl_get_captures(e_.get(), &st.raw, &p, 1, &fx), e_.get(), "get_captures");
        return tiles_of(fx.captures, st.width);
    }
    std::optional<std::vector<Tile>> detect_shieldwall(Play play, const GameState& st) const {   // logic.rs:535-569
        tafl_play p = play.to_c();
        tafl_play_effects fx;
        detail::check(tafl_detect_shieldwall(e_.get(), &st.raw, &p, 1, &fx), e_.get(), "detect_shieldwall");
        if (fx.n_captures == 0xFF) return std::nullopt;
        return tiles_of(fx.captures, st.width);
    }
    std::optional<Enclosure> find_enclosure(Tile t, PieceSet enclosed, PieceSet enclosing, bool abort_on_edge,
                                            bool abort_on_corner, const GameState& st) const {   // logic.rs:309-401
        tafl_enclosure_query q{};
        q.row = t.row; q.col = t.col;
        q.abort_on_edge = abort_on_edge; q.abort_on_corner = abort_on_corner;
        q.enclosed_pieces = enclosed; q.enclosing_pieces = enclosing;
        tafl_enclosure en;
        detail::check(tafl_find_enclosure(e_.get(), &st.raw, &q, 1, &en), e_.get(), "find_enclosure");
        if (!en.found) return std::nullopt;
        return Enclosure{tiles_of(en.occupied, st.width), tiles_of(en.unoccupied, st.width), tiles_of(en.boundary, st.width)};
    }
    // find_enclosure + enclosure_secure (logic.rs:408-463); nullopt = no enclosure
    std::optional<bool> enclosure_secure(Tile t, PieceSet enclosed, PieceSet enclosing, bool abort_on_edge, bool abort_on_corner,
                                         bool inside_safe, bool outside_safe, const GameState& st) const {
        tafl_enclosure_query q{};
        q.row = t.row; q.col = t.col;
        q.abort_on_edge = abort_on_edge; q.abort_on_corner = abort_on_corner;
        q.enclosed_pieces = enclosed; q.enclosing_pieces = enclosing;
        q.inside_safe = inside_safe; q.outside_safe = outside_safe;
        uint8_t out = 0;
        detail::check(tafl_enclosure_secure(e_.get(), &st.raw, &q, 1, &out), e_.get(), "enclosure_secure");
        if (out == 0xFF) return std::nullopt;
        return out != 0;
    }
    bool detect_exit_fort(const GameState& st) const {   // logic.rs:572-601
        uint8_t out = 0;
        detail::check(tafl_detect_exit_fort(e_.get(), &st.raw, 1, &out), e_.get(), "detect_exit_fort");
        return out != 0;
    }
    bool side_can_play(Side s, const GameState& st) const {   // logic.rs:837-846
        uint8_t sd = s == Side::Defender ? 1 : 0, out = 0;
        detail::check(tafl_side_can_play(e_.get(), &st.raw, &sd, 1, &out), e_.get(), "side_can_play");
        return out != 0;
    }
};

// ---- game/mod.rs:76-116: Game (rules + state + history) ----------------------------------------------
struct PlayRecord {   // play.rs:106-113
    Side side;
    Play play;
    PlayEffects effects;
};

// `n_games` resident games advanced in lockstep on one GPU: the batched form of Game.  do_play takes one
// Play per game and returns one Result<GameStatus, PlayInvalid> per game (game/mod.rs:95-101).
class Games {
protected:
    std::unique_ptr<tafl_engine, detail::EngineDeleter> e_;
    uint32_t n_ = 0;
    int width_ = 0;
    tafl_play* h_plays_ = nullptr;   // page-locked: tafl_step_status reads / writes these in place
    uint8_t *h_status_ = nullptr, *h_err_ = nullptr;

public:
    // Game::new(rules, starting_board) (game/mod.rs:86-91) x n_games; history_capacity > 0 keeps that many
    // (state, PlayRecord) pairs per game for undo_last_play (the reference's Vecs are unbounded).
    static Result<std::unique_ptr<Games>, ParseError> create(const Ruleset& rules, const std::string& starting_board, uint32_t n_games,
                                                             int width_bits = 0, int device = 0, uint32_t history_capacity = 0) {
        tafl_state probe;   // parse errors are values, as in the reference
        int w = width_bits ? width_bits : 128;
        int rc = tafl_state_from_fen(starting_board.c_str(), w, rules.starting_side, &probe);
        if (rc == TAFL_ERR_UNSUPPORTED && !width_bits) rc = 0;
        if (rc) return Result<std::unique_ptr<Games>, ParseError>::Err((ParseError)rc);
        std::unique_ptr<Games> g(new Games());
        tafl_engine* e = nullptr;
        rc = tafl_engine_create(&rules, starting_board.c_str(), width_bits, n_games, device, &e);
        if (rc <= TAFL_ERR_PARSE_BAD_LINE_LEN && rc >= TAFL_ERR_UNSUPPORTED) return Result<std::unique_ptr<Games>, ParseError>::Err((ParseError)rc);
        if (rc) throw EngineError(rc, std::string("tafl_engine_create: ") + tafl_last_error(nullptr));
        g->e_.reset(e);
        int sl, dev, sm;
        tafl_engine_info(e, &g->n_, &g->width_, &sl, &dev, &sm);
        void* p = nullptr;
        detail::check(tafl_host_alloc((size_t)n_games * (sizeof(tafl_play) + 2), &p), e, "tafl_host_alloc");
        g->h_plays_ = (tafl_play*)p;
        g->h_status_ = (uint8_t*)(g->h_plays_ + n_games);
        g->h_err_ = g->h_status_ + n_games;
        if (history_capacity) detail::check(tafl_history_enable(e, history_capacity), e, "tafl_history_enable");
        return g;
    }
    virtual ~Games() { if (h_plays_) tafl_host_free(h_plays_); }
    Games(const Games&) = delete;
    Games& operator=(const Games&) = delete;

    uint32_t len() const { return n_; }
    int width() const { return width_; }
    tafl_engine* handle() const { return e_.get(); }

    // restart_finished: a game whose status is Over is first replaced by a new one (Game::new, what a host does when it
    // sees a finished status) and the play is applied to that — same launch (tafl_step_status_restart)
    std::vector<Result<GameStatus, PlayInvalid>> do_play(const std::vector<Play>& plays, bool restart_finished = false) {
        if (plays.size() != n_) throw std::logic_error("Games::do_play: one Play per game");
        for (uint32_t i = 0; i < n_; ++i) h_plays_[i] = plays[i].to_c();
        detail::check((restart_finished ? tafl_step_status_restart : tafl_step_status)(e_.get(), h_plays_, h_status_, h_err_), e_.get(), "do_play");
        std::vector<Result<GameStatus, PlayInvalid>> out;
        out.reserve(n_);
        for (uint32_t i = 0; i < n_; ++i) {
            if (h_err_[i]) out.push_back(Result<GameStatus, PlayInvalid>::Err((PlayInvalid)h_err_[i]));
            else out.push_back(GameStatus{h_status_[i]});
        }
        return out;
    }
    uint32_t restart_finished() {   // Game::new for every game whose status is Over; returns how many
        uint32_t n = 0;
        detail::check(tafl_reset_finished(e_.get(), &n), e_.get(), "restart_finished");
        return n;
    }
    void undo_last_play(const std::vector<uint8_t>* mask = nullptr) {   // game/mod.rs:103-108
        detail::check(tafl_undo_last_play(e_.get(), mask ? mask->data() : nullptr), e_.get(), "undo_last_play");
    }
    std::vector<GameState> states() const {
        std::vector<tafl_state> raw(n_);
        detail::check(tafl_get_states(e_.get(), raw.data(), 0, n_), e_.get(), "states");
        std::vector<GameState> out(n_);
        for (uint32_t i = 0; i < n_; ++i) { out[i].raw = raw[i]; out[i].width = width_; }
        return out;
    }
    std::vector<PlayRecord> play_history(uint32_t game) const {   // Game.play_history (game/mod.rs:80)
        std::vector<uint32_t> len(n_);
        detail::check(tafl_history_len(e_.get(), len.data()), e_.get(), "history_len");
        std::vector<tafl_play_record> rec(len[game] ? len[game] : 1);
        uint32_t n = 0;
        detail::check(tafl_play_history(e_.get(), game, rec.data(), (uint32_t)rec.size(), &n), e_.get(), "play_history");
        std::vector<PlayRecord> out;
        for (uint32_t i = 0; i < n; ++i)
            out.push_back(PlayRecord{(Side)rec[i].side, Play::from_c(rec[i].play),
                                     PlayEffects{tiles_of(rec[i].effects.captures, width_), GameStatus{rec[i].effects.status}}});
        return out;
    }
    // get_all_possible_moves (game/main.rs:33-43) of every game, canonical order
    std::vector<std::vector<Play>> legal_plays() const {
        std::vector<tafl_move> mv((size_t)n_ * TAFL_MAX_MOVES);
        std::vector<uint32_t> cnt(n_);
        detail::check(tafl_legal_moves_host(e_.get(), mv.data(), cnt.data(), 0, n_), e_.get(), "legal_plays");
        std::vector<std::vector<Play>> out(n_);
        for (uint32_t g = 0; g < n_; ++g)
            for (uint32_t i = 0; i < cnt[g]; ++i) {
                tafl_play p;
                tafl_move_to_play(mv[(size_t)g * TAFL_MAX_MOVES + i], width_, &p);
                out[g].push_back(Play::from_c(p));
            }
        return out;
    }
    // the fused random-play step of the benchmark (count -> pick -> apply -> outcome), n_steps plies
    void step_random(uint64_t seed, uint32_t n_steps, bool auto_reset = true) {
        detail::check(tafl_step_random(e_.get(), seed, n_steps, auto_reset, 1), e_.get(), "step_random");
    }
    tafl_stats stats() const {
        tafl_stats s;
        detail::check(tafl_stats_get(e_.get(), &s), e_.get(), "stats");
        return s;
    }

protected:
    Games() = default;
};

// One game with the reference's exact surface: Game::new / do_play / undo_last_play / iter_plays / state.
class Game {
    std::unique_ptr<Games> g_;
    Ruleset rules_;

public:
    static Result<Game, ParseError> create(const Ruleset& rules, const std::string& starting_board, int width_bits = 0, int device = 0,
                                           uint32_t history_capacity = 1024) {
        auto r = Games::create(rules, starting_board, 1, width_bits, device, history_capacity);
        if (r.is_err()) return Result<Game, ParseError>::Err(r.unwrap_err());
        Game g;
        g.g_ = r.take();
        g.rules_ = rules;
        return g;
    }
    Game(Game&&) = default;
    Game& operator=(Game&&) = default;

    Result<GameStatus, PlayInvalid> do_play(Play play) { return g_->do_play({play})[0]; }   // game/mod.rs:95-101
    void undo_last_play() { g_->undo_last_play(); }                                            // game/mod.rs:103-108
    GameState state() const { return g_->states()[0]; }
    std::vector<PlayRecord> play_history() const { return g_->play_history(0); }
    std::vector<Play> get_all_possible_moves() const { return g_->legal_plays()[0]; }
    Result<std::vector<Play>, BoardError> iter_plays(Tile t) const {   // game/mod.rs:112-115
        GameState st = state();
        uint8_t rc2[2] = {t.row, t.col};
        tafl_move mv[32];
        int32_t n = 0;
        detail::check(tafl_piece_plays(g_->handle(), &st.raw, rc2, 1, mv, &n), g_->handle(), "iter_plays");
        if (n < 0) return Result<std::vector<Play>, BoardError>::Err(BoardError::NoPiece);
        std::vector<Play> out;
        for (int i = 0; i < n; ++i) {
            tafl_play p;
            tafl_move_to_play(mv[i], st.width, &p);
            out.push_back(Play::from_c(p));
        }
        return out;
    }

private:
    Game() = default;
};

}  // namespace hnefatafl

#endif  // TAFL_HPP_
