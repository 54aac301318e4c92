/*
 * tafl.h — C ABI of libtafl_b200.so, the B200-native lockstep Hnefatafl engine.
 *
 * Every entry point is the batched equivalent of one public item of the reference
 * `hnefatafl` crate (paths relative to the reference tree) or of `src/mcts.py`; the item each
 * one replaces is cited beside it.  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions
 *  - every call returns TAFL_OK (0) or a negative tafl_err; no exceptions cross the ABI.
 *  - buffers named `h_*` are HOST pointers (copied inside the call), `d_*` are DEVICE pointers
 *    on the engine's device (no copy).  The caller owns every buffer.
 *  - an engine is thread-compatible: one caller at a time per handle.  All work is enqueued on
 *    the engine's own CUDA stream; blocking calls return after the stream is synchronised,
 *    `_async` variants return after enqueue.
 *  - there is no CPU fallback: if no CUDA device is usable, tafl_engine_create fails with
 *    TAFL_ERR_CUDA and nothing else can be called.
 */
#ifndef TAFL_B200_H
#define TAFL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * Enumerations (values follow the reference's declaration order)
 * ---------------------------------------------------------------------------------------- */

/* pieces.rs:13-17  `enum Side { Attacker = 0, Defender = 8 }` */
enum { TAFL_ATTACKER = 0, TAFL_DEFENDER = 8 };

/* pieces.rs:31-38  `enum PieceType` bit flags; PieceSet (pieces.rs:157) = type << side */
enum { TAFL_KING = 1, TAFL_SOLDIER = 2, TAFL_KNIGHT = 4, TAFL_COMMANDER = 8, TAFL_GUARD = 16,
       TAFL_MERCENARY = 32 };
#define TAFL_PIECESET_NONE 0x0000u
#define TAFL_PIECESET_ALL  0xFFFFu
#define TAFL_PIECESET_TYPE(t) ((uint16_t)((t) | ((t) << 8)))     /* PieceSet::from_piece_type */
#define TAFL_PIECESET_SIDE(s) ((uint16_t)(0x00FFu << (s)))       /* PieceSet::from(Side)      */
#define TAFL_PIECESET_PIECE(t, s) ((uint16_t)((t) << (s)))       /* PieceSet::from_piece      */

/* tiles.rs:167-170  `enum Axis { Vertical = 0, Horizontal = 0x80 }` */
enum { TAFL_VERTICAL = 0, TAFL_HORIZONTAL = 0x80 };

/* rules.rs:6-17 */
enum { TAFL_THRONE_NOTHRONE = 0, TAFL_THRONE_NOPASS = 1, TAFL_THRONE_KINGPASS = 2,
       TAFL_THRONE_NOENTRY = 3, TAFL_THRONE_KINGENTRY = 4 };
/* rules.rs:22-30 */
enum { TAFL_KING_STRONG = 0, TAFL_KING_STRONG_BY_THRONE = 1, TAFL_KING_WEAK = 2 };
/* rules.rs:34-42 */
enum { TAFL_KING_ARMED = 0, TAFL_KING_ANVIL = 1, TAFL_KING_HAMMER = 2 };
/* rules.rs:64-69; 0 = Option::None */
enum { TAFL_ENCL_NONE = 0, TAFL_ENCL_WITH_EDGE_ACCESS = 1, TAFL_ENCL_WITHOUT_EDGE_ACCESS = 2 };

/* error.rs:50-71  `enum PlayInvalid`, declaration order, 0 = Ok */
enum { TAFL_PLAY_OK = 0, TAFL_WRONG_PLAYER = 1, TAFL_NO_PIECE = 2, TAFL_OUT_OF_BOUNDS = 3,
       TAFL_NO_COMMON_AXIS = 4, TAFL_BLOCKED_BY_PIECE = 5, TAFL_MOVE_THROUGH_BLOCKED_TILE = 6,
       TAFL_MOVE_ONTO_BLOCKED_TILE = 7, TAFL_TOO_FAR = 8, TAFL_GAME_OVER = 9 };

/* game/mod.rs:17-33  `enum WinReason`, declaration order */
enum { TAFL_WIN_KING_ESCAPED = 0, TAFL_WIN_EXIT_FORT = 1, TAFL_WIN_KING_CAPTURED = 2,
       TAFL_WIN_ALL_CAPTURED = 3, TAFL_WIN_ENCLOSED = 4, TAFL_WIN_NO_PLAYS = 5,
       TAFL_WIN_REPETITION = 6 };
/* game/mod.rs:37-42  `enum DrawReason` */
enum { TAFL_DRAW_REPETITION = 0, TAFL_DRAW_NO_PLAYS = 1 };

/* GameStatus (game/mod.rs:64-70) packed in one byte:
 *   0                      Ongoing
 *   0x80 | reason<<1 | w   Over(Win(reason, side)),  w = 1 iff side == Defender
 *   0xC0 | reason<<1       Over(Draw(reason))                                              */
#define TAFL_STATUS_ONGOING 0u
#define TAFL_STATUS_WIN(reason, winner_is_defender) \
    ((uint8_t)(0x80u | ((reason) << 1) | ((winner_is_defender) ? 1u : 0u)))
#define TAFL_STATUS_DRAW(reason) ((uint8_t)(0xC0u | ((reason) << 1)))
#define TAFL_STATUS_IS_OVER(s) (((s) & 0x80u) != 0)
#define TAFL_STATUS_IS_DRAW(s) (((s) & 0xC0u) == 0xC0u)
#define TAFL_STATUS_REASON(s) (((s) >> 1) & 7u)
#define TAFL_STATUS_WINNER_IS_DEFENDER(s) (((s) & 1u) != 0)

/* library error codes */
typedef enum {
    TAFL_OK = 0,
    TAFL_ERR_INVALID_ARG = -1,
    TAFL_ERR_CUDA = -2,          /* no device / CUDA runtime failure: message in tafl_last_error */
    TAFL_ERR_PARSE_BAD_LINE_LEN = -3,   /* error.rs:11  ParseError::BadLineLen */
    TAFL_ERR_PARSE_BAD_CHAR = -4,       /* error.rs:13  ParseError::BadChar    */
    TAFL_ERR_PARSE_EMPTY = -5,          /* error.rs:15  ParseError::EmptyString */
    TAFL_ERR_PARSE_BAD_INT = -6,        /* error.rs:18  ParseError::BadInt     */
    TAFL_ERR_PARSE_BAD_PLAY = -7,       /* error.rs:20  ParseError::BadPlay(DisjointTiles) */
    TAFL_ERR_PARSE_BAD_STRING = -8,     /* error.rs:22  ParseError::BadString  */
    TAFL_ERR_UNSUPPORTED = -9,   /* board does not fit u64 (<=7) / u128 (<=11) bitfields */
    TAFL_ERR_CAPACITY = -10,     /* MCTS node/edge pool exhausted                       */
    TAFL_ERR_NO_PIECE = -11      /* error.rs:44  BoardError::NoPiece                    */
} tafl_err;

/* ------------------------------------------------------------------------------------------
 * POD mirrors of the reference types
 * ---------------------------------------------------------------------------------------- */

/* play.rs:23-27  `struct Play { from: Tile, movement: AxisOffset }` */
typedef struct {
    uint8_t from_row, from_col;
    uint8_t axis;   /* TAFL_VERTICAL | TAFL_HORIZONTAL */
    int8_t  disp;   /* signed displacement along the axis */
} tafl_play;

/* rules.rs:83-117  `struct Ruleset` (+ HostilityRules :46-50, ShieldwallRules :54-59,
 * RepetitionRule :73-79).  Option<..> fields carry an `_enabled` flag / 0 value. */
typedef struct {
    uint8_t  edge_escape;
    uint8_t  king_strength;          /* TAFL_KING_STRONG ..           */
    uint8_t  king_attack;            /* TAFL_KING_ARMED ..            */
    uint8_t  shieldwall_enabled;     /* Option<ShieldwallRules>       */
    uint8_t  sw_corners_may_close;
    uint8_t  exit_fort;
    uint8_t  throne_movement;        /* TAFL_THRONE_*                 */
    uint8_t  starting_side;          /* TAFL_ATTACKER | TAFL_DEFENDER */
    uint16_t sw_captures;            /* PieceSet                      */
    uint16_t may_enter_corners;      /* PieceSet                      */
    uint16_t hostility_throne;       /* PieceSet                      */
    uint16_t hostility_corners;      /* PieceSet                      */
    uint16_t hostility_edge;         /* PieceSet                      */
    uint16_t slow_pieces;            /* PieceSet                      */
    uint8_t  enclosure_win;          /* TAFL_ENCL_*                   */
    uint8_t  repetition_enabled;     /* Option<RepetitionRule>        */
    uint8_t  repetition_is_loss;
    uint8_t  draw_on_no_plays;
    uint8_t  linnaean_capture;
    uint8_t  _pad[3];
    uint32_t n_repetitions;
} tafl_rules;

/* One game: `GameState<BitfieldBoardState<T>>` (game/state.rs:120-133) packed to 64 bytes.
 * This is also the HBM-resident form inside the engine.
 *
 * att/def are the two bitfield words of BitfieldBoardState (board/state.rs:117-121) in the
 * REFERENCE bit layout: bit (row*ROW_WIDTH + col), ROW_WIDTH = 7 for u64 boards and 11 for u128
 * boards (bitfield.rs:178-179); the king's row lives in the top nibble of `def`, its column in
 * the top nibble of `att` (board/state.rs:127-147).  For u128 boards [0] is the low 64 bits and
 * the nibble is bits 60..63 of [1]; for u64 boards only [0] is used (nibble = bits 60..63 of
 * [0]) and [1] must be 0. */
typedef struct {
    uint64_t att[2];
    uint64_t def[2];
    /* RepetitionTracker.recent_plays (game/state.rs:49, utils.rs:30-81): ring of 4 records.
     * byte0 from_row, byte1 from_col, byte2 disp (int8), byte3 flags:
     *   bit0 Some(..), bit1 captures, bit2 side==Defender, bit3 axis==Horizontal */
    uint32_t recent[4];
    uint8_t  att_reps, def_reps;     /* RepetitionTracker.{attacker,defender}_reps (saturating) */
    uint8_t  flags;                  /* bit0 attacker_mid_pair, bit1 defender_mid_pair,
                                        bit2 side_to_play == Defender                        */
    uint8_t  status;                 /* packed GameStatus, see TAFL_STATUS_*                 */
    uint32_t turn;                   /* GameState.turn                                       */
    uint32_t plays_since_capture;    /* GameState.plays_since_capture (write-only upstream)  */
    uint8_t  ring_head;              /* FixedSizeQueue.first_i                               */
    uint8_t  side_len;               /* BitfieldBoardState.side_len                          */
    uint8_t  _pad[2];
} tafl_state;

#define TAFL_FLAG_ATT_MID 1u
#define TAFL_FLAG_DEF_MID 2u
#define TAFL_FLAG_DEF_TO_PLAY 4u

#define TAFL_REC_SOME 1u
#define TAFL_REC_CAPTURES 2u
#define TAFL_REC_DEFENDER 4u
#define TAFL_REC_HORIZONTAL 8u

/* Effects of one play: `PlayRecord.effects` (play.rs:106-113, game/mod.rs:56-61).  The capture
 * set is a tile bitboard in the board's own bit layout (no nibbles). */
typedef struct {
    uint64_t captures[2];
    uint8_t  status;        /* game outcome after the play, packed GameStatus */
    uint8_t  err;           /* PlayInvalid code, 0 if the play was applied     */
    uint8_t  n_captures;
    uint8_t  _pad[5];
} tafl_play_effects;

/* `Enclosure` (logic.rs:25-32) as three tile bitboards; found == 0 mirrors Option::None. */
typedef struct {
    uint64_t occupied[2];
    uint64_t unoccupied[2];
    uint64_t boundary[2];
    uint8_t  found;
    uint8_t  _pad[7];
} tafl_enclosure;

/* Compact move: from_bit | to_bit << 7, bit indices in the board's own layout.
 * TAFL_MAX_MOVES bounds a legal-move list (<= pieces * 2(N-1); 512 is safe for 11x11). */
typedef uint16_t tafl_move;
#define TAFL_MOVE_FROM(m) ((m) & 127u)
#define TAFL_MOVE_TO(m) (((m) >> 7) & 127u)
#define TAFL_MAX_MOVES 512

/* presets (preset.rs) */
enum { TAFL_PRESET_COPENHAGEN = 0, TAFL_PRESET_BRANDUBH = 1, TAFL_PRESET_MAGPIE = 2,
       TAFL_PRESET_TABLUT = 3, TAFL_PRESET_KOCH = 4 };

/* ------------------------------------------------------------------------------------------
 * Host-only helpers (no device needed)
 * ---------------------------------------------------------------------------------------- */

/* preset::rules::{COPENHAGEN,BRANDUBH,MAGPIE,TABLUT,KOCH}  (preset.rs:12-124) */
int tafl_preset_rules(int preset_id, tafl_rules* out);
/* preset::boards::{COPENHAGEN,BRANDUBH,MAGPIE,TABLUT} (preset.rs:127-134); KOCH has no board of
 * its own upstream — the demo pairs it with the Brandubh board (game/main.rs:137-140). */
int tafl_preset_board_fen(int preset_id, const char** out_fen);

/* BoardState::from_fen (board/state.rs:225-250) + GameState::new (game/state.rs:136-145).
 * width_bits = 64 or 128 selects BitfieldBoardState<u64|u128>. */
int tafl_state_from_fen(const char* fen, int width_bits, int side_to_play, tafl_state* out);
/* BoardState::to_fen (board/state.rs:271-295); buf must hold >= 256 bytes. */
int tafl_state_to_fen(const tafl_state* st, int width_bits, char* buf, size_t buf_len);
/* Play::from_tiles (play.rs:36-49) / Play::to (play.rs:59-62) */
int tafl_play_from_tiles(uint8_t from_row, uint8_t from_col, uint8_t to_row, uint8_t to_col,
                         tafl_play* out);
/* Play: FromStr / Display (play.rs:70-92, tiles.rs:131-157), e.g. "d6-f6" */
int tafl_play_from_str(const char* s, tafl_play* out);
int tafl_play_to_str(const tafl_play* p, char* buf, size_t buf_len);
/* compact move <-> Play for a given bitfield width */
int tafl_move_to_play(tafl_move m, int width_bits, tafl_play* out);

/* BoardState::from_display_str / to_display_str (board/state.rs:252-269, 297-311) */
int tafl_state_from_display_str(const char* s, int width_bits, int side_to_play, tafl_state* out);
int tafl_state_to_display_str(const tafl_state* st, int width_bits, char* buf, size_t buf_len);

/* `trait BoardState` accessors (board/state.rs:13-81) on one packed state */
int tafl_board_get_king(const tafl_state* st, int width_bits, uint8_t* row, uint8_t* col);   /* :17,127 */
int tafl_board_set_king(tafl_state* st, int width_bits, uint8_t row, uint8_t col);           /* :20,136 */
int tafl_board_is_king(const tafl_state* st, int width_bits, uint8_t row, uint8_t col);      /* :23     */
int tafl_board_set_piece(tafl_state* st, int width_bits, uint8_t row, uint8_t col,
                         int piece_type, int side);                                           /* :28,149 */
int tafl_board_clear_tile(tafl_state* st, int width_bits, uint8_t row, uint8_t col);         /* :31,167 */
/* returns 0 and *piece_type = 0 for an empty tile (Option::None) */
int tafl_board_get_piece(const tafl_state* st, int width_bits, uint8_t row, uint8_t col,
                         int* piece_type, int* side);                                         /* :34,173 */
int tafl_board_tile_occupied(const tafl_state* st, int width_bits, uint8_t row, uint8_t col);/* :37,189 */
int tafl_board_count_pieces(const tafl_state* st, int width_bits, int side);                 /* :41,195 */
/* tiles in ascending bit order; out_rc = [row, col] pairs (>= 2*128 bytes); returns the count */
int tafl_board_iter_occupied(const tafl_state* st, int width_bits, int side, uint8_t* out_rc);/* :45,202 */
/* TAFL_ERR_NO_PIECE mirrors the `expect("No piece to move.")` panic (board/state.rs:219) */
int tafl_board_move_piece(tafl_state* st, int width_bits, uint8_t from_row, uint8_t from_col,
                          uint8_t to_row, uint8_t to_col);                                    /* :50,218 */
int tafl_board_swap_pieces(tafl_state* st, int width_bits, uint8_t r1, uint8_t c1, uint8_t r2,
                           uint8_t c2);                                                       /* :69     */
/* Tile: FromStr / Display (tiles.rs:131-157), e.g. "a8" <-> (7, 0) */
int tafl_tile_from_str(const char* s, uint8_t* row, uint8_t* col);
int tafl_tile_to_str(uint8_t row, uint8_t col, char* buf, size_t buf_len);
/* RepetitionTracker::track_play (game/state.rs:92-113) on one packed state */
int tafl_track_play(tafl_state* st, int side, const tafl_play* play, int captures);
/* board_to_matrix (game/main.rs:55-83): side_len*side_len bytes, row-major */
int tafl_board_to_matrix(const tafl_state* st, int width_bits, uint8_t* out, size_t out_len);

const char* tafl_version(void);

/* ------------------------------------------------------------------------------------------
 * Engine: `GameLogic` (logic.rs:61-72) + n_games resident `GameState`s on one GPU
 * ---------------------------------------------------------------------------------------- */
typedef struct tafl_engine tafl_engine;

/* Game::new(rules, starting_board) (game/mod.rs:86-91) x n_games, on CUDA device `device`.
 * width_bits: 64 | 128 | 0 (= smallest that fits the FEN's side length). */
int tafl_engine_create(const tafl_rules* rules, const char* start_fen, int width_bits,
                       uint32_t n_games, int device, tafl_engine** out);
int tafl_engine_destroy(tafl_engine* e);
const char* tafl_last_error(const tafl_engine* e);   /* e may be NULL: last create() error */
int tafl_engine_info(const tafl_engine* e, uint32_t* n_games, int* width_bits, int* side_len,
                     int* device, int* sm_count);
/* cudaStream_t of the engine (for callers that want to order their own work / events) */
void* tafl_engine_stream(const tafl_engine* e);
int tafl_engine_sync(tafl_engine* e);
/* Kernel specialisation in use: 0 = generic (rule flags read at run time), 1 = Copenhagen 11x11,
 * 2 = Tablut 9x9, 3 = Brandubh 7x7 on u64 — chosen at create time when the Ruleset and board equal
 * the preset (preset.rs:12-134); results are identical either way (tests run both).
 * tafl_engine_force_generic(e, 1) switches the generic kernels on, (e, 0) back. */
int tafl_engine_spec(const tafl_engine* e);
int tafl_engine_force_generic(tafl_engine* e, int on);
/* number of kernels this engine has launched so far (bench.py's gpu_launches) */
uint64_t tafl_launch_count(const tafl_engine* e);

/* reset games to the start position (ids == NULL: all games) */
int tafl_reset(tafl_engine* e, const uint32_t* h_game_ids, uint32_t n);
/* copy resident states in/out (reference layout) */
int tafl_set_states(tafl_engine* e, const tafl_state* h_states, uint32_t first, uint32_t n);
int tafl_get_states(tafl_engine* e, tafl_state* h_states, uint32_t first, uint32_t n);
/* device pointer to the resident [n_games] tafl_state array */
tafl_state* tafl_states_device(tafl_engine* e);

/* get_all_possible_moves (game/main.rs:33-43): every legal play of the side to move of every
 * resident game, canonical order = pieces ascending bit index x (V+,V-,H+,H-) x distance
 * ascending (play.rs:157,166-183,189-225).  d_moves is [n_games][TAFL_MAX_MOVES] tafl_move,
 * d_counts [n_games]; both DEVICE buffers (NULL = engine-owned, see tafl_moves_device). */
int tafl_legal_moves(tafl_engine* e, tafl_move* d_moves, uint32_t* d_counts);
int tafl_legal_moves_async(tafl_engine* e, tafl_move* d_moves, uint32_t* d_counts);   /* enqueue only, no stream sync */
int tafl_moves_device(tafl_engine* e, tafl_move** d_moves, uint32_t** d_counts);
/* same, results copied to HOST buffers */
int tafl_legal_moves_host(tafl_engine* e, tafl_move* h_moves, uint32_t* h_counts,
                          uint32_t first, uint32_t n);

/* Game::do_play (game/mod.rs:95-101) for every resident game: one tafl_play per game (HOST).
 * h_effects[i].err != 0 leaves game i unchanged (logic.rs:827-834). h_effects may be NULL. */
int tafl_step(tafl_engine* e, const tafl_play* h_plays, tafl_play_effects* h_effects);
/* The same call with the reference's exact return type, Result<GameStatus, PlayInvalid>
 * (game/mod.rs:95): per game one status byte and one PlayInvalid byte come back instead of the
 * 24-byte effects record.  HOST buffers; h_err may be NULL.  When the buffers are page-locked
 * (tafl_host_alloc / tafl_host_register, cudaHostAlloc, a pinned torch tensor) the kernel reads the
 * plays and writes the result bytes in place over PCIe -- one launch and one sync per call instead of
 * three copy-engine hand-offs (26.5 vs 41.4 us per 65 536-game ply on a B200); pageable buffers are
 * staged through the copy engine.  Same results either way. */
int tafl_step_status(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err);
/* tafl_reset_finished + tafl_step_status in ONE launch: a game that had ended before this call is first
 * replaced by a new one (`Game::new`, game/mod.rs:84-93: start position, episode counter + 1, history
 * forgotten) and the play is then applied to it.  What a host driving many games does anyway when it
 * sees a finished status; fusing it saves a launch per ply on the host-buffer path. */
int tafl_step_status_restart(tafl_engine* e, const tafl_play* h_plays, uint8_t* h_status, uint8_t* h_err);
/* Page-locked host memory for the HOST-buffer calls, so a host language needs no CUDA binding of its
 * own: tafl_host_alloc returns `bytes` of pinned, device-mapped memory (free with tafl_host_free);
 * tafl_host_register pins an existing allocation in place (e.g. a Rust Vec's buffer) until
 * tafl_host_unregister.  No engine handle: the memory is usable with every engine of the process. */
int tafl_host_alloc(size_t bytes, void** out);
int tafl_host_free(void* p);
int tafl_host_register(void* p, size_t bytes);
int tafl_host_unregister(void* p);
/* Game<T>::play_history / state_history / undo_last_play (game/mod.rs:77-108) for every
 * resident game.  After tafl_history_enable, every VALID play of tafl_step / tafl_step_status
 * pushes (state before the play, PlayRecord) for its game — an invalid play pushes nothing,
 * as in Game::do_play (mod.rs:96: the `?` returns before the push).  The reference Vecs are
 * unbounded; here the last `capacity` plies per game are held (older ones are dropped).
 * tafl_reset / tafl_set_states clear the history of the games they touch. */
typedef struct {
    tafl_play play;
    uint8_t   side;              /* TAFL_ATTACKER | TAFL_DEFENDER: PlayRecord.side (play.rs:109) */
    uint8_t   _pad[3];
    tafl_play_effects effects;   /* captures + outcome (PlayEffects, game/mod.rs:54-61)          */
} tafl_play_record;              /* 32 bytes */
int tafl_history_enable(tafl_engine* e, uint32_t capacity);
/* undo_last_play (mod.rs:103-108) in the games with h_mask[i] != 0 (NULL = every game); a game
 * with no recorded play is left unchanged, as in the reference */
int tafl_undo_last_play(tafl_engine* e, const uint8_t* h_mask);
/* play_history.len() per game / the recorded plays of one game, oldest held first */
int tafl_history_len(tafl_engine* e, uint32_t* h_len);
int tafl_play_history(tafl_engine* e, uint32_t game, tafl_play_record* h_out, uint32_t cap,
                      uint32_t* n_out);
/* do_valid_play (logic.rs:782-820) on DEVICE-resident compact moves, no validation, no copy. */
int tafl_step_moves_device(tafl_engine* e, const tafl_move* d_moves, uint8_t* d_status);

/* Random playouts (BASELINE.json configs 1-2): n_steps lockstep plies for every
 * resident game, move index = mulhi64(splitmix64(seed ^ game_id*0x9E3779B97F4A7C15 ^
 * ply*0xBF58476D1CE4E5B9), n_legal) into the canonical list; `ply` is the engine-wide step
 * counter (tafl_random_ply) and game_id = first_game_id + resident index.  With auto_reset a
 * game that is Over at the start of a step is first reset to the start position.  Games that
 * are Over (and not reset) do not step.  steps_per_launch <= 0 picks the default (1 = state
 * round-trips HBM every ply). */
int tafl_step_random(tafl_engine* e, uint64_t seed, uint32_t n_steps, int auto_reset,
                     int steps_per_launch);
int tafl_step_random_async(tafl_engine* e, uint64_t seed, uint32_t n_steps, int auto_reset,
                           int steps_per_launch);
/* The same random plies by value with HOST buffers (GameLogic::do_play is by value, logic.rs:827-834):
 * h_in[n_games] -> resident states -> n_steps plies -> h_out[n_games]; equivalent to tafl_set_states +
 * tafl_step_random + tafl_get_states, with the copies and kernels of successive game chunks
 * overlapped on several streams.  Pinned host memory is needed for the overlap (pageable works,
 * serially).  h_in and h_out may be the same buffer.  Rows are validated on the device: a row whose
 * side_len is not the engine's is not applied (that game keeps its resident state, which is what
 * h_out then holds for it) and the call returns TAFL_ERR_INVALID_ARG after the other games stepped. */
int tafl_step_random_host(tafl_engine* e, const tafl_state* h_in, tafl_state* h_out, uint64_t seed,
                          uint32_t n_steps, int auto_reset);
/* One word per game describing the most recent ply of tafl_step_random* (the "action" and "status"
 * outputs of an env step): bits 0..15 the tafl_move played, bits 16..23 the game's packed GameStatus
 * after it, bit 24 set iff a play was made (clear for a game that was Over and not reset). */
typedef uint32_t tafl_action;
#define TAFL_ACTION_MOVE(a) ((tafl_move)((a) & 0xFFFFu))
#define TAFL_ACTION_STATUS(a) ((uint8_t)(((a) >> 16) & 0xFFu))
#define TAFL_ACTION_STEPPED(a) ((((a) >> 24) & 1u) != 0)
tafl_action* tafl_actions_device(tafl_engine* e);                    /* [n_games], DEVICE */
int tafl_actions_host(tafl_engine* e, tafl_action* h_actions, uint32_t first, uint32_t n);
int tafl_set_game_id_base(tafl_engine* e, uint64_t first_game_id);
int tafl_set_random_ply(tafl_engine* e, uint64_t ply);
uint64_t tafl_random_ply(const tafl_engine* e);

/* counters accumulated on device by tafl_step_random since the last tafl_stats_reset */
typedef struct {
    uint64_t steps;              /* do_play's applied                                   */
    uint64_t legal_moves;        /* sum of legal-move counts over stepped positions     */
    uint64_t captures;           /* pieces captured                                     */
    uint64_t games_finished;     /* games that reached Over                             */
    uint64_t finished_plies;     /* sum of GameState.turn over finished games           */
    uint64_t win_reason[2][8];   /* [winner: 0 att, 1 def][WinReason]                   */
    uint64_t draw_reason[2];     /* [DrawReason]                                        */
    uint64_t state_checksum;     /* filled by tafl_stats_get: xor-fold of all states    */
} tafl_stats;
int tafl_stats_reset(tafl_engine* e);
int tafl_stats_get(tafl_engine* e, tafl_stats* out);

/* ------------------------------------------------------------------------------------------
 * Functional (by-value) batch API: GameLogic methods on caller-supplied states (HOST buffers)
 * ---------------------------------------------------------------------------------------- */

/* GameLogic::validate_play (logic.rs:219-222): out_err[i] = PlayInvalid code or 0 */
int tafl_validate_plays(tafl_engine* e, const tafl_state* h_states, const tafl_play* h_plays,
                        uint32_t n, uint8_t* h_out_err);
/* GameLogic::do_play (logic.rs:827-834): h_out_states[i] = new state (or the input state when
 * the play is invalid), h_out_effects[i] = captures + outcome + err */
int tafl_do_plays(tafl_engine* e, const tafl_state* h_states, const tafl_play* h_plays,
                  uint32_t n, tafl_state* h_out_states, tafl_play_effects* h_out_effects);
/* get_all_possible_moves (game/main.rs:33-43) for arbitrary states */
int tafl_legal_moves_of(tafl_engine* e, const tafl_state* h_states, uint32_t n,
                        tafl_move* h_moves /* [n][TAFL_MAX_MOVES] */, uint32_t* h_counts);
/* tafl_legal_moves_of with NO device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments, same results (rows of
 * h_moves past h_counts[i] are zero).  The device staging (states in, move rows and counts out) is a
 * file-scope `static __device__` array of TAFL_STATIC_SCRATCH_BYTES that every GPU which loads the
 * library holds until the process exits; the call makes no cudaMalloc / cudaFree / allocating
 * library call.  n > TAFL_STATIC_MAX_GAMES -> TAFL_ERR_CAPACITY with a message in tafl_last_error.
 * All calls on one device share the array: calls that overlap on the GPU (two engines or two
 * threads on the same device) race, so callers serialise them. */
#define TAFL_STATIC_MAX_GAMES 2048u
#define TAFL_STATIC_SCRATCH_BYTES \
    (TAFL_STATIC_MAX_GAMES * (sizeof(tafl_state) + TAFL_MAX_MOVES * sizeof(tafl_move) + sizeof(uint32_t)))
uint32_t tafl_static_max_games(void);
int tafl_legal_moves_of_static(tafl_engine* e, const tafl_state* h_states, uint32_t n,
                               tafl_move* h_moves /* [n][TAFL_MAX_MOVES] */, uint32_t* h_counts);
/* GameLogic::iter_plays(tile) (logic.rs:850-856, play.rs:139-226): plays of the piece on one
 * tile (either side); h_counts[i] = -1 mirrors Err(BoardError::NoPiece) */
int tafl_piece_plays(tafl_engine* e, const tafl_state* h_states, const uint8_t* h_tiles_rc,
                     uint32_t n, tafl_move* h_moves /* [n][32] */, int32_t* h_counts);
/* GameLogic::get_captures (logic.rs:604-699): the piece at play.from is moved on a copy of the
 * board first, exactly as do_valid_play does (logic.rs:789-791) */
int tafl_get_captures(tafl_engine* e, const tafl_state* h_states, const tafl_play* h_plays,
                      uint32_t n, tafl_play_effects* h_out);
/* GameLogic::detect_shieldwall (logic.rs:535-569) on the board AS GIVEN (no piece is moved);
 * h_out[i].n_captures = 0xFF mirrors None */
int tafl_detect_shieldwall(tafl_engine* e, const tafl_state* h_states, const tafl_play* h_plays,
                           uint32_t n, tafl_play_effects* h_out);
/* GameLogic::find_enclosure (logic.rs:309-401) */
typedef struct {
    uint8_t  row, col;
    uint8_t  abort_on_edge, abort_on_corner;
    uint16_t enclosed_pieces, enclosing_pieces;   /* PieceSet */
    uint8_t  inside_safe, outside_safe;           /* used by tafl_enclosure_secure */
    uint8_t  _pad[6];
} tafl_enclosure_query;
int tafl_find_enclosure(tafl_engine* e, const tafl_state* h_states,
                        const tafl_enclosure_query* h_q, uint32_t n, tafl_enclosure* h_out);
/* find_enclosure + GameLogic::enclosure_secure (logic.rs:408-463):
 * h_out[i] = 0xFF if no enclosure, else 0/1 */
int tafl_enclosure_secure(tafl_engine* e, const tafl_state* h_states,
                          const tafl_enclosure_query* h_q, uint32_t n, uint8_t* h_out);
/* GameLogic::detect_exit_fort (logic.rs:572-601) */
int tafl_detect_exit_fort(tafl_engine* e, const tafl_state* h_states, uint32_t n,
                          uint8_t* h_out);
/* GameLogic::side_can_play (logic.rs:837-846) */
int tafl_side_can_play(tafl_engine* e, const tafl_state* h_states, const uint8_t* h_sides,
                       uint32_t n, uint8_t* h_out);

/* ------------------------------------------------------------------------------------------
 * MCTS: `class MCTS` (src/mcts.py:11-136) with the stub evaluator (uniform priors over the
 * legal moves, value 0), one search tree per resident game, f64 PUCT, first-max tie-break.
 * ---------------------------------------------------------------------------------------- */
#define TAFL_MCTS_TREE_REUSE 1u   /* keep the subtree below the move played (tafl_mcts_advance) */
typedef struct {
    uint32_t num_sims;          /* args.numMCTSSims (mcts.py:37)                          */
    double   cpuct;             /* args.cpuct (mcts.py:112)                               */
    double   draw_value;        /* getGameEnded value of a draw (non-zero; 1e-4 default)  */
    uint32_t edge_capacity;     /* per-tree edge pool, 0 = default node_capacity*160      */
    uint32_t flags;             /* TAFL_MCTS_*                                            */
    uint32_t node_capacity;     /* per-tree node pool, 0 = default num_sims+1 (x4 with
                                   tree reuse; a tree that outgrows it -> TAFL_ERR_CAPACITY) */
    uint32_t vl_width;          /* virtual loss: simulations in flight per tree (0 / 1 = mcts.py's sequential search;
                                   up to 16).  Build-defined semantics, restated in
                                   oracle/mcts_oracle.hpp::run_from_root_vl: a batch of vl_width descents each sees
                                   the earlier ones as losses (n = N + vl, q = (N Q - vl) / n), then the new leaves
                                   are expanded together and the batch is backed up in order.  Applies to
                                   tafl_mcts_search (stub evaluator fused) and to the evaluator hook below, where
                                   one round then carries vl_width leaves per tree: the leaf batch has
                                   n_games * vl_width rows, row = game * vl_width + simulation, and a search takes
                                   1 + ceil((num_sims - 1) / vl_width) rounds (a fresh root is a batch of one). */
} tafl_mcts_config;

/* allocate node/edge pools for every resident game (idempotent for an equal config) */
int tafl_mcts_configure(tafl_engine* e, const tafl_mcts_config* cfg);
/* MCTS.getActionProb (mcts.py:28-53) for every resident game from a FRESH tree rooted at the
 * game's current state: runs cfg.num_sims x MCTS.search (mcts.py:55-136).  Outputs are DEVICE
 * buffers owned by the engine (tafl_mcts_results_device), layout [n_games][TAFL_MAX_MOVES]:
 * root moves (canonical order), visit counts Nsa, Q values Qsa; d_counts = number of root
 * edges; d_best = index of the first-max visit count (temp = 0 without the random tie choice). */
int tafl_mcts_search(tafl_engine* e);
int tafl_mcts_search_async(tafl_engine* e);
int tafl_mcts_results_device(tafl_engine* e, tafl_move** d_moves, uint32_t** d_visits,
                             double** d_q, uint32_t** d_counts, uint32_t** d_best);
int tafl_mcts_results_host(tafl_engine* e, uint32_t first, uint32_t n, tafl_move* h_moves,
                           uint32_t* h_visits, double* h_q, uint32_t* h_counts,
                           uint32_t* h_best);
/* The policy of getActionProb (mcts.py:40-53) from the last search's root visit counts:
 * temp == 0 one-hot on the first maximum; temp == 1 N/sum(N) (bit-exact with CPython); else
 * N^(1/temp)/sum (within 1e-12 relative).  h_probs [n_games][TAFL_MAX_MOVES] f64 may be NULL.
 * choose != 0 also replaces d_best by a move DRAWN from that policy (the Coach's
 * np.random.choice(p=probs); build-defined, the reference has no Coach): temp == 1 ->
 * k = mulhi64(splitmix64(seed ^ game_id*0x9E3779B97F4A7C15 ^ ply*0xBF58476D1CE4E5B9), sum N),
 * first action with cumsum(N) > k; other temp > 0 -> the same walk over the f64 weights.
 * tafl_mcts_play_best / tafl_pack_records then act on the drawn move. */
int tafl_mcts_policy(tafl_engine* e, double temp, int choose, uint64_t seed, double* h_probs);
/* play the chosen move (d_best) in every resident game that is Ongoing: do_valid_play */
int tafl_mcts_play_best(tafl_engine* e, uint8_t* h_status /* may be NULL */);
/* Tree reuse (cfg.flags & TAFL_MCTS_TREE_REUSE): the reference keeps ONE MCTS object per
 * episode, so the statistics below the move played survive into the next getActionProb
 * (mcts.py:20-26, the dicts are never cleared).  tafl_mcts_advance = tafl_mcts_play_best +
 * re-rooting every tree at the child reached by the move (subtree compacted in place); the
 * next tafl_mcts_search / tafl_mcts_begin then continues from the carried tree when its root
 * still equals the game's state, and starts a fresh tree otherwise.  Call tafl_pack_records
 * BEFORE advancing (the records describe the searched root). */
int tafl_mcts_advance(tafl_engine* e, uint8_t* h_status /* may be NULL */);
int tafl_mcts_reroot(tafl_engine* e);   /* the re-rooting half alone: play_best, pack_records, reroot */
/* Root exploration noise (the reference's TODO at src/mcts.rs:53 "add Dirichlet noise"; src/mcts.py has
 * none, so this is build-defined and pinned by the oracle only): in every evaluator-path search that
 * follows, the priors of each game's root become (1 - eps) * P[a] + eps * noise[game][a], a < L, once per
 * search -- when the root is expanded, or before the first round when tree reuse carried an expanded
 * root.  `noise` is caller-drawn (Dirichlet(alpha) over the root's L legal moves, zeros beyond),
 * [n_games][TAFL_MAX_MOVES] f64 in host or device memory, copied by this call and in force until the
 * next call; NULL or eps == 0 switches it off.  The fused stub search (tafl_mcts_search) has implicit
 * uniform priors and refuses to run while noise is set.  Call after tafl_mcts_configure. */
int tafl_mcts_root_noise(tafl_engine* e, const double* noise, double eps);
/* totals since configure: simulations run, nodes expanded, sum of path depths, sum of edges */
typedef struct {
    uint64_t sims, nodes, depth_sum, edges, terminal_hits, max_edges_used;
} tafl_mcts_stats;
int tafl_mcts_stats_get(tafl_engine* e, tafl_mcts_stats* out);

/* ------------------------------------------------------------------------------------------
 * Evaluator hook: `self.nnet.predict(canonicalBoard)` (src/mcts.py:85) as a
 * batched DEVICE boundary.  A search is num_sims ROUNDS; one round runs one MCTS.search
 * (mcts.py:55-136) in every tree, cut in two at the evaluator call:
 *
 *   tafl_mcts_begin(e)                    fresh trees rooted at the games' current states
 *   repeat num_sims times:
 *     tafl_mcts_select_leaves(e, &batch)  PUCT descent with per-action priors (mcts.py:104-123);
 *                                         trees that reach a terminal node finish their
 *                                         simulation here (mcts.py:77-81) and ask for nothing
 *     <evaluator>                         any work enqueued on batch.stream that fills
 *                                         batch.d_priors / batch.d_values for rows with
 *                                         d_counts > 0 (other rows are ignored)
 *     tafl_mcts_expand_backup(e)          mask + renormalise with numpy's np.sum order, the
 *                                         "all valid moves were masked" fallback (mcts.py:87-99),
 *                                         Ns = 0, backup of -v (mcts.py:101-102, 125-136)
 *   tafl_mcts_finish(e)                   root results, as after tafl_mcts_search
 *
 * Action index a == position in the leaf's canonical legal-move list d_moves[row][a]; priors
 * at a >= d_counts[row] are masked.  Priors and value cross the boundary as f32 (what a
 * network emits) and are widened to f64 before any arithmetic.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    const tafl_state* d_states;   /* [n_games] leaf positions                                    */
    const uint8_t*    d_planes;   /* [n_games][side_len^2] board_to_matrix (game/main.rs:55-83)  */
    const tafl_move*  d_moves;    /* [n_games][TAFL_MAX_MOVES] canonical legal moves of the leaf */
    const uint32_t*   d_counts;   /* [n_games] legal moves; 0 = no evaluation wanted this round  */
    float*            d_priors;   /* [n_games][TAFL_MAX_MOVES] OUT                               */
    float*            d_values;   /* [n_games] OUT, from the point of view of the side to move   */
    uint32_t          n_games, side_len;   /* n_games = ROWS of the batch: resident games x max(1, vl_width) */
    void*             stream;     /* cudaStream_t the evaluator must enqueue its work on         */
} tafl_leaf_batch;
typedef int (*tafl_evaluator_fn)(void* user, const tafl_leaf_batch* batch);   /* 0 = ok */

int tafl_mcts_begin(tafl_engine* e);
int tafl_mcts_select_leaves(tafl_engine* e, tafl_leaf_batch* out /* may be NULL */);
int tafl_mcts_expand_backup(tafl_engine* e);
int tafl_mcts_finish(tafl_engine* e);
/* begin + num_sims x (select_leaves, fn(user, &batch), expand_backup) + finish */
int tafl_mcts_search_with(tafl_engine* e, tafl_evaluator_fn fn, void* user);
/* built-in device evaluators.  TAFL_EVAL_STUB: uniform priors over the action space, value 0
 * (the north-star stub; results equal tafl_mcts_search bit for bit).  TAFL_EVAL_HASH: a
 * deterministic TEST evaluator, priors and value are a hash of (seed, position, side to move):
 *   h = splitmix64(seed ^ splitmix64(att[0] ^ splitmix64(att[1] ^ splitmix64(def[0] ^
 *       splitmix64(def[1] ^ side)))));  all priors 0 if (h & 15) == 0 (exercises mcts.py:91-99),
 *   else P[a] = ((splitmix64(h + (a+1)*0x9E3779B97F4A7C15) >> 40) + 1) * 2^-24;
 *   v = ((splitmix64(h ^ 0xD1B54A32D192ED03) >> 40) - 2^23) * 2^-24.                          */
#define TAFL_EVAL_STUB 0
#define TAFL_EVAL_HASH 1
int tafl_eval_builtin(tafl_engine* e, int evaluator, uint64_t seed);   /* fills the pending batch */
/* the pending batch copied to HOST buffers (any may be NULL) / evaluator outputs taken from HOST
 * buffers [n_games][TAFL_MAX_MOVES] f32 and [n_games] f32 — for callers without a CUDA runtime */
int tafl_leaf_batch_host(tafl_engine* e, tafl_state* h_states, uint8_t* h_planes, tafl_move* h_moves,
                         uint32_t* h_counts);
int tafl_leaf_eval_host(tafl_engine* e, const float* h_priors, const float* h_values);
int tafl_mcts_search_builtin(tafl_engine* e, int evaluator, uint64_t seed);

/* ------------------------------------------------------------------------------------------
 * Self-play records (replay-buffer rows; nearest upstream analogue: write_to_file,
 * game/main.rs:86-132, with board_to_matrix :55-83 as the position encoding)
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    uint64_t game_id;
    uint32_t ply;
    uint16_t n_moves;
    uint8_t  side_to_play;      /* TAFL_ATTACKER | TAFL_DEFENDER */
    uint8_t  status;            /* status AFTER the chosen move   */
    uint64_t att[2], def[2];    /* position the search was run on */
    uint16_t best;              /* index of the move played       */
    uint16_t _pad;
    uint32_t episode;           /* how often this game id has been restarted (tafl_reset /
                                 * tafl_reset_finished): (game_id, episode) names one episode */
} tafl_record_header;           /* 56 bytes */
/* pack one record per resident game into a DEVICE buffer:
 * header + moves[TAFL_REC_MOVES] u16 + visits[TAFL_REC_MOVES] u32, fixed stride.  A row holds
 * every legal move of its root (TAFL_REC_MOVES == TAFL_MAX_MOVES: nothing is ever truncated). */
#define TAFL_REC_MOVES TAFL_MAX_MOVES
#define TAFL_REC_STRIDE (56 + 2 * TAFL_REC_MOVES + 4 * TAFL_REC_MOVES)
int tafl_pack_records(tafl_engine* e, void* d_out /* [n_games][TAFL_REC_STRIDE] */);
/* The same rows, but only for the games that were searched (Ongoing at search time, n_moves > 0),
 * compacted in game order on the device: row k of d_out is the k-th such game.  The number of rows
 * goes to d_count (DEVICE u32, may be NULL) and to h_count[0]; h_count[1] receives the overflow
 * flags of the search the rows come from (see tafl_mcts_overflow).  h_count (HOST u32[2], may be
 * NULL) is written when the work enqueued here has run — pass pinned memory and wait for the
 * stream / an event.  No host synchronisation: the replay-record path never stalls the next search. */
int tafl_pack_records_compact(tafl_engine* e, void* d_out /* [n_games][TAFL_REC_STRIDE] */,
                              uint32_t* d_count, uint32_t* h_count);
/* Restart every finished game from the engine's start position (episode counter + 1, history of
 * those games forgotten): the self-play sweep keeps all n_games trees busy.  *h_n (may be NULL)
 * receives the number of games restarted (forces a stream sync). */
int tafl_reset_finished(tafl_engine* e, uint32_t* h_n);
/* engine-wide MCTS overflow flags of the LAST search (0 = every tree ran all its simulations;
 * bit 0 edge pool, bit 1 node pool, bit 2 piece table); syncs the engine stream.  The asynchronous
 * search calls do not report TAFL_ERR_CAPACITY themselves: callers of tafl_mcts_search_async
 * check here before trusting the results. */
int tafl_mcts_overflow(tafl_engine* e, uint32_t* h_flags);


/* ------------------------------------------------------------------------------------------
 * Multi-GPU: games shard by id range, one process per GPU, and the ONLY
 * exchange is the gather of the replay records to the replay-buffer rank — the batched
 * counterpart of write_to_file collecting positions (game/main.rs:86-132).  NCCL is resolved
 * at run time (dlopen of libnccl.so.2, a copy already in the process wins; TAFL_NCCL_LIB
 * overrides), ncclUniqueId / ncclComm_t cross this ABI as 128 opaque bytes / a void*.
 *   rank 0: tafl_nccl_unique_id(id); ship the 128 bytes to the other ranks by any means;
 *   all   : tafl_nccl_comm_create(e, id, n_ranks, rank, &comm);
 *   per move: tafl_pack_records(e, d_rows); tafl_gather_records(e, comm, n_ranks, rank, root,
 *             d_rows, n_local, d_all, capacity_rows, h_counts).
 * Ranks may hold different numbers of rows: the counts are all-gathered into h_counts
 * [n_ranks] (HOST, valid on every rank), the rows then move with grouped ncclSend/ncclRecv to
 * their final offset in d_all on the root, rank order == game-id order.  Work is enqueued on
 * the engine's stream; the root's rows are complete after tafl_engine_sync. */
int tafl_nccl_unique_id(void* out128);
int tafl_nccl_comm_create(tafl_engine* e, const void* id128, int n_ranks, int rank, void** out_comm);
int tafl_nccl_comm_destroy(void* comm);
int tafl_gather_records(tafl_engine* e, void* comm, int n_ranks, int rank, int root,
                        const void* d_send /* [n_local][TAFL_REC_STRIDE] */, uint32_t n_local,
                        void* d_recv /* root only: [recv_capacity_rows][TAFL_REC_STRIDE] */,
                        uint64_t recv_capacity_rows, uint32_t* h_counts /* [n_ranks] */);
/* The same exchange enqueued on `stream` (a cudaStream_t of the engine's device; NULL = the engine's
 * stream) — the caller orders it behind tafl_pack_records with an event, so the gather of move k
 * overlaps the search of move k+1 on the engine's stream and the one host wait this call contains
 * (the all-gathered row counts) never stalls the search.  A rank other than the root may pass
 * h_counts == NULL: it then only enqueues (count exchange + its send) and returns without waiting. */
int tafl_gather_records_on(tafl_engine* e, void* comm, int n_ranks, int rank, int root,
                           const void* d_send, uint32_t n_local, void* d_recv,
                           uint64_t recv_capacity_rows, uint32_t* h_counts, void* stream);

/* ------------------------------------------------------------------------------------------
 * Large boards: BitfieldBoardState<U256> (row stride 15, up to 15x15) and <U512> (row stride 21,
 * up to 21x21) — LargeBasicBoardState / HugeBasicBoardState, game/board/state.rs:337-340,
 * game/bitfield.rs:180-181.  The reference instantiates its width-generic logic tests on them
 * (game/game/logic.rs:1015-1021,1081-1087); no preset or benchmark configuration uses them, so
 * this is the by-value GameLogic surface (logic.rs:61-880) only: no resident games, no MCTS.
 * Same bit layout as the reference: tile (r, c) = bit r * stride + c, king row / column in the top
 * four bits of the defender / attacker word.  Width 256 uses words 0..3 of each array, 512 all 8.
 * Tile indices no longer fit a tafl_move, so move lists are lists of tafl_play.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    uint64_t att[8], def[8];
    uint32_t recent[4];
    uint8_t  att_reps, def_reps, flags, status;
    uint32_t turn, plays_since_capture;
    uint8_t  ring_head, side_len, _pad[2];
} tafl_state_w;                 /* 160 bytes */
typedef struct { uint64_t captures[8]; uint8_t status, err, n_captures, _pad[5]; } tafl_play_effects_w;   /* 72 bytes */
typedef struct { uint64_t occupied[8], unoccupied[8], boundary[8]; uint8_t found, _pad[7]; } tafl_enclosure_w;   /* 200 bytes */
#define TAFL_W_MAX_PLAYS 2048   /* capacity of one position's play list: 4 (N^2 - 1) <= 1760 for N = 21 */

/* host helpers on wide states (width_bits 256 | 512): GameState::new(fen, side), Display, move_piece */
int tafl_w_state_from_fen(const char* fen, int width_bits, int side_to_play, tafl_state_w* out);
int tafl_w_state_to_fen(const tafl_state_w* st, int width_bits, char* buf, size_t buf_len);
int tafl_w_board_move_piece(tafl_state_w* st, int width_bits, uint8_t from_row, uint8_t from_col,
                            uint8_t to_row, uint8_t to_col);

/* GameLogic::new(rules, side_len) for a wide bitfield (logic.rs:70-72); device = CUDA ordinal */
typedef struct tafl_wide tafl_wide;
int tafl_wide_create(const tafl_rules* rules, int side_len, int width_bits, int device, tafl_wide** out);
int tafl_wide_destroy(tafl_wide* w);
const char* tafl_wide_last_error(const tafl_wide* w);
/* by-value batch calls, HOST arrays in and out, one CUDA kernel each (the wide twins of the calls above) */
int tafl_w_validate_plays(tafl_wide* w, const tafl_state_w* h_states, const tafl_play* h_plays, uint32_t n, uint8_t* h_err);
int tafl_w_do_plays(tafl_wide* w, const tafl_state_w* h_states, const tafl_play* h_plays, uint32_t n,
                    tafl_state_w* h_out_states, tafl_play_effects_w* h_out_fx);
int tafl_w_legal_plays_of(tafl_wide* w, const tafl_state_w* h_states, uint32_t n,
                          tafl_play* h_plays /* [n][TAFL_W_MAX_PLAYS] */, uint32_t* h_counts);
int tafl_w_piece_plays(tafl_wide* w, const tafl_state_w* h_states, const uint8_t* h_tiles_rc, uint32_t n,
                       tafl_play* h_plays /* [n][64] */, int32_t* h_counts /* -1 == Err(NoPiece) */);
int tafl_w_get_captures(tafl_wide* w, const tafl_state_w* h_states, const tafl_play* h_plays, uint32_t n, tafl_play_effects_w* h_out);
int tafl_w_detect_shieldwall(tafl_wide* w, const tafl_state_w* h_states, const tafl_play* h_plays, uint32_t n, tafl_play_effects_w* h_out);
int tafl_w_find_enclosure(tafl_wide* w, const tafl_state_w* h_states, const tafl_enclosure_query* h_q, uint32_t n, tafl_enclosure_w* h_out);
int tafl_w_enclosure_secure(tafl_wide* w, const tafl_state_w* h_states, const tafl_enclosure_query* h_q, uint32_t n, uint8_t* h_out);
int tafl_w_detect_exit_fort(tafl_wide* w, const tafl_state_w* h_states, uint32_t n, uint8_t* h_out);
int tafl_w_side_can_play(tafl_wide* w, const tafl_state_w* h_states, const uint8_t* h_sides, uint32_t n, uint8_t* h_out);

#ifdef __cplusplus
}
#endif
#endif /* TAFL_B200_H */
