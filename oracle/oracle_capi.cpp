// oracle_capi.cpp — extern "C" surface of the CPU oracle (liboracle.so) for ctypes.
// TEST INFRASTRUCTURE ONLY (see tafl_oracle.hpp).  Function names mirror include/tafl.h with an
// `orc_` prefix and no engine handle: (rules, width_bits) are passed explicitly.
#include <atomic>
#include <cstring>
#include <thread>

#include "mcts_oracle.hpp"

using namespace orc;

namespace {

template <class F>
int dispatch(int width, F&& f) {
    if (width == 64) return f((uint64_t)0);
    if (width == 128) return f((u128)0);
    return TAFL_ERR_INVALID_ARG;
}

template <class W>
void bb_set(uint64_t out[2], const std::set<Tile>& tiles) {
    out[0] = out[1] = 0;
    for (auto& t : tiles) {
        unsigned b = (unsigned)t.row * Board<W>::RW + t.col;
        out[b >> 6] |= 1ull << (b & 63);
    }
}

void add_outcome(tafl_stats* st, uint8_t status, size_t turn) {
    st->games_finished += 1;
    st->finished_plies += turn;
    if (TAFL_STATUS_IS_DRAW(status)) st->draw_reason[TAFL_STATUS_REASON(status) & 1] += 1;
    else st->win_reason[TAFL_STATUS_WINNER_IS_DEFENDER(status) ? 1 : 0][TAFL_STATUS_REASON(status)] += 1;
}

void merge_stats(tafl_stats* a, const tafl_stats& b) {
    a->steps += b.steps; a->legal_moves += b.legal_moves; a->captures += b.captures;
    a->games_finished += b.games_finished; a->finished_plies += b.finished_plies;
    for (int i = 0; i < 2; ++i) for (int j = 0; j < 8; ++j) a->win_reason[i][j] += b.win_reason[i][j];
    for (int j = 0; j < 2; ++j) a->draw_reason[j] += b.draw_reason[j];
}

template <class W>
void playout_range(const GameLogic& logic, const GameState<W>& start, uint64_t seed, uint64_t game_id_base,
                   uint32_t g0, uint32_t g1, uint64_t ply0, uint32_t n_steps, int auto_reset,
                   tafl_state* states, tafl_stats* stats) {
    for (uint32_t g = g0; g < g1; ++g) {
        GameState<W> s;
        unpack_state(states[g], &s);
        for (uint32_t i = 0; i < n_steps; ++i) {
            if (s.status != TAFL_STATUS_ONGOING) {
                if (!auto_reset) break;
                s = start;
            }
            auto moves = logic.get_all_possible_moves(s);
            if (moves.empty()) break;
            uint32_t k = pick_index(seed, game_id_base + g, ply0 + i, (uint32_t)moves.size());
            auto res = logic.do_valid_play(moves[k], s);
            s = res.first;
            stats->steps += 1;
            stats->legal_moves += moves.size();
            stats->captures += res.second.effects.captures.size();
            if (s.status != TAFL_STATUS_ONGOING) add_outcome(stats, s.status, s.turn);
        }
        pack_state(s, &states[g]);
    }
}

}  // namespace

extern "C" {

int orc_preset_rules(int id, tafl_rules* out) { return preset_rules(id, out); }
const char* orc_preset_board_fen(int id) { return preset_board(id); }

int orc_state_from_fen(const char* fen, int width, int side, tafl_state* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        int rc = GameState<W>::create(fen, side ? Side::Defender : Side::Attacker, &s);
        if (rc) return rc;
        pack_state(s, out);
        return 0;
    });
}
int orc_state_from_display_str(const char* str, int width, int side, tafl_state* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        int rc = Board<W>::from_display_str(str, &s.board);
        if (rc) return rc;
        s.side_to_play = side ? Side::Defender : Side::Attacker;
        pack_state(s, out);
        return 0;
    });
}
int orc_state_to_fen(const tafl_state* st, int width, char* buf, size_t len) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state(*st, &s);
        std::string f = s.board.to_fen();
        if (f.size() + 1 > len) return (int)TAFL_ERR_INVALID_ARG;
        std::memcpy(buf, f.c_str(), f.size() + 1);
        return 0;
    });
}
int orc_state_to_display_str(const tafl_state* st, int width, char* buf, size_t len) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state(*st, &s);
        std::string f = s.board.to_display_str();
        if (f.size() + 1 > len) return (int)TAFL_ERR_INVALID_ARG;
        std::memcpy(buf, f.c_str(), f.size() + 1);
        return 0;
    });
}

// board accessors (board/state.rs:13-81) on a packed state; op-coded to keep the surface small
// op: 0 get_king -> out[0]=row,out[1]=col | 1 set_piece(row,col,type,side) | 2 clear_tile(row,col)
//     3 get_piece(row,col) -> out[0]=type(0 none),out[1]=side | 4 tile_occupied -> out[0]
//     5 count_pieces(side=a) -> out[0] | 6 move_piece(row,col -> a,b) | 7 swap_pieces(row,col,a,b)
//     8 set_king(row,col)
int orc_board_op(tafl_state* st, int width, int op, int row, int col, int a, int b, int* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state(*st, &s);
        Tile t{(uint8_t)row, (uint8_t)col};
        switch (op) {
            case 0: { Tile k = s.board.get_king(); out[0] = k.row; out[1] = k.col; break; }
            case 1: s.board.set_piece(t, Piece{(PieceType)a, b ? Side::Defender : Side::Attacker}); break;
            case 2: s.board.clear_tile(t); break;
            case 3: { auto p = s.board.get_piece(t); out[0] = p ? (int)p->type : 0; out[1] = p ? (int)p->side : 0; break; }
            case 4: out[0] = s.board.tile_occupied(t); break;
            case 5: out[0] = s.board.count_pieces(a ? Side::Defender : Side::Attacker); break;
            case 6: if (!s.board.get_piece(t)) return (int)TAFL_ERR_NO_PIECE;
                    s.board.move_piece(t, Tile{(uint8_t)a, (uint8_t)b}); break;
            case 7: s.board.swap_pieces(t, Tile{(uint8_t)a, (uint8_t)b}); break;
            case 8: s.board.set_king(t); break;
            default: return (int)TAFL_ERR_INVALID_ARG;
        }
        pack_state(s, st);
        return 0;
    });
}
// iter_occupied (board/state.rs:202-216): out_rc = [row,col]*, returns count
int orc_iter_occupied(const tafl_state* st, int width, int side, uint8_t* out_rc) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state(*st, &s);
        auto v = s.board.iter_occupied(side ? Side::Defender : Side::Attacker);
        for (size_t i = 0; i < v.size(); ++i) { out_rc[2 * i] = v[i].row; out_rc[2 * i + 1] = v[i].col; }
        return (int)v.size();
    });
}

int orc_play_from_str(const char* s, tafl_play* out) {
    Play p;
    int rc = play_from_str(s, &p);
    if (rc) return rc;
    *out = play_to_pod(p);
    return 0;
}
int orc_play_to_str(const tafl_play* p, char* buf, size_t len) {
    std::string s = play_to_str(play_from_pod(*p));
    if (s.size() + 1 > len) return TAFL_ERR_INVALID_ARG;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 0;
}
int orc_tile_from_str(const char* s, uint8_t* row, uint8_t* col) {
    Tile t;
    int rc = tile_from_str(s, &t);
    if (rc) return rc;
    *row = t.row; *col = t.col;
    return 0;
}

int orc_validate_plays(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                       uint8_t* out_err) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out_err[i] = (uint8_t)logic.validate_play(play_from_pod(pl[i]), s);
        }
        return 0;
    });
}
int orc_do_plays(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                 tafl_state* out_st, tafl_play_effects* out_fx) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s, ns;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            PlayRecord rec;
            int err = logic.do_play(play_from_pod(pl[i]), s, &ns, &rec);
            tafl_play_effects fx{};
            fx.err = (uint8_t)err;
            if (err) { out_st[i] = st[i]; fx.status = st[i].status; }
            else {
                pack_state(ns, &out_st[i]);
                std::set<Tile> tiles;
                for (auto& c : rec.effects.captures) tiles.insert(c.tile);
                bb_set<W>(fx.captures, tiles);
                fx.n_captures = (uint8_t)tiles.size();
                fx.status = ns.status;
            }
            if (out_fx) out_fx[i] = fx;
        }
        return 0;
    });
}
int orc_legal_moves_of(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, tafl_move* moves,
                       uint32_t* counts) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto v = logic.get_all_possible_moves(s);
            counts[i] = (uint32_t)v.size();
            for (size_t k = 0; k < v.size() && k < TAFL_MAX_MOVES; ++k)
                moves[(size_t)i * TAFL_MAX_MOVES + k] = play_to_move<W>(v[k]);
        }
        return 0;
    });
}
int orc_piece_plays(const tafl_rules* r, int width, const tafl_state* st, const uint8_t* tiles_rc, uint32_t n,
                    tafl_move* moves, int32_t* counts) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            std::vector<Play> v;
            bool ok = logic.iter_plays(Tile{tiles_rc[2 * i], tiles_rc[2 * i + 1]}, s, &v);
            counts[i] = ok ? (int32_t)v.size() : -1;
            for (size_t k = 0; k < v.size() && k < 32; ++k) moves[(size_t)i * 32 + k] = play_to_move<W>(v[k]);
        }
        return 0;
    });
}
int orc_get_captures(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                     tafl_play_effects* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            Play p = play_from_pod(pl[i]);
            tafl_play_effects fx{};
            if (!s.board.get_piece(p.from)) { fx.err = TAFL_NO_PIECE; out[i] = fx; continue; }
            Piece mp = s.board.move_piece(p.from, p.to());
            auto caps = logic.get_captures(p, mp, s);
            std::set<Tile> tiles;
            for (auto& c : caps) tiles.insert(c.tile);
            bb_set<W>(fx.captures, tiles);
            fx.n_captures = (uint8_t)tiles.size();
            out[i] = fx;
        }
        return 0;
    });
}
int orc_detect_shieldwall(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                          tafl_play_effects* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto res = logic.detect_shieldwall(play_from_pod(pl[i]), s);
            tafl_play_effects fx{};
            if (!res) fx.n_captures = 0xFF;
            else { bb_set<W>(fx.captures, *res); fx.n_captures = (uint8_t)res->size(); }
            out[i] = fx;
        }
        return 0;
    });
}
int orc_find_enclosure(const tafl_rules* r, int width, const tafl_state* st, const tafl_enclosure_query* q,
                       uint32_t n, tafl_enclosure* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto e = logic.find_enclosure(Tile{q[i].row, q[i].col}, PieceSet{q[i].enclosed_pieces},
                                          PieceSet{q[i].enclosing_pieces}, q[i].abort_on_edge, q[i].abort_on_corner,
                                          s.board);
            tafl_enclosure o{};
            if (e) {
                o.found = 1;
                bb_set<W>(o.occupied, e->occupied);
                bb_set<W>(o.unoccupied, e->unoccupied);
                bb_set<W>(o.boundary, e->boundary);
            }
            out[i] = o;
        }
        return 0;
    });
}
int orc_enclosure_secure(const tafl_rules* r, int width, const tafl_state* st, const tafl_enclosure_query* q,
                         uint32_t n, uint8_t* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto e = logic.find_enclosure(Tile{q[i].row, q[i].col}, PieceSet{q[i].enclosed_pieces},
                                          PieceSet{q[i].enclosing_pieces}, q[i].abort_on_edge, q[i].abort_on_corner,
                                          s.board);
            out[i] = e ? (uint8_t)logic.enclosure_secure(*e, q[i].inside_safe, q[i].outside_safe, s.board) : 0xFF;
        }
        return 0;
    });
}
int orc_detect_exit_fort(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, uint8_t* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out[i] = logic.detect_exit_fort(s.board);
        }
        return 0;
    });
}
int orc_side_can_play(const tafl_rules* r, int width, const tafl_state* st, const uint8_t* sides, uint32_t n,
                      uint8_t* out) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out[i] = logic.side_can_play(sides[i] ? Side::Defender : Side::Attacker, s);
        }
        return 0;
    });
}
// RepetitionTracker::track_play (game/state.rs:92-113) on a packed state
int orc_track_play(tafl_state* st, int width, int side, const tafl_play* p, int captures) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state(*st, &s);
        s.repetitions.track_play(side ? Side::Defender : Side::Attacker, play_from_pod(*p), captures != 0);
        pack_state(s, st);
        return 0;
    });
}

// Random playouts, n_threads host threads over disjoint game ranges.
int orc_random_playout(const tafl_rules* r, const char* start_fen, int width, uint64_t seed, uint64_t game_id_base,
                       uint32_t n_games, uint64_t ply0, uint32_t n_steps, int auto_reset, int n_threads,
                       tafl_state* states, tafl_stats* stats) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> start;
        int rc = GameState<W>::create(start_fen, r->starting_side ? Side::Defender : Side::Attacker, &start);
        if (rc) return rc;
        GameLogic logic(*r, start.board.side_len);
        if (n_threads < 1) n_threads = 1;
        std::vector<tafl_stats> part((size_t)n_threads);
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) {
            uint32_t g0 = (uint32_t)((uint64_t)n_games * t / n_threads), g1 = (uint32_t)((uint64_t)n_games * (t + 1) / n_threads);
            part[t] = tafl_stats{};
            th.emplace_back([&, t, g0, g1] {
                playout_range<W>(logic, start, seed, game_id_base, g0, g1, ply0, n_steps, auto_reset, states, &part[t]);
            });
        }
        for (auto& x : th) x.join();
        for (auto& p : part) merge_stats(stats, p);
        return 0;
    });
}

// MCTS.getActionProb for n root states (fresh tree each), n_threads host threads.
static int mcts_search_impl(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                            double cpuct, double draw_value, int use_eval, uint64_t eval_seed, int n_threads,
                            tafl_move* moves, uint32_t* visits, double* q, uint32_t* counts, uint32_t* best,
                            tafl_mcts_stats* stats) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        if (n_threads < 1) n_threads = 1;
        std::vector<tafl_mcts_stats> part((size_t)n_threads);
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) {
            uint32_t g0 = (uint32_t)((uint64_t)n * t / n_threads), g1 = (uint32_t)((uint64_t)n * (t + 1) / n_threads);
            part[t] = tafl_mcts_stats{};
            th.emplace_back([&, t, g0, g1] {
                for (uint32_t g = g0; g < g1; ++g) {
                    GameState<W> s;
                    unpack_state(roots[g], &s);
                    GameLogic logic(*r, s.board.side_len);
                    Mcts<W> m(logic, cpuct, draw_value);
                    if (use_eval)
                        m.eval = [eval_seed](const GameState<W>& gs, float* P, float* v) {
                            tafl_state ps;
                            pack_state(gs, &ps);
                            hash_eval(eval_seed, ps, P, v);
                        };
                    m.get_action_prob(s, num_sims);
                    auto& root = m.nodes[0];
                    size_t L = root.moves.size();
                    counts[g] = (uint32_t)L;
                    uint32_t b = 0, bestn = 0;
                    for (size_t a = 0; a < L && a < TAFL_MAX_MOVES; ++a) {
                        size_t o = (size_t)g * TAFL_MAX_MOVES + a;
                        moves[o] = play_to_move<W>(root.moves[a]);
                        visits[o] = root.N[a];
                        q[o] = root.Q[a];
                        if (root.N[a] > bestn) { bestn = root.N[a]; b = (uint32_t)a; }
                    }
                    best[g] = b;
                    auto& ps = part[t];
                    ps.sims += m.stats.sims; ps.nodes += m.stats.nodes; ps.depth_sum += m.stats.depth_sum;
                    ps.edges += m.stats.edges; ps.terminal_hits += m.stats.terminal_hits;
                    uint64_t used = 0;
                    for (auto& nd : m.nodes) used += nd.moves.size();
                    if (used > ps.max_edges_used) ps.max_edges_used = used;
                }
            });
        }
        for (auto& x : th) x.join();
        if (stats) {
            *stats = tafl_mcts_stats{};
            for (auto& p : part) {
                stats->sims += p.sims; stats->nodes += p.nodes; stats->depth_sum += p.depth_sum;
                stats->edges += p.edges; stats->terminal_hits += p.terminal_hits;
                if (p.max_edges_used > stats->max_edges_used) stats->max_edges_used = p.max_edges_used;
            }
        }
        return 0;
    });
}

int orc_mcts_search(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                    double cpuct, double draw_value, int n_threads, tafl_move* moves, uint32_t* visits, double* q,
                    uint32_t* counts, uint32_t* best, tafl_mcts_stats* stats) {
    return mcts_search_impl(r, width, roots, n, num_sims, cpuct, draw_value, 0, 0, n_threads, moves, visits, q, counts, best,
                            stats);
}
// the same search with the TAFL_EVAL_HASH test evaluator (non-uniform priors, non-zero values)
int orc_mcts_search_eval(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                         double cpuct, double draw_value, uint64_t eval_seed, int n_threads, tafl_move* moves,
                         uint32_t* visits, double* q, uint32_t* counts, uint32_t* best, tafl_mcts_stats* stats) {
    return mcts_search_impl(r, width, roots, n, num_sims, cpuct, draw_value, 1, eval_seed, n_threads, moves, visits, q, counts,
                            best, stats);
}
// TAFL_EVAL_HASH for n packed states: P [n][TAFL_MAX_MOVES] f32, v [n] f32
int orc_hash_eval(const tafl_state* st, uint32_t n, uint64_t seed, float* P, float* v) {
    for (uint32_t i = 0; i < n; ++i) hash_eval(seed, st[i], P + (size_t)i * TAFL_MAX_MOVES, v + i);
    return 0;
}
double orc_np_sum(const double* a, uint64_t n) { return np_sum(a, (size_t)n); }

// An episode prefix with ONE MCTS object per game (tree reuse, see Mcts::advance) or a fresh tree
// per move: n_moves x (num_sims simulations from the current root, play the first-max move).
// Outputs per (game, move): [n][n_moves][TAFL_MAX_MOVES] visits / q, [n][n_moves] counts / best;
// final states [n].  use_eval != 0 -> TAFL_EVAL_HASH evaluator.
// `noise` (optional): root exploration noise [n][n_moves][TAFL_MAX_MOVES] f64, mixed in with weight eps
// (Mcts::noise_root; evaluator searches only).
}  // extern "C"
static int mcts_episode_impl(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                             uint32_t n_moves, int reuse, double cpuct, double draw_value, int use_eval, uint64_t eval_seed,
                             int n_threads, uint32_t* visits, double* q, uint32_t* counts, uint32_t* best,
                             tafl_state* final_states, tafl_mcts_stats* stats, const double* noise, double eps, uint32_t vl_width) {
    return dispatch(width, [&](auto w) {
        using W = decltype(w);
        if (n_threads < 1) n_threads = 1;
        std::vector<tafl_mcts_stats> part((size_t)n_threads);
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) {
            uint32_t g0 = (uint32_t)((uint64_t)n * t / n_threads), g1 = (uint32_t)((uint64_t)n * (t + 1) / n_threads);
            part[t] = tafl_mcts_stats{};
            th.emplace_back([&, t, g0, g1] {
                for (uint32_t g = g0; g < g1; ++g) {
                    GameState<W> s;
                    unpack_state(roots[g], &s);
                    GameLogic logic(*r, s.board.side_len);
                    Mcts<W> m(logic, cpuct, draw_value);
                    m.vl_width = vl_width ? vl_width : 1;
                    if (use_eval)
                        m.eval = [eval_seed](const GameState<W>& gs, float* P, float* v) {
                            tafl_state ps;
                            pack_state(gs, &ps);
                            hash_eval(eval_seed, ps, P, v);
                        };
                    m.reset(s);
                    for (uint32_t mv = 0; mv < n_moves; ++mv) {
                        const size_t row = (size_t)g * n_moves + mv;
                        counts[row] = 0;
                        best[row] = 0;
                        if (noise) {
                            m.noise_eps = eps;
                            m.root_noise.assign(noise + row * TAFL_MAX_MOVES, noise + (row + 1) * TAFL_MAX_MOVES);
                        }
                        if (s.status != TAFL_STATUS_ONGOING) {   // getActionProb on a finished game: every search returns -Es
                            m.run_from_root(num_sims);
                            continue;
                        }
                        m.run_from_root(num_sims);
                        auto& rt = m.nodes[m.root];
                        const size_t L = rt.moves.size();
                        counts[row] = (uint32_t)L;
                        uint32_t b = 0, bestn = 0;
                        for (size_t a = 0; a < L && a < TAFL_MAX_MOVES; ++a) {
                            visits[row * TAFL_MAX_MOVES + a] = rt.N[a];
                            q[row * TAFL_MAX_MOVES + a] = rt.Q[a];
                            if (rt.N[a] > bestn) { bestn = rt.N[a]; b = (uint32_t)a; }
                        }
                        best[row] = b;
                        if (L == 0) continue;
                        auto res = logic.do_valid_play(rt.moves[b], rt.state);
                        s = res.first;
                        if (!reuse || !m.advance((int)b)) m.reset(s);
                    }
                    pack_state(s, &final_states[g]);
                    auto& ps = part[t];
                    ps.sims += m.stats.sims; ps.nodes += m.stats.nodes; ps.depth_sum += m.stats.depth_sum;
                    ps.edges += m.stats.edges; ps.terminal_hits += m.stats.terminal_hits;
                }
            });
        }
        for (auto& x : th) x.join();
        if (stats) {
            *stats = tafl_mcts_stats{};
            for (auto& p : part) {
                stats->sims += p.sims; stats->nodes += p.nodes; stats->depth_sum += p.depth_sum;
                stats->edges += p.edges; stats->terminal_hits += p.terminal_hits;
            }
        }
        return 0;
    });
}

extern "C" {
int orc_mcts_episode_noise(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                           uint32_t n_moves, int reuse, double cpuct, double draw_value, int use_eval, uint64_t eval_seed,
                           int n_threads, uint32_t* visits, double* q, uint32_t* counts, uint32_t* best,
                           tafl_state* final_states, tafl_mcts_stats* stats, const double* noise, double eps) {
    return mcts_episode_impl(r, width, roots, n, num_sims, n_moves, reuse, cpuct, draw_value, use_eval, eval_seed, n_threads, visits,
                             q, counts, best, final_states, stats, noise, eps, 1);
}
// the stub-evaluator episode with `vl_width` simulations in flight per tree (Mcts::run_from_root_vl)
int orc_mcts_episode_vl(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                        uint32_t n_moves, int reuse, double cpuct, double draw_value, uint32_t vl_width, int n_threads,
                        uint32_t* visits, double* q, uint32_t* counts, uint32_t* best, tafl_state* final_states,
                        tafl_mcts_stats* stats) {
    return mcts_episode_impl(r, width, roots, n, num_sims, n_moves, reuse, cpuct, draw_value, 0, 0, n_threads, visits, q, counts,
                             best, final_states, stats, nullptr, 0.0, vl_width);
}

// virtual loss with the TAFL_EVAL_HASH evaluator (what the evaluator hook runs with vl_width > 1) and optional root noise
int orc_mcts_episode_vl_eval(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                             uint32_t n_moves, int reuse, double cpuct, double draw_value, uint32_t vl_width, int use_eval,
                             uint64_t eval_seed, int n_threads, uint32_t* visits, double* q, uint32_t* counts, uint32_t* best,
                             tafl_state* final_states, tafl_mcts_stats* stats, const double* noise, double eps) {
    return mcts_episode_impl(r, width, roots, n, num_sims, n_moves, reuse, cpuct, draw_value, use_eval, eval_seed, n_threads, visits,
                             q, counts, best, final_states, stats, noise, eps, vl_width);
}

int orc_mcts_episode(const tafl_rules* r, int width, const tafl_state* roots, uint32_t n, uint32_t num_sims,
                     uint32_t n_moves, int reuse, double cpuct, double draw_value, int use_eval, uint64_t eval_seed,
                     int n_threads, uint32_t* visits, double* q, uint32_t* counts, uint32_t* best, tafl_state* final_states,
                     tafl_mcts_stats* stats) {
    return orc_mcts_episode_noise(r, width, roots, n, num_sims, n_moves, reuse, cpuct, draw_value, use_eval, eval_seed, n_threads,
                                  visits, q, counts, best, final_states, stats, nullptr, 0.0);
}

int orc_hw_threads(void) { return (int)std::thread::hardware_concurrency(); }

// ---- large boards (U256 / U512): the by-value calls over tafl_state_w, mirrors of include/tafl.h's tafl_w_* ----------
}  // extern "C"
namespace {
template <class F>
int dispatch_w(int width, F&& f) {
    if (width == 256) return f(U256{});
    if (width == 512) return f(U512{});
    return TAFL_ERR_INVALID_ARG;
}
template <class W>
void bb_set_w(uint64_t out[8], const std::set<Tile>& tiles) {
    for (int i = 0; i < 8; ++i) out[i] = 0;
    for (auto& t : tiles) {
        unsigned b = (unsigned)t.row * Board<W>::RW + t.col;
        out[b >> 6] |= 1ull << (b & 63);
    }
}
}  // namespace
extern "C" {
int orc_w_state_from_fen(const char* fen, int width, int side, tafl_state_w* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        int rc = GameState<W>::create(fen, side ? Side::Defender : Side::Attacker, &s);
        if (rc) return rc;
        pack_state_w(s, out);
        return 0;
    });
}
int orc_w_move_piece(tafl_state_w* st, int width, int fr, int fc, int tr, int tc) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        GameState<W> s;
        unpack_state_w(*st, &s);
        Tile t{(uint8_t)fr, (uint8_t)fc};
        if (!s.board.get_piece(t)) return (int)TAFL_ERR_NO_PIECE;
        s.board.move_piece(t, Tile{(uint8_t)tr, (uint8_t)tc});
        pack_state_w(s, st);
        return 0;
    });
}
int orc_w_validate_plays(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, uint8_t* out_err) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out_err[i] = (uint8_t)logic.validate_play(play_from_pod(pl[i]), s);
        }
        return 0;
    });
}
int orc_w_do_plays(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_state_w* out_st,
                   tafl_play_effects_w* out_fx) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s, ns;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            PlayRecord rec;
            int err = logic.do_play(play_from_pod(pl[i]), s, &ns, &rec);
            tafl_play_effects_w fx{};
            fx.err = (uint8_t)err;
            if (err) { out_st[i] = st[i]; fx.status = st[i].status; }
            else {
                pack_state_w(ns, &out_st[i]);
                std::set<Tile> tiles;
                for (auto& c : rec.effects.captures) tiles.insert(c.tile);
                bb_set_w<W>(fx.captures, tiles);
                fx.n_captures = (uint8_t)tiles.size();
                fx.status = ns.status;
            }
            if (out_fx) out_fx[i] = fx;
        }
        return 0;
    });
}
int orc_w_legal_plays_of(const tafl_rules* r, int width, const tafl_state_w* st, uint32_t n, tafl_play* plays, uint32_t* counts) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto v = logic.get_all_possible_moves(s);
            counts[i] = (uint32_t)v.size();
            for (size_t k = 0; k < v.size() && k < TAFL_W_MAX_PLAYS; ++k) plays[(size_t)i * TAFL_W_MAX_PLAYS + k] = play_to_pod(v[k]);
        }
        return 0;
    });
}
int orc_w_piece_plays(const tafl_rules* r, int width, const tafl_state_w* st, const uint8_t* tiles_rc, uint32_t n, tafl_play* plays,
                      int32_t* counts) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            std::vector<Play> v;
            bool ok = logic.iter_plays(Tile{tiles_rc[2 * i], tiles_rc[2 * i + 1]}, s, &v);
            counts[i] = ok ? (int32_t)v.size() : -1;
            for (size_t k = 0; k < v.size() && k < 64; ++k) plays[(size_t)i * 64 + k] = play_to_pod(v[k]);
        }
        return 0;
    });
}
int orc_w_get_captures(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_play_effects_w* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            Play p = play_from_pod(pl[i]);
            tafl_play_effects_w fx{};
            if (!s.board.get_piece(p.from)) { fx.err = TAFL_NO_PIECE; out[i] = fx; continue; }
            Piece mp = s.board.move_piece(p.from, p.to());
            auto caps = logic.get_captures(p, mp, s);
            std::set<Tile> tiles;
            for (auto& c : caps) tiles.insert(c.tile);
            bb_set_w<W>(fx.captures, tiles);
            fx.n_captures = (uint8_t)tiles.size();
            out[i] = fx;
        }
        return 0;
    });
}
int orc_w_detect_shieldwall(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_play_effects_w* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto res = logic.detect_shieldwall(play_from_pod(pl[i]), s);
            tafl_play_effects_w fx{};
            if (!res) fx.n_captures = 0xFF;
            else { bb_set_w<W>(fx.captures, *res); fx.n_captures = (uint8_t)res->size(); }
            out[i] = fx;
        }
        return 0;
    });
}
int orc_w_find_enclosure(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_enclosure_query* q, uint32_t n, tafl_enclosure_w* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto e = logic.find_enclosure(Tile{q[i].row, q[i].col}, PieceSet{q[i].enclosed_pieces}, PieceSet{q[i].enclosing_pieces},
                                          q[i].abort_on_edge, q[i].abort_on_corner, s.board);
            tafl_enclosure_w o{};
            if (e) {
                o.found = 1;
                bb_set_w<W>(o.occupied, e->occupied);
                bb_set_w<W>(o.unoccupied, e->unoccupied);
                bb_set_w<W>(o.boundary, e->boundary);
            }
            out[i] = o;
        }
        return 0;
    });
}
int orc_w_enclosure_secure(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_enclosure_query* q, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            auto e = logic.find_enclosure(Tile{q[i].row, q[i].col}, PieceSet{q[i].enclosed_pieces}, PieceSet{q[i].enclosing_pieces},
                                          q[i].abort_on_edge, q[i].abort_on_corner, s.board);
            out[i] = e ? (uint8_t)logic.enclosure_secure(*e, q[i].inside_safe, q[i].outside_safe, s.board) : 0xFF;
        }
        return 0;
    });
}
int orc_w_detect_exit_fort(const tafl_rules* r, int width, const tafl_state_w* st, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out[i] = logic.detect_exit_fort(s.board) ? 1 : 0;
        }
        return 0;
    });
}
int orc_w_side_can_play(const tafl_rules* r, int width, const tafl_state_w* st, const uint8_t* sides, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto w) {
        using W = decltype(w);
        for (uint32_t i = 0; i < n; ++i) {
            GameState<W> s;
            unpack_state_w(st[i], &s);
            GameLogic logic(*r, s.board.side_len);
            out[i] = logic.side_can_play(sides[i] ? Side::Defender : Side::Attacker, s) ? 1 : 0;
        }
        return 0;
    });
}

}  // extern "C"
