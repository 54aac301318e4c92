// kat_driver.cpp — runs the reference's unit tests (tests/golden/reference_kats.json, fed line by line by
// tests/test_cpp_host.py) through the C++ host mirror include/tafl.hpp, i.e. GameLogic / GameState / Game
// with the reference's names on top of the C ABI and the CUDA kernels.  Test infrastructure only.
//
//   stdin : one command per line (see the `if (op == ...)` chain)     stdout: one result line per query
//   `selftest_host` needs no GPU (text forms, presets, board accessors); everything else creates engines.
#include <algorithm>
#include <cstdio>
#include <iostream>
#include <map>
#include <sstream>

#include "tafl.hpp"

using namespace hnefatafl;

static Side side_of(const std::string& s) { return s == "D" ? Side::Defender : Side::Attacker; }

static std::string tiles_str(std::vector<Tile> t) {
    std::sort(t.begin(), t.end());
    std::string s;
    for (auto& x : t) s += (s.empty() ? "" : ",") + std::to_string(x.row) + "." + std::to_string(x.col);
    return s.empty() ? "-" : s;
}

static bool set_rule(Ruleset& r, const std::string& k, long v) {
#define F(name) if (k == #name) { r.name = (decltype(r.name))v; return true; }
    F(edge_escape) F(king_strength) F(king_attack) F(shieldwall_enabled) F(sw_corners_may_close) F(exit_fort)
    F(throne_movement) F(starting_side) F(sw_captures) F(may_enter_corners) F(hostility_throne) F(hostility_corners)
    F(hostility_edge) F(slow_pieces) F(enclosure_win) F(repetition_enabled) F(repetition_is_loss) F(draw_on_no_plays)
    F(linnaean_capture) F(n_repetitions)
#undef F
    return false;
}

static Play play_rc(std::istringstream& in) {
    int fr, fc, tr, tc;
    in >> fr >> fc >> tr >> tc;
    Play p;   // Play::from_tiles without the common-axis check, as the reference's tests build plays
    p.from = Tile{(uint8_t)fr, (uint8_t)fc};
    if (fr == tr) { p.axis = Axis::Horizontal; p.displacement = (int8_t)(tc - fc); }
    else { p.axis = Axis::Vertical; p.displacement = (int8_t)(tr - fr); }
    return p;
}

static int selftest_host() {
    auto t = Tile::from_str("a8");
    if (!t.is_ok() || !(t.unwrap() == Tile{7, 0}) || t.unwrap().to_string() != "a8") return 1;
    auto p = Play::from_str("d6-f6");
    if (!p.is_ok() || p.unwrap().to_string() != "d6-f6" || !(p.unwrap().to() == Tile{5, 5})) return 2;
    if (!Play::from_tiles(Tile{1, 1}, Tile{2, 2}).is_err()) return 3;
    if (!Play::from_str("zz").is_err()) return 4;
    auto g = GameState::create(preset::boards::BRANDUBH(), Side::Attacker, 64);
    if (!g.is_ok()) return 5;
    GameState s = g.unwrap();
    if (s.side_len() != 7 || !(s.get_king() == Tile{3, 3}) || !s.is_king(Tile{3, 3})) return 6;
    if (s.count_pieces(Side::Attacker) != 8 || s.count_pieces(Side::Defender) != 5) return 7;
    if (s.to_fen() != preset::boards::BRANDUBH()) return 8;
    auto pc = s.get_piece(Tile{3, 3});
    if (!pc || !(*pc == Piece{PieceType::King, Side::Defender}) || s.get_piece(Tile{0, 0})) return 9;
    s.move_piece(Tile{0, 3}, Tile{0, 0});
    if (!s.tile_occupied(Tile{0, 0}) || s.tile_occupied(Tile{0, 3})) return 10;
    bool panicked = false;
    try { s.move_piece(Tile{6, 6}, Tile{6, 5}); } catch (const std::logic_error&) { panicked = true; }
    if (!panicked) return 11;
    if (!GameState::create("3x3/7", Side::Attacker, 64).is_err()) return 12;
    if (GameState::create("7/6", Side::Attacker, 64).unwrap_err() != ParseError::BadLineLen) return 13;
    if (preset::rules::TABLUT().edge_escape != 1 || preset::rules::COPENHAGEN().exit_fort != 1) return 14;
    if (!(GameStatus::Win(WinReason::KingCaptured, Side::Attacker).win_reason() == WinReason::KingCaptured)) return 15;
    return 0;
}

// Game (game/mod.rs:76-116) end to end: do_play, PlayInvalid values, history, undo_last_play (the reference's test_undo)
static int selftest_game() {
    auto bad = Game::create(preset::rules::BRANDUBH(), "7/7/7/x", 64);
    if (!bad.is_err()) return 1;
    auto gr = Game::create(preset::rules::BRANDUBH(), preset::boards::BRANDUBH(), 64);
    if (!gr.is_ok()) return 2;
    Game g = gr.take();
    const GameState s0 = g.state();
    auto moves = g.get_all_possible_moves();
    if (moves.empty()) return 3;
    auto wrong = g.do_play(Play::from_tiles(Tile{3, 3}, Tile{3, 4}).unwrap());   // defender piece, attacker to move
    if (!wrong.is_err() || wrong.unwrap_err() != PlayInvalid::WrongPlayer) return 4;
    if (!(g.state() == s0) || !g.play_history().empty()) return 5;
    auto r1 = g.do_play(moves[0]);
    if (!r1.is_ok() || !r1.unwrap().is_ongoing()) return 6;
    const GameState s1 = g.state();
    if (s1 == s0 || s1.side_to_play() != Side::Defender || s1.turn() != s0.turn() + 1) return 7;
    auto r2 = g.do_play(g.get_all_possible_moves()[0]);
    if (!r2.is_ok()) return 8;
    auto hist = g.play_history();
    if (hist.size() != 2 || !(hist[0].play == moves[0]) || hist[0].side != Side::Attacker || hist[1].side != Side::Defender) return 9;
    g.undo_last_play();
    if (!(g.state() == s1) || g.play_history().size() != 1) return 10;
    g.undo_last_play();
    if (!(g.state() == s0)) return 11;
    g.undo_last_play();   // nothing left: unchanged, as in the reference
    if (!(g.state() == s0)) return 12;
    if (!g.iter_plays(Tile{0, 0}).is_err()) return 13;
    auto kp = g.iter_plays(Tile{0, 3});
    if (!kp.is_ok() || kp.unwrap().empty()) return 14;
    // batched: every game of a lockstep batch gets its own Result
    auto br = Games::create(preset::rules::COPENHAGEN(), preset::boards::COPENHAGEN(), 3, 128);
    if (!br.is_ok()) return 15;
    std::unique_ptr<Games> bp = br.take();
    Games& b = *bp;
    auto lp = b.legal_plays();
    if (lp[0].size() != 116) return 16;   // 116 entries at the Copenhagen start
    std::vector<Play> plays = {lp[0][0], Play::from_tiles(Tile{5, 5}, Tile{5, 6}).unwrap(), lp[2][115]};
    auto res = b.do_play(plays);
    if (!res[0].is_ok() || !res[1].is_err() || !res[2].is_ok()) return 17;
    b.step_random(3, 40);
    if (b.stats().steps != 120) return 18;
    if (b.restart_finished() > 3) return 19;   // Game::new for finished games: every game is Ongoing afterwards
    auto lp2 = b.legal_plays();
    if (lp2[0].empty() || lp2[1].empty() || lp2[2].empty()) return 20;
    auto res2 = b.do_play({lp2[0][0], lp2[1][0], lp2[2][0]}, /*restart_finished=*/true);
    if (!res2[0].is_ok() || !res2[1].is_ok() || !res2[2].is_ok()) return 21;
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::string(argv[1]) == "selftest_host") {
        int rc = selftest_host();
        std::printf(rc ? "selftest_host FAILED at %d\n" : "selftest_host ok\n", rc);
        return rc;
    }
    Ruleset rules = preset::rules::COPENHAGEN();
    std::map<std::string, int> presets = {{"COPENHAGEN", TAFL_PRESET_COPENHAGEN}, {"BRANDUBH", TAFL_PRESET_BRANDUBH},
                                          {"MAGPIE", TAFL_PRESET_MAGPIE}, {"TABLUT", TAFL_PRESET_TABLUT}, {"KOCH", TAFL_PRESET_KOCH}};
    std::unique_ptr<GameLogic> logic;
    GameState st;
    int width = 128;
    std::string line;
    try {
        while (std::getline(std::cin, line)) {
            std::istringstream in(line);
            std::string op;
            if (!(in >> op)) continue;
            if (op == "selftest_game") {
                int rc = selftest_game();
                if (rc) std::printf("selftest_game FAILED at %d\n", rc); else std::printf("selftest_game ok\n");
            } else if (op == "rules") {
                std::string name, kv;
                in >> name;
                rules = preset::rules::get(presets.at(name));
                while (in >> kv) {
                    auto eq = kv.find('=');
                    if (!set_rule(rules, kv.substr(0, eq), std::stol(kv.substr(eq + 1)))) { std::printf("error unknown rule %s\n", kv.c_str()); return 2; }
                }
            } else if (op == "logic") {
                int side_len;
                in >> width >> side_len;
                logic.reset(new GameLogic(rules, (uint8_t)side_len, width));
            } else if (op == "new") {
                std::string fen, side;
                in >> fen >> side;
                if (fen[0] == '@') fen = preset::boards::get(presets.at(fen.substr(1)));
                auto r = GameState::create(fen, side_of(side), width);
                if (r.is_err()) { std::printf("error parse %d\n", (int)r.unwrap_err()); return 2; }
                st = r.unwrap();
                std::printf("new side_len=%d\n", st.side_len());
            } else if (op == "set_side") {
                std::string side;
                in >> side;
                st.set_side_to_play(side_of(side));
            } else if (op == "move_piece") {
                int fr, fc, tr, tc;
                in >> fr >> fc >> tr >> tc;
                st.move_piece(Tile{(uint8_t)fr, (uint8_t)fc}, Tile{(uint8_t)tr, (uint8_t)tc});
            } else if (op == "validate") {
                auto r = logic->validate_play(play_rc(in), st);
                std::printf("validate %s\n", r.is_ok() ? "Ok" : to_string(r.unwrap_err()));
            } else if (op == "do_play" || op == "do_play_str") {
                Play p;
                if (op == "do_play_str") {
                    std::string s;
                    in >> s;
                    auto pr = Play::from_str(s);
                    if (pr.is_err()) { std::printf("error play parse\n"); return 2; }
                    p = pr.unwrap();
                } else {
                    p = play_rc(in);
                }
                auto r = logic->do_play(p, st);
                if (r.is_err()) {
                    std::printf("do_play err=%s\n", to_string(r.unwrap_err()));
                } else {
                    const DoPlayResult& d = r.unwrap();
                    std::printf("do_play status=%u state_status=%u captures=%s\n", d.effects.game_outcome.raw, d.new_state.status().raw,
                                tiles_str(d.effects.captures).c_str());
                    st = d.new_state;
                }
            } else if (op == "iter_plays") {
                int r, c;
                in >> r >> c;
                auto res = logic->iter_plays(Tile{(uint8_t)r, (uint8_t)c}, st);
                if (res.is_err()) { std::printf("iter_plays err\n"); continue; }
                std::vector<Tile> dst;
                bool from_ok = true;
                for (auto& p : res.unwrap()) { dst.push_back(p.to()); from_ok &= p.from == Tile{(uint8_t)r, (uint8_t)c}; }
                std::printf("iter_plays n=%zu from_ok=%d %s\n", dst.size(), (int)from_ok, tiles_str(dst).c_str());
            } else if (op == "get_captures") {
                std::printf("get_captures %s\n", tiles_str(logic->get_captures(play_rc(in), st)).c_str());
            } else if (op == "detect_shieldwall") {
                auto r = logic->detect_shieldwall(play_rc(in), st);
                std::printf("detect_shieldwall %s\n", r ? tiles_str(*r).c_str() : "none");
            } else if (op == "find_enclosure" || op == "enclosure_secure") {
                int r, c, enclosed, enclosing, ae, ac, ins = 0, outs = 0;
                in >> r >> c >> enclosed >> enclosing >> ae >> ac;
                if (op == "enclosure_secure") {
                    in >> ins >> outs;
                    auto s = logic->enclosure_secure(Tile{(uint8_t)r, (uint8_t)c}, (PieceSet)enclosed, (PieceSet)enclosing, ae, ac, ins, outs, st);
                    std::printf("enclosure_secure %s\n", !s ? "none" : (*s ? "1" : "0"));
                } else {
                    auto e = logic->find_enclosure(Tile{(uint8_t)r, (uint8_t)c}, (PieceSet)enclosed, (PieceSet)enclosing, ae, ac, st);
                    if (!e) std::printf("find_enclosure none\n");
                    else std::printf("find_enclosure occ=%s unocc=%s bnd=%s\n", tiles_str(e->occupied).c_str(), tiles_str(e->unoccupied).c_str(),
                                     tiles_str(e->boundary).c_str());
                }
            } else if (op == "exit_fort") {
                std::printf("exit_fort %d\n", (int)logic->detect_exit_fort(st));
            } else if (op == "can_play") {
                std::string side;
                in >> side;
                std::printf("can_play %d\n", (int)logic->side_can_play(side_of(side), st));
            } else {
                std::printf("error unknown op %s\n", op.c_str());
                return 2;
            }
            std::fflush(stdout);
        }
    } catch (const std::exception& ex) {
        std::printf("error exception %s\n", ex.what());
        return 3;
    }
    return 0;
}
