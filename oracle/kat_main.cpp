// kat_main.cpp — the reference crate's own unit tests, run against the CPU oracle.
// TEST INFRASTRUCTURE ONLY.  Each block cites the reference test it carries over (FEN strings,
// tiles and expected values verbatim).  Exit code 0 = every known-answer test passed.
#include <cstdio>
#include <string>

#include "mcts_oracle.hpp"

using namespace orc;
using u64 = uint64_t;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                              \
    do {                                                                         \
        ++g_checks;                                                              \
        if (!(cond)) { ++g_fail; std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); } \
    } while (0)

static tafl_rules RULES(int id) { tafl_rules r; preset_rules(id, &r); return r; }
static Tile T(int r, int c) { return Tile{(uint8_t)r, (uint8_t)c}; }
static Play P(Tile a, Tile b) { return *Play::from_tiles(a, b); }
static Piece SOLDIER(Side s) { return Piece{PieceType::Soldier, s}; }
template <class W> static GameState<W> GS(const char* fen, Side side) {
    GameState<W> s;
    int rc = GameState<W>::create(fen, side, &s);
    CHECK(rc == 0);
    return s;
}
static std::set<Tile> TS(std::initializer_list<Tile> l) { return std::set<Tile>(l); }
template <class W> static std::set<Tile> plays_to(const GameLogic& lg, const GameState<W>& s, Tile t, bool* ok) {
    std::vector<Play> v;
    *ok = lg.iter_plays(t, s, &v);
    std::set<Tile> out;
    for (auto& p : v) { CHECK(p.from == t); out.insert(p.to()); }
    CHECK(out.size() == v.size());
    return out;
}

// board/state.rs:354-372  test_from_str
static void test_from_str() {
    Board<u64> a, b;
    CHECK(Board<u64>::from_fen("3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3", &a) == 0);
    CHECK(Board<u64>::from_display_str("...t...\n...t...\n...T...\nttTKTtt\n...T...\n...t...\n...t...", &b) == 0);
    CHECK(a == b);
}
// board/state.rs:375-404  test_piece_movement
static void test_piece_movement() {
    Board<u64> st;
    CHECK(Board<u64>::from_fen("3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3", &st) == 0);
    CHECK(st.get_king() == T(3, 3));
    st.set_piece(T(1, 5), SOLDIER(Side::Attacker));
    st.set_piece(T(4, 1), SOLDIER(Side::Defender));
    st.move_piece(T(3, 3), T(0, 4));
    CHECK(st.get_king() == T(0, 4));
    CHECK(st.to_fen() == "3tK2/3t1t1/3T3/ttT1Ttt/1T1T3/3t3/3t3");
    for (Tile t : {T(0, 3), T(2, 3), T(0, 4)}) CHECK(st.tile_occupied(t));
    for (Tile t : {T(3, 3), T(5, 4), T(1, 1)}) CHECK(!st.tile_occupied(t));
}
// board/state.rs:408-431  test_iter_occupied
static void test_iter_occupied() {
    Board<u64> st;
    Board<u64>::from_fen("3t3/3t3/3T3/ttTKTtt/3T3/3t3/3t3", &st);
    auto av = st.iter_occupied(Side::Attacker);
    auto dv = st.iter_occupied(Side::Defender);
    CHECK(std::set<Tile>(av.begin(), av.end()) ==
          TS({T(0, 3), T(1, 3), T(5, 3), T(6, 3), T(3, 0), T(3, 1), T(3, 5), T(3, 6)}));
    CHECK(std::set<Tile>(dv.begin(), dv.end()) == TS({T(2, 3), T(3, 3), T(4, 3), T(3, 2), T(3, 4)}));
    for (size_t i = 1; i < av.size(); ++i) CHECK(av[i - 1] < av[i]);  // ascending bit order (board/state.rs:94-101)
}
// board/state.rs:434-444  test_swap_pieces
static void test_swap_pieces() {
    Board<u64> b;
    CHECK(Board<u64>::from_fen("5/1K3/5/5/3t1", &b) == 0);
    CHECK(*b.get_piece(T(1, 1)) == KING);
    CHECK(*b.get_piece(T(4, 3)) == SOLDIER(Side::Attacker));
    CHECK(b.get_king() == T(1, 1));
    b.swap_pieces(T(1, 1), T(4, 3));
    CHECK(*b.get_piece(T(4, 3)) == KING);
    CHECK(*b.get_piece(T(1, 1)) == SOLDIER(Side::Attacker));
    CHECK(b.get_king() == T(4, 3));
}
// board/state.rs:447-451  test_count_pieces
static void test_count_pieces() {
    Board<u128> b;
    CHECK(Board<u128>::from_fen(preset_board(TAFL_PRESET_COPENHAGEN), &b) == 0);
    CHECK(b.count_pieces(Side::Attacker) == 24);
    CHECK(b.count_pieces(Side::Defender) == 13);
}
// board/geometry.rs:163-198  test_neighbors
static void test_neighbors() {
    BoardGeometry geo(7);
    auto n = geo.neighbors(T(0, 0));
    CHECK(std::set<Tile>(n.begin(), n.end()) == TS({T(0, 1), T(1, 0)}) && n.size() == 2);
    n = geo.neighbors(T(3, 2));
    CHECK(std::set<Tile>(n.begin(), n.end()) == TS({T(2, 2), T(3, 1), T(3, 3), T(4, 2)}) && n.size() == 4);
    CHECK((n == std::vector<Tile>{T(2, 2), T(4, 2), T(3, 1), T(3, 3)}));  // order up,down,left,right (:5)
    auto b = geo.tiles_between(T(2, 2), T(2, 5));
    CHECK(std::set<Tile>(b.begin(), b.end()) == TS({T(2, 3), T(2, 4)}) && b.size() == 2);
    b = geo.tiles_between(T(1, 3), T(4, 3));
    CHECK(std::set<Tile>(b.begin(), b.end()) == TS({T(2, 3), T(3, 3)}) && b.size() == 2);
    CHECK(geo.tiles_between(T(1, 1), T(3, 3)).empty());
    CHECK(geo.tiles_between(T(1, 1), T(1, 1)).empty());
    CHECK(geo.tiles_between(T(1, 1), T(1, 2)).empty());
}
// pieces.rs:282-320  test_piece_set (set algebra on the u16)
static void test_piece_set() {
    auto has = [](uint16_t v, PieceType t, Side s) { return PieceSet{v}.contains(Piece{t, s}); };
    uint16_t ps = TAFL_PIECESET_TYPE(TAFL_KING) | TAFL_PIECESET_TYPE(TAFL_SOLDIER) | TAFL_PIECESET_TYPE(TAFL_GUARD);
    for (Side s : {Side::Attacker, Side::Defender}) {
        CHECK(has(ps, PieceType::King, s)); CHECK(has(ps, PieceType::Soldier, s)); CHECK(has(ps, PieceType::Guard, s));
        CHECK(!has(ps, PieceType::Commander, s)); CHECK(!has(ps, PieceType::Knight, s)); CHECK(!has(ps, PieceType::Mercenary, s));
    }
    ps &= (uint16_t)~TAFL_PIECESET_PIECE(TAFL_KING, TAFL_ATTACKER);  // unset_piece
    CHECK(has(ps, PieceType::King, Side::Defender)); CHECK(!has(ps, PieceType::King, Side::Attacker));
    ps |= TAFL_PIECESET_PIECE(TAFL_COMMANDER, TAFL_DEFENDER);  // set_piece
    CHECK(has(ps, PieceType::Commander, Side::Defender)); CHECK(!has(ps, PieceType::Commander, Side::Attacker));
    CHECK(has(ps, PieceType::King, Side::Defender)); CHECK(!has(ps, PieceType::King, Side::Attacker));
    for (Side s : {Side::Attacker, Side::Defender}) {
        CHECK(has(ps, PieceType::Soldier, s)); CHECK(has(ps, PieceType::Guard, s));
        CHECK(!has(ps, PieceType::Knight, s)); CHECK(!has(ps, PieceType::Mercenary, s));
    }
}
// utils.rs:115-125  test_fixed_queue (ring semantics: push overwrites first, first advances)
static void test_fixed_queue() {
    RepetitionTracker tr;
    Play a = P(T(0, 0), T(0, 1));
    for (int i = 0; i < 4; ++i) { CHECK(tr.first_i == (size_t)i); tr.track_play(Side::Attacker, a, true); }
    CHECK(tr.first_i == 0);
    CHECK(tr.queue[tr.first_i].has_value());
}
// game/state.rs:165-187  test_repetition_tracker
static void test_repetition_tracker() {
    RepetitionTracker tr;
    auto PS = [](const char* s) { Play p; CHECK(play_from_str(s, &p) == 0); return p; };
    for (size_t i = 0; i < 5; ++i) {
        tr.track_play(Side::Attacker, PS("a1-b1"), false); CHECK(tr.get_repetitions(Side::Attacker) == i);
        tr.track_play(Side::Defender, PS("a2-b2"), false); CHECK(tr.get_repetitions(Side::Defender) == i);
        tr.track_play(Side::Attacker, PS("b1-a1"), false); CHECK(tr.get_repetitions(Side::Attacker) == i);
        tr.track_play(Side::Defender, PS("b2-a2"), false); CHECK(tr.get_repetitions(Side::Defender) == i);
    }
    for (size_t i = 0; i < 5; ++i) {
        tr.track_play(Side::Attacker, PS("f1-g1"), false); CHECK(tr.get_repetitions(Side::Attacker) == i);
        tr.track_play(Side::Defender, PS("f2-g2"), false); CHECK(tr.get_repetitions(Side::Defender) == i);
        tr.track_play(Side::Attacker, PS("g1-f1"), false); CHECK(tr.get_repetitions(Side::Attacker) == i);
        tr.track_play(Side::Defender, PS("g2-f2"), false); CHECK(tr.get_repetitions(Side::Defender) == i);
    }
}
// game/mod.rs:137-207  test_iter_plays
static void test_iter_plays() {
    tafl_rules r = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(r, 7);
    auto s = GS<u64>(preset_board(TAFL_PRESET_BRANDUBH), Side::Attacker);
    bool ok;
    plays_to(lg, s, T(0, 0), &ok); CHECK(!ok);
    plays_to(lg, s, T(1, 0), &ok); CHECK(!ok);
    CHECK(plays_to(lg, s, T(0, 3), &ok) == TS({T(0, 1), T(0, 2), T(0, 4), T(0, 5)})); CHECK(ok);
    CHECK(plays_to(lg, s, T(1, 3), &ok) == TS({T(1, 0), T(1, 1), T(1, 2), T(1, 4), T(1, 5), T(1, 6)})); CHECK(ok);
    CHECK(plays_to(lg, s, T(2, 3), &ok) == TS({T(2, 0), T(2, 1), T(2, 2), T(2, 4), T(2, 5), T(2, 6)})); CHECK(ok);
    CHECK(plays_to(lg, s, T(3, 3), &ok).empty()); CHECK(ok);
    auto s2 = GS<u64>("1T5/7/7/1t3K1/7/7/7", Side::Attacker);
    CHECK(plays_to(lg, s2, T(3, 1), &ok) ==
          TS({T(1, 1), T(2, 1), T(4, 1), T(5, 1), T(6, 1), T(3, 0), T(3, 2), T(3, 4)}));
    CHECK(ok);
}
// game/mod.rs:210-230  test_undo (history is host-side; here: states differ and are value types)
static void test_undo() {
    tafl_rules r = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(r, 7);
    auto s0 = GS<u64>(preset_board(TAFL_PRESET_BRANDUBH), Side::Attacker);
    GameState<u64> s1, s2, s3;
    CHECK(lg.do_play(P(T(0, 3), T(0, 2)), s0, &s1, nullptr) == 0);
    CHECK(lg.do_play(P(T(2, 3), T(2, 1)), s1, &s2, nullptr) == 0);
    CHECK(lg.do_play(P(T(1, 3), T(1, 1)), s2, &s3, nullptr) == 0);
    CHECK(!(s0.board == s1.board)); CHECK(!(s0.board == s2.board)); CHECK(!(s0.board == s3.board));
    CHECK(s3.turn == 3 && s0.turn == 0);
}

static tafl_rules TEST_RULES() {  // logic.rs:904-908
    tafl_rules r = RULES(TAFL_PRESET_BRANDUBH);
    r.slow_pieces = TAFL_PIECESET_TYPE(TAFL_KING);
    r.throne_movement = TAFL_THRONE_NOPASS;
    return r;
}
// logic.rs:923-1021  test_play_validity (u64 and u128 instantiations; U256/U512 are out of scope)
template <class W> static void test_play_validity() {
    tafl_rules br = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(br, 7);
    auto st = GS<W>(preset_board(TAFL_PRESET_BRANDUBH), Side::Attacker);
    CHECK(lg.validate_play(P(T(3, 1), T(4, 1)), st) == TAFL_PLAY_OK);
    CHECK(lg.validate_play(P(T(0, 3), T(0, 0)), st) == TAFL_MOVE_ONTO_BLOCKED_TILE);
    CHECK(lg.validate_play(P(T(1, 1), T(2, 1)), st) == TAFL_NO_PIECE);
    CHECK(lg.validate_play(P(T(0, 3), T(0, 7)), st) == TAFL_OUT_OF_BOUNDS);
    CHECK(lg.validate_play(P(T(0, 3), T(2, 3)), st) == TAFL_BLOCKED_BY_PIECE);
    GameState<W> ns;
    CHECK(lg.do_play(P(T(3, 1), T(4, 1)), st, &ns, nullptr) == 0);
    st = ns;
    CHECK(lg.validate_play(P(T(3, 3), T(3, 2)), st) == TAFL_BLOCKED_BY_PIECE);
    st.board.move_piece(T(3, 2), T(4, 2));
    st.board.move_piece(T(3, 3), T(3, 2));
    CHECK(lg.validate_play(P(T(2, 3), T(3, 3)), st) == TAFL_MOVE_ONTO_BLOCKED_TILE);
    CHECK(lg.validate_play(P(T(3, 2), T(3, 3)), st) == TAFL_PLAY_OK);

    tafl_rules tr = TEST_RULES();
    GameLogic lg2(tr, 7);
    auto st2 = GS<W>("7/5Tt/2T4/2t2t1/Tt4T/2t4/2T2K1", Side::Defender);
    CHECK(lg2.validate_play(P(T(6, 5), T(6, 3)), st2) == TAFL_TOO_FAR);
    CHECK(lg2.validate_play(P(T(6, 5), T(6, 4)), st2) == TAFL_PLAY_OK);
    st2.side_to_play = Side::Attacker;
    CHECK(lg2.validate_play(P(T(3, 2), T(3, 4)), st2) == TAFL_MOVE_THROUGH_BLOCKED_TILE);
}
// logic.rs:1023-1087  test_play_outcome
template <class W> static void test_play_outcome() {
    tafl_rules tr = TEST_RULES();
    GameLogic lg(tr, 7);
    auto proto = GS<W>("4t2/5Tt/2T4/2t2t1/Tt4T/2t4/2T2K1", Side::Attacker);
    auto caps_tiles = [](const std::set<PlacedPiece>& c) { std::set<Tile> o; for (auto& p : c) o.insert(p.tile); return o; };
    {
        auto st = proto;
        Play play = P(T(0, 4), T(6, 4));
        Piece piece = st.board.move_piece(play.from, play.to());
        auto caps = lg.get_captures(play, piece, st);
        CHECK(caps.size() == 1 && caps.begin()->tile == T(6, 5) && caps.begin()->piece == KING);
        st.board.move_piece(play.to(), play.from);
        GameState<W> ns;
        CHECK(lg.do_play(play, st, &ns, nullptr) == 0);
        CHECK(ns.status == Win(TAFL_WIN_KING_CAPTURED, Side::Attacker));
    }
    {
        auto st = proto;
        st.side_to_play = Side::Defender;
        Play play = P(T(4, 6), T(4, 2));
        Piece piece = st.board.move_piece(play.from, play.to());
        auto caps = lg.get_captures(play, piece, st);
        CHECK(caps_tiles(caps) == TS({T(4, 1), T(3, 2), T(5, 2)}));
        for (auto& c : caps) CHECK(c.piece == SOLDIER(Side::Attacker));
        st.board.move_piece(play.to(), play.from);
        GameState<W> ns;
        CHECK(lg.do_play(play, st, &ns, nullptr) == 0);
        CHECK(ns.status == TAFL_STATUS_ONGOING);
    }
    {
        auto st = proto;
        st.side_to_play = Side::Defender;
        Play play = P(T(6, 5), T(6, 6));
        Piece piece = st.board.move_piece(play.from, play.to());
        CHECK(lg.get_captures(play, piece, st).empty());
        st.board.move_piece(play.to(), play.from);
        GameState<W> ns;
        CHECK(lg.do_play(play, st, &ns, nullptr) == 0);
        CHECK(ns.status == Win(TAFL_WIN_KING_ESCAPED, Side::Defender));
    }
    {
        auto st = proto;
        st.side_to_play = Side::Defender;
        Play play = P(T(6, 5), T(5, 5));
        Piece piece = st.board.move_piece(play.from, play.to());
        CHECK(lg.get_captures(play, piece, st).empty());
        st.board.move_piece(play.to(), play.from);
        GameState<W> ns;
        CHECK(lg.do_play(play, st, &ns, nullptr) == 0);
        CHECK(ns.status == TAFL_STATUS_ONGOING);
    }
}
// logic.rs:1090-1169  test_shieldwalls
static void test_shieldwalls() {
    tafl_rules cop = RULES(TAFL_PRESET_COPENHAGEN);
    tafl_rules no_corner = cop;
    no_corner.sw_corners_may_close = 0; no_corner.sw_captures = TAFL_PIECESET_TYPE(TAFL_SOLDIER);
    tafl_rules king_cap = cop;
    king_cap.sw_corners_may_close = 0; king_cap.sw_captures = TAFL_PIECESET_ALL;
    const char* corner_sw = "9/9/9/9/6t2/7tT/7tT/7tT/9";
    const char* regular_sw = "9/9/9/6t2/7tT/7tT/7tT/8t/9";
    const char* regular_sw_king = "9/9/9/6t2/7tT/7tK/7tT/8t/9";
    const char* no_sw_gap = "9/9/9/6t2/7tT/8T/7tT/8t/9";
    const char* no_sw_friend = "9/9/9/6t2/7tT/6tTT/7tT/8t/9";
    const char* no_sw_small = "9/9/9/6t2/7tT/8t/9/9/9";
    Play cm = P(T(4, 6), T(4, 8)), m = P(T(3, 6), T(3, 8)), n = P(T(3, 6), T(3, 7));
    GameLogic corner_logic(cop, 9), no_corner_logic(no_corner, 9), regular_logic(no_corner, 9), king_cap_logic(king_cap, 9);
    auto corner_state = GS<u128>(corner_sw, Side::Attacker);
    CHECK(!corner_logic.detect_shieldwall(n, corner_state));
    auto w = corner_logic.detect_shieldwall(cm, corner_state);
    CHECK(w && *w == TS({T(5, 8), T(6, 8), T(7, 8)}));
    CHECK(!no_corner_logic.detect_shieldwall(m, corner_state));
    auto regular_state = GS<u128>(regular_sw, Side::Attacker);
    w = regular_logic.detect_shieldwall(m, regular_state);
    CHECK(w && *w == TS({T(4, 8), T(5, 8), T(6, 8)}));
    auto king_state = GS<u128>(regular_sw_king, Side::Attacker);
    w = regular_logic.detect_shieldwall(m, king_state);
    CHECK(w && *w == TS({T(4, 8), T(6, 8)}));
    w = king_cap_logic.detect_shieldwall(m, king_state);
    CHECK(w && *w == TS({T(4, 8), T(5, 8), T(6, 8)}));
    CHECK(!regular_logic.detect_shieldwall(m, GS<u128>(no_sw_gap, Side::Attacker)));
    CHECK(!regular_logic.detect_shieldwall(m, GS<u128>(no_sw_friend, Side::Attacker)));
    CHECK(!regular_logic.detect_shieldwall(m, GS<u128>(no_sw_small, Side::Attacker)));
}
// logic.rs:1172-1213  test_encl_secure
static void test_encl_secure() {
    const char* setup_1 = "7/2ttt2/1t1K1t1/2ttt2/7";
    const char* setup_2 = "7/1tttt2/1t1K1t1/2tttt1/7";
    const char* setup_3 = "2t1t2/1t1t1t1/1t1K1t1/2ttt2/7";
    const char* setup_4 = "2t2t1/1t3t1/1t1K1t1/2ttt2/7";
    tafl_rules cop = RULES(TAFL_PRESET_COPENHAGEN);
    tafl_rules safe = cop;
    safe.hostility_corners = safe.hostility_edge = safe.hostility_throne = TAFL_PIECESET_NONE;
    struct C { const char* s; bool in, out, expect; tafl_rules r; };
    C cand[] = {{setup_1, false, true, true, cop},  {setup_1, false, false, false, cop},
                {setup_2, false, true, true, cop},  {setup_2, true, false, true, cop},
                {setup_3, false, true, false, cop}, {setup_4, false, true, false, cop},
                {setup_4, false, true, true, safe}, {setup_4, true, false, true, cop}};
    for (auto& c : cand) {
        GameLogic lg(c.r, 7);
        auto st = GS<u64>(c.s, Side::Attacker);
        auto e = lg.find_enclosure(T(2, 3), PieceSet::of_type(PieceType::King),
                                   PieceSet::of_piece(SOLDIER(Side::Attacker)), false, false, st.board);
        CHECK(e.has_value());
        if (e) CHECK(lg.enclosure_secure(*e, c.in, c.out, st.board) == c.expect);
    }
}
// logic.rs:1216-1233  test_exit_forts
static void test_exit_forts() {
    tafl_rules cop = RULES(TAFL_PRESET_COPENHAGEN);
    GameLogic lg(cop, 9);
    for (const char* s : {"9/9/8t/7tT/7T1/6tT1/7TK/7tT/9", "9/9/9/9/9/5TTTT/5T2K/6TTT/9"})
        CHECK(lg.detect_exit_fort(GS<u128>(s, Side::Attacker).board));
    for (const char* s : {"9/9/9/8T/7Tt/7T1/7TK/8T/9", "9/9/9/8T/7TT/7TT/7TK/8T/9", "9/9/9/8T/9/4t2T1/7TK/8T/9",
                          "9/9/9/9/9/6TTT/5T2K/6TTT/9"})
        CHECK(!lg.detect_exit_fort(GS<u128>(s, Side::Attacker).board));
}
// logic.rs:1236-1385  test_enclosures
static void test_enclosures() {
    tafl_rules br = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(br, 7);
    PieceSet K = PieceSet::of_type(PieceType::King), S = PieceSet::of_type(PieceType::Soldier);
    PieceSet AS = PieceSet::of_piece(SOLDIER(Side::Attacker));
    Board<u64> b;
    Board<u64>::from_fen("2ttt2/1t1K1t1/2tttt1/7/7/7/7", &b);
    auto e = lg.find_enclosure(T(1, 3), K, S, true, true, b);
    CHECK(e.has_value());
    if (e) {
        CHECK(e->occupied == TS({T(1, 3)}));
        CHECK(e->unoccupied == TS({T(1, 2), T(1, 4)}));
        CHECK(e->boundary == TS({T(0, 2), T(0, 3), T(0, 4), T(1, 1), T(1, 5), T(2, 2), T(2, 3), T(2, 4)}));
    }
    Board<u64>::from_fen("2t1t2/1t1K1t1/2tttt1/7/7/7/7", &b);
    CHECK(!lg.find_enclosure(T(1, 3), K, S, true, true, b));
    e = lg.find_enclosure(T(1, 3), K, S, false, true, b);
    CHECK(e.has_value());
    if (e) {
        CHECK(e->occupied == TS({T(1, 3)}));
        CHECK(e->unoccupied == TS({T(0, 3), T(1, 2), T(1, 4)}));
        CHECK(e->boundary == TS({T(0, 2), T(0, 4), T(1, 1), T(1, 5), T(2, 2), T(2, 3), T(2, 4)}));
    }
    Board<u64>::from_fen("5t1/4tK1/4ttt/7/7/7/7", &b);
    CHECK(!lg.find_enclosure(T(1, 3), K, S, false, true, b));
    e = lg.find_enclosure(T(1, 5), K, S, false, false, b);
    CHECK(e.has_value());
    if (e) {
        CHECK(e->occupied == TS({T(1, 5)}));
        CHECK(e->unoccupied == TS({T(0, 6), T(1, 6)}));
        CHECK(e->boundary == TS({T(0, 5), T(1, 4), T(2, 5), T(2, 6)}));
    }
    Board<u64>::from_fen("2ttt2/1t1KTt1/2tttt1/7/7/7/7", &b);
    CHECK(!lg.find_enclosure(T(1, 3), K, AS, true, true, b));
    PieceSet KD{(uint16_t)(TAFL_PIECESET_PIECE(TAFL_KING, TAFL_DEFENDER) | TAFL_PIECESET_PIECE(TAFL_SOLDIER, TAFL_DEFENDER))};
    e = lg.find_enclosure(T(1, 3), KD, AS, true, true, b);
    CHECK(e.has_value());
    if (e) {
        CHECK(e->occupied == TS({T(1, 3), T(1, 4)}));
        CHECK(e->unoccupied == TS({T(1, 2)}));
        CHECK(e->boundary == TS({T(0, 2), T(0, 3), T(0, 4), T(1, 1), T(1, 5), T(2, 2), T(2, 3), T(2, 4)}));
    }
    Board<u64>::from_fen("1t2t2/1t1K1t1/2tttt1/7/7/7/7", &b);
    CHECK(lg.find_enclosure(T(1, 3), K, AS, false, false, b).has_value());
}
// logic.rs:1388-1403  test_can_play
static void test_can_play() {
    tafl_rules br = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(br, 7);
    auto s = GS<u64>("2tt3/1tTKt2/2tt3/7/7/7/7", Side::Attacker);
    CHECK(lg.side_can_play(Side::Attacker, s)); CHECK(!lg.side_can_play(Side::Defender, s));
    s = GS<u64>("2tKt2/3t3/7/7/7/7/7", Side::Attacker);
    CHECK(lg.side_can_play(Side::Attacker, s)); CHECK(!lg.side_can_play(Side::Defender, s));
}
// logic.rs:1406-1421  test_repetitions
static void test_repetitions() {
    tafl_rules br = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(br, 7);
    auto s = GS<u64>(preset_board(TAFL_PRESET_BRANDUBH), Side::Attacker);
    auto doit = [&](const char* str) {
        Play p; CHECK(play_from_str(str, &p) == 0);
        GameState<u64> ns; CHECK(lg.do_play(p, s, &ns, nullptr) == 0); s = ns;
    };
    for (int i = 0; i < 3; ++i) { doit("d6-f6"); doit("d5-f5"); doit("f6-d6"); doit("f5-d5"); }
    CHECK(s.status == TAFL_STATUS_ONGOING);
    doit("d6-f6");
    CHECK(s.status == Win(TAFL_WIN_REPETITION, Side::Defender));
}
// logic.rs:1424-1462  test_strong_king_capture
static void test_strong_king_capture() {
    tafl_rules br = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(br, 7);
    struct C { Tile a, b; const char* fen; bool captured; };
    C cs[] = {{T(3, 6), T(3, 5), "1T5/7/4t2/4K1t/4t2/7/7", true}, {T(1, 4), T(2, 4), "1T5/4t2/7/4Kt1/4t2/7/7", true},
              {T(3, 6), T(3, 5), "1T5/7/7/4K1t/4t2/7/7", false},  {T(1, 4), T(2, 4), "1T5/4t2/7/4K2/4t2/7/7", false}};
    for (auto& c : cs) {
        GameState<u64> ns; PlayRecord rec;
        CHECK(lg.do_play(P(c.a, c.b), GS<u64>(c.fen, Side::Attacker), &ns, &rec) == 0);
        if (c.captured) {
            CHECK(rec.effects.captures.size() == 1 && rec.effects.captures.begin()->tile == T(3, 4) &&
                  rec.effects.captures.begin()->piece == KING);
            CHECK(rec.effects.game_outcome && *rec.effects.game_outcome == Win(TAFL_WIN_KING_CAPTURED, Side::Attacker));
        } else {
            CHECK(rec.effects.captures.empty()); CHECK(!rec.effects.game_outcome);
        }
    }
}
// logic.rs:1465-1482  test_linnaean_capture
static void test_linnaean_capture() {
    tafl_rules tb = RULES(TAFL_PRESET_TABLUT);
    GameLogic lg(tb, 9);
    GameState<u128> ns; PlayRecord rec;
    CHECK(lg.do_play(P(T(4, 0), T(4, 2)), GS<u128>("tT7/9/9/4t4/t2TKt3/4t4/9/9/9", Side::Attacker), &ns, &rec) == 0);
    CHECK(rec.effects.captures.size() == 1 && rec.effects.captures.begin()->tile == T(4, 3) &&
          rec.effects.captures.begin()->piece == SOLDIER(Side::Defender));
}
// tiles.rs:226-235 test_tile_creation, :237-273 test_moves
static void test_tiles_moves() {
    for (int r = 0; r < 16; ++r) for (int c = 0; c < 16; ++c) { Tile t = T(r, c); CHECK(t.row == r && t.col == c); }
    auto p = Play::from_tiles(T(2, 4), T(2, 6));
    CHECK(p && p->from == T(2, 4) && p->axis == Axis::Horizontal && p->disp == 2 && p->to() == T(2, 6));
    p = Play::from_tiles(T(2, 3), T(5, 3));
    CHECK(p && p->from == T(2, 3) && p->axis == Axis::Vertical && p->disp == 3 && p->to() == T(5, 3));
    p = Play::from_tiles(T(1, 4), T(1, 1));
    CHECK(p && p->axis == Axis::Horizontal && p->disp == -3 && p->distance() == 3 && p->to() == T(1, 1));
    p = Play::from_tiles(T(7, 5), T(0, 5));
    CHECK(p && p->axis == Axis::Vertical && p->disp == -7 && p->to() == T(0, 5));
    CHECK(!Play::from_tiles(T(2, 3), T(3, 6)));
}
// tiles.rs:276-292 test_parsing_tiles, :295-324 test_parsing_moves
static void test_parsing() {
    Tile t;
    CHECK(tile_from_str("a8", &t) == 0 && t == T(7, 0)); CHECK(tile_to_str(T(7, 0)) == "a8");
    CHECK(tile_from_str("f14", &t) == 0 && t == T(13, 5)); CHECK(tile_to_str(T(13, 5)) == "f14");
    CHECK(tile_from_str("", &t) == TAFL_ERR_PARSE_EMPTY);
    CHECK(tile_from_str("[53", &t) == TAFL_ERR_PARSE_BAD_CHAR);
    CHECK(tile_from_str("a!!", &t) == TAFL_ERR_PARSE_BAD_INT);
    Play m;
    CHECK(play_from_str("a8-a11", &m) == 0 && m == P(T(7, 0), T(10, 0))); CHECK(play_to_str(m) == "a8-a11");
    CHECK(play_from_str("f5-d5", &m) == 0 && m == P(T(4, 5), T(4, 3))); CHECK(play_to_str(m) == "f5-d5");
    CHECK(play_from_str("f5-d6", &m) == TAFL_ERR_PARSE_BAD_PLAY);
    CHECK(play_from_str("f5-d7-d6", &m) == TAFL_ERR_PARSE_BAD_STRING);
    CHECK(play_from_str("f5-d]", &m) == TAFL_ERR_PARSE_BAD_INT);
    CHECK(play_from_str("!5-d5", &m) == TAFL_ERR_PARSE_BAD_CHAR);
}
// SURVEY §6 sizing facts (derived from preset.rs): start-position legal-move counts
static void test_start_counts() {
    struct C { int preset; int n; int att, def; };
    for (C c : {C{TAFL_PRESET_COPENHAGEN, 11, 116, 60}, C{TAFL_PRESET_TABLUT, 9, 80, 56}}) {
        tafl_rules r = RULES(c.preset);
        GameLogic lg(r, (uint8_t)c.n);
        auto s = GS<u128>(preset_board(c.preset), Side::Attacker);
        CHECK((int)lg.get_all_possible_moves(s).size() == c.att);
        s.side_to_play = Side::Defender;
        CHECK((int)lg.get_all_possible_moves(s).size() == c.def);
    }
    tafl_rules r = RULES(TAFL_PRESET_BRANDUBH);
    GameLogic lg(r, 7);
    auto s = GS<u64>(preset_board(TAFL_PRESET_BRANDUBH), Side::Attacker);
    CHECK(lg.get_all_possible_moves(s).size() == 40);
    s.side_to_play = Side::Defender;
    CHECK(lg.get_all_possible_moves(s).size() == 24);
}

int main() {
    test_from_str(); test_piece_movement(); test_iter_occupied(); test_swap_pieces(); test_count_pieces();
    test_neighbors(); test_piece_set(); test_fixed_queue(); test_repetition_tracker(); test_iter_plays();
    test_undo();
    test_play_validity<u64>(); test_play_validity<u128>();
    test_play_outcome<u64>(); test_play_outcome<u128>();
    test_shieldwalls(); test_encl_secure(); test_exit_forts(); test_enclosures(); test_can_play();
    test_repetitions(); test_strong_king_capture(); test_linnaean_capture(); test_tiles_moves(); test_parsing();
    test_start_counts();
    std::printf("oracle KAT: %d checks, %d failed\n", g_checks, g_fail);
    return g_fail ? 1 : 0;
}
