"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the
CPU oracle on the same seeded inputs, against the committed golden vectors, and — at the
BASELINE.json sizes — through size-independent properties.  Bit-exact: this is integer work."""
import hashlib
import json
import os

import numpy as np
import pytest

import kat_runner
from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import engine as E
from alphazeroforhnefatafl_b200.engine import TaflError
from parity_cases import (PRESET_CONFIGS, board_for, compare_by_value, compare_wide, random_boards, sample_positions,
                          wide_random_boards)
from product_host import GpuImplFactory, ProductHost

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def gpu(built_lib):
    f = GpuImplFactory(0)
    yield f
    f.close()


def test_gpu_reference_kats(gpu):
    """The reference's own unit tests (golden/reference_kats.json) through the CUDA path."""
    n = kat_runner.run_all(ProductHost(), gpu)
    assert n >= 150


def test_gpu_golden_playouts(built_lib):
    """Committed oracle playout vectors: end states (sha256) and statistics."""
    with open(os.path.join(GOLDEN, "playout_vectors.json")) as f:
        vectors = json.load(f)["vectors"]
    for v in vectors:
        rules = E.preset_rules(v["preset"])
        fen = E.preset_board(A.PRESET_BRANDUBH if v["preset"] == A.PRESET_KOCH else v["preset"])
        with E.LockstepGames(rules, fen, v["n_games"], 0, v["width"]) as g:
            g.set_game_id_base(v["game_id_base"])
            g.step_random(v["seed"], v["n_steps"], auto_reset=v["auto_reset"], steps_per_launch=1)
            st = g.get_states()
            assert hashlib.sha256(st.tobytes()).hexdigest() == v["states_sha256"], v["name"]
            assert [st[i].tobytes().hex() for i in range(2)] == v["first_states_hex"]
            assert g.stats().as_dict() == v["stats"], v["name"]


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS)
def test_gpu_random_playouts_vs_oracle(orc, built_lib, name, preset, width):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    for seed, n_games, n_steps, auto_reset, spl in ((41, 300, 500, True, 1), (42, 257, 333, False, 7), (43, 64, 900, True, 64)):
        so, so_stats = orc.random_playout(rules, fen, width, seed, n_games, n_steps, auto_reset=auto_reset,
                                          game_id_base=1000, n_threads=8)
        with E.LockstepGames(rules, fen, n_games, 0, width) as g:
            g.set_game_id_base(1000)
            g.step_random(seed, n_steps, auto_reset=auto_reset, steps_per_launch=spl)
            assert g.get_states().tobytes() == so.tobytes(), (name, seed)
            assert g.stats().as_dict() == so_stats.as_dict(), (name, seed)


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS)
def test_gpu_by_value_vs_oracle(orc, gpu, name, preset, width):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    N = len(fen.split("/"))
    rng = np.random.default_rng(1234)
    states = np.concatenate([sample_positions(orc, rules, fen, width, 128, seed=51),
                             random_boards(orc, rng, N, width, 800)])
    compare_by_value(orc.by_value(rules, width), gpu(rules, width, N), states, width, rng)


def test_gpu_rule_variants_vs_oracle(orc, gpu):
    rng = np.random.default_rng(321)
    base = orc.preset_rules(A.PRESET_COPENHAGEN)
    sets = [0, 0xFFFF, A.pieceset_type(A.KING), A.pieceset_type(A.SOLDIER), A.pieceset_side(A.ATTACKER),
            A.pieceset_side(A.DEFENDER)]
    for trial in range(12):
        r = base.copy(
            throne_movement=int(rng.integers(0, 5)), king_strength=int(rng.integers(0, 3)),
            king_attack=int(rng.integers(0, 3)), edge_escape=int(rng.integers(0, 2)),
            shieldwall_enabled=int(rng.integers(0, 2)), sw_corners_may_close=int(rng.integers(0, 2)),
            sw_captures=int(rng.choice(sets)), exit_fort=int(rng.integers(0, 2)),
            may_enter_corners=int(rng.choice(sets)), hostility_throne=int(rng.choice(sets)),
            hostility_corners=int(rng.choice(sets)), hostility_edge=int(rng.choice(sets)),
            slow_pieces=int(rng.choice([0, 0, A.pieceset_type(A.KING), 0xFFFF])),
            enclosure_win=int(rng.integers(0, 3)), repetition_enabled=int(rng.integers(0, 2)),
            repetition_is_loss=int(rng.integers(0, 2)), draw_on_no_plays=int(rng.integers(0, 2)),
            linnaean_capture=int(rng.integers(0, 2)), n_repetitions=int(rng.integers(1, 4)))
        N, width = [(7, 64), (9, 128), (11, 128), (7, 128)][trial % 4]
        states = random_boards(orc, rng, N, width, 200)
        compare_by_value(orc.by_value(r, width), gpu(r, width, N), states, width, rng)


def test_gpu_resident_legal_moves_and_step(orc, built_lib):
    """tafl_legal_moves_host + tafl_step (Game::do_play with validation) on resident games, driven
    move by move from the host, against the oracle's do_play."""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    n = 96
    rng = np.random.default_rng(8)
    ref = orc.by_value(rules, 128)
    with E.LockstepGames(rules, fen, n, 0, 128) as g:
        states = g.get_states()
        for ply in range(40):
            moves, counts = g.legal_moves()
            mr, cr = ref.legal_moves_of(states)
            assert np.array_equal(counts, cr)
            pick = (rng.integers(0, 1 << 30, n) % np.maximum(counts, 1)).astype(np.int64)
            chosen = moves[np.arange(n), pick]
            plays = E.moves_to_plays(chosen, 128)
            if ply % 5 == 4:   # inject an illegal play: the game must stay unchanged
                plays[0] = (0, 0, A.VERTICAL, 1)
            fx = g.step(plays)
            exp_states, exp_fx = ref.do_plays(states, plays)
            assert fx.tobytes() == exp_fx.tobytes()
            states = g.get_states()
            assert states.tobytes() == exp_states.tobytes()


def test_gpu_reset_and_set_states(orc, built_lib):
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    with E.LockstepGames(rules, fen, 50, 0, 64) as g:
        start = g.get_states()
        g.step_random(1, 30)
        moved = g.get_states()
        assert moved.tobytes() != start.tobytes()
        g.reset([3, 7])
        st = g.get_states()
        assert st[3].tobytes() == start[3].tobytes() and st[7].tobytes() == start[7].tobytes()
        assert st[4].tobytes() == moved[4].tobytes()
        g.set_states(moved[10:20], first=0)
        assert g.get_states(0, 10).tobytes() == moved[10:20].tobytes()
        g.reset()
        assert g.get_states().tobytes() == start.tobytes()


def test_gpu_full_size_properties(orc, built_lib):
    """BASELINE config 2 size (65 536 Copenhagen games): size-independent properties.
    (1) sharding: game ids [0, n) on one engine == two engines over disjoint id ranges;
    (2) steps_per_launch does not change results; (3) the first 64 games equal the oracle;
    (4) conservation: steps == n_games * n_steps under auto-reset; checksum is reproducible."""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    n, steps, seed = 65536, 200, 2024
    with E.LockstepGames(rules, fen, n, 0, 128) as g:
        g.step_random(seed, steps, steps_per_launch=1)
        full = g.get_states()
        s_full = g.stats()
        assert s_full.steps == n * steps
        g.reset(); g.stats_reset(); g.set_random_ply(0)
        g.step_random(seed, steps, steps_per_launch=8)
        assert g.get_states().tobytes() == full.tobytes()
        assert g.stats().state_checksum == s_full.state_checksum
    half = n // 2
    parts = []
    for base in (0, half):
        with E.LockstepGames(rules, fen, half, 0, 128) as g:
            g.set_game_id_base(base)
            g.step_random(seed, steps, steps_per_launch=4)
            parts.append(g.get_states())
    assert np.concatenate(parts).tobytes() == full.tobytes()
    so, _ = orc.random_playout(rules, fen, 128, seed, 64, steps, n_threads=8)
    assert so.tobytes() == full[:64].tobytes()
    so, _ = orc.random_playout(rules, fen, 128, seed, 32, steps, game_id_base=n - 32, n_threads=8)
    assert so.tobytes() == full[n - 32:].tobytes()


@pytest.mark.parametrize("name,preset,width,spec", [("copenhagen", A.PRESET_COPENHAGEN, 128, 1), ("tablut", A.PRESET_TABLUT, 128, 2),
                                                    ("brandubh", A.PRESET_BRANDUBH, 64, 3), ("brandubh_u128", A.PRESET_BRANDUBH, 128, 0),
                                                    ("koch", A.PRESET_KOCH, 64, 3)])   # Koch differs from Brandubh only in hostility MASKS (run-time data)
def test_gpu_preset_specialisation_equals_generic(orc, built_lib, name, preset, width, spec):
    """The preset-specialised kernels (compile-time rule flags) are picked for the preset rules + board,
    and both they and the generic kernels (forced) reproduce the oracle: playouts and MCTS."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    so, so_stats = orc.random_playout(rules, fen, width, 77, 512, 700, game_id_base=5, n_threads=8)
    roots = sample_positions(orc, rules, fen, width, 32, seed=88, max_steps=200)
    ref = orc.mcts_search(rules, width, roots, 200, n_threads=8)
    for generic in (False, True):
        with E.LockstepGames(rules, fen, 512, 0, width) as g:
            assert g.spec == spec
            if generic:
                g.force_generic(True)
                assert g.spec == 0
            g.set_game_id_base(5)
            g.step_random(77, 700)
            assert g.get_states().tobytes() == so.tobytes(), (name, generic)
            assert g.stats().as_dict() == so_stats.as_dict()
        with E.LockstepGames(rules, fen, len(roots), 0, width) as g:
            if generic:
                g.force_generic(True)
            g.set_states(roots)
            g.mcts_configure(200, 1.0, 1e-4)
            g.mcts_search()
            got = g.mcts_results()
            assert np.array_equal(got["visits"], ref["visits"]) and np.array_equal(got["moves"], ref["moves"])
            assert got["q"].tobytes() == ref["q"].tobytes()


def test_gpu_step_random_host_pipeline(orc, built_lib):
    """tafl_step_random_host (chunked, multi-stream H2D / kernel / D2H) == set_states + step_random +
    get_states == the oracle, for ragged sizes that do not divide into the chunk count."""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    for n in (1, 63, 1000, 8193):
        st, _ = orc.random_playout(rules, fen, 128, 5, n, 37, game_id_base=9, n_threads=8)
        exp, exp_stats = orc.random_playout(rules, fen, 128, 5, n, 3, game_id_base=9, ply0=37, n_threads=8, states=st.copy())
        with E.LockstepGames(rules, fen, n, 0, 128) as g:
            g.set_game_id_base(9)
            g.set_random_ply(37)
            got = g.step_random_host(st, 5, 3)
            assert got.tobytes() == exp.tobytes(), n
            assert g.get_states().tobytes() == exp.tobytes()
            assert g.stats().as_dict() == exp_stats.as_dict()
            assert g.random_ply == 40
            if n >= 63:   # a row for another board size is rejected on the device: that game keeps its resident state
                bad = exp.copy()
                bad["side_len"][[5, n - 1]] = 9
                keep = g.get_states()
                with pytest.raises(TaflError):
                    g.step_random_host(bad, 5, 1)
                after = g.get_states()
                assert after[[5, n - 1]].tobytes() == keep[[5, n - 1]].tobytes()
                ok = np.ones(n, bool)
                ok[[5, n - 1]] = False
                exp2, _ = orc.random_playout(rules, fen, 128, 5, n, 1, game_id_base=9, ply0=40, n_threads=8, states=exp.copy())
                assert after[ok].tobytes() == exp2[ok].tobytes()


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS[:3])
def test_gpu_recorded_actions_replay(orc, built_lib, name, preset, width):
    """The actions the random ply reports (tafl_actions_host) are the plays it applied: replaying them
    through Game::do_play (tafl_step_status, host buffers) from the start reproduces the same states,
    the same statuses, and the oracle agrees on every successor."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    n, plies = 700, 90
    with E.LockstepGames(rules, fen, n, 0, width) as a, E.LockstepGames(rules, fen, n, 0, width) as b:
        for ply in range(plies):
            before = a.get_states()
            a.step_random(123, 1, auto_reset=False)
            moves, status, stepped = a.last_actions()
            after = a.get_states()
            assert np.array_equal(status, after["status"])
            assert np.array_equal(stepped, before["status"] == 0)
            plays = E.moves_to_plays(moves, width)
            st, err = b.game_do_play(plays)
            assert np.array_equal(err == 0, stepped)
            assert np.array_equal(err[~stepped], np.full((~stepped).sum(), A.GAME_OVER, np.uint8))
            assert np.array_equal(st, status)
            assert b.get_states().tobytes() == after.tobytes(), ply
            if ply % 30 == 0:
                exp, fx = orc.by_value(rules, width).do_plays(before[stepped], plays[stepped])
                assert exp.tobytes() == after[stepped].tobytes()


@pytest.mark.gpu
def test_gpu_do_play_pinned_buffers_in_place(orc, built_lib, monkeypatch):
    """tafl_step_status with pinned caller buffers (the kernel reads the plays and writes the status /
    PlayInvalid bytes in place over PCIe) == pageable buffers (copy engine) == TAFL_HOST_IO=0, for ragged
    game counts, invalid plays included, with and without the PlayInvalid output; oracle on the successors."""
    import torch
    from alphazeroforhnefatafl_b200._lib import lib
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), board_for(orc, A.PRESET_COPENHAGEN)
    for n in (1, 65, 4100):
        with E.LockstepGames(rules, fen, n, 0, 128) as src:
            seq = []
            for ply in range(12):
                src.step_random(11, 1, auto_reset=False)
                pl = E.moves_to_plays(src.last_actions()[0], 128)
                if ply == 4:
                    pl.view(np.uint8).reshape(n, 4)[::7, 3] = 0      # zero displacement: invalid
                seq.append(pl)
        results = {}
        for label, mode, pinned in (("in_place", "3", True), ("copy", "0", True), ("pageable", "3", False)):
            monkeypatch.setenv("TAFL_HOST_IO", mode)
            with E.LockstepGames(rules, fen, n, 0, 128) as g:
                got = []
                for ply, pl in enumerate(seq):
                    before = g.get_states()
                    if pinned:
                        h_pl = torch.from_numpy(pl.view(np.uint8).reshape(n, 4).copy()).pin_memory()
                        h_st = torch.full((n + 8,), 0xEE, dtype=torch.uint8).pin_memory()    # guard bytes after the rows
                        h_er = torch.full((n + 8,), 0xEE, dtype=torch.uint8).pin_memory()
                        want_err = ply % 2 == 0
                        rc = lib().tafl_step_status(g.handle, h_pl.data_ptr(), h_st.data_ptr(), h_er.data_ptr() if want_err else None)
                        assert rc == 0
                        assert (h_st[n:] == 0xEE).all() and (h_er[n if want_err else 0:] == 0xEE).all()
                        st, er = h_st[:n].numpy().copy(), (h_er[:n].numpy().copy() if want_err else None)
                    else:      # plain numpy memory -> the copy-engine path; the last plies on an array pinned in place
                        pl2 = np.ascontiguousarray(pl).copy()
                        st, er = np.empty(n, np.uint8), np.empty(n, np.uint8)
                        reg = ply >= 9 and lib().tafl_host_register(A.ptr(pl2), pl2.nbytes) == 0
                        assert lib().tafl_step_status(g.handle, A.ptr(pl2), A.ptr(st), A.ptr(er)) == 0
                        if reg:
                            assert lib().tafl_host_unregister(A.ptr(pl2)) == 0
                        if ply == 11:   # and the Python mirror (pinned staging inside game_do_play) on a copy of the game
                            with E.LockstepGames(rules, fen, n, 0, 128) as g2:
                                g2.set_states(before)
                                st2, er2 = g2.game_do_play(pl)
                                assert st2.tobytes() == st.tobytes() and er2.tobytes() == er.tobytes()
                                assert g2.get_states().tobytes() == g.get_states().tobytes()
                    after = g.get_states()
                    got.append((st.tobytes(), None if er is None else er.tobytes(), after.tobytes()))
                    if label == "in_place" and er is not None:
                        ok = er == 0
                        exp, _ = orc.by_value(rules, 128).do_plays(before[ok], pl[ok])
                        assert exp.tobytes() == after[ok].tobytes()
                        assert after[~ok].tobytes() == before[~ok].tobytes()
                        if ply == 4:
                            assert (~ok).any()
                results[label] = got
        for a, b, c in zip(results["in_place"], results["copy"], results["pageable"]):
            assert a[0] == b[0] == c[0] and a[2] == b[2] == c[2]
            assert a[1] == b[1] and (a[1] is None or a[1] == c[1])


@pytest.mark.gpu
def test_gpu_graph_pdl_stepping_equals_plain_launches(orc, built_lib, monkeypatch):
    """n_steps single-ply launches as one CUDA graph with programmatic (PDL) edges == plain launches ==
    the oracle, across re-captures (changed n_steps / seed / game-id base) and launch-to-launch ply carry"""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), board_for(orc, A.PRESET_COPENHAGEN)
    n = 3000
    plan = [(5, 64), (5, 64), (5, 17), (9, 17), (9, 3), (9, 40)]     # (seed, n_steps); 3 < 4 -> plain path inside the graph engine
    finals = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("TAFL_STEP_GRAPH", mode)
        with E.LockstepGames(rules, fen, n, 0, 128) as g:
            g.set_game_id_base(77)
            for k, (seed, steps) in enumerate(plan):
                if k == 4:
                    g.set_game_id_base(1000)
                g.step_random(seed, steps, steps_per_launch=1)
            finals[mode] = (g.get_states(), g.stats().as_dict(), g.launches)
    assert finals["1"][0].tobytes() == finals["0"][0].tobytes()
    assert finals["1"][1] == finals["0"][1]
    assert finals["1"][2] == finals["0"][2] >= sum(s for _, s in plan)    # the launch count is the same claim either way
    # and against the oracle for the first two calls (same seed, consecutive plies)
    monkeypatch.setenv("TAFL_STEP_GRAPH", "1")
    with E.LockstepGames(rules, fen, n, 0, 128) as g:
        g.set_game_id_base(77)
        g.step_random(5, 64, steps_per_launch=1)
        g.step_random(5, 64, steps_per_launch=1)
        got = g.get_states()
    exp, _ = orc.random_playout(rules, fen, 128, 5, n, 128, game_id_base=77, n_threads=8)
    assert got.tobytes() == exp.tobytes()


def test_gpu_chained_step_graph_equals_whole_grid_dependencies(orc, built_lib, monkeypatch):
    """The step graph with CTA-to-CTA tokens (TAFL_STEP_CHAIN; default for grids of about one wave, forced here with 2)
    == the same graph with whole-grid dependencies == plain launches == the oracle: states, statistics and the
    actions of the last ply, over several graph launches, re-captures and a batch that is not a whole number of CTAs."""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), board_for(orc, A.PRESET_COPENHAGEN)
    for n, plan in ((65536 + 17, [(5, 64), (5, 64), (7, 33), (7, 64)]), (3000, [(5, 64), (5, 64), (9, 17), (9, 40)])):
        finals = {}
        for mode, env in (("chained", {"TAFL_STEP_CHAIN": "2"}), ("grid", {"TAFL_STEP_CHAIN": "0"}), ("plain", {"TAFL_STEP_GRAPH": "0"})):
            for k in ("TAFL_STEP_CHAIN", "TAFL_STEP_GRAPH"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            with E.LockstepGames(rules, fen, n, 0, 128) as g:
                g.set_game_id_base(77)
                for seed, steps in plan:
                    g.step_random(seed, steps, steps_per_launch=1)
                moves, status, stepped = g.last_actions()
                finals[mode] = (g.get_states().tobytes(), g.stats().as_dict(), moves.tobytes(), status.tobytes(), g.launches)
        assert finals["chained"] == finals["grid"] == finals["plain"], n
    for k in ("TAFL_STEP_CHAIN", "TAFL_STEP_GRAPH"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("TAFL_STEP_CHAIN", "2")
    with E.LockstepGames(rules, fen, 3000, 0, 128) as g:
        g.set_game_id_base(77)
        g.step_random(5, 64, steps_per_launch=1)
        g.step_random(5, 64, steps_per_launch=1)
        got = g.get_states()
    exp, _ = orc.random_playout(rules, fen, 128, 5, 3000, 128, game_id_base=77, n_threads=8)
    assert got.tobytes() == exp.tobytes()


def test_gpu_chained_step_graph_reports_a_lost_token(built_lib, monkeypatch):
    """A CTA that never sees its predecessor's token must not hang the GPU: the poll gives up, the other CTAs stop
    waiting, and the host call that synchronises next returns an error (the test hook makes CTA 0 of node 3 publish
    a wrong token); the engine works again afterwards."""
    import time
    rules, fen = E.preset.rules.COPENHAGEN, E.preset.boards.COPENHAGEN
    monkeypatch.setenv("TAFL_STEP_CHAIN", "2")
    monkeypatch.setenv("TAFL_STEP_CHAIN_LOSE", "3")
    with E.LockstepGames(rules, fen, 4096, 0, 128) as g:
        t0 = time.perf_counter()
        with pytest.raises(TaflError) as ei:
            g.step_random(1, 16, steps_per_launch=1)
        assert "token" in str(ei.value) and time.perf_counter() - t0 < 60.0
        g.reset()
        g.step_random(1, 3, steps_per_launch=1)          # fewer than four plies: plain launches, no chain
        assert g.stats().as_dict()["steps"] > 0


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS[:3])
def test_gpu_legal_move_kernels_agree(orc, built_lib, monkeypatch, name, preset, width):
    """get_all_possible_moves three ways — thread per game (TAFL_LM_G=1) and the warp-cooperative kernel with 8 and 16
    lanes per game — against the oracle: same counts, same moves, same canonical order; finished games list nothing;
    rows are written in whole 16-byte vectors whose unused tail is zero."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    n = 777                                   # not a multiple of any CTA's game count
    rc, st0 = orc.state_from_fen(fen, width, A.ATTACKER)
    states = np.concatenate([sample_positions(orc, rules, fen, width, 300, seed=13, max_steps=240),
                             random_boards(orc, np.random.default_rng(8), len(fen.split("/")), width, n - 300)])
    em, ec = orc.by_value(rules, width).legal_moves_of(states)
    assert (ec == 0).any() and (ec > 0).any()
    for G in ("1", "8", "16"):
        monkeypatch.setenv("TAFL_LM_G", G)
        with E.LockstepGames(rules, fen, n, 0, width) as g:
            g.set_states(states)
            mv, cnt = g.legal_moves()
            assert np.array_equal(cnt, ec), G
            for i in range(n):
                assert np.array_equal(mv[i, :cnt[i]], em[i, :ec[i]]), (G, i)
                if G != "1":
                    assert not mv[i, cnt[i]:(cnt[i] + 7) // 8 * 8].any(), (G, i)
            m2, c2 = g.legal_moves_of(states[:100])          # the by-value entry point runs the same kernel
            assert np.array_equal(c2, ec[:100])


def test_gpu_reference_kats_on_large_bitfields(gpu):
    """SURVEY §8f.4 / VERDICT r01 item 9: the reference's unit tests on BitfieldBoardState<U256> / <U512> (the widths its
    generic tests are instantiated at, game/game/logic.rs:1015-1021,1081-1087) through the CUDA path (tafl_w_*)."""
    n = kat_runner.run_all(ProductHost(), gpu, widths=(256, 512))
    assert n >= 160


@pytest.mark.parametrize("width,N", [(256, 13), (256, 15), (512, 19), (512, 21), (512, 11)])
def test_gpu_large_boards_vs_oracle(orc, built_lib, width, N):
    """15x15 / 21x21-class boards through the C ABI's tafl_w_* kernels against the oracle: every by-value query on
    random boards and on their successors, Copenhagen rules and rule variants; FEN text round trip on the host."""
    rng = np.random.default_rng(2000 + width + N)
    base = orc.preset_rules(A.PRESET_COPENHAGEN)
    for trial in range(3):
        rules = base if trial == 0 else base.copy(throne_movement=int(rng.integers(0, 5)), king_strength=int(rng.integers(0, 3)),
                                                   edge_escape=int(rng.integers(0, 2)), hostility_edge=int(rng.choice([0, 0xFFFF])),
                                                   linnaean_capture=int(rng.integers(0, 2)), enclosure_win=int(rng.integers(0, 3)))
        states = wide_random_boards(orc, rng, N, width, 200)
        with E.WideLogic(rules, N, width) as w:
            nxt = compare_wide(orc.by_value(rules, width), w, states, N, rng)
            if len(nxt):
                compare_wide(orc.by_value(rules, width), w, nxt, N, rng)
            for st in states[:10]:
                fen = E.WideLogic.state_to_fen(st, width)
                back = E.WideLogic.state_from_fen(fen, width, A.DEFENDER if st["flags"] & A.FLAG_DEF_TO_PLAY else A.ATTACKER)
                if N <= 16:   # a king beyond row / column 15 does not survive the four-bit nibbles (reference behaviour)
                    assert np.array_equal(back["att"][0], st["att"]) and np.array_equal(back["def"][0], st["def"])
    with pytest.raises(TaflError):
        E.WideLogic(base, 16, 256)          # a 16x16 board does not fit the 256-bit field
    with E.WideLogic(base, N, width) as w:
        bad = wide_random_boards(orc, rng, N - 1, width, 1)
        with pytest.raises(TaflError):
            w.validate_plays(bad, np.zeros(1, A.PLAY_DTYPE))     # side_len mismatch


def test_gpu_do_play_with_restart_equals_reset_then_do_play(orc, built_lib):
    """tafl_step_status_restart == tafl_reset_finished + tafl_step_status, ply for ply, over games that keep ending:
    states, status / PlayInvalid bytes, and the episode counters the replay records carry."""
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    n, plies = 512, 160
    rng = np.random.default_rng(17)
    with E.LockstepGames(rules, fen, n, 0, 64) as a, E.LockstepGames(rules, fen, n, 0, 64) as b:
        ended = 0
        for ply in range(plies):
            st = a.get_states()
            fresh = np.repeat(orc.state_from_fen(fen, 64, A.ATTACKER)[1], n)
            cur = np.where((st["status"] != 0)[:, None], fresh.view(np.uint8).reshape(n, 64), st.view(np.uint8).reshape(n, 64))
            cur = np.ascontiguousarray(cur).view(A.STATE_DTYPE).reshape(n)      # the positions the plays will meet
            mv, cnt = orc.by_value(rules, 64).legal_moves_of(cur)
            pick = mv[np.arange(n), rng.integers(0, 1 << 30, n) % np.maximum(cnt, 1)]
            plays = E.moves_to_plays(pick, 64)
            plays["disp"][::37] = 0                                              # a few invalid plays
            sa, ea = a.game_do_play(plays, restart_finished=True)
            b.reset_finished()
            sb, eb = b.game_do_play(plays)
            assert np.array_equal(sa, sb) and np.array_equal(ea, eb), ply
            assert a.get_states().tobytes() == b.get_states().tobytes(), ply
            ended += int((sa != 0).sum())
        assert ended > n        # every game ended at least once on average: the restart path was exercised
