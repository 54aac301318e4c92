"""CPU differential tests: the kernels' __host__ __device__ game logic (csrc/tafl_core.cuh,
tafl_select.cuh) compiled with g++ (tests/hostsim) against the oracle.  This is how the device
math is checked on a machine without a GPU; the GPU tests repeat the same comparisons through
the C ABI."""
import numpy as np
import pytest

import kat_runner
import oracle_ffi as F
from alphazeroforhnefatafl_b200 import _abi as A
from parity_cases import (PRESET_CONFIGS, board_for, compare_by_value, compare_wide, random_boards, sample_positions,
                          wide_random_boards)


def test_hostsim_reference_kats(orc, sim):
    n = kat_runner.run_all(F.OracleHost(), lambda rules, width, side_len: sim.by_value(rules, width))
    assert n >= 150


def test_hostsim_reference_kats_on_large_bitfields(orc, sim):
    """The reference's own unit tests on BitfieldBoardState<U256> / <U512> (row strides 15 / 21): the oracle first
    (its restatement of the reference at those widths must give the reference's answers), then the kernels' device
    functions (csrc/tafl_wide.cuh over tafl_core.cuh) compiled for the host."""
    n_orc = kat_runner.run_all(F.OracleHost(), lambda rules, width, side_len: orc.by_value(rules, width), widths=(256, 512))
    n_sim = kat_runner.run_all(F.OracleHost(), lambda rules, width, side_len: sim.by_value(rules, width), widths=(256, 512))
    assert n_orc == n_sim and n_orc >= 160


@pytest.mark.parametrize("width,N", [(256, 13), (256, 15), (512, 19), (512, 21), (512, 11)])
def test_hostsim_large_boards_vs_oracle(orc, sim, width, N):
    """15x15 / 21x21-class boards (and an 11x11 board in the 512-bit layout) under Copenhagen rules and a few rule
    variants: every by-value query of the device functions against the oracle, on random boards and on the
    positions reached by playing one legal play from each."""
    rng = np.random.default_rng(1000 + width + N)
    base = orc.preset_rules(A.PRESET_COPENHAGEN)
    for trial in range(3):
        rules = base if trial == 0 else base.copy(throne_movement=int(rng.integers(0, 5)), king_strength=int(rng.integers(0, 3)),
                                                   edge_escape=int(rng.integers(0, 2)), hostility_edge=int(rng.choice([0, 0xFFFF])),
                                                   linnaean_capture=int(rng.integers(0, 2)), enclosure_win=int(rng.integers(0, 3)))
        states = wide_random_boards(orc, rng, N, width, 60)
        nxt = compare_wide(orc.by_value(rules, width), sim.by_value(rules, width), states, N, rng)
        if len(nxt):
            compare_wide(orc.by_value(rules, width), sim.by_value(rules, width), nxt, N, rng)


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS)
def test_hostsim_random_playouts(orc, sim, name, preset, width):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    rc, st0 = orc.state_from_fen(fen, width, A.ATTACKER)
    for seed, n_games, n_steps, auto_reset in ((11, 96, 700, True), (12, 64, 500, False)):
        so, so_stats = orc.random_playout(rules, fen, width, seed, n_games, n_steps, auto_reset=auto_reset, n_threads=8)
        sh, sh_stats = sim.random_playout(rules, st0, width, seed, n_games, n_steps, auto_reset=auto_reset)
        assert so.tobytes() == sh.tobytes()
        assert so_stats.as_dict() == sh_stats.as_dict()


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS)
def test_hostsim_by_value_on_played_positions(orc, sim, name, preset, width):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    states = sample_positions(orc, rules, fen, width, 160, seed=21)
    rng = np.random.default_rng(5)
    assert compare_by_value(orc.by_value(rules, width), sim.by_value(rules, width), states, width, rng) > 0


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS)
def test_hostsim_by_value_on_synthetic_boards(orc, sim, name, preset, width):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    N = len(fen.split("/"))
    rng = np.random.default_rng(77)
    states = random_boards(orc, rng, N, width, 600)
    compare_by_value(orc.by_value(rules, width), sim.by_value(rules, width), states, width, rng)


def test_hostsim_rule_variants_on_synthetic_boards(orc, sim):
    """Non-preset Ruleset combinations: throne rules, king strength/attack, hostile
    edges, slow pieces, enclosure/edge-escape/shield-wall variants."""
    rng = np.random.default_rng(123)
    base = orc.preset_rules(A.PRESET_COPENHAGEN)
    sets = [0, 0xFFFF, A.pieceset_type(A.KING), A.pieceset_type(A.SOLDIER), A.pieceset_side(A.ATTACKER),
            A.pieceset_side(A.DEFENDER)]
    for trial in range(24):
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
        states = random_boards(orc, rng, N, width, 120)
        compare_by_value(orc.by_value(r, width), sim.by_value(r, width), states, width, rng)
        # and whole random games under these rules
        fen = orc.preset_board({7: A.PRESET_BRANDUBH, 9: A.PRESET_TABLUT, 11: A.PRESET_COPENHAGEN}[N])
        rc, st0 = orc.state_from_fen(fen, width, A.DEFENDER if r.starting_side else A.ATTACKER)
        so, so_stats = orc.random_playout(r, fen, width, trial, 24, 300, n_threads=8)
        sh, sh_stats = sim.random_playout(r, st0, width, trial, 24, 300)
        assert so.tobytes() == sh.tobytes(), trial
        assert so_stats.as_dict() == sh_stats.as_dict()


@pytest.mark.parametrize("name,preset,width", PRESET_CONFIGS[:3])
def test_hostsim_select_kth_move(orc, sim, name, preset, width):
    """The bit-sliced count/select (tafl_select.cuh) returns exactly the k-th entry of the
    canonical move list (game/main.rs:33-43 order)."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    states = sample_positions(orc, rules, fen, width, 200, seed=31)
    states = states[states["status"] == 0]
    moves, counts = orc.by_value(rules, width).legal_moves_of(states)
    rng = np.random.default_rng(9)
    for rep in range(6):
        ks = rng.integers(0, 1 << 30, len(states)).astype(np.uint32)
        got, got_counts = sim.select_moves(rules, width, states, ks)
        assert np.array_equal(got_counts, counts)
        nz = counts > 0
        exp = moves[np.arange(len(states))[nz], (ks[nz] % counts[nz])]
        assert np.array_equal(got[nz], exp)


@pytest.mark.parametrize("name,preset,width,spec", [("copenhagen", A.PRESET_COPENHAGEN, 128, 1),
                                                    ("tablut", A.PRESET_TABLUT, 128, 2),
                                                    ("brandubh", A.PRESET_BRANDUBH, 64, 3)])
def test_hostsim_preset_specialisation_equals_generic(orc, sim, name, preset, width, spec):
    """The compile-time preset specialisations (SpecTraits, tafl_core.cuh) are selected for the preset
    rules + board and give the same results as the generic run-time-flag code and as the oracle."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    rc, st0 = orc.state_from_fen(fen, width, A.ATTACKER)
    so, so_stats = orc.random_playout(rules, fen, width, 33, 128, 900, n_threads=8)
    used0, gen0 = sim.spec_used(spec), sim.spec_used(0)
    a, a_stats = sim.random_playout(rules, st0, width, 33, 128, 900)
    assert sim.spec_used(spec) == used0 + 1 and sim.spec_used(0) == gen0
    sim.force_generic(1)
    try:
        b, b_stats = sim.random_playout(rules, st0, width, 33, 128, 900)
        assert sim.spec_used(0) == gen0 + 1
        rng = np.random.default_rng(3)
        N = len(fen.split("/"))
        states = random_boards(orc, rng, N, width, 400)
        compare_by_value(orc.by_value(rules, width), sim.by_value(rules, width), states, width, rng)
    finally:
        sim.force_generic(0)
    assert a.tobytes() == b.tobytes() == so.tobytes()
    assert a_stats.as_dict() == b_stats.as_dict() == so_stats.as_dict()
    # a preset Ruleset on another board size must fall back to the generic code
    if preset == A.PRESET_COPENHAGEN:
        g0 = sim.spec_used(0)
        st7 = orc.state_from_fen(orc.preset_board(A.PRESET_BRANDUBH), 128, A.ATTACKER)[1]
        sim.random_playout(rules, st7, 128, 1, 4, 20)
        assert sim.spec_used(0) == g0 + 1


def test_hostsim_piece_table_equals_move_list(orc, sim):
    """The compressed per-piece move table of the MCTS node pools (pt_pack / pt_decode) expands to
    exactly the canonical legal-move list, for the presets and for random rule variants (throne
    rules, corner rights, slow pieces) on played and synthetic positions."""
    rng = np.random.default_rng(2024)
    def check(rules, width, states):
        mo, co = orc.by_value(rules, width).legal_moves_of(states)
        mt, ct = sim.piece_table_moves(rules, width, states)
        me, ce = sim.emit_moves(rules, width, states)   # the per-piece tile arithmetic of k_legal_moves_g
        assert np.array_equal(co, ct) and np.array_equal(co, ce)
        for i in range(len(states)):
            assert np.array_equal(mo[i, :co[i]], mt[i, :ct[i]]), i
            assert np.array_equal(mo[i, :co[i]], me[i, :ce[i]]), i
    for name, preset, width in PRESET_CONFIGS:
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        check(rules, width, sample_positions(orc, rules, fen, width, 150, seed=5))
        check(rules, width, random_boards(orc, rng, len(fen.split("/")), width, 300))
    base = orc.preset_rules(A.PRESET_COPENHAGEN)
    sets = [0, 0xFFFF, A.pieceset_type(A.KING), A.pieceset_type(A.SOLDIER), A.pieceset_side(A.ATTACKER), A.pieceset_side(A.DEFENDER)]
    for trial in range(40):
        r = base.copy(throne_movement=int(rng.integers(0, 5)), may_enter_corners=int(rng.choice(sets)),
                      slow_pieces=int(rng.choice([0, 0, A.pieceset_type(A.KING), 0xFFFF, A.pieceset_side(A.DEFENDER)])))
        N, width = [(7, 64), (9, 128), (11, 128), (7, 128)][trial % 4]
        check(r, width, random_boards(orc, rng, N, width, 150))


def test_hostsim_under_sanitizers():
    """The differential tests of this file once more, with the kernel math built with ASan + UBSan
    (-fno-sanitize-recover): out-of-range shifts, bad extract32 / bit() arguments, out-of-bounds walks
    abort the run.  compute-sanitizer is closed on the GPU pool, so this is the memory-safety check
    the shared __host__ __device__ code gets."""
    import os
    import subprocess
    import sys
    if os.environ.get("TAFL_HOSTSIM_SAN") == "1":
        pytest.skip("this IS the sanitizer run")
    rt = F.sanitizer_runtime()
    if rt is None:
        pytest.skip("g++ has no libasan.so on this machine")
    F.build_hostsim(sanitize=True)
    env = dict(os.environ, TAFL_HOSTSIM_SAN="1", LD_PRELOAD=rt,
               ASAN_OPTIONS="detect_leaks=0:abort_on_error=1", UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_hostsim.py"), "-x", "-q", "-p", "no:cacheprovider"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-4000:] + "\n" + r.stderr[-4000:]
