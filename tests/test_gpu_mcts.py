"""GPU parity tests of the batched MCTS (tafl_mcts.cu) against the CPU MCTS oracle
(oracle/mcts_oracle.hpp, a restatement of the reference's src/mcts.py with the stub evaluator).
Visit counts, move lists, best move and search statistics are bit-exact; Q within 1e-6 relative
(BASELINE.json north_star) — in practice also bit-exact because the f64 operation order is the
oracle's and FMA contraction is disabled."""
import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import engine as E
from alphazeroforhnefatafl_b200._lib import TaflError
from parity_cases import board_for, sample_positions

pytestmark = pytest.mark.gpu
Q_RTOL = 1e-6


def check_search(orc, rules, fen, width, roots, sims, cpuct=1.0, draw_value=1e-4):
    n = len(roots)
    ref = orc.mcts_search(rules, width, roots, sims, cpuct=cpuct, draw_value=draw_value, n_threads=8)
    with E.LockstepGames(rules, fen, n, 0, width) as g:
        g.set_states(roots)
        g.mcts_configure(sims, cpuct, draw_value)
        g.mcts_search()
        got = g.mcts_results()
        stats = g.mcts_stats().as_dict()
        assert np.array_equal(got["counts"], ref["counts"])
        for i in range(n):
            c = int(ref["counts"][i])
            assert np.array_equal(got["moves"][i, :c], ref["moves"][i, :c]), i
            assert np.array_equal(got["visits"][i, :c], ref["visits"][i, :c]), i
            assert np.allclose(got["q"][i, :c], ref["q"][i, :c], rtol=Q_RTOL, atol=0.0), i
            assert not got["visits"][i, c:].any()
        assert np.array_equal(got["best"], ref["best"])
        assert stats == ref["stats"].as_dict()
        q_exact = all(np.array_equal(got["q"][i, :ref["counts"][i]], ref["q"][i, :ref["counts"][i]]) for i in range(n))
        # play the chosen move: equals the oracle's do_play of moves[best]
        status = g.mcts_play_best()
        after = g.get_states()
        ongoing = (roots["status"] == 0) & (ref["counts"] > 0)
        plays = E.moves_to_plays(ref["moves"][np.arange(n), ref["best"]], width)
        exp, fx = orc.by_value(rules, width).do_plays(roots[ongoing], plays[ongoing])
        assert not fx["err"].any()
        assert after[ongoing].tobytes() == exp.tobytes()
        assert after[~ongoing].tobytes() == roots[~ongoing].tobytes()
        assert np.array_equal(status, after["status"])
    return ref, q_exact


@pytest.mark.parametrize("name,preset,width,sims", [
    ("copenhagen", A.PRESET_COPENHAGEN, 128, 800), ("tablut", A.PRESET_TABLUT, 128, 400),
    ("brandubh", A.PRESET_BRANDUBH, 64, 800), ("brandubh_u128", A.PRESET_BRANDUBH, 128, 150),
    ("koch", A.PRESET_KOCH, 64, 300), ("magpie", A.PRESET_MAGPIE, 64, 100)])
def test_gpu_mcts_vs_oracle(orc, built_lib, name, preset, width, sims):
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    roots = sample_positions(orc, rules, fen, width, 96, seed=61, max_steps=240)
    ref, q_exact = check_search(orc, rules, fen, width, roots, sims)
    assert q_exact, "Q values differ in the last bits"
    # mcts.py:83-102: a fresh Ongoing root spends its first simulation on expansion
    og = (roots["status"] == 0) & (ref["counts"] > 0)
    assert np.array_equal(ref["visits"][og].sum(1), np.full(og.sum(), sims - 1))


def test_gpu_mcts_terminal_values(orc, built_lib):
    """Small boards late in the game: terminal leaves give non-zero Q (wins, losses, draws)."""
    rules, fen = orc.preset_rules(A.PRESET_TABLUT), orc.preset_board(A.PRESET_TABLUT)
    roots = sample_positions(orc, rules, fen, 128, 128, seed=71, max_steps=400)
    ref, q_exact = check_search(orc, rules, fen, 128, roots, 500, cpuct=1.7, draw_value=0.25)
    assert q_exact
    assert ref["stats"].terminal_hits > 0
    assert np.abs(ref["q"]).max() > 0


def test_gpu_mcts_capacity_error(orc, built_lib):
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    with E.LockstepGames(rules, fen, 8, 0, 128) as g:
        g.mcts_configure(64, 1.0, 1e-4, edge_capacity=500)
        with pytest.raises(TaflError) as ei:
            g.mcts_search()
        assert ei.value.code == A.ERR_CAPACITY


def test_gpu_overflow_flag_is_per_search(orc, built_lib):
    """the engine-wide overflow word used to stay set for ever.  A search that exhausts a pool
    reports TAFL_ERR_CAPACITY with the pool named; the next search on the same engine, with room, is clean."""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    with E.LockstepGames(rules, fen, 8, 0, 128) as g:
        roots = g.get_states()
        g.mcts_configure(64, 1.0, 1e-4, edge_capacity=500)
        with pytest.raises(TaflError) as ei:
            g.mcts_search()
        assert ei.value.code == A.ERR_CAPACITY and "edge pool" in str(ei.value)
        g.mcts_search(sync=False)                     # the asynchronous call does not report ...
        assert g.mcts_overflow() == 1                 # ... tafl_mcts_overflow does (bit 0: edge pool)
        g.mcts_configure(64, 1.0, 1e-4, edge_capacity=500, tree_reuse=False, node_capacity=65)   # same pools: no reallocation
        with pytest.raises(TaflError):
            g.mcts_search()
        g.mcts_configure(64, 1.0, 1e-4)               # pools with room
        g.mcts_search()                               # TAFL_OK
        assert g.mcts_overflow() == 0
        exp = orc.mcts_search(rules, 128, roots, 64)
        assert np.array_equal(g.mcts_results()["visits"], exp["visits"])
    # node pool: a carried subtree + num_sims new nodes outgrow node_capacity (Brandubh: few moves, deep trees)
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    with E.LockstepGames(rules, fen, 8, 0, 64) as g:
        g.mcts_configure(400, 1.0, 1e-4, tree_reuse=True, node_capacity=401)
        with pytest.raises(TaflError) as ei:
            for _ in range(8):
                g.mcts_search()
                g.mcts_advance()
        assert ei.value.code == A.ERR_CAPACITY and "node pool" in str(ei.value)
        g.reset()
        g.mcts_configure(400, 1.0, 1e-4)
        g.mcts_search()
        assert g.mcts_overflow() == 0


def test_gpu_records_compact_and_episodes(orc, built_lib):
    """tafl_pack_records_compact == the rows of tafl_pack_records with n_moves > 0, in game order, with the row count
    and overflow word delivered to pinned host memory; tafl_reset_finished restarts finished games and bumps the
    episode counter the record header carries."""
    import torch
    from alphazeroforhnefatafl_b200 import selfplay as SP
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    n = 300
    with E.LockstepGames(rules, fen, n, 0, 64) as g:
        g.step_random(9, 70, auto_reset=False)        # a good part of the games is over
        over0 = g.get_states()["status"] != 0
        assert 0 < over0.sum() < n
        g.mcts_configure(80)
        g.mcts_search()
        g.mcts_play_best(False)
        full = torch.zeros((n, A.REC_STRIDE), dtype=torch.uint8, device="cuda")
        comp = torch.full((n, A.REC_STRIDE), 0xEE, dtype=torch.uint8, device="cuda")
        pin = E.PinnedArray(2, np.uint32)
        pin.a[:] = 77
        g.pack_records(full.data_ptr())
        g.pack_records_compact(comp.data_ptr(), A.ptr(pin.a))
        g.sync()
        rf = SP.records_to_numpy(full)
        keep = rf["h"]["n_moves"] > 0
        assert np.array_equal(keep, ~over0)
        assert int(pin.a[0]) == keep.sum() and int(pin.a[1]) == 0
        rc = SP.records_to_numpy(comp[:int(pin.a[0])])
        assert rc.tobytes() == rf[keep].tobytes()
        assert (comp[int(pin.a[0]):].cpu().numpy() == 0xEE).all()      # nothing written behind the last row
        assert not rf["h"]["episode"].any()
        # restart the finished games: new episode, start position; the others are untouched
        before = g.get_states()
        fin = before["status"] != 0
        assert g.reset_finished(want_count=True) == fin.sum()
        after = g.get_states()
        start = orc.state_from_fen(fen, 64, A.ATTACKER)[1][0]
        assert after[fin].tobytes() == np.repeat(start, fin.sum()).tobytes() and after[~fin].tobytes() == before[~fin].tobytes()
        g.mcts_search()
        g.mcts_play_best(False)
        g.pack_records(full.data_ptr())
        g.sync()
        rf2 = SP.records_to_numpy(full)
        assert np.array_equal(rf2["h"]["episode"], fin.astype(np.uint32))
        assert (rf2["h"]["n_moves"] > 0).all()
        # outcomes are back-filled per (game id, episode): an episode-0 result never leaks into episode 1
        z = SP.backfill_outcomes(np.concatenate([rf[keep], rf2]))
        z2 = z[keep.sum():]
        ended2 = (rf2["h"]["status"] & 0x80) != 0
        assert np.isnan(z2[fin & ~ended2]).all() and not np.isnan(z2[ended2]).any()
        pin.close()


def test_gpu_record_holds_a_high_mobility_root(orc, built_lib):
    """a root with more than 256 legal moves used to be truncated silently.  A record row now
    holds TAFL_MAX_MOVES moves; checked on a synthetic 11x11 position with > 256 plays."""
    import torch
    from alphazeroforhnefatafl_b200 import selfplay as SP
    rules = orc.preset_rules(A.PRESET_COPENHAGEN)
    # 24 attackers spread so that almost every one has long rays both ways (found by hill climbing on the oracle)
    rc, best_st = orc.state_from_fen("3tt3t2/1Kt6t1/1t8t/t5t4/t6t3/1t8t/5t5/t8t1/1t8t/2t1t3t2/3t2tt3", 128, A.ATTACKER)
    assert rc == 0
    best = int(orc.by_value(rules, 128).legal_moves_of(best_st)[1][0])
    assert best > 256
    with E.LockstepGames(rules, orc.preset_board(A.PRESET_COPENHAGEN), 2, 0, 128) as g:
        g.set_states(np.repeat(best_st, 2))
        g.mcts_configure(best + 40, 1.0, 1e-4)
        g.mcts_search()
        res = g.mcts_results()
        g.mcts_play_best(False)
        buf = torch.zeros((2, A.REC_STRIDE), dtype=torch.uint8, device="cuda")
        g.pack_records(buf.data_ptr())
        g.sync()
        rec = SP.records_to_numpy(buf)
        exp = orc.mcts_search(rules, 128, np.repeat(best_st, 2), best + 40)
        for i in range(2):
            assert rec["h"]["n_moves"][i] == best == exp["counts"][i]
            assert np.array_equal(rec["moves"][i, :best], exp["moves"][i, :best]) and not rec["moves"][i, best:].any()
            assert np.array_equal(rec["visits"][i, :best], exp["visits"][i, :best]) and rec["visits"][i, :best].sum() == best + 39
            assert rec["h"]["best"][i] == exp["best"][i] == res["best"][i]


def test_gpu_pack_records(orc, built_lib):
    import torch
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    n = 40
    roots = sample_positions(orc, rules, fen, 64, n, seed=81, max_steps=60)
    with E.LockstepGames(rules, fen, n, 0, 64) as g:
        g.set_states(roots)
        g.set_game_id_base(5000)
        g.mcts_configure(120)
        g.mcts_search()
        res = g.mcts_results()
        g.mcts_play_best(False)
        buf = torch.zeros(n * A.REC_STRIDE, dtype=torch.uint8, device="cuda")
        g.pack_records(buf.data_ptr())
        g.sync()
        after = g.get_states()
    raw = buf.cpu().numpy().reshape(n, A.REC_STRIDE)
    hdr = np.dtype([("game_id", "<u8"), ("ply", "<u4"), ("n_moves", "<u2"), ("side", "u1"), ("status", "u1"),
                    ("att", "<u8", (2,)), ("def", "<u8", (2,)), ("best", "<u2"), ("_pad", "<u2"), ("episode", "<u4")])
    assert hdr.itemsize == 56
    for i in range(n):
        h = raw[i, :56].view(hdr)[0]
        assert h["game_id"] == 5000 + i and h["ply"] == roots[i]["turn"]
        assert h["n_moves"] == res["counts"][i] and h["best"] == res["best"][i]
        assert h["side"] == (A.DEFENDER if roots[i]["flags"] & A.FLAG_DEF_TO_PLAY else A.ATTACKER)
        assert h["status"] == after[i]["status"]
        assert np.array_equal(h["att"], roots[i]["att"]) and np.array_equal(h["def"], roots[i]["def"])
        c = int(res["counts"][i])
        mv = raw[i, 56:56 + 2 * A.REC_MOVES].view("<u2")
        vs = raw[i, 56 + 2 * A.REC_MOVES:].view("<u4")
        assert np.array_equal(mv[:c], res["moves"][i, :c]) and not mv[c:].any()
        assert np.array_equal(vs[:c], res["visits"][i, :c]) and not vs[c:].any()


def test_gpu_selfplay_episode_vs_oracle(orc, built_lib):
    """MCTS self-play (search -> first-max move -> pack -> gather) for several moves on Tablut
    (BASELINE configs[3] rules): records and game states follow the oracle move for move."""
    from alphazeroforhnefatafl_b200 import selfplay as SP
    rules, fen = orc.preset_rules(A.PRESET_TABLUT), orc.preset_board(A.PRESET_TABLUT)
    n, sims, moves = 48, 96, 10
    ref_bv = orc.by_value(rules, 128)
    with E.LockstepGames(rules, fen, n, 0, 128) as g:
        g.set_game_id_base(777)
        g.step_random(5, 30, auto_reset=False)       # diversify the starting positions
        states = g.get_states()
        sp = SP.SelfPlay(g, num_sims=sims)
        all_recs = []
        for mv in range(moves):
            out = sp.move()
            g.sync()
            recs = SP.records_to_numpy(out)
            ref = orc.mcts_search(rules, 128, states, sims, n_threads=8)
            ongoing = (states["status"] == 0) & (ref["counts"] > 0)
            assert len(recs) == int(ongoing.sum())
            idx = np.nonzero(ongoing)[0]
            plays = E.moves_to_plays(ref["moves"][np.arange(n), ref["best"]], 128)
            nxt, fx = ref_bv.do_plays(states[ongoing], plays[ongoing])
            for r, i, s2 in zip(recs, idx, nxt):
                h = r["h"]
                c = int(ref["counts"][i])
                assert h["game_id"] == 777 + i and h["ply"] == states[i]["turn"] and h["n_moves"] == c
                assert h["best"] == ref["best"][i] and h["status"] == s2["status"]
                assert np.array_equal(r["moves"][:c], ref["moves"][i, :c])
                assert np.array_equal(r["visits"][:c], ref["visits"][i, :c])
            states[ongoing] = nxt
            assert g.get_states().tobytes() == states.tobytes()
            all_recs.append(recs)
        z = SP.backfill_outcomes(np.concatenate(all_recs))
        assert z.shape[0] == sum(len(r) for r in all_recs)


def test_gpu_selfplay_root_noise_vs_oracle(orc, built_lib):
    """SelfPlay(root_noise=(alpha, eps)) with the hash evaluator and tree reuse: the Dirichlet rows the
    driver draws per move, fed to the oracle, give the same visit counts and the same games"""
    from alphazeroforhnefatafl_b200 import selfplay as SP
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    n, sims, moves = 40, 80, 6
    with E.LockstepGames(rules, fen, n, 0, 64) as g:
        g.set_game_id_base(300)
        g.step_random(8, 12, auto_reset=False)
        roots = g.get_states()
        with pytest.raises(ValueError):
            SP.SelfPlay(g, num_sims=sims, root_noise=(0.3, 0.25))
        sp = SP.SelfPlay(g, num_sims=sims, tree_reuse=True, evaluator=("hash", 4), root_noise=(0.3, 0.25))
        noise = np.zeros((n, moves, A.MAX_MOVES))
        visits = np.zeros((n, moves, A.MAX_MOVES), np.uint32)
        for mv in range(moves):
            noise[:, mv] = g.dirichlet_root_noise(0.3, (21, 300, mv))      # what move() draws: (seed, id base, move index)
            L = g.legal_moves()[1]
            live = g.get_states()["status"] == 0
            assert np.allclose(noise[live, mv].sum(1), 1.0) and (noise[np.arange(n), mv, np.maximum(L, 1) - 1][live] > 0).all()
            sp.move(gather=False, seed=21)
            g.sync()
            visits[:, mv] = g.mcts_results()["visits"]
        final = g.get_states()
    exp = orc.mcts_episode(rules, 64, roots, sims, moves, True, eval_seed=4, n_threads=8, noise=noise, eps=0.25)
    assert np.array_equal(visits, exp["visits"])
    assert final.tobytes() == exp["final"].tobytes()


def test_gpu_mcts_group_sizes_are_bit_identical(orc, built_lib, monkeypatch):
    """k_mcts_sims_g with 32 / 16 / 8 lanes per tree (2 / 4 trees per warp, converged create + expand)
    == the first-generation warp-per-tree kernel (TAFL_MCTS_G=0) == the oracle, incl. tree reuse over
    several moves, trees that are not a multiple of the trees-per-CTA, and the capacity error"""
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), orc.preset_board(A.PRESET_COPENHAGEN)
    roots = sample_positions(orc, rules, fen, 128, 203, seed=61, max_steps=220)     # 203: ragged last CTA
    sims, n_moves = 150, 4
    exp = orc.mcts_episode(rules, 128, roots, sims, n_moves, True, n_threads=8)
    for G in ("0", "32", "16", "8"):
        monkeypatch.setenv("TAFL_MCTS_G", G)
        with E.LockstepGames(rules, fen, len(roots), 0, 128) as g:
            g.set_states(roots)
            g.mcts_configure(sims, 1.0, 1e-4, tree_reuse=True)
            for j in range(n_moves):
                g.mcts_search()
                r = g.mcts_results()
                assert np.array_equal(r["counts"], exp["counts"][:, j]), (G, j)
                assert np.array_equal(r["visits"], exp["visits"][:, j]), (G, j)
                assert r["q"].tobytes() == exp["q"][:, j].tobytes(), (G, j)
                g.mcts_advance()
            assert g.get_states().tobytes() == exp["final"].tobytes()
            st = g.mcts_stats().as_dict()
            for k in ("sims", "nodes", "edges", "terminal_hits", "depth_sum"):
                assert st[k] == getattr(exp["stats"], k), (G, k)
        with E.LockstepGames(rules, fen, 5, 0, 128) as g:      # edge pool too small -> TAFL_ERR_CAPACITY
            g.mcts_configure(64, 1.0, 1e-4, edge_capacity=500)
            with pytest.raises(TaflError):
                g.mcts_search()


def test_gpu_tree_reuse_dropped_when_the_search_kind_changes(orc, built_lib):
    """k_mcts_sims_g writes a node's edges lazily (only [0, Ns] is ever read under uniform priors); the
    evaluator-hook kernels need every edge initialised.  A carried tree of the other kind is dropped."""
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    roots = sample_positions(orc, rules, fen, 64, 48, seed=71, max_steps=60)
    with E.LockstepGames(rules, fen, len(roots), 0, 64) as g:
        g.set_states(roots)
        g.mcts_configure(120, 1.0, 1e-4, tree_reuse=True)
        g.mcts_search()
        g.mcts_advance()
        after = g.get_states()
        g.mcts_search_builtin(A.EVAL_STUB)          # other kind: fresh trees
        got = g.mcts_results()
        exp = orc.mcts_search(rules, 64, after, 120)
        assert np.array_equal(got["visits"], exp["visits"]) and got["q"].tobytes() == exp["q"].tobytes()
        g.mcts_advance()
        after2 = g.get_states()
        g.mcts_search()                              # and back
        got = g.mcts_results()
        exp = orc.mcts_search(rules, 64, after2, 120)
        assert np.array_equal(got["visits"], exp["visits"]) and got["q"].tobytes() == exp["q"].tobytes()


@pytest.mark.parametrize("name,preset,width,sims", [("copenhagen", A.PRESET_COPENHAGEN, 128, 130), ("brandubh", A.PRESET_BRANDUBH, 64, 300),
                                                    ("tablut", A.PRESET_TABLUT, 128, 200)])
def test_gpu_virtual_loss_matches_oracle(orc, built_lib, name, preset, width, sims):
    """k_mcts_sims_vl (W simulations in flight per tree) against oracle/mcts_oracle.hpp::run_from_root_vl at W = 1, 2, 4,
    8, 16: visit counts, Q, statistics and the games played, over several moves with tree reuse, from positions that
    include finished games and endgames (terminal leaves, collisions on narrow nodes).  W = 1 is the sequential search."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    roots = sample_positions(orc, rules, fen, width, 203, seed=41, max_steps=260)
    n_moves = 3
    seq = orc.mcts_episode(rules, width, roots, sims, n_moves, True, n_threads=8)
    for W in (1, 2, 4, 8, 16):
        exp = orc.mcts_episode_vl(rules, width, roots, sims, n_moves, True, W, n_threads=8)
        if W == 1:
            assert np.array_equal(exp["visits"], seq["visits"]) and exp["q"].tobytes() == seq["q"].tobytes()
        with E.LockstepGames(rules, fen, len(roots), 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, 1.0, 1e-4, tree_reuse=True, vl_width=W)
            for j in range(n_moves):
                g.mcts_search()
                r = g.mcts_results()
                assert np.array_equal(r["counts"], exp["counts"][:, j]), (W, j)
                assert np.array_equal(r["visits"], exp["visits"][:, j]), (W, j)
                assert r["q"].tobytes() == exp["q"][:, j].tobytes(), (W, j)
                g.mcts_advance()
            assert g.get_states().tobytes() == exp["final"].tobytes(), W
            st = g.mcts_stats().as_dict()
            for k in ("sims", "nodes", "edges", "terminal_hits", "depth_sum"):
                assert st[k] == getattr(exp["stats"], k), (W, k)
    with E.LockstepGames(rules, fen, 4, 0, width) as g:
        with pytest.raises(TaflError):
            g.mcts_configure(sims, 1.0, 1e-4, vl_width=17)


@pytest.mark.parametrize("name,preset,width,sims", [("copenhagen", A.PRESET_COPENHAGEN, 128, 110), ("brandubh", A.PRESET_BRANDUBH, 64, 260),
                                                    ("tablut", A.PRESET_TABLUT, 128, 150)])
def test_gpu_virtual_loss_through_the_evaluator_hook(orc, built_lib, monkeypatch, name, preset, width, sims):
    """k_mcts_hook_vl — W leaves per tree and round, the evaluator sees n_trees x W rows — against the oracle's
    run_from_root_vl with the hash evaluator at W = 2, 4, 8, 16: visits, Q, statistics, games played over several moves
    with tree reuse and root noise, fused and split launches; with the stub evaluator it equals the fused virtual-loss
    kernel k_mcts_sims_vl bit for bit; the split C-ABI calls with host copies of the leaf batch drive the same kernel."""
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    roots = sample_positions(orc, rules, fen, width, 203, seed=43, max_steps=260)
    n, n_moves = len(roots), 3
    rng = np.random.default_rng(5)
    noise = rng.dirichlet(np.full(A.MAX_MOVES, 0.3), size=(n, n_moves))
    for W, fuse, eps in ((2, "1", 0.0), (4, "1", 0.25), (8, "0", 0.0), (16, "1", 0.0)):
        monkeypatch.setenv("TAFL_HOOK_FUSE", fuse)
        exp = orc.mcts_episode_vl_eval(rules, width, roots, sims, n_moves, True, W, eval_seed=31, cpuct=1.3, draw_value=0.2,
                                       n_threads=8, noise=noise if eps else None, eps=eps)
        with E.LockstepGames(rules, fen, n, 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, 1.3, 0.2, tree_reuse=True, vl_width=W)
            for j in range(n_moves):
                if eps:
                    g.mcts_root_noise(noise[:, j], eps)
                g.mcts_search_builtin(A.EVAL_HASH, 31)
                r = g.mcts_results()
                assert np.array_equal(r["counts"], exp["counts"][:, j]), (W, j)
                assert np.array_equal(r["visits"], exp["visits"][:, j]), (W, j)
                assert r["q"].tobytes() == exp["q"][:, j].tobytes(), (W, j)
                g.mcts_advance()
            assert g.get_states().tobytes() == exp["final"].tobytes(), W
            st = g.mcts_stats().as_dict()
            for k in ("sims", "nodes", "edges", "terminal_hits", "depth_sum"):
                assert st[k] == getattr(exp["stats"], k), (W, k)
    monkeypatch.delenv("TAFL_HOOK_FUSE", raising=False)
    # the stub evaluator through the hook == the fused virtual-loss kernel
    for W in (4, 16):
        res = {}
        for mode in ("fused", "hook"):
            with E.LockstepGames(rules, fen, n, 0, width) as g:
                g.set_states(roots)
                g.mcts_configure(sims, 1.0, 1e-4, vl_width=W)
                if mode == "fused":
                    g.mcts_search()
                else:
                    g.mcts_search_builtin(A.EVAL_STUB)
                r = g.mcts_results()
                res[mode] = (r["counts"].tobytes(), r["visits"].tobytes(), r["q"].tobytes(),
                             {k: v for k, v in g.mcts_stats().as_dict().items() if k in ("sims", "nodes", "edges", "terminal_hits", "depth_sum")})
        assert res["fused"] == res["hook"], W
    # split calls, evaluator on the host: the batch has n * W rows, row = game * W + simulation
    W = 4
    exp = orc.mcts_episode_vl_eval(rules, width, roots[:40], 30, 1, False, W, eval_seed=None, n_threads=8)
    with E.LockstepGames(rules, fen, 40, 0, width) as g:
        g.set_states(roots[:40])
        g.mcts_configure(30, 1.0, 1e-4, vl_width=W, edge_capacity=31 * 320)   # endgame roots: more than 160 plays per node
        assert g.leaf_rows == 40 * W
        g.mcts_begin()
        rounds = 0
        while True:
            try:
                g.mcts_select_leaves()
            except TaflError:
                break
            b = g.leaf_batch_host()
            assert b["counts"].shape == (40 * W,)
            live = b["counts"].reshape(40, W) > 0
            assert not live[:, 1:].any() if rounds == 0 else True        # the first batch of a fresh tree is its root alone
            pri = np.full((40 * W, A.MAX_MOVES), 1.0 / A.MAX_MOVES, np.float32)
            g.leaf_eval_host(pri, np.zeros(40 * W, np.float32))
            g.mcts_expand_backup()
            rounds += 1
        assert rounds == 1 + (30 - 1 + W - 1) // W
        g.mcts_finish()
        r = g.mcts_results()
        assert np.array_equal(r["visits"], exp["visits"][:, 0]) and r["q"].tobytes() == exp["q"][:, 0].tobytes()
