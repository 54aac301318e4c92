"""Evaluator-hook MCTS against outputs of the UNMODIFIED reference `src/mcts.py`
run with a NON-UNIFORM evaluator (tests/golden/mcts_eval_vectors.json, generator
tests/golden/make_mcts_eval_vectors.py): pins mcts.py:85-99 (mask, renormalise with np.sum,
"all valid moves were masked" fallback), per-action priors in PUCT and non-zero leaf values.

* CPU: the C++ oracle in evaluator mode reproduces visit counts exactly and Q bit for bit; its
  np.sum restatement equals numpy's; its TAFL_EVAL_HASH equals the numpy one of the generator.
* GPU: the split-phase CUDA search (select -> evaluator -> expand/backup, through the C ABI)
  reproduces the same vectors with the built-in device evaluator AND with priors computed on
  the host and handed in per simulation (the real evaluator boundary).
"""
import json
import os

import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "mcts_eval_vectors.json")) as f:
    GOLDEN = json.load(f)
CASES = GOLDEN["cases"]
IDS = ["%s-%d" % (c["name"], i) for i, c in enumerate(CASES)]
Q_RTOL = 1e-6


def root_of(case):
    return np.frombuffer(bytes.fromhex(case["root_hex"]), A.STATE_DTYPE).copy()


def expected(case):
    return (np.array(case["visits"], np.uint32), np.array([float.fromhex(x) for x in case["q_hex"]], np.float64))


def test_golden_file_shape():
    assert len(CASES) >= 15
    assert "UNMODIFIED reference" in GOLDEN["provenance"]
    assert sum(c["masked_evals"] for c in CASES) > 50      # the all-masked fallback is exercised
    mixed = sum(1 for c in CASES for x in c["q_hex"] if 0.0 < abs(float.fromhex(x)) < 1.0)
    assert mixed > 200


def test_np_sum_restatement_matches_numpy(orc):
    rng = np.random.default_rng(5)
    for n in (1, 5, 7, 8, 9, 31, 64, 127, 128, 129, 200, 511, 512, 513, 1000):
        for _ in range(20):
            a = rng.random(n) * rng.choice([1.0, 1e-6, 1e6])
            if rng.random() < 0.5:
                a[rng.integers(0, n):] = 0.0
            assert orc.np_sum(a) == float(np.sum(a)), n


def test_hash_evaluator_known_answers(orc):
    for k in GOLDEN["evaluator_kat"]:
        st = np.frombuffer(bytes.fromhex(k["state_hex"]), A.STATE_DTYPE).copy()
        P, v = orc.hash_eval(st, k["seed"])
        assert [float(x) for x in P[0, :8]] == k["p_first"]
        assert float(P[0, -1]) == k["p_last"]
        assert float(P[0].astype(np.float64).sum()) == k["p_sum_f64"]
        assert float(v[0]) == k["v"]


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_eval_oracle_matches_reference_mcts_py(orc, case):
    rules = orc.preset_rules(case["preset"])
    got = orc.mcts_search_eval(rules, case["width"], root_of(case), case["sims"], case["eval_seed"], cpuct=case["cpuct"],
                               draw_value=case["draw_value"])
    n, q = expected(case)
    c = case["count"]
    assert int(got["counts"][0]) == c
    assert np.array_equal(got["visits"][0, :c], n)
    assert got["q"][0, :c].tobytes() == q.tobytes(), "Q differs from the reference in the last bits"
    assert got["stats"].nodes == case["n_nodes"]


EPISODES = GOLDEN["episodes"]


def _check_episode(ep, visits, q, counts, best, final, i=0):
    for j, mv in enumerate(ep["moves"]):
        L = len(mv["visits"])
        assert int(counts[i, j]) == L, (ep["name"], j)
        assert np.array_equal(visits[i, j, :L], np.array(mv["visits"], np.uint32)), (ep["name"], j)
        ref_q = np.array([float.fromhex(x) for x in mv["q_hex"]])
        assert q[i, j, :L].tobytes() == ref_q.tobytes(), (ep["name"], j)
        assert int(best[i, j]) == mv["best"]
    for j in range(len(ep["moves"]), ep["n_moves"]):
        assert int(counts[i, j]) == 0          # the game ended before move j
    assert final[i:i + 1].tobytes().hex() == ep["final_hex"]


@pytest.mark.parametrize("ep", EPISODES, ids=["%s-%d" % (e["name"], i) for i, e in enumerate(EPISODES)])
def test_oracle_tree_reuse_matches_reference_mcts_object_kept_across_moves(orc, ep):
    """one reference MCTS object per episode (mcts.py:20-26 dicts persist) == Mcts::advance"""
    rules = orc.preset_rules(ep["preset"])
    root = np.frombuffer(bytes.fromhex(ep["root_hex"]), A.STATE_DTYPE).copy()
    got = orc.mcts_episode(rules, ep["width"], root, ep["sims"], ep["n_moves"], True, cpuct=ep["cpuct"],
                           draw_value=ep["draw_value"], eval_seed=ep["eval_seed"])
    assert any(sum(mv["visits"]) > ep["sims"] for mv in ep["moves"][1:])   # reuse really carries visits over
    _check_episode(ep, got["visits"], got["q"], got["counts"], got["best"], got["final"])


def _groups():
    groups = {}
    for case in CASES:
        key = (case["preset"], case["width"], case["sims"], case["cpuct"], case["draw_value"], case["eval_seed"])
        groups.setdefault(key, []).append(case)
    return groups


def _check(got, cases):
    for i, case in enumerate(cases):
        n, q = expected(case)
        c = case["count"]
        assert int(got["counts"][i]) == c
        assert np.array_equal(got["visits"][i, :c], n), case["name"]
        assert np.allclose(got["q"][i, :c], q, rtol=Q_RTOL, atol=0.0), case["name"]
        assert got["q"][i, :c].tobytes() == q.tobytes(), case["name"]


@pytest.mark.gpu
def test_gpu_eval_search_matches_reference_mcts_py(orc, built_lib):
    """built-in device evaluator, whole search on the device"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for
    checked = 0
    for (preset, width, sims, cpuct, dv, eseed), cases in _groups().items():
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        roots = np.concatenate([root_of(c) for c in cases])
        with E.LockstepGames(rules, fen, len(cases), 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, cpuct, dv)
            g.mcts_search_builtin(A.EVAL_HASH, eseed)
            _check(g.mcts_results(), cases)
            checked += len(cases)
    assert checked == len(CASES)


@pytest.mark.gpu
def test_gpu_eval_boundary_with_host_priors(orc, built_lib):
    """the real boundary: leaf batch out, priors / values computed OUTSIDE the library (here: on the
    host by the oracle's copy of the hash evaluator, from the batch's leaf states), handed back"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for
    done = 0
    for (preset, width, sims, cpuct, dv, eseed), cases in _groups().items():
        if sims > 400:
            continue
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        bv = orc.by_value(rules, width)
        roots = np.concatenate([root_of(c) for c in cases])
        with E.LockstepGames(rules, fen, len(cases), 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, cpuct, dv)
            g.mcts_begin()
            for r in range(sims):
                g.mcts_select_leaves()
                b = g.leaf_batch_host()
                live = b["counts"] > 0
                if r == 0:
                    assert live.all() and b["states"].tobytes() == roots.tobytes()
                if r % 37 == 0 and live.any():   # the batch rows describe the leaf: moves and planes
                    em, ec = bv.legal_moves_of(b["states"][live])
                    assert np.array_equal(ec, b["counts"][live])
                    for k, i in enumerate(np.nonzero(live)[0]):
                        assert np.array_equal(b["moves"][i, :ec[k]], em[k, :ec[k]])
                        assert np.array_equal(b["planes"][i], E.board_to_matrix(b["states"][i:i + 1], width).reshape(-1))
                P, v = orc.hash_eval(b["states"], eseed)
                P[~live] = np.nan      # rows that asked for nothing must be ignored
                v[~live] = np.nan
                g.leaf_eval_host(P, v)
                g.mcts_expand_backup()
            g.mcts_finish()
            _check(g.mcts_results(), cases)
            done += len(cases)
    assert done >= 10


@pytest.mark.gpu
def test_gpu_stub_through_the_hook_equals_fused_search(orc, built_lib):
    """TAFL_EVAL_STUB through select / evaluate / expand == the fused k_mcts_sims search, bit for bit"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for, sample_positions
    for preset, width, sims in ((A.PRESET_COPENHAGEN, 128, 300), (A.PRESET_BRANDUBH, 64, 500)):
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        roots = sample_positions(orc, rules, fen, width, 96, seed=41, max_steps=150)
        with E.LockstepGames(rules, fen, len(roots), 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, 1.25, 1e-4)
            g.mcts_search()
            a = g.mcts_results()
            sa = g.mcts_stats().as_dict()
            g.mcts_search_builtin(A.EVAL_STUB)
            b = g.mcts_results()
            sb = g.mcts_stats().as_dict()
        for k in ("moves", "visits", "q", "counts", "best"):
            assert a[k].tobytes() == b[k].tobytes(), k
        for k in ("sims", "nodes", "depth_sum", "edges", "terminal_hits"):
            assert sb[k] == 2 * sa[k], (k, sa[k], sb[k])


@pytest.mark.gpu
def test_gpu_eval_phase_errors(orc, built_lib):
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    with E.LockstepGames(rules, fen, 4, 0, 64) as g:
        g.mcts_configure(3, 1.0, 1e-4)
        with pytest.raises(E.TaflError):
            g.mcts_select_leaves()          # not begun
        g.mcts_begin()
        with pytest.raises(E.TaflError):
            g.mcts_expand_backup()          # nothing pending
        for _ in range(3):
            g.mcts_select_leaves()
            with pytest.raises(E.TaflError):
                g.mcts_finish()             # evaluation pending
            g.eval_builtin(A.EVAL_STUB)
            g.mcts_expand_backup()
        with pytest.raises(E.TaflError):
            g.mcts_select_leaves()          # node pool is sized for num_sims rounds
        g.mcts_finish()
        assert g.mcts_results()["visits"].sum() == 4 * 2


POLICY = GOLDEN["policy"]


def _host_draw(visits, seed, gid, ply):
    """include/tafl.h tafl_mcts_policy, temp == 1: k = mulhi64(splitmix64(..), sum N), first cumsum > k"""
    M = (1 << 64) - 1

    def sm(x):
        x = (x + 0x9E3779B97F4A7C15) & M
        x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
        x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
        return x ^ (x >> 31)
    x = sm(seed ^ ((gid * 0x9E3779B97F4A7C15) & M) ^ ((ply * 0xBF58476D1CE4E5B9) & M))
    k = (x * int(visits.sum())) >> 64
    return int(np.searchsorted(np.cumsum(visits.astype(np.int64)), k, side="right"))


@pytest.mark.gpu
def test_gpu_policy_matches_reference_get_action_prob(orc, built_lib):
    """MCTS.getActionProb (mcts.py:28-53) of the unmodified reference: temp 1 bit-exact, temp 0.5 / 2
    within 1e-12, temp 0 one-hot inside the reference's tie set; the drawn move follows the ABI's rule"""
    from types import SimpleNamespace
    from alphazeroforhnefatafl_b200 import engine as E
    from alphazeroforhnefatafl_b200.mcts import MCTS
    from parity_cases import board_for
    for name in sorted({c["name"] for c in POLICY}):
        cases = [c for c in POLICY if c["name"] == name]
        c0 = cases[0]
        rules, fen = orc.preset_rules(c0["preset"]), board_for(orc, c0["preset"])
        roots = np.concatenate([root_of(c) for c in cases])
        with E.LockstepGames(rules, fen, len(cases), 0, c0["width"]) as g:
            g.set_states(roots)
            g.set_game_id_base(1000)
            m = MCTS(g, SimpleNamespace(numMCTSSims=c0["sims"], cpuct=c0["cpuct"]), nnet=("hash", c0["eval_seed"]))
            for temp in (1, 0.5, 2, 0):
                moves, probs, counts = m.getActionProb(temp=temp)
                for i, c in enumerate(cases):
                    L = len(c["visits"])
                    assert int(counts[i]) == L and not probs[i, L:].any()
                    assert np.array_equal(m.last["visits"][i, :L], np.array(c["visits"], np.uint32))
                    if temp == 0:
                        assert probs[i, :L].sum() == 1.0 and int(np.argmax(probs[i, :L])) == c["ties_t0"][0]
                        continue
                    ref = np.array([float.fromhex(x) for x in c["probs"][str(temp)]])
                    if temp == 1:
                        assert probs[i, :L].tobytes() == ref.tobytes()
                    else:
                        assert np.allclose(probs[i, :L], ref, rtol=1e-12, atol=0.0)
            # the draw: temp == 1 is an integer walk, reproducible on the host
            for seed in (1, 2, 3, 99):
                m.choose(temp=1, seed=seed)
                best = g.mcts_results()["best"]
                for i, c in enumerate(cases):
                    ply = int(roots["turn"][i])
                    assert int(best[i]) == _host_draw(np.array(c["visits"]), seed, 1000 + i, ply)
            m.choose(temp=0)   # back to the first maximum
            assert [int(b) for b in g.mcts_results()["best"]] == [int(np.argmax(c["visits"])) for c in cases]
            # a drawn move is played and recorded
            m.choose(temp=0.5, seed=5)
            best = g.mcts_results()["best"]
            for i, c in enumerate(cases):
                assert c["visits"][int(best[i])] > 0
            st = m.play_best()
            assert len(st) == len(cases)


def _run_episode_gpu(g, sims, n_moves, eval_seed, reuse=True):
    n = g.n_games
    visits = np.zeros((n, n_moves, A.MAX_MOVES), np.uint32)
    q = np.zeros((n, n_moves, A.MAX_MOVES), np.float64)
    counts = np.zeros((n, n_moves), np.uint32)
    best = np.zeros((n, n_moves), np.uint32)
    for j in range(n_moves):
        if eval_seed is None:
            g.mcts_search()
        else:
            g.mcts_search_builtin(A.EVAL_HASH, eval_seed)
        r = g.mcts_results()
        visits[:, j], q[:, j], counts[:, j], best[:, j] = r["visits"], r["q"], r["counts"], r["best"]
        if reuse:
            g.mcts_advance()
        else:
            g.mcts_play_best()
    return visits, q, counts, best, g.get_states()


@pytest.mark.gpu
def test_gpu_tree_reuse_matches_reference_mcts_object_kept_across_moves(orc, built_lib):
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for
    for name in sorted({e["name"] for e in EPISODES}):
        eps = [e for e in EPISODES if e["name"] == name]
        e0 = eps[0]
        rules, fen = orc.preset_rules(e0["preset"]), board_for(orc, e0["preset"])
        roots = np.concatenate([np.frombuffer(bytes.fromhex(e["root_hex"]), A.STATE_DTYPE) for e in eps])
        with E.LockstepGames(rules, fen, len(eps), 0, e0["width"]) as g:
            g.set_states(roots)
            g.mcts_configure(e0["sims"], e0["cpuct"], e0["draw_value"], tree_reuse=True)
            visits, q, counts, best, final = _run_episode_gpu(g, e0["sims"], e0["n_moves"], e0["eval_seed"])
        for i, ep in enumerate(eps):
            _check_episode(ep, visits, q, counts, best, final, i)


@pytest.mark.gpu
@pytest.mark.parametrize("preset,width,sims,n_moves,eval_seed", [
    (A.PRESET_COPENHAGEN, 128, 160, 6, None), (A.PRESET_COPENHAGEN, 128, 120, 5, 77),
    (A.PRESET_BRANDUBH, 64, 300, 12, None), (A.PRESET_TABLUT, 128, 200, 8, 78)])
def test_gpu_tree_reuse_matches_oracle_at_scale(orc, built_lib, preset, width, sims, n_moves, eval_seed):
    """many games, mid-game roots, games ending inside the episode: visits / Q / final states bit-exact"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for, sample_positions
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    roots = sample_positions(orc, rules, fen, width, 192, seed=51, max_steps=250)
    exp = orc.mcts_episode(rules, width, roots, sims, n_moves, True, eval_seed=eval_seed, n_threads=8)
    with E.LockstepGames(rules, fen, len(roots), 0, width) as g:
        g.set_states(roots)
        g.mcts_configure(sims, 1.0, 1e-4, tree_reuse=True)
        visits, q, counts, best, final = _run_episode_gpu(g, sims, n_moves, eval_seed)
        st = g.mcts_stats().as_dict()
    assert np.array_equal(counts, exp["counts"])
    assert np.array_equal(visits, exp["visits"])
    assert q.tobytes() == exp["q"].tobytes()
    assert np.array_equal(best, exp["best"])
    assert final.tobytes() == exp["final"].tobytes()
    assert (visits.sum(2)[:, 1:] > sims).any()       # statistics were carried over
    for k in ("nodes", "edges", "terminal_hits", "depth_sum"):
        assert st[k] == getattr(exp["stats"], k), k


@pytest.mark.gpu
def test_gpu_tree_reuse_dropped_when_states_are_replaced(orc, built_lib):
    """a carried tree is only used while its root equals the game's state"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for, sample_positions
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    roots = sample_positions(orc, rules, fen, 64, 32, seed=52, max_steps=60)
    other = sample_positions(orc, rules, fen, 64, 32, seed=53, max_steps=60)
    with E.LockstepGames(rules, fen, len(roots), 0, 64) as g:
        g.mcts_configure(150, 1.0, 1e-4, tree_reuse=True)
        g.set_states(roots)
        g.mcts_search()
        g.mcts_advance()
        g.set_states(other)          # the caller replaced the positions: carried trees are stale
        g.mcts_search()
        got = g.mcts_results()
    exp = orc.mcts_search(rules, 64, other, 150)
    assert np.array_equal(got["visits"], exp["visits"]) and got["q"].tobytes() == exp["q"].tobytes()


# ---- root exploration noise (src/mcts.rs:53 TODO; build-defined, pinned by the oracle) ------------------
def _noise_rows(n, n_moves, seed):
    rng = np.random.default_rng(seed)
    z = rng.gamma(0.3, size=(n, n_moves, A.MAX_MOVES))
    return z / z[:, :, :64].sum(2, keepdims=True)


def test_oracle_root_noise_semantics(orc):
    """eps = 0 leaves the search untouched; eps = 1 with a one-hot row forces every simulation through
    that root action; the noise is applied once per search, also to a root carried by tree reuse."""
    from parity_cases import board_for, sample_positions
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    roots = sample_positions(orc, rules, fen, 64, 6, seed=61, max_steps=40)
    base = orc.mcts_episode(rules, 64, roots, 60, 3, True, eval_seed=9)
    z = _noise_rows(len(roots), 3, 1)
    same = orc.mcts_episode(rules, 64, roots, 60, 3, True, eval_seed=9, noise=z, eps=0.0)
    assert np.array_equal(base["visits"], same["visits"]) and base["q"].tobytes() == same["q"].tobytes()
    mixed = orc.mcts_episode(rules, 64, roots, 60, 3, True, eval_seed=9, noise=z, eps=0.25)
    assert not np.array_equal(base["visits"], mixed["visits"])
    onehot = np.zeros_like(z)
    onehot[:, :, 2] = 1.0
    forced = orc.mcts_episode(rules, 64, roots, 60, 3, True, eval_seed=9, noise=onehot, eps=1.0)
    for g in range(len(roots)):     # the simulation after the root's expansion sees u > 0 on action 2 only
        if forced["counts"][g, 0] > 2:
            assert forced["visits"][g, 0, 2] >= 1
    # a root carried by tree reuse takes that move's noise too: rows that differ at move 1 only
    z2 = z.copy()
    z2[:, 1] = _noise_rows(len(roots), 1, 5)[:, 0]
    other = orc.mcts_episode(rules, 64, roots, 60, 3, True, eval_seed=9, noise=z2, eps=0.25)
    assert np.array_equal(other["visits"][:, 0], mixed["visits"][:, 0])
    assert not np.array_equal(other["visits"][:, 1], mixed["visits"][:, 1])


@pytest.mark.gpu
@pytest.mark.parametrize("preset,width,sims,n_moves,reuse", [
    (A.PRESET_COPENHAGEN, 128, 100, 4, True), (A.PRESET_BRANDUBH, 64, 200, 8, True), (A.PRESET_TABLUT, 128, 120, 5, False)])
def test_gpu_root_noise_matches_oracle(orc, built_lib, preset, width, sims, n_moves, reuse):
    """tafl_mcts_root_noise: visits / Q / chosen moves / final states bit-exact with the oracle's noise_root,
    fresh trees and carried trees, through the graph-launched built-in evaluator and the split-phase calls"""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for, sample_positions
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    roots = sample_positions(orc, rules, fen, width, 96, seed=62, max_steps=200)
    n = len(roots)
    z = _noise_rows(n, n_moves, 2)
    exp = orc.mcts_episode(rules, width, roots, sims, n_moves, reuse, eval_seed=31, n_threads=8, noise=z, eps=0.25)
    for split in (False, True):
        with E.LockstepGames(rules, fen, n, 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, 1.0, 1e-4, tree_reuse=reuse)
            visits = np.zeros((n, n_moves, A.MAX_MOVES), np.uint32)
            q = np.zeros((n, n_moves, A.MAX_MOVES), np.float64)
            best = np.zeros((n, n_moves), np.uint32)
            for j in range(n_moves):
                g.mcts_root_noise(z[:, j], 0.25)
                if split:
                    g.mcts_begin()
                    for _ in range(sims):
                        g.mcts_select_leaves()
                        g.eval_builtin(A.EVAL_HASH, 31)
                        g.mcts_expand_backup()
                    g.mcts_finish()
                else:
                    g.mcts_search_builtin(A.EVAL_HASH, 31)
                r = g.mcts_results()
                visits[:, j], q[:, j], best[:, j] = r["visits"], r["q"], r["best"]
                g.mcts_advance() if reuse else g.mcts_play_best()
            final = g.get_states()
            # noise is refused by the fused stub search, and can be switched off again
            with pytest.raises(Exception):
                g.mcts_search()
            g.mcts_root_noise(None, 0.0)
            g.mcts_search()
        assert np.array_equal(visits, exp["visits"]), split
        assert q.tobytes() == exp["q"].tobytes(), split
        assert np.array_equal(best, exp["best"])
        assert final.tobytes() == exp["final"].tobytes()
    plain = orc.mcts_episode(rules, width, roots, sims, n_moves, reuse, eval_seed=31, n_threads=8)
    assert not np.array_equal(plain["visits"], exp["visits"])


@pytest.mark.gpu
def test_gpu_hook_generations_are_bit_identical(orc, built_lib, monkeypatch):
    """The evaluator-hook search three ways — first-generation warp-per-tree kernel pair (TAFL_HOOK_GEN=1), the
    second-generation G-lanes-per-tree kernel with its two phases in separate launches (TAFL_HOOK_FUSE=0) and fused
    (expand + next descent in one launch), G = 8 / 16 / 32 — against the oracle with the hash evaluator: visits, Q,
    statistics; a tree count that is not a multiple of the trees per CTA; games that end inside the search."""
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for, sample_positions
    for preset, width, sims in ((A.PRESET_COPENHAGEN, 128, 90), (A.PRESET_BRANDUBH, 64, 260)):
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        roots = sample_positions(orc, rules, fen, width, 203, seed=97, max_steps=230)
        exp = orc.mcts_search_eval(rules, width, roots, sims, 5, cpuct=1.3, draw_value=0.2, n_threads=8)
        for gen, fuse, G in (("1", "1", "16"), ("2", "0", "16"), ("2", "1", "16"), ("2", "1", "8"), ("2", "1", "32")):
            monkeypatch.setenv("TAFL_HOOK_GEN", gen)
            monkeypatch.setenv("TAFL_HOOK_FUSE", fuse)
            monkeypatch.setenv("TAFL_MCTS_G", G)
            with E.LockstepGames(rules, fen, len(roots), 0, width) as g:
                g.set_states(roots)
                g.mcts_configure(sims, 1.3, 0.2)
                g.mcts_search_builtin(A.EVAL_HASH, 5)
                got = g.mcts_results()
                st = g.mcts_stats().as_dict()
                key = (preset, gen, fuse, G)
                assert np.array_equal(got["counts"], exp["counts"]), key
                assert np.array_equal(got["visits"], exp["visits"]), key
                assert got["q"].tobytes() == exp["q"].tobytes(), key
                for k in ("sims", "nodes", "edges", "terminal_hits", "depth_sum"):
                    assert st[k] == getattr(exp["stats"], k), (key, k)
                # the split C-ABI calls drive the same kernels: one more search, phase by phase, stub evaluator
                g.mcts_begin()
                for _ in range(12):
                    g.mcts_select_leaves()
                    g.eval_builtin(A.EVAL_STUB)
                    g.mcts_expand_backup()
                g.mcts_finish()
                r12 = g.mcts_results()
                assert (r12["visits"].sum(1)[roots["status"] == 0] == 11).all(), key
