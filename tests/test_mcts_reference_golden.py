"""MCTS parity against outputs of the UNMODIFIED reference `src/mcts.py` (class MCTS), frozen in
tests/golden/mcts_reference_vectors.json by tests/golden/make_mcts_reference_vectors.py (which
imports the reference in the build container; the reference tree does not exist on the GPU box).

* CPU: the C++ MCTS oracle (oracle/mcts_oracle.hpp) reproduces the reference's root visit counts
  exactly and its Q values bit for bit  -> the oracle is pinned by reference outputs.
* GPU: the CUDA search (tafl_mcts.cu through the C ABI) reproduces the same vectors: visit counts
  bit-exact, Q within 1e-6 relative (BASELINE.json north_star) — asserted bit-exact as well.
"""
import json
import os

import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "mcts_reference_vectors.json")) as f:
    GOLDEN = json.load(f)
CASES = GOLDEN["cases"]
IDS = ["%s-%d" % (c["name"], i) for i, c in enumerate(CASES)]
Q_RTOL = 1e-6


def root_of(case):
    return np.frombuffer(bytes.fromhex(case["root_hex"]), A.STATE_DTYPE).copy()


def expected(case):
    return (np.array(case["visits"], np.uint32), np.array([float.fromhex(x) for x in case["q_hex"]], np.float64))


def test_golden_file_shape():
    assert len(CASES) >= 20
    assert "UNMODIFIED reference" in GOLDEN["provenance"]
    # the fixture exercises non-trivial arithmetic: some Q are non-zero and not +-1
    mixed = sum(1 for c in CASES for x in c["q_hex"] if 0.0 < abs(float.fromhex(x)) < 1.0)
    assert mixed > 100
    for c in CASES:
        assert sum(c["visits"]) == c["sims"] - 1          # mcts.py:83-102: first simulation expands the root


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_mcts_oracle_matches_reference_mcts_py(orc, case):
    rules = orc.preset_rules(case["preset"])
    got = orc.mcts_search(rules, case["width"], root_of(case), case["sims"], cpuct=case["cpuct"],
                          draw_value=case["draw_value"])
    n, q = expected(case)
    c = case["count"]
    assert int(got["counts"][0]) == c
    assert np.array_equal(got["visits"][0, :c], n)
    assert got["q"][0, :c].tobytes() == q.tobytes(), "Q differs from the reference in the last bits"
    assert int(got["best"][0]) == int(np.argmax(n))       # first-max
    # the oracle counts the root among the nodes; mcts.py's Es dict has one entry per node met
    assert got["stats"].nodes == case["n_nodes"]


@pytest.mark.gpu
def test_gpu_mcts_matches_reference_mcts_py(orc, built_lib):
    from alphazeroforhnefatafl_b200 import engine as E
    from parity_cases import board_for
    groups = {}
    for case in CASES:
        key = (case["preset"], case["width"], case["sims"], case["cpuct"], case["draw_value"])
        groups.setdefault(key, []).append(case)
    checked = 0
    for (preset, width, sims, cpuct, dv), cases in groups.items():
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        roots = np.concatenate([root_of(c) for c in cases])
        with E.LockstepGames(rules, fen, len(cases), 0, width) as g:
            g.set_states(roots)
            g.mcts_configure(sims, cpuct, dv)
            g.mcts_search()
            got = g.mcts_results()
        for i, case in enumerate(cases):
            n, q = expected(case)
            c = case["count"]
            assert int(got["counts"][i]) == c
            assert np.array_equal(got["visits"][i, :c], n), case["name"]
            assert np.allclose(got["q"][i, :c], q, rtol=Q_RTOL, atol=0.0), case["name"]
            assert got["q"][i, :c].tobytes() == q.tobytes(), case["name"]
            assert int(got["best"][i]) == int(np.argmax(n))
            checked += 1
    assert checked == len(CASES)


def test_oracle_virtual_loss_width_1_is_the_sequential_search(orc):
    """The virtual-loss restatement (Mcts::run_from_root_vl) with one simulation in flight is mcts.py's search bit for
    bit — visits, Q, final states and statistics over several moves with and without tree reuse; wider batches keep
    the simulation count and differ only in where the visits go."""
    import numpy as np
    from alphazeroforhnefatafl_b200 import _abi as A
    from parity_cases import sample_positions
    for preset, width, sims in ((A.PRESET_BRANDUBH, 64, 150), (A.PRESET_COPENHAGEN, 128, 90), (A.PRESET_TABLUT, 128, 120)):
        rules, fen = orc.preset_rules(preset), orc.preset_board(A.PRESET_BRANDUBH if preset == A.PRESET_KOCH else preset)
        roots = sample_positions(orc, rules, fen, width, 20, seed=3, max_steps=200)
        for reuse in (False, True):
            a = orc.mcts_episode(rules, width, roots, sims, 3, reuse, n_threads=4)
            b = orc.mcts_episode_vl(rules, width, roots, sims, 3, reuse, 1, n_threads=4)
            assert np.array_equal(a["visits"], b["visits"]) and a["q"].tobytes() == b["q"].tobytes()
            assert a["final"].tobytes() == b["final"].tobytes() and a["stats"].as_dict() == b["stats"].as_dict()
            for W in (4, 16):
                c = orc.mcts_episode_vl(rules, width, roots, sims, 3, reuse, W, n_threads=4)
                assert c["stats"].sims == a["stats"].sims
                live = roots["status"] == 0
                assert (c["visits"][:, 0].sum(1)[live] == sims - 1).all()      # the first simulation expands the root


def test_oracle_virtual_loss_with_an_evaluator(orc):
    """run_from_root_vl with the hash evaluator and root noise (what the evaluator hook runs with vl_width > 1): one
    simulation in flight is the sequential evaluator search bit for bit; the stub through this entry point is the stub
    virtual-loss search; wider batches keep the simulation count, differ in where the visits go, and a collision
    (two simulations of one batch on the same fresh edge) takes the opener's leaf value instead of evaluating twice."""
    import numpy as np
    from alphazeroforhnefatafl_b200 import _abi as A
    from parity_cases import sample_positions
    for preset, width, sims in ((A.PRESET_BRANDUBH, 64, 150), (A.PRESET_COPENHAGEN, 128, 80)):
        rules, fen = orc.preset_rules(preset), orc.preset_board(preset)
        roots = sample_positions(orc, rules, fen, width, 16, seed=5, max_steps=200)
        noise = np.random.default_rng(2).dirichlet(np.full(A.MAX_MOVES, 0.3), size=(len(roots), 2))
        for reuse in (False, True):
            for nz, eps in ((None, 0.0), (noise, 0.25)):
                a = orc.mcts_episode(rules, width, roots, sims, 2, reuse, eval_seed=9, n_threads=4, noise=nz, eps=eps)
                b = orc.mcts_episode_vl_eval(rules, width, roots, sims, 2, reuse, 1, eval_seed=9, n_threads=4, noise=nz, eps=eps)
                assert np.array_equal(a["visits"], b["visits"]) and a["q"].tobytes() == b["q"].tobytes()
                assert a["final"].tobytes() == b["final"].tobytes() and a["stats"].as_dict() == b["stats"].as_dict()
            for W in (2, 16):
                c = orc.mcts_episode_vl_eval(rules, width, roots, sims, 2, reuse, W, eval_seed=9, n_threads=4)
                assert c["stats"].sims == a["stats"].sims
                assert c["stats"].nodes <= a["stats"].nodes          # collisions open no node
                s = orc.mcts_episode_vl(rules, width, roots, sims, 2, reuse, W, n_threads=4)
                t = orc.mcts_episode_vl_eval(rules, width, roots, sims, 2, reuse, W, eval_seed=None, n_threads=4)
                assert np.array_equal(s["visits"], t["visits"]) and s["q"].tobytes() == t["q"].tobytes()
