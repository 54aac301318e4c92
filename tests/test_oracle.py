"""CPU tests of the oracle (oracle/): the reference's own unit tests as known answers, and the
committed golden playout vectors."""
import hashlib
import json
import os
import subprocess

import numpy as np

import kat_runner
import oracle_ffi as F
from alphazeroforhnefatafl_b200 import _abi as A
from parity_cases import board_for

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_oracle_kat_binary(orc):
    """oracle/kat_main.cpp: all 25 reference tests restated natively against the oracle types."""
    exe = os.path.join(F.ORACLE_DIR, "oracle_kat")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "0 failed" in r.stdout


def test_oracle_reference_kats(orc):
    n = kat_runner.run_all(F.OracleHost(), lambda rules, width, side_len: orc.by_value(rules, width))
    assert n >= 150


def test_oracle_golden_playouts(orc):
    with open(os.path.join(GOLDEN, "playout_vectors.json")) as f:
        vectors = json.load(f)["vectors"]
    for v in vectors:
        rules, fen = orc.preset_rules(v["preset"]), board_for(orc, v["preset"])
        st, stats = orc.random_playout(rules, fen, v["width"], v["seed"], v["n_games"], v["n_steps"],
                                       auto_reset=v["auto_reset"], game_id_base=v["game_id_base"], n_threads=4)
        assert stats.as_dict() == v["stats"], v["name"]
        assert hashlib.sha256(st.tobytes()).hexdigest() == v["states_sha256"], v["name"]


def test_oracle_start_move_counts(orc):
    """legal moves at the start — Copenhagen 116, Tablut 80, Brandubh 40 (attackers)."""
    for preset, width, n in ((A.PRESET_COPENHAGEN, 128, 116), (A.PRESET_TABLUT, 128, 80), (A.PRESET_BRANDUBH, 64, 40)):
        rules = orc.preset_rules(preset)
        rc, st = orc.state_from_fen(orc.preset_board(preset), width, A.ATTACKER)
        assert rc == 0
        _, counts = orc.by_value(rules, width).legal_moves_of(st)
        assert int(counts[0]) == n


def test_oracle_threads_deterministic(orc):
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    a, sa = orc.random_playout(rules, fen, 64, 3, 40, 300, n_threads=1)
    b, sb = orc.random_playout(rules, fen, 64, 3, 40, 300, n_threads=5)
    assert a.tobytes() == b.tobytes() and sa.as_dict() == sb.as_dict()
    # game-id sharding: two half ranges == the full range (multi-GPU contract)
    c, _ = orc.random_playout(rules, fen, 64, 3, 20, 300, game_id_base=0)
    d, _ = orc.random_playout(rules, fen, 64, 3, 20, 300, game_id_base=20)
    assert np.concatenate([c, d]).tobytes() == a.tobytes()
