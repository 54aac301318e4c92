#!/usr/bin/env python
"""Writes tests/golden/mcts_reference_vectors.json by running the UNMODIFIED reference MCTS
(`/root/reference/src/mcts.py`, class MCTS) on tafl positions.

The reference ships no Game/NNet adapter, so this script supplies one, served by the CPU game oracle
(oracle/liboracle.so) — the adapter is the build-defined part:

  * board object        = (packed 64-byte state, action path from the root)
  * stringRepresentation = root-relative action path  -> node identity is the path (tree semantics)
  * action index         = position in the canonical legal-move list (game/main.rs:33-43)
  * getGameEnded(b, 1)   = +1 / -1 from the side to move, `draw_value` for a draw, 0 if Ongoing
  * nnet.predict         = uniform prior over the action space, value 0 (the stub evaluator)

Everything else — PUCT arithmetic, expansion, masking/renormalisation, backup, visit counting — is
the reference's own code.  Needs /root/reference, so it runs in the build container only; the
vectors travel to the GPU box as a committed fixture.

    python tests/golden/make_mcts_reference_vectors.py
"""
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import oracle_ffi as F  # noqa: E402
from alphazeroforhnefatafl_b200 import _abi as A  # noqa: E402
from parity_cases import board_for, sample_positions  # noqa: E402

REF = os.environ.get("TAFL_REFERENCE", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_mcts", os.path.join(REF, "src", "mcts.py"))
ref_mcts = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_mcts)

ACTION_SIZE = A.MAX_MOVES


class Board:
    __slots__ = ("state", "path", "moves", "count")

    def __init__(self, state, path):
        self.state, self.path, self.moves, self.count = state, path, None, None


class TaflGame:
    """alpha-zero-general style Game interface expected by mcts.py (lines 40-41, 75-86, 122-123)."""

    def __init__(self, orc, rules, width, draw_value):
        self.bv, self.width, self.draw_value = orc.by_value(rules, width), width, draw_value

    def _moves(self, b):
        if b.moves is None:
            m, c = self.bv.legal_moves_of(b.state)
            b.moves, b.count = m[0], int(c[0])
        return b.moves, b.count

    def stringRepresentation(self, b):
        return ",".join(map(str, b.path))

    def getActionSize(self):
        return ACTION_SIZE

    def getGameEnded(self, b, player):
        s = int(b.state["status"][0])
        if s == 0:
            return 0
        if (s & 0xC0) == 0xC0:
            return self.draw_value
        winner_def = (s & 1) != 0
        to_play_def = (int(b.state["flags"][0]) & 4) != 0
        return 1.0 if winner_def == to_play_def else -1.0

    def getValidMoves(self, b, player):
        _, c = self._moves(b)
        v = np.zeros(ACTION_SIZE)
        v[:c] = 1
        return v

    def getNextState(self, b, player, a):
        m, c = self._moves(b)
        assert 0 <= a < c
        from alphazeroforhnefatafl_b200 import engine as E
        plays = E.moves_to_plays(np.array([m[a]], np.uint16), self.width)
        nxt, fx = self.bv.do_plays(b.state, plays)
        assert not fx["err"].any()
        return Board(nxt, b.path + (a,)), -player

    def getCanonicalForm(self, b, player):
        return b


class StubNet:
    def predict(self, b):
        return np.full(ACTION_SIZE, 1.0 / ACTION_SIZE), 0.0


def run_case(orc, rules, width, root, sims, cpuct, draw_value):
    game = TaflGame(orc, rules, width, draw_value)
    args = types.SimpleNamespace(numMCTSSims=sims, cpuct=cpuct)
    m = ref_mcts.MCTS(game, StubNet(), args)
    b = Board(root, ())
    for _ in range(sims):                      # getActionProb's loop (mcts.py:37-38)
        m.search(b)
    s = game.stringRepresentation(b)
    _, count = game._moves(b) if int(root["status"][0]) == 0 else (None, 0)
    visits = [int(m.Nsa.get((s, a), 0)) for a in range(count)]
    q = [float(m.Qsa.get((s, a), 0.0)) for a in range(count)]
    # whole-tree digest: (path, a) -> (N, Q) for every visited edge
    edges = sorted((k[0], k[1], int(n), float(m.Qsa[k]).hex()) for k, n in m.Nsa.items())
    terminal = sum(1 for v in m.Es.values() if v != 0)
    return dict(count=count, visits=visits, q_hex=[x.hex() for x in q], n_nodes=len(m.Es), n_expanded=len(m.Ps),
                n_terminal_nodes=terminal, n_edges_visited=len(edges),
                depth2_sample=[e for e in edges if e[0] != ""][:40])


def main():
    orc = F.oracle()
    cases = []
    plan = [  # (name, preset, width, n_positions, seed, max_steps, sims, cpuct, draw_value)
        ("copenhagen", A.PRESET_COPENHAGEN, 128, 4, 11, 200, 160, 1.0, 1e-4),
        ("tablut", A.PRESET_TABLUT, 128, 4, 12, 300, 200, 1.0, 1e-4),
        ("tablut_cpuct", A.PRESET_TABLUT, 128, 3, 13, 400, 200, 1.7, 0.25),
        ("brandubh", A.PRESET_BRANDUBH, 64, 5, 14, 120, 300, 1.0, 1e-4),
        ("brandubh_u128", A.PRESET_BRANDUBH, 128, 2, 15, 120, 150, 2.5, 1e-4),
        ("koch", A.PRESET_KOCH, 64, 2, 16, 150, 200, 1.0, 1e-4),
        # deep searches on small boards: many terminal leaves, Q values mix through the incremental mean
        ("brandubh_deep", A.PRESET_BRANDUBH, 64, 3, 17, 200, 2500, 1.0, 1e-4),
        ("tablut_deep", A.PRESET_TABLUT, 128, 2, 18, 500, 1500, 1.3, 0.05),
        ("magpie_deep", A.PRESET_MAGPIE, 64, 2, 19, 200, 1200, 0.7, 1e-4),
    ]
    for name, preset, width, npos, seed, max_steps, sims, cpuct, dv in plan:
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        # over-sample, prefer roots whose search meets terminal nodes (non-zero Q exercises the arithmetic)
        pool = sample_positions(orc, rules, fen, width, 48, seed=seed, max_steps=max_steps)
        pool = pool[pool["status"] == 0]
        ref = orc.mcts_search(rules, width, pool, sims, cpuct=cpuct, draw_value=dv, n_threads=8)
        score = np.abs(ref["q"]).sum(1)
        order = np.argsort(-score, kind="stable")
        chosen = list(order[: max(1, npos - 1)]) + [int(order[-1])]
        for i in chosen[:npos]:
            root = pool[i:i + 1].copy()
            out = run_case(orc, rules, width, root, sims, cpuct, dv)
            out.update(name=name, preset=preset, width=width, sims=sims, cpuct=cpuct, draw_value=dv,
                       root_hex=root.tobytes().hex())
            cases.append(out)
            print(name, "L=%d" % out["count"], "sum N=%d" % sum(out["visits"]), "terminal nodes=%d" % out["n_terminal_nodes"],
                  "nonzero Q=%d" % sum(1 for x in out["q_hex"] if float.fromhex(x) != 0.0))
    path = os.path.join(HERE, "mcts_reference_vectors.json")
    with open(path, "w") as f:
        json.dump(dict(provenance="generated by tests/golden/make_mcts_reference_vectors.py: UNMODIFIED reference "
                                  "src/mcts.py (class MCTS) over a Game/NNet adapter served by the CPU game oracle",
                       adapter="node = action path from the root; action = canonical move index; stub evaluator "
                               "P uniform, v = 0; getGameEnded +-1 / draw_value",
                       cases=cases), f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", path, len(cases))


if __name__ == "__main__":
    main()
