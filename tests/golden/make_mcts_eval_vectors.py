#!/usr/bin/env python
"""Writes tests/golden/mcts_eval_vectors.json: the UNMODIFIED reference MCTS
(`/root/reference/src/mcts.py`) driven with a NON-UNIFORM evaluator, so that the reference's own
mask / renormalise / "all valid moves were masked" code (mcts.py:85-99), per-action priors in PUCT
(mcts.py:112-115) and non-zero leaf values in the backup (mcts.py:102,127-133) are pinned.

The evaluator is TAFL_EVAL_HASH (include/tafl.h): priors and value are a hash of (seed, position,
side to move).  It is implemented HERE in numpy, independently of the C++ oracle and the CUDA
kernel, and handed to mcts.py through ``nnet.predict`` as (P.astype(float64), float(v)).  The
Game adapter is the one of make_mcts_reference_vectors.py (build-defined part).

Needs /root/reference, so it runs in the build container only.

    python tests/golden/make_mcts_eval_vectors.py
"""
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import make_mcts_reference_vectors as base  # noqa: E402  (loads the reference module + adapter)
from make_mcts_reference_vectors import A, F, Board, TaflGame, board_for, ref_mcts, sample_positions  # noqa: E402

M64 = (1 << 64) - 1


def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def hash_eval(seed, state):
    """TAFL_EVAL_HASH in plain Python integers; returns (P float32[MAX_MOVES], v float32)."""
    att, de = [int(x) for x in state["att"][0]], [int(x) for x in state["def"][0]]
    side = 1 if int(state["flags"][0]) & 4 else 0
    h = splitmix64(de[1] ^ side)
    h = splitmix64(de[0] ^ h)
    h = splitmix64(att[1] ^ h)
    h = splitmix64(att[0] ^ h)
    h = splitmix64(seed ^ h)
    P = np.zeros(A.MAX_MOVES, np.float32)
    if (h & 15) != 0:
        for a in range(A.MAX_MOVES):
            x = splitmix64((h + (a + 1) * 0x9E3779B97F4A7C15) & M64)
            P[a] = np.float32((x >> 40) + 1) * np.float32(2.0 ** -24)
    y = splitmix64(h ^ 0xD1B54A32D192ED03)
    v = (np.float32((y >> 40)) - np.float32(8388608.0)) * np.float32(2.0 ** -24)
    return P, np.float32(v)


class HashNet:
    def __init__(self, seed):
        self.seed, self.calls, self.masked = seed, 0, 0

    def predict(self, b):
        P, v = hash_eval(self.seed, b.state)
        self.calls += 1
        self.masked += int(not P.any())
        return P.astype(np.float64), float(v)


def run_case(orc, rules, width, root, sims, cpuct, draw_value, seed):
    game = TaflGame(orc, rules, width, draw_value)
    net = HashNet(seed)
    m = ref_mcts.MCTS(game, net, types.SimpleNamespace(numMCTSSims=sims, cpuct=cpuct))
    b = Board(root, ())
    for _ in range(sims):
        m.search(b)
    s = game.stringRepresentation(b)
    _, count = game._moves(b)
    visits = [int(m.Nsa.get((s, a), 0)) for a in range(count)]
    q = [float(m.Qsa.get((s, a), 0.0)) for a in range(count)]
    priors = [float(x).hex() for x in m.Ps[s][:count]]
    return dict(count=count, visits=visits, q_hex=[x.hex() for x in q], root_prior_hex=priors, n_nodes=len(m.Es),
                n_expanded=len(m.Ps), evals=net.calls, masked_evals=net.masked)


def main():
    orc = F.oracle()
    cases = []
    plan = [  # (name, preset, width, n_positions, pos_seed, max_steps, sims, cpuct, draw_value, eval_seed)
        ("copenhagen", A.PRESET_COPENHAGEN, 128, 4, 21, 200, 200, 1.0, 1e-4, 101),
        ("copenhagen_cpuct", A.PRESET_COPENHAGEN, 128, 2, 22, 300, 300, 2.5, 1e-4, 102),
        ("tablut", A.PRESET_TABLUT, 128, 4, 23, 300, 300, 1.0, 1e-4, 103),
        ("brandubh", A.PRESET_BRANDUBH, 64, 5, 24, 120, 400, 1.0, 1e-4, 104),
        ("brandubh_deep", A.PRESET_BRANDUBH, 64, 2, 25, 200, 2000, 1.3, 0.05, 105),
        ("koch", A.PRESET_KOCH, 64, 2, 26, 150, 300, 0.7, 1e-4, 106),
    ]
    for name, preset, width, npos, pseed, max_steps, sims, cpuct, dv, eseed in plan:
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        pool = sample_positions(orc, rules, fen, width, 32, seed=pseed, max_steps=max_steps)
        pool = pool[pool["status"] == 0]
        for i in range(npos):
            root = pool[i:i + 1].copy()
            out = run_case(orc, rules, width, root, sims, cpuct, dv, eseed)
            out.update(name=name, preset=preset, width=width, sims=sims, cpuct=cpuct, draw_value=dv, eval_seed=eseed,
                       root_hex=root.tobytes().hex())
            cases.append(out)
            print(name, "L=%d" % out["count"], "sum N=%d" % sum(out["visits"]), "evals=%d masked=%d" % (out["evals"], out["masked_evals"]))
    # getActionProb itself (mcts.py:28-53), temperatures 1, 0.5, 2 and 0, on fresh MCTS objects
    policy = []
    for name, preset, width, pseed, sims, eseed in (("copenhagen", A.PRESET_COPENHAGEN, 128, 27, 120, 107),
                                                     ("brandubh", A.PRESET_BRANDUBH, 64, 28, 250, 108)):
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        pool = sample_positions(orc, rules, fen, width, 16, seed=pseed, max_steps=120)
        pool = pool[pool["status"] == 0]
        for i in range(3):
            root = pool[i:i + 1].copy()
            game = TaflGame(orc, rules, width, 1e-4)
            out = dict(name=name, preset=preset, width=width, sims=sims, cpuct=1.0, draw_value=1e-4, eval_seed=eseed,
                       root_hex=root.tobytes().hex(), probs={})
            for temp in (1, 0.5, 2, 0):
                m = ref_mcts.MCTS(game, HashNet(eseed), types.SimpleNamespace(numMCTSSims=sims, cpuct=1.0))
                np.random.seed(0)
                probs = m.getActionProb(Board(root, ()), temp=temp)
                _, count = game._moves(Board(root, ()))
                assert all(p == 0 for p in probs[count:])
                if temp == 0:   # the reference picks randomly among ties; keep the tie set
                    s = game.stringRepresentation(Board(root, ()))
                    cnt = [int(m.Nsa.get((s, a), 0)) for a in range(count)]
                    out["ties_t0"] = [a for a in range(count) if cnt[a] == max(cnt)]
                    out["visits"] = cnt
                else:
                    out["probs"][str(temp)] = [float(x).hex() for x in probs[:count]]
            policy.append(out)
            print("policy", name, i)
    # tree reuse: ONE reference MCTS object per episode (its dicts persist, mcts.py:20-26), first-max move played
    episodes = []
    for name, preset, width, pseed, sims, n_moves, eseed in (("copenhagen", A.PRESET_COPENHAGEN, 128, 29, 100, 6, 109),
                                                              ("brandubh", A.PRESET_BRANDUBH, 64, 30, 200, 10, 110),
                                                              ("brandubh_stub", A.PRESET_BRANDUBH, 64, 32, 200, 10, None),
                                                              ("tablut_stub", A.PRESET_TABLUT, 128, 33, 150, 6, None)):
        rules, fen = orc.preset_rules(preset), board_for(orc, preset)
        pool = sample_positions(orc, rules, fen, width, 16, seed=pseed, max_steps=100)
        pool = pool[pool["status"] == 0]
        for i in range(2):
            root = pool[i:i + 1].copy()
            game = TaflGame(orc, rules, width, 1e-4)
            net = HashNet(eseed) if eseed is not None else base.StubNet()
            m = ref_mcts.MCTS(game, net, types.SimpleNamespace(numMCTSSims=sims, cpuct=1.0))
            b = Board(root, ())
            moves = []
            for _ in range(n_moves):
                if game.getGameEnded(b, 1) != 0:
                    break
                for _ in range(sims):
                    m.search(b)
                s = game.stringRepresentation(b)
                _, count = game._moves(b)
                visits = [int(m.Nsa.get((s, a), 0)) for a in range(count)]
                q = [float(m.Qsa.get((s, a), 0.0)).hex() for a in range(count)]
                a = int(np.argmax(visits))
                moves.append(dict(visits=visits, q_hex=q, best=a))
                b, _ = game.getNextState(b, 1, a)
            episodes.append(dict(name=name, preset=preset, width=width, sims=sims, cpuct=1.0, draw_value=1e-4, eval_seed=eseed,
                                 root_hex=root.tobytes().hex(), n_moves=n_moves, moves=moves, final_hex=b.state.tobytes().hex()))
            print("episode", name, i, [sum(mv["visits"]) for mv in moves])
    # evaluator known answers: P[:8] and v for a few positions (pins the numpy / C++ / CUDA copies)
    kat = []
    rules, fen = orc.preset_rules(A.PRESET_COPENHAGEN), board_for(orc, A.PRESET_COPENHAGEN)
    pool = sample_positions(orc, rules, fen, 128, 24, seed=31, max_steps=100)
    for i in range(len(pool)):
        P, v = hash_eval(777, pool[i:i + 1])
        kat.append(dict(state_hex=pool[i:i + 1].tobytes().hex(), seed=777, p_first=[float(x) for x in P[:8]],
                        p_last=float(P[-1]), p_sum_f64=float(P.astype(np.float64).sum()), v=float(v)))
    path = os.path.join(HERE, "mcts_eval_vectors.json")
    with open(path, "w") as f:
        json.dump(dict(provenance="generated by tests/golden/make_mcts_eval_vectors.py: UNMODIFIED reference src/mcts.py "
                                  "(class MCTS) with the TAFL_EVAL_HASH evaluator implemented in numpy; numpy %s" % np.__version__,
                       cases=cases, evaluator_kat=kat, policy=policy, episodes=episodes), f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", path, len(cases))


if __name__ == "__main__":
    main()
