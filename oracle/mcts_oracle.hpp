// mcts_oracle.hpp — CPU oracle for the MCTS path: a C++ restatement of the reference's
// `class MCTS` (reference tree: src/mcts.py) over the game oracle in tafl_oracle.hpp.
//
// TEST INFRASTRUCTURE ONLY (see the header of tafl_oracle.hpp).
//
// PARITY PIN: src/mcts.py has no tests of its own, but it is importable in the build container, so
// tests/golden/make_mcts_reference_vectors.py runs the UNMODIFIED reference class over a Game/NNet
// adapter served by the game oracle and freezes its visit counts / Q values / node counts in
// tests/golden/mcts_reference_vectors.json; tests/test_mcts_reference_golden.py checks this
// restatement against them bit for bit.  What the reference leaves undefined is the adapter
// (action indexing, stringRepresentation, getGameEnded values, canonical form); the build-defined
// choices, used identically by the adapter, are:
//   * node identity = path from the root (tree, as in src/mcts.rs:9-32), not the transposition-
//     merged dict of mcts.py:20-26 — a tafl state also carries side-to-move and repetition
//     history, so "same board string" is not "same state";
//   * action index = position in the canonical legal-move list (game/main.rs:33-43);
//   * evaluator stub: P(a) = 1/L exactly for each of the L legal moves, v = 0;
//   * getGameEnded: +1 if the side to move at the node has won, -1 if it has lost,
//     `draw_value` (default 1e-4, must be non-zero: mcts.py:79) for a draw;
//   * temp = 0 move choice: first maximum instead of np.random.choice among ties (mcts.py:44-45).
// The arithmetic of select / expand / backup follows mcts.py line by line, in f64, with no FMA
// contraction (compile with -ffp-contract=off).
#pragma once
#include <cmath>
#include <limits>

#include "tafl_oracle.hpp"

namespace orc {

static constexpr double MCTS_EPS = 1e-8;  // mcts.py:6

template <class W>
struct Mcts {
    struct Node {
        GameState<W> state;
        double Es = 0.0;        // mcts.py:77-78  self.Es[s]
        bool expanded = false;  // "s in self.Ps"
        std::vector<Play> moves;      // self.Vs[s] as the list of valid actions
        std::vector<double> Q;        // self.Qsa[(s,a)]
        std::vector<uint32_t> N;      // self.Nsa[(s,a)]; 0 == "(s,a) not in Qsa"
        std::vector<int32_t> child;   // node reached by (s,a), -1 = not created yet
        uint32_t Ns = 0;              // self.Ns[s]
        double P = 0.0;               // self.Ps[s][a], uniform over the valid moves
    };
    const GameLogic& logic;
    double cpuct, draw_value;
    std::vector<Node> nodes;
    tafl_mcts_stats stats{};

    Mcts(const GameLogic& l, double c, double dv) : logic(l), cpuct(c), draw_value(dv) {}

    // game.getGameEnded(canonicalBoard, 1) — build-defined, see header
    double game_ended(const GameState<W>& s) const {
        if (s.status == TAFL_STATUS_ONGOING) return 0.0;
        if (TAFL_STATUS_IS_DRAW(s.status)) return draw_value;
        Side winner = TAFL_STATUS_WINNER_IS_DEFENDER(s.status) ? Side::Defender : Side::Attacker;
        return winner == s.side_to_play ? 1.0 : -1.0;
    }
    int new_node(const GameState<W>& s) {
        Node n;
        n.state = s;
        n.Es = game_ended(s);
        nodes.push_back(std::move(n));
        stats.nodes += 1;
        return (int)nodes.size() - 1;
    }
    void reset(const GameState<W>& root) {
        nodes.clear();
        new_node(root);
    }

    // mcts.py:55-136
    double search(int idx, int depth) {
        if (nodes[idx].Es != 0.0) {  // mcts.py:79-81
            stats.terminal_hits += 1;
            stats.depth_sum += (uint64_t)depth;
            return -nodes[idx].Es;
        }
        if (!nodes[idx].expanded) {  // mcts.py:83-102
            Node& n = nodes[idx];
            n.moves = logic.get_all_possible_moves(n.state);
            size_t L = n.moves.size();
            n.P = L ? 1.0 / (double)L : 0.0;
            n.Q.assign(L, 0.0);
            n.N.assign(L, 0);
            n.child.assign(L, -1);
            n.Ns = 0;
            n.expanded = true;
            stats.edges += L;
            stats.depth_sum += (uint64_t)depth;
            double v = 0.0;
            return -v;
        }
        int best_act = -1;
        {  // mcts.py:104-119
            Node& n = nodes[idx];
            double cur_best = -std::numeric_limits<double>::infinity();
            for (size_t a = 0; a < n.moves.size(); ++a) {
                double u;
                if (n.N[a] > 0) {
                    u = n.Q[a] + ((cpuct * n.P) * std::sqrt((double)n.Ns)) / (double)(1 + n.N[a]);
                } else {
                    u = (cpuct * n.P) * std::sqrt((double)n.Ns + MCTS_EPS);
                }
                if (u > cur_best) { cur_best = u; best_act = (int)a; }
            }
        }
        if (best_act < 0) return 0.0;  // Ongoing node without legal moves: unreachable via do_play
        int a = best_act;
        int ch = nodes[idx].child[a];
        if (ch < 0) {  // mcts.py:122-123  getNextState + getCanonicalForm
            auto res = logic.do_valid_play(nodes[idx].moves[a], nodes[idx].state);
            ch = new_node(res.first);
            nodes[idx].child[a] = ch;
        }
        double v = search(ch, depth + 1);  // mcts.py:125
        Node& n = nodes[idx];
        if (n.N[a] > 0) {  // mcts.py:127-129
            n.Q[a] = ((double)n.N[a] * n.Q[a] + v) / (double)(n.N[a] + 1);
            n.N[a] += 1;
        } else {  // mcts.py:131-133
            n.Q[a] = v;
            n.N[a] = 1;
        }
        n.Ns += 1;  // mcts.py:135
        return -v;
    }

    // mcts.py:28-53 with temp = 0 and first-max tie-break
    void get_action_prob(const GameState<W>& root, uint32_t num_sims) {
        reset(root);
        for (uint32_t i = 0; i < num_sims; ++i) {
            search(0, 0);
            stats.sims += 1;
        }
    }
};

}  // namespace orc
