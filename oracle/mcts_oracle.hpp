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
#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <utility>

#include "tafl_oracle.hpp"

namespace orc {

static constexpr double MCTS_EPS = 1e-8;  // mcts.py:6

// np.sum over a contiguous float64 vector, as mcts.py:88,99 call it.  numpy is a third-party
// dependency of the reference that is not part of /root/reference (pinned here: numpy 2.3.5, the
// version the golden vectors were generated with); its add-reduce is the published pairwise
// summation: < 8 elements sequential; <= 128 elements eight stride-8 accumulators combined as
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) plus a sequential tail; larger inputs split at n/2 rounded
// down to a multiple of 8.  The result is added to the identity 0.0.  Checked against np.sum in
// tests/test_mcts_eval_golden.py.
inline double np_pairwise_sum(const double* a, size_t n) {
    if (n < 8) {
        double res = 0.0;
        for (size_t i = 0; i < n; ++i) res += a[i];
        return res;
    }
    if (n <= 128) {
        double r[8];
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        size_t i = 8;
        for (; i < n - (n % 8); i += 8)
            for (int j = 0; j < 8; ++j) r[j] += a[i + j];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    }
    size_t n2 = n / 2;
    n2 -= n2 % 8;
    return np_pairwise_sum(a, n2) + np_pairwise_sum(a + n2, n - n2);
}
inline double np_sum(const double* a, size_t n) { return 0.0 + np_pairwise_sum(a, n); }

// The deterministic TEST evaluator behind TAFL_EVAL_HASH (include/tafl.h): priors and value are a
// hash of (seed, position, side to move), exactly representable in f32.  One state in 16 returns
// all-zero priors so that the "all valid moves were masked" branch (mcts.py:91-99) is exercised.
inline void hash_eval(uint64_t seed, const tafl_state& st, float* P /*[TAFL_MAX_MOVES]*/, float* v) {
    const uint64_t side = (st.flags & TAFL_FLAG_DEF_TO_PLAY) ? 1u : 0u;
    uint64_t h = splitmix64(st.def[1] ^ side);
    h = splitmix64(st.def[0] ^ h);
    h = splitmix64(st.att[1] ^ h);
    h = splitmix64(st.att[0] ^ h);
    h = splitmix64(seed ^ h);
    const bool masked = (h & 15u) == 0;
    for (uint32_t a = 0; a < TAFL_MAX_MOVES; ++a) {
        const uint64_t x = splitmix64(h + (uint64_t)(a + 1) * 0x9E3779B97F4A7C15ull);
        P[a] = masked ? 0.0f : (float)((x >> 40) + 1) * 5.9604644775390625e-08f;   // (0, 1], step 2^-24
    }
    const uint64_t y = splitmix64(h ^ 0xD1B54A32D192ED03ull);
    *v = ((float)(int64_t)(y >> 40) - 8388608.0f) * 5.9604644775390625e-08f;        // [-0.5, 0.5)
}

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
        std::vector<uint32_t> VL;     // virtual-loss mode: simulations of the current batch that hold (s,a) on their path
        uint32_t vln = 0;             //                    ... that pass through s
        double v_leaf = 0.0;          //                    the evaluator's value when s was expanded (a collision takes it)
        double P = 0.0;               // self.Ps[s][a], uniform over the valid moves
        std::vector<double> Pa;       // evaluator mode: self.Ps[s][a] per action (masked, renormalised)
    };
    // evaluator mode (nnet.predict, mcts.py:85): fills P[TAFL_MAX_MOVES] and v for a position.  The
    // boundary is f32 (what a network emits), widened to f64 before masking (adapter: predict
    // returns P.astype(float64), float(v)).  Unset = the uniform-prior / value-0 stub.
    std::function<void(const GameState<W>&, float*, float*)> eval;
    // Root exploration noise (the reference's TODO at src/mcts.rs:53, absent from src/mcts.py; build-defined):
    // the priors of the node a search starts from become (1 - eps) * P[a] + eps * noise[a], once per
    // search, when that node is expanded or -- with a retained tree -- before the first simulation.
    std::vector<double> root_noise;
    double noise_eps = 0.0;
    bool root_noised = false;
    void noise_root() {
        Node& n = nodes[root];
        if (noise_eps == 0.0 || root_noised || !n.expanded || !eval) return;
        for (size_t a = 0; a < n.Pa.size(); ++a) n.Pa[a] = (1.0 - noise_eps) * n.Pa[a] + noise_eps * root_noise[a];
        root_noised = true;
    }
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
    void reset(const GameState<W>& root_state) {
        nodes.clear();
        root = 0;
        new_node(root_state);
    }

    // mcts.py:55-136
    double search(int idx, int depth) {
        if (nodes[idx].Es != 0.0) {  // mcts.py:79-81
            stats.terminal_hits += 1;
            stats.depth_sum += (uint64_t)depth;
            return -nodes[idx].Es;
        }
        if (!nodes[idx].expanded) {  // mcts.py:83-102
            stats.depth_sum += (uint64_t)depth;
            return -expand(idx);
        }
        return search_expanded(idx, depth);
    }
    // mcts.py:83-101: Ps, Vs, Ns of a leaf; returns the evaluator's v (0 for the stub)
    double expand(int idx) {
        {
            Node& n = nodes[idx];
            n.moves = logic.get_all_possible_moves(n.state);
            size_t L = n.moves.size();
            double v_eval = 0.0;
            if (eval) {  // mcts.py:85-99 with a real evaluator: mask, renormalise, uniform fallback
                std::vector<float> pf(TAFL_MAX_MOVES);
                float vf = 0.0f;
                eval(n.state, pf.data(), &vf);
                std::vector<double> ps(TAFL_MAX_MOVES), valids(TAFL_MAX_MOVES, 0.0);
                for (size_t a = 0; a < L && a < TAFL_MAX_MOVES; ++a) valids[a] = 1.0;
                for (size_t a = 0; a < TAFL_MAX_MOVES; ++a) ps[a] = (double)pf[a] * valids[a];   // :87
                double sum = np_sum(ps.data(), ps.size());                                       // :88
                if (sum > 0) {
                    for (auto& x : ps) x /= sum;                                                 // :90
                } else {
                    for (size_t a = 0; a < TAFL_MAX_MOVES; ++a) ps[a] = ps[a] + valids[a];       // :98
                    double s2 = np_sum(ps.data(), ps.size());
                    for (auto& x : ps) x /= s2;                                                  // :99
                }
                n.Pa.assign(ps.begin(), ps.begin() + (std::ptrdiff_t)L);
                v_eval = (double)vf;
            }
            n.P = L ? 1.0 / (double)L : 0.0;
            n.Q.assign(L, 0.0);
            n.N.assign(L, 0);
            n.VL.assign(L, 0);
            n.child.assign(L, -1);
            n.Ns = 0;
            n.vln = 0;
            n.v_leaf = v_eval;
            n.expanded = true;
            stats.edges += L;
            if (idx == root) noise_root();
            return v_eval;
        }
    }
    double search_expanded(int idx, int depth) {
        int best_act = -1;
        {  // mcts.py:104-119
            Node& n = nodes[idx];
            double cur_best = -std::numeric_limits<double>::infinity();
            for (size_t a = 0; a < n.moves.size(); ++a) {
                double u;
                const double P = eval ? n.Pa[a] : n.P;
                if (n.N[a] > 0) {
                    u = n.Q[a] + ((cpuct * P) * std::sqrt((double)n.Ns)) / (double)(1 + n.N[a]);
                } else {
                    u = (cpuct * P) * std::sqrt((double)n.Ns + MCTS_EPS);
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
        root_noised = false;
        for (uint32_t i = 0; i < num_sims; ++i) {
            search(0, 0);
            stats.sims += 1;
        }
    }

    // Tree reuse: the reference keeps ONE MCTS object per episode, so the statistics of the
    // subtree below the move played survive into the next getActionProb (its dicts are never
    // cleared, mcts.py:20-26).  `root` walks down the retained tree.
    int root = 0;
    void run_from_root(uint32_t num_sims) {
        if (vl_width > 1) { run_from_root_vl(num_sims); return; }
        root_noised = false;
        noise_root();
        for (uint32_t i = 0; i < num_sims; ++i) {
            search(root, 0);
            stats.sims += 1;
        }
    }

    // -----------------------------------------------------------------------------------------------------------
    // Virtual-loss mode (north star: "select (PUCT with virtual loss)"; intent at src/mcts.rs:49-55; NOT in mcts.py,
    // so build-defined and pinned only by this restatement): `vl_width` simulations are in flight per tree.
    // One batch = w <= vl_width simulations, run in three sweeps, each in simulation-index order:
    //   select   every simulation descends from the root with PUCT over the statistics AS IF each simulation that
    //            currently holds an edge on its path had already lost there: n = N + vl, q = (N Q - vl) / n,
    //            Ns_eff = Ns + vl(node); it then adds itself to vl along its own path.  A simulation stops at a
    //            terminal node, at the (unexpanded) root, at the first edge nobody has followed yet (it creates the
    //            child), or — collision — at an edge whose child another simulation of the same batch is creating.
    //   expand   the new leaves are expanded (mcts.py:83-101 with the evaluator in use; stub: uniform priors, v = 0); a
    //            collision takes the same value as the simulation that created the child.
    //   backup   every simulation removes its virtual losses and applies the real update of mcts.py:125-136.
    // While the root is unexpanded a batch holds one simulation.  With vl_width == 1 no simulation ever sees a
    // virtual loss, and the result is mcts.py's, bit for bit (the golden vectors guard that).
    // With an evaluator the batch is what one evaluator call sees per tree: vl_width leaves (the hook's leaf batch is
    // then n_trees x vl_width rows).
    // -----------------------------------------------------------------------------------------------------------
    uint32_t vl_width = 1;
    void run_from_root_vl(uint32_t num_sims) {
        enum { K_TERM = 1, K_ROOT, K_EMPTY, K_NEW, K_COLL };
        struct Sim { std::vector<std::pair<int, int>> path; int leaf = 0, kind = 0; uint32_t depth = 0; double v = 0.0; };
        uint32_t done = 0;
        root_noised = false;
        noise_root();
        while (done < num_sims) {
            uint32_t w = std::min<uint32_t>(vl_width, num_sims - done);
            if (!nodes[root].expanded && nodes[root].Es == 0.0) w = 1;
            std::vector<Sim> sims(w);
            for (uint32_t i = 0; i < w; ++i) {   // ---- select
                Sim& S = sims[i];
                int idx = root;
                while (true) {
                    if (nodes[idx].Es != 0.0) { S.kind = K_TERM; S.leaf = idx; S.v = -nodes[idx].Es; break; }
                    if (!nodes[idx].expanded) { S.kind = K_ROOT; S.leaf = idx; break; }
                    const size_t L = nodes[idx].moves.size();
                    if (L == 0) { S.kind = K_EMPTY; S.leaf = idx; S.v = 0.0; break; }
                    int a = -1;
                    {
                        Node& n = nodes[idx];
                        const uint32_t ns_eff = n.Ns + n.vln;
                        const double sq_vis = std::sqrt((double)ns_eff), sq_new = std::sqrt((double)ns_eff + MCTS_EPS);
                        double cur_best = -std::numeric_limits<double>::infinity();
                        for (size_t k = 0; k < L; ++k) {
                            const uint32_t nn = n.N[k] + n.VL[k];
                            const double c = cpuct * (eval ? n.Pa[k] : n.P);
                            double u;
                            if (nn > 0) {
                                const double qv = n.VL[k] ? ((double)n.N[k] * n.Q[k] - (double)n.VL[k]) / (double)nn : n.Q[k];
                                u = qv + (c * sq_vis) / (double)(1u + nn);
                            } else {
                                u = c * sq_new;
                            }
                            if (u > cur_best) { cur_best = u; a = (int)k; }
                        }
                    }
                    const uint32_t vl_before = nodes[idx].VL[a];
                    nodes[idx].VL[a] += 1;
                    nodes[idx].vln += 1;
                    S.path.push_back({idx, a});
                    S.depth += 1;
                    if (nodes[idx].N[a] == 0) {
                        if (vl_before > 0) { S.kind = K_COLL; S.leaf = nodes[idx].child[a]; break; }
                        auto res = logic.do_valid_play(nodes[idx].moves[a], nodes[idx].state);
                        const int ch = new_node(res.first);
                        nodes[idx].child[a] = ch;
                        S.kind = K_NEW;
                        S.leaf = ch;
                        break;
                    }
                    idx = nodes[idx].child[a];
                }
            }
            for (uint32_t i = 0; i < w; ++i) {   // ---- expand
                Sim& S = sims[i];
                Node& n = nodes[S.leaf];
                if (S.kind == K_EMPTY) continue;
                stats.depth_sum += S.depth;
                if (S.kind == K_TERM) { stats.terminal_hits += 1; continue; }
                if (n.Es != 0.0) { S.v = -n.Es; stats.terminal_hits += 1; continue; }   // new / collided child is terminal
                if (S.kind == K_COLL) { S.v = -n.v_leaf; continue; }   // the value the simulation that opened the child got
                S.v = -expand(S.leaf);
            }
            for (uint32_t i = 0; i < w; ++i) {   // ---- backup (mcts.py:125-136), virtual losses removed
                Sim& S = sims[i];
                double v = S.v;
                for (size_t k = S.path.size(); k-- > 0;) {
                    Node& n = nodes[S.path[k].first];
                    const int a = S.path[k].second;
                    n.VL[a] -= 1;
                    n.vln -= 1;
                    if (n.N[a] > 0) {
                        n.Q[a] = ((double)n.N[a] * n.Q[a] + v) / (double)(n.N[a] + 1);
                        n.N[a] += 1;
                    } else {
                        n.Q[a] = v;
                        n.N[a] = 1;
                    }
                    n.Ns += 1;
                    v = -v;
                }
            }
            stats.sims += w;
            done += w;
        }
    }
    // play action a at the root; false if (root, a) was never traversed (no child to descend to)
    bool advance(int a) {
        if (a < 0 || (size_t)a >= nodes[root].child.size() || nodes[root].child[a] < 0) return false;
        root = nodes[root].child[a];
        return true;
    }
};

}  // namespace orc
