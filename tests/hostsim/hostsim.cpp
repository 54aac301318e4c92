// hostsim.cpp — TEST HARNESS ONLY.  Compiles the kernels' __host__ __device__ game logic
// (alphazeroforhnefatafl_b200/csrc/tafl_core.cuh, tafl_select.cuh) with g++ so the exact device
// math can be differentially tested against the CPU oracle on this GPU-less machine.  It is
// not part of the product and is never loaded by the package.
// Signatures mirror oracle/oracle_capi.cpp (`sim_` instead of `orc_`).
#include <cstring>
#include <vector>

#include "../../alphazeroforhnefatafl_b200/csrc/tafl_ctx.hpp"
#include "../../alphazeroforhnefatafl_b200/csrc/tafl_select.cuh"
#include "../../alphazeroforhnefatafl_b200/csrc/tafl_wide.cuh"

using namespace tafl;

namespace {
template <class F>
int dispatch(int width, F&& f) {
    if (width == 64) return f(BB64{});
    if (width == 128) return f(BB128{});
    return TAFL_ERR_INVALID_ARG;
}
template <class BB> void put(uint64_t o[2], BB b) { o[0] = b.lo64(); o[1] = b.hi64(); }

// Build the Ctx and hand it to f — as the preset SPECIALISATION when the rules and board match one
// (exactly the engine's choice, tafl_engine_create), else as the generic run-time Ctx.
int g_force_generic = 0;
int g_spec_used[SPEC_COUNT] = {0, 0, 0, 0};
template <class BB, class F>
void with_ctx(const tafl_rules& r, int N, F&& f) {
    const Ctx<BB> cx = make_ctx<BB>(r, N);
    if (!g_force_generic) {
        if constexpr (BB::BITS == 64) {
            if (spec_matches<SPEC_BRANDUBH>(cx)) { g_spec_used[SPEC_BRANDUBH]++; f(spec_cast<SPEC_BRANDUBH>(cx)); return; }
        } else {
            if (spec_matches<SPEC_COPENHAGEN>(cx)) { g_spec_used[SPEC_COPENHAGEN]++; f(spec_cast<SPEC_COPENHAGEN>(cx)); return; }
            if (spec_matches<SPEC_TABLUT>(cx)) { g_spec_used[SPEC_TABLUT]++; f(spec_cast<SPEC_TABLUT>(cx)); return; }
        }
    }
    g_spec_used[SPEC_GENERIC]++;
    f(cx);
}
}  // namespace

extern "C" {

void sim_force_generic(int on) { g_force_generic = on; }
int sim_spec_used(int spec) { return (spec >= 0 && spec < SPEC_COUNT) ? g_spec_used[spec] : -1; }

int sim_validate_plays(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                       uint8_t* out_err) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            out_err[i] = (uint8_t)validate_play(cx, st[i], pl[i], (st[i].flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
            });
        }
        return 0;
    });
}
int sim_do_plays(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                 tafl_state* out_st, tafl_play_effects* out_fx) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            tafl_play_effects fx{};
            int err = validate_play(cx, st[i], pl[i], (st[i].flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
            fx.err = (uint8_t)err;
            out_st[i] = st[i];
            if (!err) {
                BB caps = do_valid_play(cx, out_st[i], pl[i].from_row, pl[i].from_col, pl[i].axis == TAFL_HORIZONTAL,
                                        pl[i].disp);
                put(fx.captures, caps);
                fx.n_captures = (uint8_t)caps.popc();
            }
            fx.status = out_st[i].status;
            if (out_fx) out_fx[i] = fx;
            });
        }
        return 0;
    });
}
int sim_legal_moves_of(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, tafl_move* moves,
                       uint32_t* counts) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            counts[i] = (uint32_t)gen_moves(cx, st[i], moves + (size_t)i * TAFL_MAX_MOVES, TAFL_MAX_MOVES);
            });
        }
        return 0;
    });
}
int sim_piece_plays(const tafl_rules* r, int width, const tafl_state* st, const uint8_t* tiles_rc, uint32_t n,
                    tafl_move* moves, int32_t* counts) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            counts[i] = piece_plays(cx, st[i], tiles_rc[2 * i], tiles_rc[2 * i + 1], moves + (size_t)i * 32);
            });
        }
        return 0;
    });
}
int sim_get_captures(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                     tafl_play_effects* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            out[i] = query_captures(cx, st[i], pl[i]);
            });
        }
        return 0;
    });
}
int sim_detect_shieldwall(const tafl_rules* r, int width, const tafl_state* st, const tafl_play* pl, uint32_t n,
                          tafl_play_effects* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            out[i] = query_shieldwall(cx, st[i], pl[i]);
            });
        }
        return 0;
    });
}
int sim_find_enclosure(const tafl_rules* r, int width, const tafl_state* st, const tafl_enclosure_query* q,
                       uint32_t n, tafl_enclosure* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            uint8_t sec;
            out[i] = query_enclosure(cx, st[i], q[i], &sec);
            });
        }
        return 0;
    });
}
int sim_enclosure_secure(const tafl_rules* r, int width, const tafl_state* st, const tafl_enclosure_query* q,
                         uint32_t n, uint8_t* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            query_enclosure(cx, st[i], q[i], &out[i]);
            });
        }
        return 0;
    });
}
int sim_detect_exit_fort(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, uint8_t* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            out[i] = detect_exit_fort(cx, load_pos(cx, st[i]));
            });
        }
        return 0;
    });
}
int sim_side_can_play(const tafl_rules* r, int width, const tafl_state* st, const uint8_t* sides, uint32_t n,
                      uint8_t* out) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            out[i] = query_can_play(cx, st[i], sides[i] != 0);
            });
        }
        return 0;
    });
}

// legal-move list rebuilt from the compressed piece table the MCTS node pools store (tafl_select.cuh)
int sim_piece_table_moves(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, tafl_move* moves,
                          uint32_t* counts) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
                counts[i] = (uint32_t)pt_gen_moves(cx, st[i], moves + (size_t)i * TAFL_MAX_MOVES, TAFL_MAX_MOVES);
            });
        }
        return 0;
    });
}

// ---- large boards (BB256 / BB512, tafl_state_w): the device functions behind tafl_w_* (csrc/tafl_wide.cuh) -------
}  // extern "C"
namespace {
template <class F>
int dispatch_w(int width, F&& f) {
    if (width == 256) return f(BB256{});
    if (width == 512) return f(BB512{});
    return TAFL_ERR_INVALID_ARG;
}
}  // namespace
extern "C" {
int sim_w_validate_plays(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            const Ctx<BB> cx = make_ctx<BB>(*r, st[i].side_len);
            out[i] = (uint8_t)validate_play(cx, st[i], pl[i], (st[i].flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
        }
        return 0;
    });
}
int sim_w_do_plays(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_state_w* out_st,
                   tafl_play_effects_w* out_fx) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            const Ctx<BB> cx = make_ctx<BB>(*r, st[i].side_len);
            out_st[i] = st[i];
            const tafl_play_effects_w fx = do_play_w(cx, out_st[i], pl[i]);
            if (out_fx) out_fx[i] = fx;
        }
        return 0;
    });
}
int sim_w_legal_plays_of(const tafl_rules* r, int width, const tafl_state_w* st, uint32_t n, tafl_play* plays, uint32_t* counts) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            const Ctx<BB> cx = make_ctx<BB>(*r, st[i].side_len);
            counts[i] = (uint32_t)gen_plays(cx, st[i], plays + (size_t)i * TAFL_W_MAX_PLAYS, TAFL_W_MAX_PLAYS);
        }
        return 0;
    });
}
int sim_w_piece_plays(const tafl_rules* r, int width, const tafl_state_w* st, const uint8_t* rc, uint32_t n, tafl_play* plays, int32_t* counts) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            const Ctx<BB> cx = make_ctx<BB>(*r, st[i].side_len);
            counts[i] = piece_plays_w(cx, st[i], rc[2 * i], rc[2 * i + 1], plays + (size_t)i * 64, 64);
        }
        return 0;
    });
}
int sim_w_get_captures(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_play_effects_w* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) out[i] = query_captures_w(make_ctx<BB>(*r, st[i].side_len), st[i], pl[i]);
        return 0;
    });
}
int sim_w_detect_shieldwall(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_play* pl, uint32_t n, tafl_play_effects_w* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) out[i] = query_shieldwall_w(make_ctx<BB>(*r, st[i].side_len), st[i], pl[i]);
        return 0;
    });
}
int sim_w_find_enclosure(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_enclosure_query* q, uint32_t n, tafl_enclosure_w* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) { uint8_t sec; out[i] = query_enclosure_w(make_ctx<BB>(*r, st[i].side_len), st[i], q[i], &sec); }
        return 0;
    });
}
int sim_w_enclosure_secure(const tafl_rules* r, int width, const tafl_state_w* st, const tafl_enclosure_query* q, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) query_enclosure_w(make_ctx<BB>(*r, st[i].side_len), st[i], q[i], &out[i]);
        return 0;
    });
}
int sim_w_detect_exit_fort(const tafl_rules* r, int width, const tafl_state_w* st, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            const Ctx<BB> cx = make_ctx<BB>(*r, st[i].side_len);
            out[i] = detect_exit_fort(cx, load_pos(cx, st[i])) ? 1 : 0;
        }
        return 0;
    });
}
int sim_w_side_can_play(const tafl_rules* r, int width, const tafl_state_w* st, const uint8_t* sides, uint32_t n, uint8_t* out) {
    return dispatch_w(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) out[i] = query_can_play_w(make_ctx<BB>(*r, st[i].side_len), st[i], sides[i] != 0);
        return 0;
    });
}

// the per-piece tile-arithmetic emission of k_legal_moves_g (emit_piece_plays), pieces ascending
int sim_emit_moves(const tafl_rules* r, int width, const tafl_state* st, uint32_t n, tafl_move* moves, uint32_t* counts) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
                counts[i] = (uint32_t)emit_gen_moves(cx, st[i], moves + (size_t)i * TAFL_MAX_MOVES, TAFL_MAX_MOVES);
            });
        }
        return 0;
    });
}

// Random playouts through the fused count+select+apply path the step kernel uses.
int sim_random_playout(const tafl_rules* r, const tafl_state* start, int width, uint64_t seed, uint64_t game_id_base,
                       uint32_t n_games, uint64_t ply0, uint32_t n_steps, int auto_reset, tafl_state* states,
                       tafl_stats* stats) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        StepStats acc{};
        with_ctx<BB>(*r, start->side_len, [&](const auto& cx) {
            for (uint32_t g = 0; g < n_games; ++g) {
                tafl_state s = states[g];
                for (uint32_t i = 0; i < n_steps; ++i)
                    random_step(cx, s, *start, seed, game_id_base + g, ply0 + i, auto_reset != 0, acc);
                states[g] = s;
            }
        });
        fold_stats(acc, stats);
        return 0;
    });
}
// k-th move of the canonical list through the bit-sliced selector (tafl_select.cuh)
int sim_select_moves(const tafl_rules* r, int width, const tafl_state* st, const uint32_t* ks, uint32_t n,
                     tafl_move* out_moves, uint32_t* out_counts) {
    return dispatch(width, [&](auto z) {
        using BB = decltype(z);
        for (uint32_t i = 0; i < n; ++i) {
            with_ctx<BB>(*r, st[i].side_len, [&](const auto& cx) {
            Pos<BB> p = load_pos(cx, st[i]);
            MoveSel<BB> ms = count_moves(cx, p, (st[i].flags & TAFL_FLAG_DEF_TO_PLAY) != 0);
            out_counts[i] = (uint32_t)ms.total;
            out_moves[i] = (ms.total > 0) ? select_move(cx, p, ms, ks[i] % (uint32_t)ms.total) : (tafl_move)0xFFFF;
            });
        }
        return 0;
    });
}

}  // extern "C"
