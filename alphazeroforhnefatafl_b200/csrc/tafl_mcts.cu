// tafl_mcts.cu — placeholder until the MCTS kernels land (next commit).
#include "tafl_engine.hpp"
namespace tafl { void mcts_free(tafl_engine*) {} }
extern "C" {
int tafl_mcts_configure(tafl_engine*, const tafl_mcts_config*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_search(tafl_engine*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_search_async(tafl_engine*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_results_device(tafl_engine*, tafl_move**, uint32_t**, double**, uint32_t**, uint32_t**) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_results_host(tafl_engine*, uint32_t, uint32_t, tafl_move*, uint32_t*, double*, uint32_t*, uint32_t*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_play_best(tafl_engine*, uint8_t*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_mcts_stats_get(tafl_engine*, tafl_mcts_stats*) { return TAFL_ERR_UNSUPPORTED; }
int tafl_pack_records(tafl_engine*, void*) { return TAFL_ERR_UNSUPPORTED; }
}
