"""ctypes loader for ``libtafl_b200.so`` (the nvcc-built sm_100a library behind ``include/tafl.h``).

There is no CPU fallback: if the shared library is missing this module raises, and if no CUDA
device is usable ``tafl_engine_create`` fails with ``TAFL_ERR_CUDA``.
"""
import ctypes as C
import os

from . import _abi as A

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TAFL_LIB") or os.path.join(HERE, "libtafl_b200.so")   # TAFL_LIB: an A/B build variant (build.py --suffix)

_lib = None


class TaflError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"tafl error {code}: {msg}" if msg else f"tafl error {code}")
        self.code = code


def lib():
    """Load (once) and return the C-ABI library.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m alphazeroforhnefatafl_b200.build` "
            "(nvcc, sm_100a).  This package has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, u8p = C.c_void_p, C.POINTER(C.c_uint8)
    sig = {
        "tafl_version": (C.c_char_p, []),
        "tafl_preset_rules": (C.c_int, [C.c_int, C.POINTER(A.Rules)]),
        "tafl_preset_board_fen": (C.c_int, [C.c_int, C.POINTER(C.c_char_p)]),
        "tafl_state_from_fen": (C.c_int, [C.c_char_p, C.c_int, C.c_int, vp]),
        "tafl_state_to_fen": (C.c_int, [vp, C.c_int, C.c_char_p, C.c_size_t]),
        "tafl_state_from_display_str": (C.c_int, [C.c_char_p, C.c_int, C.c_int, vp]),
        "tafl_state_to_display_str": (C.c_int, [vp, C.c_int, C.c_char_p, C.c_size_t]),
        "tafl_play_from_tiles": (C.c_int, [C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint8, C.POINTER(A.Play)]),
        "tafl_play_from_str": (C.c_int, [C.c_char_p, C.POINTER(A.Play)]),
        "tafl_play_to_str": (C.c_int, [C.POINTER(A.Play), C.c_char_p, C.c_size_t]),
        "tafl_move_to_play": (C.c_int, [C.c_uint16, C.c_int, C.POINTER(A.Play)]),
        "tafl_board_get_king": (C.c_int, [vp, C.c_int, u8p, u8p]),
        "tafl_board_set_king": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8]),
        "tafl_board_is_king": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8]),
        "tafl_board_set_piece": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8, C.c_int, C.c_int]),
        "tafl_board_clear_tile": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8]),
        "tafl_board_get_piece": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "tafl_board_tile_occupied": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8]),
        "tafl_board_count_pieces": (C.c_int, [vp, C.c_int, C.c_int]),
        "tafl_board_iter_occupied": (C.c_int, [vp, C.c_int, C.c_int, vp]),
        "tafl_board_move_piece": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint8]),
        "tafl_board_swap_pieces": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint8]),
        "tafl_tile_from_str": (C.c_int, [C.c_char_p, u8p, u8p]),
        "tafl_tile_to_str": (C.c_int, [C.c_uint8, C.c_uint8, C.c_char_p, C.c_size_t]),
        "tafl_track_play": (C.c_int, [vp, C.c_int, C.POINTER(A.Play), C.c_int]),
        "tafl_board_to_matrix": (C.c_int, [vp, C.c_int, vp, C.c_size_t]),
        "tafl_engine_create": (C.c_int, [C.POINTER(A.Rules), C.c_char_p, C.c_int, C.c_uint32, C.c_int, C.POINTER(vp)]),
        "tafl_engine_destroy": (C.c_int, [vp]),
        "tafl_last_error": (C.c_char_p, [vp]),
        "tafl_engine_info": (C.c_int, [vp, C.POINTER(C.c_uint32), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "tafl_engine_stream": (vp, [vp]),
        "tafl_engine_sync": (C.c_int, [vp]),
        "tafl_launch_count": (C.c_uint64, [vp]),
        "tafl_engine_spec": (C.c_int, [vp]),
        "tafl_engine_force_generic": (C.c_int, [vp, C.c_int]),
        "tafl_reset": (C.c_int, [vp, vp, C.c_uint32]),
        "tafl_set_states": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32]),
        "tafl_get_states": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32]),
        "tafl_states_device": (vp, [vp]),
        "tafl_legal_moves": (C.c_int, [vp, vp, vp]),
        "tafl_legal_moves_async": (C.c_int, [vp, vp, vp]),
        "tafl_moves_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp)]),
        "tafl_legal_moves_host": (C.c_int, [vp, vp, vp, C.c_uint32, C.c_uint32]),
        "tafl_step": (C.c_int, [vp, vp, vp]),
        "tafl_step_status": (C.c_int, [vp, vp, vp, vp]),
        "tafl_step_status_restart": (C.c_int, [vp, vp, vp, vp]),
        "tafl_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
        "tafl_host_free": (C.c_int, [vp]),
        "tafl_host_register": (C.c_int, [vp, C.c_size_t]),
        "tafl_host_unregister": (C.c_int, [vp]),
        "tafl_actions_device": (vp, [vp]),
        "tafl_actions_host": (C.c_int, [vp, vp, C.c_uint32, C.c_uint32]),
        "tafl_step_moves_device": (C.c_int, [vp, vp, vp]),
        "tafl_step_random": (C.c_int, [vp, C.c_uint64, C.c_uint32, C.c_int, C.c_int]),
        "tafl_step_random_async": (C.c_int, [vp, C.c_uint64, C.c_uint32, C.c_int, C.c_int]),
        "tafl_step_random_host": (C.c_int, [vp, vp, vp, C.c_uint64, C.c_uint32, C.c_int]),
        "tafl_set_game_id_base": (C.c_int, [vp, C.c_uint64]),
        "tafl_set_random_ply": (C.c_int, [vp, C.c_uint64]),
        "tafl_random_ply": (C.c_uint64, [vp]),
        "tafl_stats_reset": (C.c_int, [vp]),
        "tafl_stats_get": (C.c_int, [vp, C.POINTER(A.Stats)]),
        "tafl_validate_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_do_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp, vp]),
        "tafl_legal_moves_of": (C.c_int, [vp, vp, C.c_uint32, vp, vp]),
        "tafl_legal_moves_of_static": (C.c_int, [vp, vp, C.c_uint32, vp, vp]),
        "tafl_static_max_games": (C.c_uint32, []),
        "tafl_piece_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp, vp]),
        "tafl_get_captures": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_detect_shieldwall": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_find_enclosure": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_enclosure_secure": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_detect_exit_fort": (C.c_int, [vp, vp, C.c_uint32, vp]),
        "tafl_side_can_play": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_history_enable": (C.c_int, [vp, C.c_uint32]),
        "tafl_undo_last_play": (C.c_int, [vp, vp]),
        "tafl_history_len": (C.c_int, [vp, vp]),
        "tafl_play_history": (C.c_int, [vp, C.c_uint32, vp, C.c_uint32, C.POINTER(C.c_uint32)]),
        "tafl_nccl_unique_id": (C.c_int, [vp]),
        "tafl_nccl_comm_create": (C.c_int, [vp, vp, C.c_int, C.c_int, C.POINTER(vp)]),
        "tafl_nccl_comm_destroy": (C.c_int, [vp]),
        "tafl_gather_records": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp, C.c_uint32, vp, C.c_uint64, vp]),
        "tafl_gather_records_on": (C.c_int, [vp, vp, C.c_int, C.c_int, C.c_int, vp, C.c_uint32, vp, C.c_uint64, vp, vp]),
        "tafl_mcts_configure": (C.c_int, [vp, C.POINTER(A.MctsConfig)]),
        "tafl_mcts_search": (C.c_int, [vp]),
        "tafl_mcts_search_async": (C.c_int, [vp]),
        "tafl_mcts_results_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                               C.POINTER(vp)]),
        "tafl_mcts_results_host": (C.c_int, [vp, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp]),
        "tafl_mcts_play_best": (C.c_int, [vp, vp]),
        "tafl_mcts_advance": (C.c_int, [vp, vp]),
        "tafl_mcts_reroot": (C.c_int, [vp]),
        "tafl_mcts_root_noise": (C.c_int, [vp, vp, C.c_double]),
        "tafl_mcts_stats_get": (C.c_int, [vp, C.POINTER(A.MctsStats)]),
        "tafl_mcts_policy": (C.c_int, [vp, C.c_double, C.c_int, C.c_uint64, vp]),
        "tafl_mcts_begin": (C.c_int, [vp]),
        "tafl_mcts_select_leaves": (C.c_int, [vp, C.POINTER(A.LeafBatch)]),
        "tafl_mcts_expand_backup": (C.c_int, [vp]),
        "tafl_mcts_finish": (C.c_int, [vp]),
        "tafl_mcts_search_with": (C.c_int, [vp, A.EVALUATOR_FN, vp]),
        "tafl_eval_builtin": (C.c_int, [vp, C.c_int, C.c_uint64]),
        "tafl_mcts_search_builtin": (C.c_int, [vp, C.c_int, C.c_uint64]),
        "tafl_leaf_batch_host": (C.c_int, [vp, vp, vp, vp, vp]),
        "tafl_leaf_eval_host": (C.c_int, [vp, vp, vp]),
        "tafl_w_state_from_fen": (C.c_int, [C.c_char_p, C.c_int, C.c_int, vp]),
        "tafl_w_state_to_fen": (C.c_int, [vp, C.c_int, C.c_char_p, C.c_size_t]),
        "tafl_w_board_move_piece": (C.c_int, [vp, C.c_int, C.c_uint8, C.c_uint8, C.c_uint8, C.c_uint8]),
        "tafl_wide_create": (C.c_int, [C.POINTER(A.Rules), C.c_int, C.c_int, C.c_int, C.POINTER(vp)]),
        "tafl_wide_destroy": (C.c_int, [vp]),
        "tafl_wide_last_error": (C.c_char_p, [vp]),
        "tafl_w_validate_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_w_do_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp, vp]),
        "tafl_w_legal_plays_of": (C.c_int, [vp, vp, C.c_uint32, vp, vp]),
        "tafl_w_piece_plays": (C.c_int, [vp, vp, vp, C.c_uint32, vp, vp]),
        "tafl_w_get_captures": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_w_detect_shieldwall": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_w_find_enclosure": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_w_enclosure_secure": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_w_detect_exit_fort": (C.c_int, [vp, vp, C.c_uint32, vp]),
        "tafl_w_side_can_play": (C.c_int, [vp, vp, vp, C.c_uint32, vp]),
        "tafl_pack_records": (C.c_int, [vp, vp]),
        "tafl_pack_records_compact": (C.c_int, [vp, vp, vp, vp]),
        "tafl_reset_finished": (C.c_int, [vp, C.POINTER(C.c_uint32)]),
        "tafl_mcts_overflow": (C.c_int, [vp, C.POINTER(C.c_uint32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    L._tafl_signatures = sig
    _lib = L
    return L


def declared_symbols():
    """Every entry point ``include/tafl.h`` declares (used by the ABI test)."""
    import re
    hdr = os.path.join(os.path.dirname(HERE), "include", "tafl.h")
    with open(hdr) as f:
        text = f.read()
    return sorted(set(re.findall(r"\b(tafl_[a-z0-9_]+)\s*\(", text)))
