"""tafl_legal_moves_of_static (no device allocation: staging in the library's file-scope __device__ array) against
tafl_legal_moves_of, through the C ABI, on two calls in a row.  Bit-exact: this is integer work.

Collected last on purpose; passes alone in a fresh process and after the other tests.
The entry point has no counterpart in the reference (include/tafl.h says why it exists)."""
import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A
from alphazeroforhnefatafl_b200 import engine as E
from alphazeroforhnefatafl_b200._lib import lib
from alphazeroforhnefatafl_b200.engine import TaflError

pytestmark = pytest.mark.gpu


def positions(rules, fen, width, n, seed, plies):
    """n distinct mid-game positions: seeded random play on the GPU (finished games restart, so all are live or just over)"""
    with E.LockstepGames(rules, fen, n, 0, width) as g:
        g.step_random(seed, plies, auto_reset=True, steps_per_launch=1)
        return g.get_states().copy()


def same_as_existing(g, states):
    m_ref, c_ref = g.legal_moves_of(states)
    m_new, c_new = g.legal_moves_of_static(states)
    assert np.array_equal(c_ref, c_new)
    assert int(c_ref.max()) > 0
    live = np.arange(A.MAX_MOVES)[None, :] < c_ref[:, None]
    assert np.array_equal(m_ref[live], m_new[live])          # same plays in the same order
    assert not m_new[~live].any()                            # rows past the count are cleared, whatever ran before
    return int(c_ref.sum())


@pytest.mark.parametrize("preset,width", [(A.PRESET_COPENHAGEN, 128), (A.PRESET_BRANDUBH, 64)])
def test_gpu_legal_moves_static_scratch_matches_allocating_path(built_lib, preset, width):
    rules, fen = E.preset_rules(preset), E.preset_board(preset)
    cap = lib().tafl_static_max_games()
    assert cap == 2048
    first = positions(rules, fen, width, cap, 9001, 23)      # the largest batch the array holds
    second = positions(rules, fen, width, 777, 9002, 41)     # then a smaller one over the same (now dirty) array
    with E.LockstepGames(rules, fen, 8, 0, width) as g:
        t1 = same_as_existing(g, first)
        t2 = same_as_existing(g, second)
        assert t1 > t2 > 0
        # a second engine on the same device shares the array and the cached symbol address
        with E.LockstepGames(rules, fen, 4, 0, width) as g2:
            same_as_existing(g2, second[:100])
        m0, c0 = g.legal_moves_of_static(first[:0])          # empty batch: nothing to do, no error
        assert m0.shape == (0, A.MAX_MOVES) and c0.shape == (0,)


def test_gpu_legal_moves_static_scratch_rejects_oversize(built_lib):
    rules, fen = E.preset_rules(A.PRESET_COPENHAGEN), E.preset_board(A.PRESET_COPENHAGEN)
    cap = lib().tafl_static_max_games()
    with E.LockstepGames(rules, fen, 8, 0, 128) as g:
        st = np.repeat(g.get_states()[:1], cap + 1)
        with pytest.raises(TaflError) as ei:
            g.legal_moves_of_static(st)
        assert ei.value.code == A.ERR_CAPACITY
        assert "TAFL_STATIC_MAX_GAMES" in str(ei.value)
        m, c = g.legal_moves_of_static(st[:cap])             # the limit itself is fine, and the engine still works
        assert (c == c[0]).all() and c[0] > 0
