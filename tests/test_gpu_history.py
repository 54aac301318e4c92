"""Game<T> history / undo (game/mod.rs:77-108) on the device: the reference's own `test_undo`
(game/mod.rs:209-231) replayed through the C ABI, plus a differential run against a host model of
`play_history` / `state_history` with invalid plays, masked undos and a history ring that overflows."""
import numpy as np
import pytest

from alphazeroforhnefatafl_b200 import _abi as A
from parity_cases import board_for

pytestmark = pytest.mark.gpu


def _play(fr, fc, tr, tc):
    hz = fr == tr
    return (fr, fc, A.HORIZONTAL if hz else A.VERTICAL, (tc - fc) if hz else (tr - fr))


def test_reference_test_undo(orc, built_lib):
    from alphazeroforhnefatafl_b200 import engine as E
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), board_for(orc, A.PRESET_BRANDUBH)
    with E.LockstepGames(rules, fen, 1, 0, 64) as g:
        g.history_enable(16)
        states = [g.get_states().tobytes()]
        for mv in (_play(0, 3, 0, 2), _play(2, 3, 2, 1), _play(1, 3, 1, 1)):      # mod.rs:213-220
            fx = g.step(E.plays_array([mv]))
            assert fx["err"][0] == 0
            states.append(g.get_states().tobytes())
            assert states[-1] != states[0]
        assert list(g.history_len()) == [3]
        for want in (states[2], states[1], states[0], states[0]):                  # mod.rs:221-228
            g.undo_last_play()
            assert g.get_states().tobytes() == want
        assert list(g.history_len()) == [0]


@pytest.mark.parametrize("preset,width,cap", [(A.PRESET_COPENHAGEN, 128, 64), (A.PRESET_BRANDUBH, 64, 5)])
def test_history_differential(orc, built_lib, preset, width, cap):
    from alphazeroforhnefatafl_b200 import engine as E
    rules, fen = orc.preset_rules(preset), board_for(orc, preset)
    n, plies = 200, 40
    rng = np.random.default_rng(9)
    with E.LockstepGames(rules, fen, n, 0, width) as g:
        g.history_enable(cap)
        stacks = [[] for _ in range(n)]          # host model: (state bytes, play, side, effects bytes)
        floor = [0] * n                          # oldest entry the device ring still holds
        for ply in range(plies):
            cur = g.get_states()
            moves, counts = g.legal_moves()
            pick = (rng.integers(0, 1 << 30, n) % np.maximum(counts, 1)).astype(np.int64)
            plays = E.moves_to_plays(moves[np.arange(n), pick], width)
            bad = rng.random(n) < 0.15           # an invalid play: nothing may be pushed for that game
            plays["disp"][bad] = 0
            plays["from_row"][bad] = 0
            plays["from_col"][bad] = 0
            fx = g.step(plays)
            for i in range(n):
                if fx["err"][i] == 0:
                    side = A.DEFENDER if cur["flags"][i] & 4 else A.ATTACKER
                    stacks[i].append((cur[i:i + 1].tobytes(), plays[i:i + 1].tobytes(), side, fx[i:i + 1].tobytes()))
                    floor[i] = max(floor[i], len(stacks[i]) - cap)
            assert (fx["err"][bad] != 0).all()
            if ply % 7 == 6:                       # masked undo
                mask = (rng.random(n) < 0.5).astype(np.uint8)
                before = g.get_states()
                g.undo_last_play(mask)
                after = g.get_states()
                for i in range(n):
                    if mask[i] and len(stacks[i]) > floor[i]:
                        assert after[i:i + 1].tobytes() == stacks[i][-1][0], i
                        stacks[i].pop()
                    else:
                        assert after[i:i + 1].tobytes() == before[i:i + 1].tobytes(), i
            assert np.array_equal(g.history_len(), np.array([len(s) for s in stacks], np.uint32))
        for i in range(0, n, 17):                 # play_history: the records still held, oldest first
            recs = g.play_history(i)
            exp = stacks[i][floor[i]:]
            assert len(recs) == len(exp)
            for r, (_, pl, side, fxb) in zip(recs, exp):
                assert r["play"].tobytes() == pl and int(r["side"]) == side and r["effects"].tobytes() == fxb
        # undo everything that is still held, newest first
        for _ in range(cap + 2):
            g.undo_last_play()
        final = g.get_states()
        assert np.array_equal(g.history_len(), np.array(floor, np.uint32))
        for i in range(n):
            if len(stacks[i]) > floor[i]:
                assert final[i:i + 1].tobytes() == stacks[i][floor[i]][0]
        # set_states / reset forget the history of the games they touch
        g.reset()
        assert not g.history_len().any()


def test_history_is_forgotten_by_mutators_that_do_not_record(orc, built_lib):
    """only Game::do_play records (state, play).  The batch mutators without a reference counterpart
    (random stepping, the MCTS move) forget the history of the games they advance, so undo_last_play never
    jumps back over unrecorded plies and play_history never has holes."""
    from alphazeroforhnefatafl_b200 import engine as E
    rules, fen = orc.preset_rules(A.PRESET_BRANDUBH), orc.preset_board(A.PRESET_BRANDUBH)
    n = 16
    with E.LockstepGames(rules, fen, n, 0, 64) as g:
        g.history_enable(8)
        mv, cnt = g.legal_moves()
        g.game_do_play(E.moves_to_plays(mv[:, 0], 64))
        assert (g.history_len() == 1).all()
        g.step_random(3, 2, auto_reset=False)
        assert (g.history_len() == 0).all()
        before = g.get_states()
        g.undo_last_play()                                   # nothing recorded: no-op, as after set_states
        assert g.get_states().tobytes() == before.tobytes()
        mv, cnt = g.legal_moves()
        g.game_do_play(E.moves_to_plays(mv[:, 0], 64))
        assert (g.history_len() == 1).all()
        g.mcts_configure(30)
        g.mcts_search()
        g.mcts_play_best(False)
        assert (g.history_len() == 0).all()
