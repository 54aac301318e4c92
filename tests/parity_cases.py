"""Differential checks shared by the CPU (hostsim vs oracle) and GPU (CUDA vs oracle) parity tests."""
import numpy as np

from alphazeroforhnefatafl_b200 import _abi as A

PRESET_CONFIGS = [  # (name, preset id, width)
    ("copenhagen", A.PRESET_COPENHAGEN, 128), ("brandubh", A.PRESET_BRANDUBH, 64), ("tablut", A.PRESET_TABLUT, 128),
    ("magpie", A.PRESET_MAGPIE, 64), ("koch", A.PRESET_KOCH, 64), ("brandubh_u128", A.PRESET_BRANDUBH, 128),
]


def board_for(orc, preset):
    return orc.preset_board(A.PRESET_BRANDUBH if preset == A.PRESET_KOCH else preset)


def sample_positions(orc, rules, fen, width, n_games, seed, max_steps=160, auto_reset=False):
    """States visited by seeded oracle random play: game g is advanced (g * 7919) % max_steps plies."""
    rc, st0 = orc.state_from_fen(fen, width, A.DEFENDER if rules.starting_side else A.ATTACKER)
    assert rc == 0
    out = np.repeat(st0, n_games)
    steps = (np.arange(n_games) * 7919) % max_steps
    for s in np.unique(steps):
        idx = np.nonzero(steps == s)[0]
        if s == 0:
            continue
        # advance each selected game separately with its own game id so streams differ
        for g in idx:
            sub, _ = orc.random_playout(rules, fen, width, seed, 1, int(s), auto_reset=auto_reset, game_id_base=int(g))
            out[g] = sub[0]
    return out


def compare_by_value(ref, test, states, width, rng):
    """ref/test: ByValueImpl-like objects.  Compares every by-value entry point on `states`."""
    n = len(states)
    checks = 0
    # legal moves: set AND order
    mr, cr = ref.legal_moves_of(states)
    mt, ct = test.legal_moves_of(states)
    assert np.array_equal(cr, ct)
    for i in range(n):
        assert np.array_equal(mr[i, :cr[i]], mt[i, :ct[i]]), i
    checks += n
    # one random legal play per ongoing position + a few illegal ones
    plays = np.zeros(n, A.PLAY_DTYPE)
    stride = A.stride_for(width)
    for i in range(n):
        if cr[i] and rng.random() < 0.85:
            m = int(mr[i, rng.integers(cr[i])])
            f, t = m & 127, (m >> 7) & 127
            fr, fc, tr, tc = f // stride, f % stride, t // stride, t % stride
        else:  # arbitrary (mostly invalid) play
            fr, fc = rng.integers(0, 12, 2)
            if rng.random() < 0.5:
                tr, tc = fr, rng.integers(0, 12)
            else:
                tr, tc = rng.integers(0, 12), fc
        plays[i]["from_row"], plays[i]["from_col"] = fr, fc
        if fr == tr:
            plays[i]["axis"], plays[i]["disp"] = A.HORIZONTAL, tc - fc
        else:
            plays[i]["axis"], plays[i]["disp"] = A.VERTICAL, tr - fr
    assert np.array_equal(ref.validate_plays(states, plays), test.validate_plays(states, plays))
    sr, fr_ = ref.do_plays(states, plays)
    st_, ft = test.do_plays(states, plays)
    assert sr.tobytes() == st_.tobytes()
    assert fr_.tobytes() == ft.tobytes()
    checks += 2 * n
    ok = fr_["err"] == 0
    if ok.any():
        a, b = ref.get_captures(states[ok], plays[ok]), test.get_captures(states[ok], plays[ok])
        assert a.tobytes() == b.tobytes()
        a, b = ref.detect_shieldwall(sr[ok], plays[ok]), test.detect_shieldwall(sr[ok], plays[ok])
        assert a.tobytes() == b.tobytes()
    # queries on the successor positions
    for side in (0, 1):
        sides = np.full(n, side, np.uint8)
        assert np.array_equal(ref.side_can_play(sr, sides), test.side_can_play(sr, sides))
    assert np.array_equal(ref.detect_exit_fort(sr), test.detect_exit_fort(sr))
    # enclosure around the king nibble tile, attackers enclosing (the enclosure-win query) and
    # around random tiles with random piece sets / abort flags
    q = np.zeros(n, A.ENCL_QUERY_DTYPE)
    hi = sr["def"][:, 1] if width == 128 else sr["def"][:, 0]
    ha = sr["att"][:, 1] if width == 128 else sr["att"][:, 0]
    q["row"], q["col"] = (hi >> np.uint64(60)).astype(np.uint8), (ha >> np.uint64(60)).astype(np.uint8)
    q["enclosed_pieces"], q["enclosing_pieces"] = A.pieceset_side(A.DEFENDER), A.pieceset_side(A.ATTACKER)
    q["abort_on_edge"], q["abort_on_corner"] = 1, 1
    q["outside_safe"] = 1
    a, b = ref.find_enclosure(sr, q), test.find_enclosure(sr, q)
    assert a.tobytes() == b.tobytes()
    assert np.array_equal(ref.enclosure_secure(sr, q), test.enclosure_secure(sr, q))
    sets = [A.pieceset_type(A.KING), A.pieceset_type(A.SOLDIER), A.pieceset_side(A.DEFENDER),
            A.pieceset_side(A.ATTACKER), A.pieceset_piece(A.SOLDIER, A.ATTACKER),
            A.pieceset_piece(A.SOLDIER, A.DEFENDER), 0, 0xFFFF]
    N = int(states["side_len"][0])
    q["row"], q["col"] = rng.integers(0, N, n), rng.integers(0, N, n)
    q["enclosed_pieces"] = rng.choice(sets, n)
    q["enclosing_pieces"] = rng.choice(sets, n)
    q["abort_on_edge"], q["abort_on_corner"] = rng.integers(0, 2, n), rng.integers(0, 2, n)
    q["inside_safe"], q["outside_safe"] = rng.integers(0, 2, n), rng.integers(0, 2, n)
    a, b = ref.find_enclosure(sr, q), test.find_enclosure(sr, q)
    assert a.tobytes() == b.tobytes()
    assert np.array_equal(ref.enclosure_secure(sr, q), test.enclosure_secure(sr, q))
    # plays of single tiles (iter_plays), including empty tiles
    tiles = np.stack([rng.integers(0, N, n), rng.integers(0, N, n)], 1).astype(np.uint8)
    (ma, ca), (mb, cb) = ref.piece_plays(sr, tiles), test.piece_plays(sr, tiles)
    assert np.array_equal(ca, cb)
    for i in range(n):
        k = max(int(ca[i]), 0)
        assert np.array_equal(ma[i, :k], mb[i, :k])
    checks += 8 * n
    return checks


def random_boards(orc, rng, N, width, n, density=None, edge_bias=True):
    """Synthetic positions (not reachable play): random soldiers + one king, biased towards the
    edges so shield walls, exit forts and enclosures actually occur."""
    out = np.zeros(n, A.STATE_DTYPE)
    for i in range(n):
        d = density if density is not None else rng.uniform(0.1, 0.6)
        grid = [["." for _ in range(N)] for _ in range(N)]
        for r in range(N):
            for c in range(N):
                p = d * (1.6 if edge_bias and (r in (0, 1, N - 2, N - 1) or c in (0, 1, N - 2, N - 1)) else 0.7)
                if rng.random() < p:
                    grid[r][c] = "t" if rng.random() < 0.55 else "T"
        if rng.random() < 0.95:
            if rng.random() < 0.5:  # king near an edge
                r, c = rng.integers(0, N), rng.choice([0, 1, N - 2, N - 1])
                if rng.random() < 0.5:
                    r, c = c, r
            else:
                r, c = rng.integers(0, N, 2)
            grid[r][c] = "K"
        rows = []
        for r in range(N):
            s, run = "", 0
            for ch in grid[r]:
                if ch == ".":
                    run += 1
                else:
                    if run:
                        s += str(run)
                        run = 0
                    s += ch
            if run:
                s += str(run)
            rows.append(s)
        side = A.DEFENDER if rng.random() < 0.5 else A.ATTACKER
        rc, st = orc.state_from_fen("/".join(rows), width, side)
        assert rc == 0
        out[i] = st[0]
    return out


def wide_random_boards(orc, rng, N, width, n):
    """random NxN positions on a wide bitfield: both sides scattered, a king somewhere, either side to move"""
    out = np.zeros(n, A.STATE_W_DTYPE)
    for i in range(n):
        d = rng.uniform(0.08, 0.45)
        rows = []
        kr, kc = rng.integers(0, N, 2)
        for r in range(N):
            s, run = "", 0
            for c in range(N):
                ch = "."
                if (r, c) == (kr, kc) and rng.random() < 0.95:
                    ch = "K"
                elif rng.random() < d * (1.5 if r in (0, 1, N - 2, N - 1) or c in (0, 1, N - 2, N - 1) else 0.8):
                    ch = "t" if rng.random() < 0.55 else "T"
                if ch == ".":
                    run += 1
                else:
                    s += (str(run) if run else "") + ch
                    run = 0
            rows.append(s + (str(run) if run else ""))
        rc, st = orc.state_from_fen("/".join(rows), width, A.DEFENDER if rng.random() < 0.5 else A.ATTACKER)
        assert rc == 0
        out[i] = st[0]
    return out


def compare_wide(ref, test, states, N, rng):
    """every wide by-value entry point of `test` against `ref` on `states`"""
    n = len(states)
    pr, cr = ref.legal_plays_of(states)
    pt, ct = test.legal_plays_of(states)
    assert np.array_equal(cr, ct)
    for i in range(n):
        assert pr[i, :cr[i]].tobytes() == pt[i, :ct[i]].tobytes(), i
    plays = np.zeros(n, A.PLAY_DTYPE)
    for i in range(n):
        if cr[i] and rng.random() < 0.85:
            plays[i] = pr[i, rng.integers(cr[i])]
        else:
            fr, fc = rng.integers(0, N + 1, 2)
            hz = rng.random() < 0.5
            plays[i] = (fr, fc, A.HORIZONTAL if hz else A.VERTICAL, int(rng.integers(-N, N + 1)))
    assert np.array_equal(ref.validate_plays(states, plays), test.validate_plays(states, plays))
    sr, fr_ = ref.do_plays(states, plays)
    st_, ft = test.do_plays(states, plays)
    assert fr_.tobytes() == ft.tobytes()
    assert sr.tobytes() == st_.tobytes()
    legal = fr_["err"] == 0
    if legal.any():
        assert ref.get_captures(states[legal], plays[legal]).tobytes() == test.get_captures(states[legal], plays[legal]).tobytes()
    assert ref.detect_shieldwall(states, plays).tobytes() == test.detect_shieldwall(states, plays).tobytes()
    tiles = rng.integers(0, N, (n, 2)).astype(np.uint8)
    a, ca = ref.piece_plays(states, tiles)
    b, cb = test.piece_plays(states, tiles)
    assert np.array_equal(ca, cb)
    for i in range(n):
        if ca[i] > 0:
            assert a[i, :ca[i]].tobytes() == b[i, :cb[i]].tobytes(), i
    q = np.zeros(n, A.ENCL_QUERY_DTYPE)
    q["row"], q["col"] = tiles[:, 0], tiles[:, 1]
    q["abort_on_edge"], q["abort_on_corner"] = rng.integers(0, 2, n), rng.integers(0, 2, n)
    sets = [0xFFFF, A.pieceset_side(A.DEFENDER), A.pieceset_side(A.ATTACKER), A.pieceset_type(A.KING), A.pieceset_type(A.SOLDIER), 0]
    q["enclosed_pieces"], q["enclosing_pieces"] = rng.choice(sets, n), rng.choice(sets, n)
    q["inside_safe"], q["outside_safe"] = rng.integers(0, 2, n), rng.integers(0, 2, n)
    assert ref.find_enclosure(states, q).tobytes() == test.find_enclosure(states, q).tobytes()
    assert np.array_equal(ref.enclosure_secure(states, q), test.enclosure_secure(states, q))
    assert np.array_equal(ref.detect_exit_fort(states), test.detect_exit_fort(states))
    sides = rng.integers(0, 2, n).astype(np.uint8)
    assert np.array_equal(ref.side_can_play(states, sides), test.side_can_play(states, sides))
    return sr[legal]
