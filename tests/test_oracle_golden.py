"""The oracle pinned against the reference's own known-answer vectors and against fixtures
produced by running the unmodified reference (tests/golden/make_golden.py).  CPU only."""
import glob
import os

import numpy as np
import pytest

import oracle


# ---- pyastar2d/tests/test_astar.py known answers (4-connected fork) ---------------------------
def test_kat_small_4connected():
    # pyastar2d/tests/test_astar.py:10-14,22-26
    weights = np.array([[1, 3, 3, 3, 3],
                        [2, 1, 3, 3, 3],
                        [2, 2, 1, 3, 3],
                        [2, 2, 2, 1, 3],
                        [2, 2, 2, 2, 1]], dtype=np.float32)
    path = oracle.astar_weights(weights, (0, 0), (4, 4))
    expected = np.array([[0, 0], [1, 0], [1, 1], [2, 1], [2, 2], [3, 2], [3, 3], [4, 3], [4, 4]])
    assert np.array_equal(path, expected)


def test_kat_no_solution():
    # pyastar2d/tests/test_astar.py:29-42
    for axis in (0, 1):
        weights = np.ones((5, 5), dtype=np.float32)
        if axis == 0:
            weights[:, 2] = np.inf
        else:
            weights[2, :] = np.inf
        assert oracle.astar_weights(weights, (0, 0), (4, 4)) is None


def test_kat_narrow():
    # pyastar2d/tests/test_astar.py:63-78
    p = oracle.astar_weights(np.ones((5, 1), np.float32), (0, 0), (4, 0))
    assert np.array_equal(p, [[0, 0], [1, 0], [2, 0], [3, 0], [4, 0]])
    p = oracle.astar_weights(np.ones((1, 5), np.float32), (0, 0), (0, 4))
    assert np.array_equal(p, [[0, 0], [0, 1], [0, 2], [0, 3], [0, 4]])


def test_kat_errors():
    # pyastar2d/tests/test_astar.py:81-127
    rng = np.random.default_rng(0)
    w = (1. + 5. * rng.random((10, 10))).astype(np.float32)
    bad = w.copy()
    bad[4, 4] = 0.25
    with pytest.raises(ValueError):
        oracle.astar_weights(bad, (0, 0), (9, 9))
    for s in [(-1, 0), (10, 0), (0, -1), (0, 10)]:
        with pytest.raises(ValueError, match='Start'):
            oracle.astar_weights(w, s, (9, 9))
        with pytest.raises(ValueError, match='Goal'):
            oracle.astar_weights(w, (0, 0), s)
    with pytest.raises(AssertionError, match='float64'):
        oracle.astar_weights(np.ones((3, 3)), (0, 0), (2, 2))


def test_match_reverse_cost():
    # pyastar2d/tests/test_astar.py:45-60 (path cost is direction independent)
    rng = np.random.default_rng(1)
    w = (1. + 5. * rng.random((25, 25))).astype(np.float32)
    f, cf = oracle.astar_weights(w, (0, 0), (24, 24), return_cost=True)
    r, cr = oracle.astar_weights(w, (24, 24), (0, 0), return_cost=True)
    assert f is not None and r is not None
    assert abs((cf - w[24, 24]) - (cr - w[0, 0])) < 1e-3


def test_max_path_length_cutoff():
    # 1x20 row, start 0, goal 19: max=19 -> None, max=20 -> L=20
    w = np.ones((1, 20), np.float32)
    assert oracle.astar_weights(w, (0, 0), (0, 19), 19) is None
    assert oracle.astar_weights(w, (0, 0), (0, 19), 20).shape[0] == 20


# ---- compiled reference (oracle/_ref) vs restatement ------------------------------------------------
@pytest.mark.skipif(not oracle.ref_available(), reason='oracle/_ref not built')
def test_restatement_equals_compiled_reference():
    rng = np.random.default_rng(5)
    n = 0
    for it in range(120):
        H, W = int(rng.integers(20, 120)), int(rng.integers(20, 120))
        m = rng.random((H, W)) < rng.uniform(0.3, 0.9)
        if it % 3 == 0:
            m[:] = False
            m[H // 2 - 1:H // 2 + 2, :] = True
            m[:, W // 2 - 1:W // 2 + 2] = True
            m[rng.integers(0, H, 10), rng.integers(0, W, 10)] ^= True
        max_px = int(rng.choice([8, 20, 45, 90]))
        src = np.stack([rng.integers(0, H, 4), rng.integers(0, W, 4)], 1)
        dst = np.stack([rng.integers(0, H, 6), rng.integers(0, W, 6)], 1)
        a = oracle.cost_matrix_astar(m, src, dst, max_px)
        pl, un = oracle.ref_cost_matrix(m, src, dst, max_px, threads=False)
        assert np.array_equal(a[0], pl) and np.array_equal(a[1], un)
        n += a[0].size
    assert n > 2000


@pytest.mark.skipif(not oracle.ref_available(), reason='oracle/_ref not built')
def test_compiled_reference_kat():
    w = np.ones((1, 20), np.float32)
    assert oracle.ref_astar_path(w, (0, 0), (0, 19), 19) is None
    assert oracle.ref_astar_path(w, (0, 0), (0, 19), 20).shape == (20, 2)


@pytest.mark.skipif(not oracle.ref_available(), reason='oracle/_ref not built')
@pytest.mark.parametrize('H,n_dets,n_min', [(2048, 250, 300), (4096, 1000, 100)])
def test_rule_equals_compiled_reference_at_baseline_shapes(H, n_dets, n_min):
    """The closed form pinned to the COMPILED reference at BASELINE's real sizes and max_px = 500
    (VERDICT r1 next #1b): >= 300 in-range pairs of the configs[1] workload (2048^2, ~200 dets) and
    >= 100 of configs[2] (4096^2, ~800 dets); 14 / 75 ms per reference call."""
    from axtrack_b200 import synth
    mask = synth.make_mask(H, H, 0)
    a, b = synth.make_tracks(mask, n_dets, 2, 0)
    d2 = ((a[:, None, :].astype(np.int64) - b[None, :, :]) ** 2).sum(-1)
    inr = (d2 < 500 * 500).sum(1)
    ns = int(np.searchsorted(np.cumsum(inr), n_min) + 1)
    pl, un = oracle.ref_cost_matrix(mask, a[:ns], b, 500, threads=False)
    r = oracle.cost_matrix_rule_ex(mask, a[:ns], b, 500)
    assert int(inr[:ns].sum()) >= n_min and int((un == 0).sum()) >= n_min // 4
    assert r[3].sum() == 0                                  # no pair of these workloads is in the heap-tie regime
    assert np.array_equal(pl, r[0]) and np.array_equal(un, r[1])


# ---- closed form vs per-pair restatement --------------------------------------------------------------
def _adversarial_case(rng):
    H, W = int(rng.integers(20, 140)), int(rng.integers(20, 140))
    m = np.zeros((H, W), bool)
    for _ in range(int(rng.integers(2, 7))):
        y = int(rng.integers(0, H)); w_ = int(rng.integers(1, 6)); m[y:y + w_, :] = True
        x = int(rng.integers(0, W)); w_ = int(rng.integers(1, 6)); m[:, x:x + w_] = True
    for _ in range(int(rng.integers(0, 8))):
        y, x, g = int(rng.integers(0, H)), int(rng.integers(0, W)), int(rng.integers(1, 5))
        m[y:y + g, x:x + g] = False
    max_px = int(rng.choice([12, 25, 40, 80, 150]))
    fg = np.argwhere(m)

    def pick(n):
        p = fg[rng.integers(0, len(fg), n)] + rng.integers(-2, 3, (n, 2)) * (rng.random((n, 1)) < 0.3)
        p[:, 0] = np.clip(p[:, 0], 0, H - 1)
        p[:, 1] = np.clip(p[:, 1], 0, W - 1)
        return p
    return m, pick(5), pick(9), max_px


def test_rule_equals_astar_random():
    """The closed form against the restated per-pair A* on adversarial small grids: EXACT on every
    pair outside the heap-tie regime (``tie_regime`` of cost_matrix_rule_ex: the reference's answer
    there depends on the pop order of equal-priority std::priority_queue entries);
    a 6.4 M-pair sweep of this generator family found 49 differing pairs, all of them flagged
    (tests/golden/make_heap_tie_cases.py)."""
    rng = np.random.default_rng(11)
    tot = flagged = 0
    for it in range(300):
        m, src, dst, max_px = _adversarial_case(rng)
        a = oracle.cost_matrix_astar(m, src, dst, max_px)
        r = oracle.cost_matrix_rule_ex(m, src, dst, max_px)
        clear = r[3] == 0
        assert np.array_equal(a[0][clear], r[0][clear]) and np.array_equal(a[1][clear], r[1][clear]), it
        tot += a[0].size
        flagged += int(r[3].sum())
        ok = (a[1] == 0) & (r[1] == 0) & (a[2] < 2 ** 24)
        assert np.array_equal(a[2][ok], r[2][ok])
    assert tot == 300 * 45 and flagged < tot // 50, (flagged, tot)      # the regime stays a corner


@pytest.mark.skipif(not oracle.ref_available(), reason='oracle/_ref not built')
def test_rule_equals_compiled_reference_random():
    """Same family against the COMPILED reference (oracle/_ref, astar.cpp unmodified)."""
    rng = np.random.default_rng(12)
    n = 0
    for it in range(60):
        m, src, dst, max_px = _adversarial_case(rng)
        pl, un = oracle.ref_cost_matrix(m, src, dst, max_px, threads=False)
        r = oracle.cost_matrix_rule_ex(m, src, dst, max_px)
        clear = r[3] == 0
        assert np.array_equal(pl[clear], r[0][clear]) and np.array_equal(un[clear], r[1][clear]), it
        n += int(clear.sum())
    assert n > 2500


def test_heap_tie_fixture(golden_dir):
    """The concrete counter-examples: pairs where the reference returns None
    because a stale heap entry with path_length max_px+1 pops on an f-tie, while the closed form
    finds the target.  The fixture holds the COMPILED reference's answers; the restated A* (same
    std::priority_queue) reproduces them, the closed form flags every one of them as tie regime."""
    g = np.load(os.path.join(golden_dir, 'heap_tie_cases.npz'))
    n = int(g['n'])
    assert n >= 12
    for i in range(n):
        H, W = g[f'shape_{i}']
        m = np.unpackbits(g[f'mask_{i}'])[:H * W].reshape(H, W)
        src, dst, max_px = g[f'src_{i}'][None], g[f'dst_{i}'][None], int(g[f'max_px_{i}'])
        a = oracle.cost_matrix_astar(m, src, dst, max_px)
        assert (int(a[0][0, 0]), int(a[1][0, 0])) == tuple(int(v) for v in g[f'ref_{i}']), i
        r = oracle.cost_matrix_rule_ex(m, src, dst, max_px)
        assert (int(r[0][0, 0]), int(r[1][0, 0])) == tuple(int(v) for v in g[f'rule_{i}']), i
        assert r[3][0, 0] == 1 and g[f'ref_{i}'][1] == 1 and r[1][0, 0] == 0, i
        if oracle.ref_available():
            pl, un = oracle.ref_cost_matrix(m, src, dst, max_px, threads=False)
            assert (int(pl[0, 0]), int(un[0, 0])) == tuple(int(v) for v in g[f'ref_{i}']), i


# ---- fixtures generated from the unmodified reference ---------------------------------------------------
def _astar_fixtures(golden_dir):
    return sorted(glob.glob(os.path.join(golden_dir, 'astar_s*.npz')))


def test_golden_astar_fixtures_exist(golden_dir):
    assert len(_astar_fixtures(golden_dir)) >= 4
    assert len(glob.glob(os.path.join(golden_dir, 'decode_nms_s*.npz'))) >= 3


@pytest.mark.parametrize('which', ['astar', 'rule'])
def test_golden_astar(golden_dir, which):
    fn = oracle.cost_matrix_astar if which == 'astar' else oracle.cost_matrix_rule
    for f in _astar_fixtures(golden_dir):
        g = np.load(f)
        H, W, max_px = int(g['H']), int(g['W']), int(g['max_px'])
        mask = np.unpackbits(g['mask_bits'])[:H * W].reshape(H, W)
        for t, tb in g['keys']:
            pl, un, wc = fn(mask, g[f'dets_yx_{tb}'], g[f'dets_yx_{t}'], max_px)
            assert np.array_equal(pl, g[f'dist_{t}_{tb}']), (f, t, tb)
            assert np.array_equal(un, g[f'none_{t}_{tb}']), (f, t, tb)
            ref_c = g[f'wcost_{t}_{tb}']
            ok = un == 0
            assert np.allclose(wc[ok], ref_c[ok], rtol=1e-5, atol=0)
            tc = oracle.transition_cost(pl, max_px, 0.6, int(t - tb))
            ref_t = g[f'tcost_{t}_{tb}']
            assert np.array_equal(np.isinf(tc), np.isinf(ref_t))
            fin = ~np.isinf(tc)
            assert np.allclose(tc[fin], ref_t[fin], rtol=1e-5, atol=0)


def test_golden_decode_nms(golden_dir):
    for f in sorted(glob.glob(os.path.join(golden_dir, 'decode_nms_s*.npz'))):
        g = np.load(f)
        o = oracle.decode_stitch_nms(g['Y'], g['tile_yx'], int(g['tilesize']), float(g['conf_thr']), int(g['nms_dist']))
        for k in ('pre_conf', 'pre_x', 'pre_y', 'conf', 'x', 'y', 'keep_idx'):
            assert np.array_equal(o[k], g[k]), (f, k)
