"""K4: path pixels of selected pairs.  Pixels are heap-tie dependent in the
reference, so parity = a valid 4-connected path from source to target whose length and weighted
cost equal the reference's (oracle closed form / restated A*), None exactly where the reference
returns None."""
import numpy as np
import pytest

import oracle
from axtrack_b200 import synth

pytestmark = pytest.mark.gpu


def _check_paths(mask, src, dst, max_px):
    import torch
    from axtrack_b200 import ops
    bits = ops.pack_mask(torch.from_numpy(mask.astype(np.uint8)).cuda())
    ii, jj = np.meshgrid(np.arange(len(src)), np.arange(len(dst)), indexing='ij')
    req = np.concatenate([np.zeros((ii.size, 1), np.int32), src[ii.ravel()], dst[jj.ravel()]], 1).astype(np.int32)
    yx, ln = ops.trace_paths(bits, mask.shape[1], req, max_px)
    ops.raise_status()
    yx, ln = yx.cpu().numpy(), ln.cpu().numpy().reshape(len(src), len(dst))
    pl, un, wc = oracle.cost_matrix_rule(mask, src, dst, max_px)
    w = np.where(mask == 1, 1.0, 65536.0)
    n_paths = 0
    for i in range(len(src)):
        for j in range(len(dst)):
            L = ln[i, j]
            if un[i, j]:
                assert L == 0, (i, j, L)
                continue
            assert L == pl[i, j], (i, j, L, pl[i, j])
            p = yx[i * len(dst) + j, :L]
            assert tuple(p[0]) == tuple(src[i]) and tuple(p[-1]) == tuple(dst[j])
            assert (np.abs(np.diff(p, axis=0)).sum(1) == 1).all()                 # 4-connected steps
            assert len({(int(a), int(b)) for a, b in p}) == L                      # no pixel twice
            assert w[p[1:, 0], p[1:, 1]].sum() == wc[i, j]                         # the reference's path cost
            n_paths += 1
    return n_paths


def test_paths_on_corridor_mask():
    rng = np.random.default_rng(5)
    mask = synth.make_mask(900, 1000, seed=2, pitch=(80, 170), width=(6, 14))
    fg = np.argwhere(mask)
    src = fg[rng.integers(0, len(fg), 6)].astype(np.int32)
    dst = fg[rng.integers(0, len(fg), 40)].astype(np.int32)
    dst[3] = src[0]                                   # start == goal: a path of one pixel
    assert _check_paths(mask, src, dst, 500) > 20


def test_paths_wide_radius():
    """max_px beyond 511: the wide all-levels kernel records and walks back the entry directions."""
    rng = np.random.default_rng(6)
    mask = synth.make_mask(1300, 1400, seed=4, pitch=(80, 170), width=(6, 14))
    fg = np.argwhere(mask)
    src = fg[rng.integers(0, len(fg), 4)].astype(np.int32)
    dst = fg[rng.integers(0, len(fg), 30)].astype(np.int32)
    assert _check_paths(mask, src, dst, 900) > 20


def test_paths_through_background_and_short_cutoffs():
    """Sources / targets off the mask (paths that pay 65536 per background pixel), gaps, small max_px."""
    rng = np.random.default_rng(9)
    n = 0
    for it in range(6):
        H, W = int(rng.integers(40, 140)), int(rng.integers(40, 140))
        m = np.zeros((H, W), bool)
        for _ in range(4):
            y = int(rng.integers(0, H)); m[y:y + int(rng.integers(1, 5)), :] = True
            x = int(rng.integers(0, W)); m[:, x:x + int(rng.integers(1, 5))] = True
        for _ in range(5):
            y, x = int(rng.integers(0, H)), int(rng.integers(0, W)); m[y:y + 3, x:x + 3] = False
        fg = np.argwhere(m)
        pick = lambda k: np.clip(fg[rng.integers(0, len(fg), k)] + rng.integers(-2, 3, (k, 2)) * (rng.random((k, 1)) < 0.4),
                                 0, [H - 1, W - 1]).astype(np.int32)
        n += _check_paths(m, pick(4), pick(8), int(rng.choice([25, 60, 150, 500])))
    assert n > 30


def test_compute_astar_path_mirror():
    from axtrack_b200.paths import compute_astar_path, compute_astar_paths
    m = np.ones((5, 5), np.uint8)
    coo, L = compute_astar_path((0, 0), (4, 4), m, 500)        # pyastar2d/tests/test_astar.py:22-26: 9 pixels
    assert L == 9 and coo.getnnz() == 9 and coo.shape == (5, 5) and coo.dtype == bool
    m2 = np.ones((1, 20), np.uint8)
    assert compute_astar_path((0, 0), (0, 19), m2, 19) == (None, None)
    coo, L = compute_astar_path((0, 0), (0, 19), m2, 20)
    assert L == 20 and sorted(coo.col.tolist()) == list(range(20))
    res = compute_astar_paths(m, [(0, 0), (2, 2)], [(0, 4), (2, 2)])
    assert res[0][1] == 5 and res[1][1] == 1
