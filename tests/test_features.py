"""K5: appearance features = mincostflow_models.feature_model
(mincostflow_models.py:30-64).  Floating point, but every operation is determined (integer counts,
one float scale, one fused multiply-add), so the bar is BITWISE equality with the unmodified
reference's cv2 output (tests/golden/feature_hist_s*.npz, tests/golden/make_golden_features.py)."""
import glob
import os

import numpy as np
import pytest

import oracle


def _fixtures(golden_dir):
    files = sorted(glob.glob(os.path.join(golden_dir, 'feature_hist_s*.npz')))
    assert len(files) >= 3
    return [np.load(f) for f in files]


def test_oracle_matches_reference_fixtures(golden_dir):
    for g in _fixtures(golden_dir):
        got = oracle.feature_hist(g['image'], g['boxes'])
        assert got.dtype == np.float32 and np.array_equal(got, g['features'])


def test_oracle_matches_cv2_when_present():
    cv2 = pytest.importorskip('cv2')
    rng = np.random.default_rng(4)
    img = rng.random((120, 150), dtype=np.float32)
    boxes = np.array([[10, 20, 70, 70], [-5, -8, 70, 70], [100, 90, 70, 70], [30, 30, 1, 1]], np.int32)
    want = []
    for x, y, w, h in boxes:
        x1, x2 = max(y, 0), max(x, 0)
        c = img[x1:max(x1 + 1, x1 + h), x2:max(x2 + 1, x2 + w)]
        hist = cv2.calcHist([c], [0], None, [180], [0, 1])
        cv2.normalize(hist, hist, alpha=0, beta=1, norm_type=cv2.NORM_MINMAX)
        want.append(hist[:, 0])
    assert np.array_equal(oracle.feature_hist(img, boxes), np.array(want))


@pytest.mark.gpu
def test_gpu_matches_reference_fixtures(golden_dir):
    import torch
    from axtrack_b200 import ops, mincostflow_models
    for g in _fixtures(golden_dir):
        got = ops.feature_hist(torch.from_numpy(g['image']).cuda(), torch.from_numpy(g['boxes']).cuda())
        assert np.array_equal(got.cpu().numpy(), g['features'])
        fm = mincostflow_models.feature_model(image=g['image'], boxes=g['boxes'])      # the drop-in
        assert fm.shape == (len(g['boxes']), 180, 1) and fm.dtype == np.float32
        assert np.array_equal(fm[:, :, 0], g['features'])
    assert mincostflow_models.feature_model(image=g['image'], boxes=np.zeros((0, 4))).shape == (0,)


@pytest.mark.gpu
def test_gpu_matches_oracle_full_frame():
    """2048x2048 frame, ~200 boxes of 70x70 around detections (config 2 shape) plus other bin counts
    and ranges; bitwise against the oracle."""
    import torch
    from axtrack_b200 import ops
    rng = np.random.default_rng(8)
    img = (rng.random((2048, 2048), dtype=np.float32) ** 2 * 1.1).astype(np.float32)
    img[rng.random(img.shape) < 0.5] = 0
    xy = rng.integers(-10, 2040, (210, 2))
    boxes = np.concatenate([xy - 35, np.full((210, 2), 70)], 1).astype(np.int32)
    d_img, d_boxes = torch.from_numpy(img).cuda(), torch.from_numpy(boxes).cuda()
    assert np.array_equal(ops.feature_hist(d_img, d_boxes).cpu().numpy(), oracle.feature_hist(img, boxes))
    for bins, lo, hi in ((16, 0.0, 1.0), (256, -0.25, 1.25), (1000, 0.0, 0.5)):
        got = ops.feature_hist(d_img, d_boxes[:40], bins, lo, hi).cpu().numpy()
        assert np.array_equal(got, oracle.feature_hist(img, boxes[:40], bins, lo, hi)), bins
