"""Golden vectors for K5 (feature histograms), produced by the UNMODIFIED reference function
axtrack/mincostflow_models.py:30-64 (feature_model: cv2.calcHist + cv2.normalize) imported from
/root/reference in the build container (cv2 4.13).  Run from the repo root:

    python -m tests.golden.make_golden_features

Writes tests/golden/feature_hist_s*.npz: image (float32), boxes (int32 x, y, w, h), features (N, 180).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness as rh   # noqa: E402


def synth_frame(seed, H, W):
    """Microscopy-like frame: smooth structures, zero background (Timelapse zeroes the image off the
    mask, Timelapse.py:217), noise, and the values that decide bins: exact bin edges k/180, 0, 1,
    values below 0 and above 1, a NaN."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    img = 0.5 + 0.5 * np.sin(yy / 17.0) * np.cos(xx / 23.0)
    img = (img * rng.random((H, W), dtype=np.float32) ** 0.5).astype(np.float32)
    img[rng.random((H, W)) < 0.35] = 0.0
    edge = rng.integers(0, 181, (H, W)).astype(np.float32) / np.float32(180.0)
    sel = rng.random((H, W)) < 0.05
    img[sel] = edge[sel]
    sel = rng.random((H, W)) < 0.02
    img[sel] = rng.uniform(-0.5, 1.5, int(sel.sum())).astype(np.float32)
    img[3, 5] = np.nan
    if seed % 2:                      # a frame whose crops fill every bin (min count > 0: the shift term)
        img = rng.random((H, W), dtype=np.float32)
    return img


def main():
    assert rh.available()
    rh.load()
    fm = sys.modules['axref.axtrack.mincostflow_models'].feature_model
    for seed, (H, W, n) in enumerate([(260, 340, 40), (200, 300, 30), (90, 75, 12)]):
        rng = np.random.default_rng(100 + seed)
        img = synth_frame(seed, H, W)
        xy = np.stack([rng.integers(-40, W - 5, n), rng.integers(-40, H - 5, n)], 1)
        wh = np.full((n, 2), 70)
        wh[::7] = rng.integers(0, 120, (len(wh[::7]), 2))       # other sizes, zero width / height
        boxes = np.concatenate([xy, wh], 1).astype(np.int32)
        feats = fm(image=img, boxes=boxes)
        out = os.path.join(HERE, f'feature_hist_s{seed}.npz')
        np.savez_compressed(out, image=img, boxes=boxes, features=np.asarray(feats, np.float32)[:, :, 0])
        print(out, feats.shape, 'min count > 0 in', int((feats.min(1) > 0).sum()), 'boxes')


if __name__ == '__main__':
    main()
