"""Golden vectors for K7 (visual similarity) and the vis_sim branch of transition_model, produced by the
UNMODIFIED reference functions axtrack/mincostflow_models.py:30-64 (feature_model) and :67-119
(transition_model: cv2.compareHist HISTCMP_BHATTACHARYYA double loop, nan_to_num, -log) imported from
/root/reference in the build container (cv2 4.13, numpy 2.3).  Run from the repo root:

    python -m tests.golden.make_golden_vis

Writes tests/golden/vis_sim_s*.npz: feats_src / feats_dst (float32 [n, 180]), dists (int64 [n_src, n_dst]),
vis (float64, the reference's vis_sim matrix), cost_w* (float64, transition_model output) per weight,
miss_rate, time_gap, max_px.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden import ref_harness as rh                      # noqa: E402
from tests.golden.make_golden_features import synth_frame       # noqa: E402

WEIGHTS = (0.25, 0.6, 1.0)


def main():
    import cv2
    assert rh.available()
    rh.load()
    mm = sys.modules['axref.axtrack.mincostflow_models']
    for seed, (H, W, n_src, n_dst, gap) in enumerate([(260, 340, 23, 37, 1), (200, 300, 9, 70, 2), (120, 90, 33, 5, 1)]):
        rng = np.random.default_rng(300 + seed)
        img = synth_frame(seed, H, W)

        def boxes(n):
            xy = np.stack([rng.integers(-30, W - 5, n), rng.integers(-30, H - 5, n)], 1)
            return np.concatenate([xy, np.full((n, 2), 70)], 1).astype(np.int32)
        fs = mm.feature_model(image=img, boxes=boxes(n_src))
        fd = mm.feature_model(image=np.ascontiguousarray(img[::-1]), boxes=boxes(n_dst))
        if seed == 2:
            fs[3] = 0                     # an all-zero histogram: s1 * s2 = 0 -> distance 1 -> similarity 0
            fd[1] = fs[5]                 # identical histograms: similarity 1 (up to rounding)
        max_px = 500
        dists = rng.integers(1, max_px + 1, (n_src, n_dst)).astype(np.int64)
        dists[rng.random(dists.shape) < 0.5] = max_px
        name = 'gold'
        key = f'{name}_t:{5:0>3}-t:{5 - gap:0>3}'
        vis = np.nan_to_num(np.array([[1 - cv2.compareHist(f1, f2, cv2.HISTCMP_BHATTACHARYYA) for f2 in fd] for f1 in fs]))
        out = dict(feats_src=fs[:, :, 0], feats_dst=fd[:, :, 0], dists=dists, vis=vis, miss_rate=0.6, time_gap=gap, max_px=max_px)
        for w in WEIGHTS:
            out[f'cost_w{w}'] = mm.transition_model(miss_rate=0.6, time_gap=gap, boxes=None, predecessor_boxes=None, features=fd,
                                                    predecessor_features=fs, frame_idx=5, dataset_name=name, astar_dists={key: dists},
                                                    max_px_assoc_dist=max_px, vis_sim_weight=w)
        path = os.path.join(HERE, f'vis_sim_s{seed}.npz')
        np.savez_compressed(path, **out)
        print(path, vis.shape, 'vis range', float(vis.min()), float(vis.max()))


if __name__ == '__main__':
    main()
