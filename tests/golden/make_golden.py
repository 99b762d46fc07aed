"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

    python tests/golden/make_golden.py

What is run, all from /root/reference through tests/golden/ref_harness.py:
  * AxonDetections._yolo_Y2pandas_det + Timelapse.stitch_tiles + AxonDetections._non_max_supression
    (AxonDetections.py:122-128) on seeded synthetic detector-head tensors;
  * AxonDetections._compute_detections_astar_paths (AxonDetections.py:526-585) -> compiled
    pyastar2d fork, then _get_astar_path_distances (AxonDetections.py:717-752);
  * mincostflow_models.transition_model (mincostflow_models.py:67-119) on those distances.
The fixtures hold inputs and reference outputs; tests compare the oracle and the CUDA path to
them without needing /root/reference.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from axtrack_b200 import synth            # noqa: E402
from tests.golden import ref_harness as rh   # noqa: E402


def gold_decode_nms(seed, H, W, n):
    import torch
    mask = synth.make_mask(H, W, seed)
    dets = synth.make_tracks(mask, n, 1, seed)[0]
    Y, _, _ = synth.make_head_tensor(dets, H, W, seed, dup_frac=0.35)
    ds = rh.make_dataset_stub(f'gold{seed}', mask, 1)
    ad = rh.make_axon_detections(ds)
    thr = ad.all_conf_thrs.min()
    tiled = ad._yolo_Y2pandas_det(torch.from_numpy(Y.copy()), conf_thr=thr)
    frame, _ = ds.stitch_tiles(tiled, reset_index=True)
    nms = ad._non_max_supression(frame)
    pre = np.stack([frame.conf.values.astype(np.float64), frame.anchor_x.values.astype(np.float64),
                    frame.anchor_y.values.astype(np.float64)], 1)
    post = np.stack([nms.conf.values.astype(np.float64), nms.anchor_x.values.astype(np.float64),
                     nms.anchor_y.values.astype(np.float64)], 1)
    # survivor indices = positions of the NMS survivors in the stitched pre-NMS frame
    keep = []
    for c, x, y in post:
        hit = np.nonzero((pre[:, 0] == c) & (pre[:, 1] == x) & (pre[:, 2] == y))[0]
        assert len(hit) == 1
        keep.append(hit[0])
    # the tile-wise frames (_pandas_tiled_dets, AxonDetections.py:131-132): tile, Axon_### label, conf, x, y
    tl = np.array([[i, int(lbl.split('_')[1]), float(r.conf), float(r.anchor_x), float(r.anchor_y)]
                   for i, df in enumerate(tiled) for lbl, r in df.iterrows()], np.float64).reshape(-1, 5)
    out = os.path.join(HERE, f'decode_nms_s{seed}.npz')
    np.savez_compressed(out, tiled=tl, Y=Y, tile_yx=synth.tile_table(H, W), H=H, W=W, conf_thr=float(thr),
                        nms_dist=int(ad.nms_min_dist), tilesize=int(ad.tilesize),
                        pre_conf=pre[:, 0].astype(np.float32), pre_x=pre[:, 1].astype(np.int64),
                        pre_y=pre[:, 2].astype(np.int64), conf=post[:, 0].astype(np.float32),
                        x=post[:, 1].astype(np.int64), y=post[:, 2].astype(np.int64),
                        keep_idx=np.array(keep, np.int64))
    print(out, 'pre', len(pre), 'post', len(post))


def gold_astar(seed, H, W, n, T, max_px, off_mask_frac=0.1, all_ones=False):
    import pandas as pd
    mask = np.ones((H, W), bool) if all_ones else synth.make_mask(H, W, seed, pitch=(70, 160), width=(5, 14))
    tracks = synth.make_tracks(mask, n, T, seed, off_mask_frac=off_mask_frac)
    name = f'gold{seed}'
    ds = rh.make_dataset_stub(name, mask, T)
    ad = rh.make_axon_detections(ds)
    ad.max_px_assoc_dist = max_px
    rng = np.random.default_rng(seed)
    dets = []
    for t in range(T):
        yx = tracks[t]
        conf = np.sort(rng.uniform(0.56, 1.1, len(yx)))[::-1]
        df = pd.DataFrame({'conf': conf.astype(np.float32), 'anchor_x': yx[:, 1].astype(np.int64),
                           'anchor_y': yx[:, 0].astype(np.int64)},
                          index=[f'Axon_{i:0>3}' for i in range(len(yx))]).convert_dtypes()
        dets.append(df)
    ad._detections = dets
    paths = ad._compute_detections_astar_paths(cache=None)
    dists = ad._get_astar_path_distances(paths)
    weights = ad._get_maskweights(0)
    tm = sys.modules['axref.axtrack.mincostflow_models'].transition_model
    out = dict(mask_bits=np.packbits(mask), H=H, W=W, T=T, max_px=max_px, name=name)
    for t in range(T):
        out[f'dets_yx_{t}'] = tracks[t].astype(np.int32)
    keys = []
    for key, rows in paths.items():
        t, tb = int(key.split('_t:')[1][:3]), int(key.split('-t:')[1][:3])
        keys.append((t, tb))
        none = np.array([[p is None for p in r] for r in rows], np.uint8)
        cost = np.full(none.shape, np.inf, np.float64)
        for i, r in enumerate(rows):
            for j, p in enumerate(r):
                if p is not None:
                    sy, sx = tracks[tb][i]
                    cost[i, j] = float(weights[p.row, p.col].astype(np.float64).sum() - weights[sy, sx])
        d = np.asarray(dists[key]).reshape(none.shape)
        out[f'dist_{t}_{tb}'] = d.astype(np.int64)
        out[f'none_{t}_{tb}'] = none
        out[f'wcost_{t}_{tb}'] = cost
        n_pred, n_cur = none.shape
        # transition_model with the deployed vis_sim_weight = 0: the cv2 double loop only feeds a
        # term multiplied by 0; run the reference function with 1-bin dummy histograms
        f_pred = [np.ones((1, 1), np.float32)] * n_pred
        f_cur = [np.ones((1, 1), np.float32)] * n_cur
        tc = tm(miss_rate=0.6, time_gap=t - tb, boxes=None, predecessor_boxes=None, features=f_cur,
                predecessor_features=f_pred, frame_idx=t, dataset_name=name,
                astar_dists={key: d}, max_px_assoc_dist=max_px, vis_sim_weight=0)
        out[f'tcost_{t}_{tb}'] = np.asarray(tc, np.float64).reshape(none.shape)
    out['keys'] = np.array(keys, np.int32)
    fn = os.path.join(HERE, f'astar_s{seed}.npz')
    np.savez_compressed(fn, **out)
    tot = sum(out[f'none_{t}_{tb}'].size for t, tb in keys)
    reach = sum(int((out[f'none_{t}_{tb}'] == 0).sum()) for t, tb in keys)
    print(fn, 'pairs', tot, 'reachable', reach)


if __name__ == '__main__':
    assert rh.available()
    gold_decode_nms(0, 1024, 1024, 60)
    gold_decode_nms(1, 700, 1300, 45)           # ragged: last tile row/col partly outside the image
    gold_decode_nms(2, 2048, 2048, 230)
    gold_astar(10, 640, 832, 28, 3, 500)
    gold_astar(11, 500, 500, 24, 3, 120, off_mask_frac=0.3)   # cut-off active
    gold_astar(12, 300, 420, 16, 2, 60, off_mask_frac=0.5)    # multi-level searches
    gold_astar(13, 200, 260, 12, 2, 500, all_ones=True)       # mask None -> all ones (Timelapse.py:210-213)
