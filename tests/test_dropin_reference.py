"""Where /root/reference is mounted (the build container): the mixin composed with the REAL
reference AxonDetections class.  CPU only: checks the override surface and that the reference's
unchanged consumers (_get_astar_path_distances recursion, transition_model) accept the
AStarPaths object; matrices here come from the oracle (the checker), not from the GPU."""
import sys

import numpy as np
import pytest

from tests.golden import ref_harness as rh

pytestmark = pytest.mark.skipif(not rh.available(), reason='/root/reference not mounted')


def test_mixin_overrides_exist_in_reference_and_compose():
    import oracle
    from axtrack_b200 import synth
    from axtrack_b200.axon_detections import B200AssociationMixin, make_drop_in, AStarPaths
    rh.load()
    Ref = sys.modules['axref.axtrack.AxonDetections'].AxonDetections
    for name in ('detect_dataset', '_compute_detections_astar_paths', '_get_astar_path_distances',
                 'get_frame_and_truedets', '_assign_IDs_to_detections'):
        assert callable(getattr(Ref, name)), name            # the reference still has what we override
        assert getattr(make_drop_in(Ref), name) is not getattr(Ref, name)
    for name in ('get_frame_dets', 'to_cache', 'from_cache', 'assign_ids', '_agg_all_IDed_dets', 'libmot_det2det', 'det2libmot_det'):
        assert getattr(make_drop_in(Ref), name) is getattr(Ref, name)      # inherited unchanged

    # the real class, real __init__ / get_frame_dets; matrices from the oracle
    import pandas as pd
    H, W, T, max_px = 300, 420, 3, 60
    mask = synth.make_mask(H, W, 12, pitch=(70, 160), width=(5, 14))
    tracks = synth.make_tracks(mask, 16, T, 12, off_mask_frac=0.5)
    ds = rh.make_dataset_stub('cmp', mask, T)
    import pickle, os
    with open(os.path.join(rh.REF, 'deployed_model/params.pkl'), 'rb') as f:
        params = pickle.load(f)
    params['DEVICE'] = 'cpu'
    ad = make_drop_in(Ref)(None, ds, params, None)
    ref = Ref(None, ds, params, None)
    assert isinstance(ad, B200AssociationMixin) and isinstance(ad, Ref)
    ad.max_px_assoc_dist = ref.max_px_assoc_dist = max_px
    dets = [pd.DataFrame({'conf': np.linspace(1, .6, len(yx)).astype(np.float32), 'anchor_x': yx[:, 1].astype(np.int64),
                          'anchor_y': yx[:, 0].astype(np.int64)}, index=[f'Axon_{i:0>3}' for i in range(len(yx))]).convert_dtypes()
            for yx in tracks]
    ad._detections = ref._detections = dets
    ref_paths = ref._compute_detections_astar_paths(cache=None)             # compiled pyastar2d fork
    ref_dists = ref._get_astar_path_distances(ref_paths)
    dists, unreach = {}, {}
    for key in ref_paths:
        t, tb = int(key.split('_t:')[1][:3]), int(key.split('-t:')[1][:3])
        pl, un, _ = oracle.cost_matrix_rule(mask, tracks[tb], tracks[t], max_px)
        dists[key], unreach[key] = pl.astype(np.int64), un
    ours = AStarPaths(dists, unreach, mask.shape)
    assert list(ours.keys()) == list(ref_paths.keys())
    fast = ad._get_astar_path_distances(ours)                                # mixin shortcut
    generic = Ref._get_astar_path_distances(ad, ours)                        # reference recursion over our leaves
    tm = sys.modules['axref.axtrack.mincostflow_models'].transition_model
    for key in ref_paths:
        want = np.asarray(ref_dists[key]).reshape(dists[key].shape)
        assert np.array_equal(fast[key], want) and np.array_equal(np.asarray(generic[key]).reshape(want.shape), want), key
        none_ref = np.array([[p is None for p in r] for r in ref_paths[key]], np.uint8).reshape(want.shape)
        none_our = np.array([[p is None for p in r] for r in ours[key]], np.uint8).reshape(want.shape)
        assert np.array_equal(none_ref, none_our), key
        t, tb = int(key.split('_t:')[1][:3]), int(key.split('-t:')[1][:3])
        n_pred, n_cur = want.shape
        args = dict(miss_rate=0.6, time_gap=t - tb, boxes=None, predecessor_boxes=None, features=[np.ones((1, 1), np.float32)] * n_cur,
                    predecessor_features=[np.ones((1, 1), np.float32)] * n_pred, frame_idx=t, dataset_name='cmp',
                    max_px_assoc_dist=max_px, vis_sim_weight=0)
        a = np.asarray(tm(astar_dists={key: np.asarray(ref_dists[key])}, **args), np.float64)
        b = np.asarray(tm(astar_dists={key: fast[key]}, **args), np.float64)
        assert np.array_equal(a, b), key                                     # bit-identical final cost matrix
