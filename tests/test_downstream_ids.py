"""north_star "the downstream assigned IDs are identical" (VERDICT r1 missing #6, row N1).

The reference's own chain  assign_ids -> _assign_IDs_to_detections -> libmot MinCostFlowTracker.process /
compute_trajectories -> libmot_det2det -> _agg_all_IDed_dets  (AxonDetections.py:505-524, 631-715, 786-842) runs
on the real reference classes with a deterministic stand-in for ortools' SimpleMinCostFlow (tests/mcf_standin.py):

* arm A: the reference end to end (compiled pyastar2d fork, coo_matrix leaves);
* arm B: the drop-in class with OUR astar_dets_paths object (AStarPaths of integer matrices; here filled by the
  closed-form oracle, which the -m gpu tests pin bit for bit to the CUDA kernels);
* arm C: arm B with libmot_bridge's arc tracker fed from arrays (the arc source is test-side numpy here; the K3
  source is checked against the same numpy on the GPU box, tests/test_vis_sim.py).

IDed_dets_all and the solver's input (every AddArc call, in order) must be identical across the arms.
CPU only, needs /root/reference."""
import sys

import numpy as np
import pandas as pd
import pytest
import torch

from tests.golden import ref_harness as rh
from tests import mcf_standin

pytestmark = pytest.mark.skipif(not rh.available(), reason='/root/reference not mounted')


class _NumpyArcs:
    """arc source for libmot_bridge's tracker: the arc table looked up with numpy (test side)."""

    def __init__(self, dists, name, max_px, miss_rate, max_num_misses, thr):
        from axtrack_b200 import mincostflow_models as mm
        self.d, self.name, self.mm = dists, name, mm
        _, self.arc_tab = mm.transition_tables(max_px, miss_rate, max_num_misses, 0.0, thr)

    def arcs(self, frame_idx, gap):
        a = self.arc_tab[gap - 1][np.asarray(self.d[f'{self.name}_t:{frame_idx:0>3}-t:{frame_idx - gap:0>3}'])]
        i, j = np.nonzero(a != self.mm.NO_ARC)
        return i, j, a[i, j]


def _setup(T=6, n=16, seed=31):
    import oracle
    from axtrack_b200 import synth
    from axtrack_b200.axon_detections import AStarPaths, make_drop_in
    rh.load()
    mcf = sys.modules['libmot.data_association.offline_mincostflow']
    mcf.pywrapgraph.SimpleMinCostFlow = mcf_standin.SimpleMinCostFlow
    Ref = sys.modules['axref.axtrack.AxonDetections'].AxonDetections
    H, W, max_px = 380, 420, 90
    mask = synth.make_mask(H, W, seed, pitch=(60, 140), width=(5, 12))
    tracks = synth.make_tracks(mask, n, T, seed, step=12, off_mask_frac=0.15, birth_death=0.05)
    ds = rh.make_dataset_stub('ids', mask, T)
    rng = np.random.default_rng(seed)
    dets = [pd.DataFrame({'conf': np.sort(rng.uniform(0.6, 1.2, len(yx)))[::-1].astype(np.float32), 'anchor_x': yx[:, 1].astype(np.int64),
                          'anchor_y': yx[:, 0].astype(np.int64)}, index=[f'Axon_{i:0>3}' for i in range(len(yx))]).convert_dtypes()
            for yx in tracks]
    img = torch.from_numpy(rng.random((1, H, W)).astype(np.float32))

    def make(cls):
        ad = cls(None, ds, dict(_params()), None)
        ad.max_px_assoc_dist = max_px
        ad.MCF_min_flow, ad.MCF_max_flow = 2, 40
        ad._detections = [d.copy() for d in dets]
        ad.get_frame_and_truedets = lambda t, unstitched=False: (img, None)      # the image only feeds feature_model (weight 0)
        return ad
    ours_paths = {}
    ref = make(Ref)
    ref.astar_dets_paths = ref._compute_detections_astar_paths(cache=None)      # the reference: compiled pyastar2d fork
    dists, unreach = {}, {}
    for key in ref.astar_dets_paths:
        t, tb = int(key.split('_t:')[1][:3]), int(key.split('-t:')[1][:3])
        pl, un, _ = oracle.cost_matrix_rule(mask, tracks[tb], tracks[t], max_px)
        dists[key], unreach[key] = pl.astype(np.int64), un
    ours = make(make_drop_in(Ref))
    ours.astar_dets_paths = AStarPaths(dists, unreach, mask.shape)
    return ref, ours, mcf, max_px


def _params():
    import os
    import pickle
    with open(os.path.join(rh.REF, 'deployed_model/params.pkl'), 'rb') as f:
        params = pickle.load(f)
    params['DEVICE'] = 'cpu'
    return params


def _ids(ad):
    graphs = []
    mcf = sys.modules['libmot.data_association.offline_mincostflow']
    orig = mcf.pywrapgraph.SimpleMinCostFlow

    class Rec(orig):
        def __init__(self):
            super().__init__()
            graphs.append(self)
    mcf.pywrapgraph.SimpleMinCostFlow = Rec
    try:
        ad._IDed_detections = ad._assign_IDs_to_detections(cache=None)
    finally:
        mcf.pywrapgraph.SimpleMinCostFlow = orig
    assert ad._IDed_detections is not None, 'the stand-in solver found no flow'
    ad.IDed_dets_all = ad._agg_all_IDed_dets()          # as assign_ids does (AxonDetections.py:522-524)
    return ad.IDed_dets_all, graphs[-1].calls


def test_assigned_ids_identical_reference_vs_drop_in():
    ref, ours, mcf, max_px = _setup()
    ours.arcs_from_arrays = False               # arm B: the reference's own arc loop on OUR distances
    ids_ref, calls_ref = _ids(ref)
    ids_our, calls_our = _ids(ours)
    assert calls_ref == calls_our and len(calls_ref) > 3 * sum(len(d) for d in ref._detections)
    assert len(ids_ref) > 0
    pd.testing.assert_frame_equal(ids_ref, ids_our)
    # arm C: transition arcs from arrays (libmot_bridge), numpy arc source
    from axtrack_b200 import libmot_bridge
    ours2_ref, ours2, _, _ = _setup()
    real = libmot_bridge.ArcTables
    libmot_bridge.ArcTables = lambda astar_dists, name, mp, mr, mnm, thr: _NumpyArcs(astar_dists, name, mp, mr, mnm, thr)
    try:
        ids_arr, calls_arr = _ids(ours2)
    finally:
        libmot_bridge.ArcTables = real
    assert calls_arr == calls_ref                # same AddArc calls in the same order
    pd.testing.assert_frame_equal(ids_ref, ids_arr)
