"""transition_model mirror: host tables vs the reference's golden outputs on CPU,
the K3 gather kernel vs the same fixtures on the GPU (bitwise)."""
import glob
import os

import numpy as np
import pytest

import oracle
from axtrack_b200 import mincostflow_models as mm

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), 'golden', 'astar_s1*.npz')))


def _cases():
    for f in GOLD:
        z = np.load(f, allow_pickle=True)
        for k in z.keys():
            if k.startswith('tcost_'):
                t, tb = map(int, k.split('_')[1:])
                yield os.path.basename(f), int(z['max_px']), t, tb, z[f'dist_{t}_{tb}'], z[k]


def test_tables_reproduce_reference_transition_model():
    """table[d] gathered over the reference's own distance matrices == the UNMODIFIED reference's
    transition_model output (tests/golden/make_golden.py), bit for bit, and == the oracle port."""
    n = 0
    for name, mp, t, tb, d, want in _cases():
        got = mm.transition_table(mp, 0.6, t - tb)[d]
        assert np.array_equal(got.view(np.int64), want.view(np.int64)), (name, t, tb)
        assert np.array_equal(got.view(np.int64), oracle.transition_cost(d, mp, 0.6, t - tb).view(np.int64))
        n += 1
    assert n >= 6


def test_arc_table_matches_libmot_loop():
    """offline_mincostflow.py:192-198: arcs only where cost <= cost_threshold, int(1e5 * cost)."""
    cost, arc = mm.transition_tables(500, 0.6, 1, 0.0, cost_threshold=0.7)
    assert cost.shape == arc.shape == (2, 501)
    for g in range(2):
        for d in (0, 1, 17, 250, 251, 252, 253, 499, 500):
            c = cost[g, d]
            assert arc[g, d] == (mm.NO_ARC if c > 0.7 else int(1e5 * c))
    assert np.isinf(cost[:, 500]).all() and (arc[:, 500] == mm.NO_ARC).all()
    # a per-distance TABLE exists for weight 0 only; other weights go through transition_model (K7, tests/test_vis_sim.py)
    with pytest.raises(ValueError):
        mm.transition_table(500, 0.6, 1, vis_sim_weight=0.2)


@pytest.mark.gpu
def test_k3_gather_bitwise_vs_golden():
    import torch
    from axtrack_b200 import ops
    from axtrack_b200.ops_pairs import make_pairs
    all_cases = list(_cases())
    for mp in sorted({c[1] for c in all_cases}):
        cases = [c for c in all_cases if c[1] == mp]
        pairs, total, _ = make_pairs([(0, 0, c[4].shape[0], 0, c[4].shape[1], c[2] - c[3]) for c in cases])
        pl = np.concatenate([c[4].reshape(-1) for c in cases]).astype(np.int32)
        ct, at = mm.transition_tables(mp, 0.6, 1, 0.0, cost_threshold=0.7)
        cost, arc = ops.transition_cost_batch(torch.from_numpy(pl).cuda(), pairs, ct, at)
        ops.raise_status()
        cost, arc = cost.cpu().numpy(), arc.cpu().numpy()
        for (name, _, t, tb, d, want), pd in zip(cases, pairs):
            o = int(pd['out_off'])
            got = cost[o:o + d.size].reshape(d.shape)
            assert np.array_equal(got.view(np.int64), want.view(np.int64)), (name, t, tb)
            a = arc[o:o + d.size].reshape(d.shape)
            ref = np.array([[mm.NO_ARC if c > 0.7 else int(1e5 * c) for c in row] for row in want], np.int64).reshape(d.shape)
            assert np.array_equal(a, ref)
    cases = all_cases
    mp = cases[0][1]
    ct, at = mm.transition_tables(mp, 0.6, 1, 0.0, cost_threshold=0.7)
    # drop-in signature of the reference's transition_model
    name, _, t, tb, d, want = cases[0]
    got = mm.transition_model(miss_rate=0.6, time_gap=t - tb, boxes=None, predecessor_boxes=None, features=[],
                              predecessor_features=[], frame_idx=t, dataset_name='ds', vis_sim_weight=0.0,
                              astar_dists={f'ds_t:{t:0>3}-t:{tb:0>3}': d}, max_px_assoc_dist=mp)
    assert np.array_equal(got.view(np.int64), want.view(np.int64))
    # a path length outside the table is reported, not read out of bounds
    bad = torch.full((4,), mp + 7, dtype=torch.int32, device='cuda')
    p1, _, _ = make_pairs([(0, 0, 2, 0, 2, 1)])
    ops.transition_cost_batch(bad, p1, ct, None)
    from axtrack_b200 import _lib
    with pytest.raises(_lib.AxbError):
        ops.raise_status()
