"""SURVEY 8f rank 1, finished (VERDICT r1 missing #1, #2): the visual-similarity term of transition_model (K7) and
the arc tables the tracker consumes (K3), against fixtures made by the UNMODIFIED reference
(tests/golden/make_golden_vis.py: mincostflow_models.transition_model with cv2 4.13) -- bitwise."""
import glob
import os

import numpy as np
import pytest

import oracle

WEIGHTS = (0.25, 0.6, 1.0)


def _fixtures(golden_dir):
    fs = sorted(glob.glob(os.path.join(golden_dir, 'vis_sim_s*.npz')))
    assert len(fs) >= 3
    return [np.load(f) for f in fs]


def _same_bits(a, b):
    return np.array_equal(np.asarray(a, np.float64).view(np.int64), np.asarray(b, np.float64).view(np.int64))


def test_oracle_matches_reference_fixtures(golden_dir):
    for g in _fixtures(golden_dir):
        vis = oracle.vis_sim(g['feats_src'], g['feats_dst'])
        assert _same_bits(vis, g['vis'])
        for w in WEIGHTS:
            c = oracle.transition_cost(g['dists'], int(g['max_px']), float(g['miss_rate']), int(g['time_gap']), w, vis)
            assert _same_bits(c, g[f'cost_w{w}']), w


def test_oracle_matches_cv2_when_present():
    cv2 = pytest.importorskip('cv2')
    rng = np.random.default_rng(3)
    for n in (180, 179, 181, 7, 64, 1):
        for _ in range(20):
            h1 = (rng.random((n, 1)) * (rng.random((n, 1)) < 0.6)).astype(np.float32)
            h2 = (rng.random((n, 1)) * (rng.random((n, 1)) < 0.6)).astype(np.float32)
            assert oracle.bhattacharyya(h1, h2) == cv2.compareHist(h1, h2, cv2.HISTCMP_BHATTACHARYYA)


@pytest.fixture(scope='module')
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('-m gpu tests need a CUDA device (no CPU fallback exists)')
    torch.cuda.set_device(0)
    return torch


@pytest.mark.gpu
def test_k7_bitwise_vs_reference_fixtures(cuda, golden_dir):
    from axtrack_b200 import ops
    for g in _fixtures(golden_dir):
        fs, fd = cuda.from_numpy(g['feats_src']).cuda(), cuda.from_numpy(g['feats_dst']).cuda()
        vis = ops.vis_sim(fs, fd).cpu().numpy()
        ops.raise_status()
        assert _same_bits(vis, g['vis'])
        # with the distances: entries at max_px are skipped (their cost is inf whatever the similarity)
        d = cuda.from_numpy(g['dists'].astype(np.int32)).cuda()
        vis2 = ops.vis_sim(fs, fd, d, int(g['max_px'])).cpu().numpy()
        far = g['dists'] == int(g['max_px'])
        assert _same_bits(vis2[~far], g['vis'][~far]) and (vis2[far] == 0).all() and far.any() and (~far).any()
    # other histogram lengths (the len % 4 tail), an empty side
    rng = np.random.default_rng(5)
    for bins in (7, 181, 64):
        a = (rng.random((11, bins)) * (rng.random((11, bins)) < 0.7)).astype(np.float32)
        b = (rng.random((50, bins)) * (rng.random((50, bins)) < 0.7)).astype(np.float32)
        got = ops.vis_sim(cuda.from_numpy(a).cuda(), cuda.from_numpy(b).cuda()).cpu().numpy()
        assert _same_bits(got, oracle.vis_sim(a, b)), bins
    assert ops.vis_sim(cuda.zeros((0, 180)).cuda(), cuda.zeros((4, 180)).cuda()).shape == (0, 4)


@pytest.mark.gpu
def test_transition_model_with_vis_sim_weight(cuda, golden_dir):
    """The drop-in transition_model for vis_sim_weight != 0: K7 + the reference's own float64 expression."""
    from axtrack_b200 import mincostflow_models as mm
    for g in _fixtures(golden_dir):
        gap = int(g['time_gap'])
        key = f'gold_t:{5:0>3}-t:{5 - gap:0>3}'
        for w in WEIGHTS:
            c = mm.transition_model(miss_rate=float(g['miss_rate']), time_gap=gap, boxes=None, predecessor_boxes=None,
                                    features=g['feats_dst'][:, :, None], predecessor_features=g['feats_src'][:, :, None], frame_idx=5,
                                    dataset_name='gold', astar_dists={key: g['dists']}, max_px_assoc_dist=int(g['max_px']), vis_sim_weight=w)
            assert _same_bits(c, g[f'cost_w{w}']), w


@pytest.mark.gpu
def test_arc_tables_and_arc_tracker(cuda):
    """libmot_bridge: K3 arc tables of all frame pairs in one launch == the table looked up with numpy, in the
    reference's loop order; the tracker subclass issues the same AddArc calls as the per-pair Python loop
    (offline_mincostflow.py:166-198), checked here against a minimal stand-in base tracker (the real libmot class is
    exercised by tests/test_downstream_ids.py in the build container)."""
    from axtrack_b200 import libmot_bridge, mincostflow_models as mm
    rng = np.random.default_rng(9)
    max_px, miss, mnm, thr = 120, 0.6, 1, 0.7
    n = [7, 0, 9, 12, 5]
    dists = {}
    for t in range(len(n)):
        for tb in (t - 1, t - 2):
            if tb >= 0:
                d = rng.integers(1, max_px + 1, (n[tb], n[t])).astype(np.int64)
                d[rng.random(d.shape) < 0.4] = max_px
                dists[f'seq_t:{t:0>3}-t:{tb:0>3}'] = d
    tabs = libmot_bridge.ArcTables(dists, 'seq', max_px, miss, mnm, thr)
    _, arc_tab = mm.transition_tables(max_px, miss, mnm, 0.0, thr)
    for key, d in dists.items():
        t, tb = int(key[6:9]), int(key[-3:])
        i, j, a = tabs.arcs(t, t - tb)
        want = arc_tab[t - tb - 1][d]
        wi, wj = np.nonzero(want != mm.NO_ARC)
        assert np.array_equal(i, wi) and np.array_equal(j, wj) and np.array_equal(a, want[wi, wj]), key

    class Graph:
        def __init__(self):
            self.calls = []

        def AddArcWithCapacityAndUnitCost(self, *a):
            self.calls.append(tuple(int(x) for x in a))

    class Base:                      # the node bookkeeping of MinCostFlowTracker.process, nothing else
        def __init__(self, loop):
            self.graph, self.nodes_in_timestep, self.node, self.node_idx = Graph(), [], {}, 2
            self.max_num_misses, self.miss_rate, self.cost_threshold, self.vis_sim_weight = mnm, miss, thr, 0
            self.astar_dists, self.dataset_name, self.max_px_assoc_dist = dists, 'seq', max_px
            self.transition_model = loop

        def process(self, boxes, scores, image=None, features=None, **kw):
            ids = []
            for _ in range(len(boxes)):
                ids.append(self.node_idx)
                self.node_idx += 2
            for k, pred in enumerate(self.nodes_in_timestep[-(1 + self.max_num_misses):]):
                if not pred or not ids:
                    continue
                gap = len(self.nodes_in_timestep[-(1 + self.max_num_misses):]) - k
                for i, costs in enumerate(self.transition_model(time_gap=gap, frame_idx=kw['frame_idx'])):
                    for j, cost in enumerate(costs):
                        if cost > self.cost_threshold:
                            continue
                        self.graph.AddArcWithCapacityAndUnitCost(pred[i] + 1, ids[j], 1, int(mm.MULTI_FACTOR * cost))
            self.nodes_in_timestep.append(ids)

    def loop_model(time_gap, frame_idx):
        return oracle.transition_cost(dists[f'seq_t:{frame_idx:0>3}-t:{frame_idx - time_gap:0>3}'], max_px, miss, time_gap)
    ref, arr = Base(loop_model), libmot_bridge.make_arc_tracker(Base)(loop_model)
    for t in range(len(n)):
        for tr in (ref, arr):
            tr.process(np.zeros((n[t], 4)), np.zeros(n[t]), frame_idx=t)
    assert ref.graph.calls == arr.graph.calls and len(ref.graph.calls) > 20


@pytest.mark.gpu
def test_feature_hist_of_a_block_of_frames(cuda):
    """K5 for all frames of a block in one launch == one launch per frame (how libmot calls feature_model)."""
    from axtrack_b200 import ops
    rng = np.random.default_rng(2)
    F, H, W = 5, 130, 170
    imgs = rng.random((F, H, W), dtype=np.float32)
    imgs[rng.random(imgs.shape) < 0.4] = 0
    boxes, frames = [], []
    for f in range(F):
        k = int(rng.integers(0, 30)) if f != 2 else 0
        xy = np.stack([rng.integers(-30, W - 5, k), rng.integers(-30, H - 5, k)], 1)
        boxes.append(np.concatenate([xy, np.full((k, 2), 70)], 1).astype(np.int32))
        frames.append(np.full(k, f, np.int32))
    d = cuda.from_numpy(imgs).cuda()
    got = ops.feature_hist_frames(d, cuda.from_numpy(np.concatenate(boxes)).cuda(), cuda.from_numpy(np.concatenate(frames)).cuda()).cpu().numpy()
    ops.raise_status()
    want = np.concatenate([oracle.feature_hist(imgs[f], boxes[f]) if len(boxes[f]) else np.zeros((0, 180), np.float32) for f in range(F)])
    assert np.array_equal(got.view(np.int32), want.view(np.int32))
    per = np.concatenate([ops.feature_hist(d[f], cuda.from_numpy(boxes[f]).cuda()).cpu().numpy() for f in range(F)])
    assert np.array_equal(got.view(np.int32), per.view(np.int32))
