"""The host-side mirror of AxonDetections (axtrack_b200/axon_detections.py) on the GPU, driven
through the reference's method names and compared with the golden fixtures of the unmodified
reference.  /root/reference does not exist on the GPU box, so the mixin sits on a minimal
stand-in base that provides only what AxonDetections.__init__ / get_frame_dets / to_cache /
from_cache provide (AxonDetections.py:22-79, 141-176, 320-323); tests/test_dropin_reference.py
composes the mixin with the real class where the reference is mounted."""
import glob
import os
import pickle

import numpy as np
import pytest

import oracle
from axtrack_b200 import synth

pytestmark = pytest.mark.gpu


class _Base:
    """What the mixin expects from the reference class."""
    def __init__(self, model, dataset, directory, max_px=500):
        self.model, self.dataset, self.name, self.dir = model, dataset, dataset.name, directory
        self.timepoint_subset = range(dataset.sizet)
        self.device = 'cpu'
        self.Sx = self.Sy = 12
        self.tilesize = 512
        self.MCF_max_num_misses = 1
        self.nms_min_dist = 23
        self.conf_thr = 0.7
        self.all_conf_thrs = np.sort(np.append(np.arange(0.55, 1, .04), self.conf_thr)).round(2)
        self.max_px_assoc_dist = max_px

    def __len__(self):
        return len(self.timepoint_subset)

    def get_frame_dets(self, which, t):
        return self._detections[t].copy()

    def to_cache(self, which, dat):
        with open(f'{self.dir}/{self.dataset.name}_{which}.pkl', 'wb') as f:
            pickle.dump(dat, f)

    def from_cache(self, which):
        with open(f'{self.dir}/{self.dataset.name}_{which}.pkl', 'rb') as f:
            return pickle.load(f)

    def _get_astar_path_distances(self, astar_paths):       # AxonDetections.py:717-752
        def rec(p):
            if isinstance(p, list):
                return [rec(e) for e in p]
            return self.max_px_assoc_dist if p is None else p.getnnz()
        keys, vals = astar_paths.keys(), list(astar_paths.values())
        return dict(zip(keys, [np.array(d) for d in rec(vals)]))


class _Dataset:
    def __init__(self, name, mask, heads):
        import torch
        from scipy.sparse import coo_matrix
        self.name, self.sizet = name, len(heads)
        self.sizey, self.sizex = mask.shape
        self.ytiles, self.xtiles = -(-mask.shape[0] // 512), -(-mask.shape[1] // 512)
        self.tile_info = torch.ones((self.ytiles, self.xtiles, self.sizet, 2))
        self.mask = [coo_matrix(mask.astype(bool))] * self.sizet
        self.heads = heads

    def construct_tiles(self, *a, **k):
        pass

    def get_frametiles_stack(self, t, device=None):
        import torch
        return torch.tensor([t]), torch.zeros(1)


class _Model:
    def __init__(self, heads):
        self.heads = heads

    def detect_axons(self, X):
        import torch
        return torch.from_numpy(self.heads[int(X[0])])


def _make(tmp_path, name, mask, heads, max_px=500):
    from axtrack_b200.axon_detections import make_drop_in
    ds = _Dataset(name, mask, heads)
    return make_drop_in(_Base)(_Model(heads), ds, str(tmp_path), max_px)


def test_detect_dataset_matches_reference_fixtures(tmp_path, golden_dir):
    import torch
    assert torch.cuda.is_available()
    for f in sorted(glob.glob(os.path.join(golden_dir, 'decode_nms_s*.npz'))):
        g = np.load(f)
        H, W = int(g['H']), int(g['W'])
        ad = _make(tmp_path, 'gold', np.ones((H, W), bool), [g['Y'], g['Y']])
        ad.detect_dataset(cache='to')
        for t in range(2):
            d = ad._detections[t]
            assert list(d.columns) == ['conf', 'anchor_x', 'anchor_y']
            assert str(d.conf.dtype) == 'Float32' and str(d.anchor_x.dtype) == 'Int64'
            assert list(d.index) == [f'Axon_{i:0>3}' for i in range(len(g['conf']))]
            assert np.array_equal(d.conf.to_numpy(np.float32), g['conf'])
            assert np.array_equal(d.anchor_x.to_numpy(np.int64), g['x']) and np.array_equal(d.anchor_y.to_numpy(np.int64), g['y'])
            tl = np.array([[i, int(lbl.split('_')[1]), float(r.conf), float(r.anchor_x), float(r.anchor_y)]
                           for i, df in enumerate(ad._pandas_tiled_dets[t]) for lbl, r in df.iterrows()], np.float64).reshape(-1, 5)
            assert np.array_equal(tl, g['tiled']), f
        again = _make(tmp_path, 'gold', np.ones((H, W), bool), [g['Y'], g['Y']])
        again.detect_dataset(cache='from')
        assert again._detections[1].equals(ad._detections[1])


def test_astar_paths_contract_and_cache(tmp_path, golden_dir):
    g = np.load(os.path.join(golden_dir, 'astar_s11.npz'))
    H, W, T, max_px = int(g['H']), int(g['W']), int(g['T']), int(g['max_px'])
    mask = np.unpackbits(g['mask_bits'])[:H * W].reshape(H, W).astype(bool)
    from axtrack_b200.axon_detections import _frame, _labels, AStarPaths
    ad = _make(tmp_path, str(g['name']), mask, [None] * T, max_px)
    ad._detections = [_frame(np.linspace(1, .6, len(g[f'dets_yx_{t}'])), g[f'dets_yx_{t}'][:, 1], g[f'dets_yx_{t}'][:, 0],
                             _labels(len(g[f'dets_yx_{t}']))) for t in range(T)]
    paths = ad._compute_detections_astar_paths(cache='to')
    assert isinstance(paths, AStarPaths)
    assert list(paths.keys()) == [f'{g["name"]}_t:{t:0>3}-t:{tb:0>3}' for t, tb in g['keys']]
    dists = ad._get_astar_path_distances(paths)
    generic = _Base._get_astar_path_distances(ad, paths)            # the reference's recursion over leaves
    for (t, tb), k in zip(g['keys'], paths.keys()):
        assert np.array_equal(dists[k], g[f'dist_{t}_{tb}']) and np.array_equal(generic[k], g[f'dist_{t}_{tb}'])
        rows = paths[k]
        none = np.array([[leaf is None for leaf in r] for r in rows], np.uint8).reshape(g[f'none_{t}_{tb}'].shape)
        assert np.array_equal(none, g[f'none_{t}_{tb}'])
        tc = oracle.transition_cost(dists[k], max_px, 0.6, int(t - tb))
        ref = g[f'tcost_{t}_{tb}']
        assert np.array_equal(np.isinf(tc), np.isinf(ref)) and np.allclose(tc[~np.isinf(tc)], ref[~np.isinf(ref)], rtol=1e-5, atol=0)
    back = ad._compute_detections_astar_paths(cache='from')
    assert list(back.keys()) == list(paths.keys())
    assert all(np.array_equal(back.dists[k], paths.dists[k]) for k in paths.keys())


def test_materialized_paths_cache_is_the_references_format(tmp_path, golden_dir):
    """materialize_paths = True (VERDICT r1 missing #4): the astar_dets_paths cache holds real
    scipy.sparse.coo_matrix(bool, (H, W)) leaves / None in a plain dict of nested lists (AxonDetections.py:578-579,
    utils.py:381-384), and a process WITHOUT axtrack_b200 reads it with the reference's own recursion
    (AxonDetections.py:717-752) and gets the reference's distances."""
    import subprocess
    import sys
    from scipy.sparse import coo_matrix
    g = np.load(os.path.join(golden_dir, 'astar_s11.npz'))
    H, W, T, max_px = int(g['H']), int(g['W']), int(g['T']), int(g['max_px'])
    mask = np.unpackbits(g['mask_bits'])[:H * W].reshape(H, W).astype(bool)
    from axtrack_b200.axon_detections import _frame, _labels
    ad = _make(tmp_path, str(g['name']), mask, [None] * T, max_px)
    ad.materialize_paths = True
    ad._detections = [_frame(np.linspace(1, .6, len(g[f'dets_yx_{t}'])), g[f'dets_yx_{t}'][:, 1], g[f'dets_yx_{t}'][:, 0],
                             _labels(len(g[f'dets_yx_{t}']))) for t in range(T)]
    paths = ad._compute_detections_astar_paths(cache='to')
    assert type(paths) is dict and list(paths.keys()) == [f'{g["name"]}_t:{t:0>3}-t:{tb:0>3}' for t, tb in g['keys']]
    w = np.where(mask, 1.0, 65536.0)
    n_paths = 0
    for (t, tb), k in zip(g['keys'], paths.keys()):
        rows = paths[k]
        assert type(rows) is list and all(type(r) is list for r in rows)
        src, dst = g[f'dets_yx_{tb}'], g[f'dets_yx_{t}']
        for i, r in enumerate(rows):
            for j, leaf in enumerate(r):
                assert (leaf is None) == bool(g[f'none_{t}_{tb}'][i, j])
                if leaf is None:
                    continue
                assert isinstance(leaf, coo_matrix) and leaf.shape == (H, W) and leaf.dtype == bool
                assert leaf.getnnz() == g[f'dist_{t}_{tb}'][i, j]
                px = set(zip(leaf.row.tolist(), leaf.col.tolist()))
                assert len(px) == leaf.getnnz() and tuple(src[i]) in px and tuple(dst[j]) in px
                # a 4-connected chain: every pixel but the two ends has exactly two path neighbours
                deg = [sum((y + dy, x + dx) in px for dy, dx in ((1, 0), (-1, 0), (0, 1), (0, -1))) for y, x in px]
                assert leaf.getnnz() == 1 or (sorted(deg)[:2] == [1, 1] and min(deg) >= 1)
                # the reference's cost of the path it returned (float32 g of the goal): sum of weights over path[1:]
                assert np.isclose(w[leaf.row, leaf.col].sum() - w[tuple(src[i])], g[f'wcost_{t}_{tb}'][i, j], rtol=1e-5)
                n_paths += 1
    assert n_paths > 50
    code = (
        "import sys, pickle, numpy as np\n"
        "sys.modules['axtrack_b200'] = None\n"
        f"paths = pickle.load(open({str(tmp_path)!r} + '/' + {str(g['name'])!r} + '_astar_dets_paths.pkl', 'rb'))\n"
        "def rec(p):\n"
        "    if isinstance(p, list): return [rec(e) for e in p]\n"
        f"    return {max_px} if p is None else p.getnnz()\n"
        "np.savez(sys.argv[1], **{k: np.array(rec(v)) for k, v in paths.items()})\n")
    out = str(tmp_path / 'dists.npz')
    subprocess.run([sys.executable, '-c', code, out], check=True, cwd=str(tmp_path), env={k: v for k, v in os.environ.items() if k != 'PYTHONPATH'})
    d = np.load(out)
    for (t, tb), k in zip(g['keys'], paths.keys()):
        assert np.array_equal(d[k], g[f'dist_{t}_{tb}'])


def test_out_of_grid_detection_raises_like_reference(tmp_path):
    from axtrack_b200.axon_detections import _frame, _labels
    mask = np.ones((60, 80), bool)
    ad = _make(tmp_path, 'oob', mask, [None, None])
    ad._detections = [_frame([.9, .8], [5, 80], [5, 7], _labels(2)), _frame([.9], [6], [6], _labels(1))]
    with pytest.raises(ValueError, match=r'Start of \(7, 80\) lies outside grid'):
        ad._compute_detections_astar_paths(cache=None)
