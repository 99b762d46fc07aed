"""Host-side mirror of the reference's ``AxonDetections`` interface for the association hot path.

``B200AssociationMixin`` overrides exactly the methods SURVEY.md 8(b) lists and nothing else:

    detect_dataset                      /root/reference/axtrack/AxonDetections.py:87-139
        (the per-frame body _yolo_Y2pandas_det + stitch_tiles + _non_max_supression, :122-128,
         becomes ONE fused kernel launch over all frames; the model forward stays PyTorch)
    _compute_detections_astar_paths     AxonDetections.py:526-585
    _get_astar_path_distances           AxonDetections.py:717-752

Same names, arguments, cache files ({dir}/{name}_{which}.pkl, AxonDetections.py:141-176), dict
keys (f'{name}_t:{t:0>3}-t:{t_bef:0>3}', :554,563) and error behaviour (ValueError texts of
astar_wrapper.py:35-45).  Use it on top of the reference's class:

    from axtrack.AxonDetections import AxonDetections as Ref
    class AxonDetections(B200AssociationMixin, Ref): pass          # or make_drop_in(Ref)

    get_frame_and_truedets              AxonDetections.py:355-376 (the frame from the dense X, no tile round trip)
    _assign_IDs_to_detections           AxonDetections.py:631-715 (unchanged code; only the tracker CLASS it builds is
                                        swapped for libmot_bridge's, which adds the transition arcs from arrays)

Everything else (get_frame_dets, libmot's solve, ...) is inherited unchanged.
The mixin only touches: self.model.detect_axons, self.dataset.{name, mask, sizey, sizex, ytiles,
xtiles, tile_info, construct_tiles, get_frametiles_stack}, self.{Sx, Sy, tilesize, nms_min_dist,
all_conf_thrs, max_px_assoc_dist, MCF_max_num_misses, timepoint_subset, device, dir}.
"""
import numpy as np
import pandas as pd
import torch

from . import ops
from .ops_pairs import make_pairs


class PathLength:
    """Leaf of ``astar_dets_paths``: the reference only ever calls ``.getnnz()`` on a leaf
    (AxonDetections.py:736) -- the number of pixels on the A* path.  The pixels themselves are
    heap-tie dependent and never read back (AxonDetections.py:924-935 is 'not implemented')."""
    __slots__ = ('n', 'shape')

    def __init__(self, n, shape):
        self.n, self.shape = int(n), tuple(shape)

    def getnnz(self):
        return self.n

    def __repr__(self):
        return f'<A* path of {self.n} pixels on a {self.shape[0]}x{self.shape[1]} frame>'


class AStarPaths(dict):
    """``astar_dets_paths``: key -> nested list [i_pred][j_cur] of PathLength | None, built lazily
    from the integer matrices the GPU returned (``.dists`` / ``.unreachable``, same keys)."""

    def __init__(self, dists, unreachable, frame_shape):
        super().__init__()
        self.dists, self.unreachable, self.frame_shape = dists, unreachable, tuple(frame_shape)
        for k in dists:
            dict.__setitem__(self, k, None)

    def _rows(self, k):
        d, u = self.dists[k], self.unreachable[k]
        return [[None if u[i, j] else PathLength(d[i, j], self.frame_shape) for j in range(d.shape[1])]
                for i in range(d.shape[0])]

    def __getitem__(self, k):
        v = dict.__getitem__(self, k)
        if v is None:
            v = self._rows(k)
            dict.__setitem__(self, k, v)
        return v

    def values(self):
        return [self[k] for k in self.keys()]

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def __reduce__(self):     # pickles as the matrices (the 'to' cache), not 8M leaf objects
        return (AStarPaths, (self.dists, self.unreachable, self.frame_shape))


def tile_table(dataset):
    """int32 [ntiles, 2] (tile_y, tile_x) of the compact non-empty tiles, row-major
    (Timelapse.flat_tile_idx2yx_tile_idx, Timelapse.py:141-148)."""
    ne = np.asarray(torch.as_tensor(dataset.tile_info)[..., 0].any(-1).cpu().numpy(), bool)
    flat = np.arange(dataset.ytiles * dataset.xtiles).reshape(dataset.ytiles, dataset.xtiles)[ne].reshape(-1)
    ty, tx = np.divmod(flat, dataset.xtiles)
    return np.stack([ty, tx], 1).astype(np.int32)


def _labels(n):
    return [f'Axon_{i:0>3}' for i in range(n)]


def _frame(conf, x, y, index):
    # dtypes as after the reference's .convert_dtypes() (AxonDetections.py:235): Float32 / Int64
    df = pd.DataFrame({'conf': pd.array(np.asarray(conf, np.float32), dtype='Float32'),
                       'anchor_x': pd.array(np.asarray(x, np.int64), dtype='Int64'),
                       'anchor_y': pd.array(np.asarray(y, np.int64), dtype='Int64')}, index=index)
    return df


class B200AssociationMixin:
    """See the module docstring.  Must precede the reference class in the MRO."""

    # ---- detect_dataset: AxonDetections.py:87-139 ------------------------------------------
    def detect_dataset(self, cache=None):
        self.dataset.construct_tiles(self.device, force_no_transformation=True)
        if cache == 'from':
            self._yolo_targets = self.from_cache('_yolo_targets')
            self._pandas_tiled_dets = self.from_cache('_pandas_tiled_dets')
            self._detections = self.from_cache('_detections')
        else:
            self._yolo_targets, heads = [], []
            print(f'Detecting axons in {self.dataset.name} data: ', end='\n')
            fpf = getattr(self, 'frames_per_forward', None)
            if fpf:
                # opt-in K6 front end: tiles cut on the device from the dense X, `fpf` frames per forward
                # (the forward's cuDNN algorithms may differ with the batch size, so head values are
                # not guaranteed bit-identical to the one-frame-per-forward default below)
                from .tiling import FrameTiler
                ds = self.dataset
                tiler = FrameTiler(ds.X.to_dense().to(self.device, torch.float32), ds.tilesize, ds.temporal_context,
                                   ds.use_motion_filtered, ds.timepoints_indices)
                Y = tiler.detect_frames(self.model, self.timepoint_subset, fpf)
                self._yolo_targets = [ds.target_tiled[ds.timepoints_indices[t]].to(self.device) for t in self.timepoint_subset]
            else:
                for t in self.timepoint_subset:
                    X, yolo_target = self.dataset.get_frametiles_stack(t, self.device)
                    heads.append(self.model.detect_axons(X))          # PyTorch forward, unchanged (model.py:119-125)
                    self._yolo_targets.append(yolo_target)
                Y = torch.stack(heads)
            self._pandas_tiled_dets, self._detections = self._decode_frames(Y)
            print('Done.\n', flush=True)
        if cache == 'to':
            self.to_cache('_yolo_targets', self._yolo_targets)
            self.to_cache('_pandas_tiled_dets', self._pandas_tiled_dets)
            self.to_cache('_detections', self._detections)

    def _decode_frames(self, Y):
        """Y [T, ntiles, Sx, Sy, 3] -> (list[T] of list[ntiles] of tile DataFrames, list[T] of frame
        DataFrames after NMS), as _yolo_Y2pandas_det + stitch_tiles + _non_max_supression produce
        them (AxonDetections.py:122-128): ONE launch of the fused kernel for all frames."""
        dev = Y.device if Y.is_cuda else torch.device('cuda', torch.cuda.current_device())
        Y = Y.detach().to(dev, torch.float32).contiguous()
        tt = tile_table(self.dataset)
        T, ntiles, Sx, Sy, _ = Y.shape
        assert len(tt) == ntiles, 'head tensor and non-empty tile table disagree'
        o = ops.decode_nms(Y, torch.from_numpy(tt).to(dev), self.tilesize, float(self.all_conf_thrs.min()),
                           int(self.nms_min_dist), want_pre=True)
        ops.raise_status(device=dev)
        h = {k: v.cpu().numpy() for k, v in o.items()}
        cpt = Sx * Sy
        tiled_all, frames = [], []
        for t in range(T):
            npre, nout = int(h['n_pre'][t]), int(h['n_out'][t])
            cell, conf, yx = h['pre_cell'][t, :npre], h['pre_conf'][t, :npre], h['pre_yx'][t, :npre]
            tile = cell // cpt
            tiled = []
            for i in range(ntiles):
                sel = np.nonzero(tile == i)[0]                      # contiguous, conf ascending
                rank = np.argsort(np.argsort(cell[sel], kind='stable'), kind='stable')   # Axon_### = rank in cell order (:222-230)
                tiled.append(_frame(conf[sel], yx[sel, 1] - tt[i, 1] * self.tilesize, yx[sel, 0] - tt[i, 0] * self.tilesize,
                                    [f'Axon_{r:0>3}' for r in rank]))
            tiled_all.append(tiled)
            frames.append(_frame(h['conf'][t, :nout], h['yx'][t, :nout, 1], h['yx'][t, :nout, 0], _labels(nout)))
        return tiled_all, frames

    # ---- _compute_detections_astar_paths: AxonDetections.py:526-585 ------------------------------
    def _frame_masks(self):
        """(uint8 [n_masks, H, W], mask_of_frame [T]): dataset.mask[t] exactly as the reference
        indexes it (AxonDetections.py:598), identical masks stored once (a 2-D mask file is stacked
        T times, Timelapse.py:213-214)."""
        uniq, first, of_frame = [], {}, []
        for t in range(len(self)):
            m = self.dataset.mask[t]
            key = id(m)
            if key not in first:
                dense = (np.asarray(m.todense()) == 1).astype(np.uint8)
                for j, u in enumerate(uniq):
                    if u.shape == dense.shape and np.array_equal(u, dense):
                        first[key] = j
                        break
                else:
                    first[key] = len(uniq)
                    uniq.append(dense)
            of_frame.append(first[key])
        return np.stack(uniq), np.asarray(of_frame, np.int32)

    def _compute_detections_astar_paths(self, cache='to'):
        if cache == 'from':
            return self.from_cache('astar_dets_paths')
        print('\nComputing A* detection paths between detections...', end='')
        T, name = len(self), self.dataset.name
        dets = [self.get_frame_dets('all', t) for t in range(T)]
        yx = [np.stack([d.anchor_y.to_numpy(dtype=np.int64), d.anchor_x.to_numpy(dtype=np.int64)], 1).astype(np.int32).reshape(-1, 2)
              for d in dets]
        offs = np.concatenate([[0], np.cumsum([len(a) for a in yx])]).astype(np.int64)
        masks, of_frame = self._frame_masks()
        keys = [(t, tb) for t in range(T) for tb in range(t - 1, t - (self.MCF_max_num_misses + 2), -1) if tb >= 0]
        pairs, total, _ = make_pairs([(int(of_frame[t]), int(offs[tb]), len(yx[tb]), int(offs[t]), len(yx[t]), t - tb) for t, tb in keys])
        dev = torch.device('cuda', torch.cuda.current_device())
        bits = ops.pack_mask(torch.from_numpy(masks).to(dev))
        all_yx = np.concatenate(yx) if T else np.zeros((0, 2), np.int32)
        pl, un, _, _ = ops.cost_matrix_batch(bits, masks.shape[2], torch.from_numpy(all_yx).to(dev), pairs,
                                             int(self.max_px_assoc_dist), want_cost=False)
        ops.raise_status(pairs, all_yx, device=dev)      # ValueError as astar_wrapper.py:39-45
        pl, un = pl.cpu().numpy(), un.cpu().numpy()
        dists, unreach = {}, {}
        for (t, tb), pd_ in zip(keys, pairs):
            o, ns, nd = int(pd_['out_off']), int(pd_['n_src']), int(pd_['n_dst'])
            lbl = f'{name}_t:{t:0>3}-t:{tb:0>3}'
            dists[lbl] = pl[o:o + ns * nd].reshape(ns, nd).astype(np.int64)
            unreach[lbl] = un[o:o + ns * nd].reshape(ns, nd)
        if getattr(self, 'materialize_paths', False):
            astar_dets_paths = self._materialize_paths(dists, unreach, keys, yx, masks, of_frame, bits)
        else:
            astar_dets_paths = AStarPaths(dists, unreach, masks.shape[1:])
        print('Done.')
        if cache == 'to':
            self.to_cache('astar_dets_paths', astar_dets_paths)
        return astar_dets_paths

    def _materialize_paths(self, dists, unreach, keys, yx, masks, of_frame, bits, batch=32768):
        """``materialize_paths = True``: the cache object exactly as the reference builds it (AxonDetections.py:578-579,
        utils.py:381-384) -- a plain dict of nested lists whose leaves are ``scipy.sparse.coo_matrix(bool, (H, W))`` of
        the path pixels or ``None`` -- so that the STOCK reference can ``from_cache('astar_dets_paths')`` it without
        this package on its path.  The pixels come from K4 (``axb_trace_paths``): a minimum-cost path of exactly the
        reference's length (which of the equal-cost paths is heap-order dependent in the reference)."""
        from scipy.sparse import coo_matrix
        name, shape = self.dataset.name, tuple(masks.shape[1:])
        max_px = int(self.max_px_assoc_dist)
        out, req, where = {}, [], []
        for (t, tb) in keys:
            lbl = f'{name}_t:{t:0>3}-t:{tb:0>3}'
            un = unreach[lbl]
            out[lbl] = [[None] * un.shape[1] for _ in range(un.shape[0])]
            ii, jj = np.nonzero(un == 0)
            for i, j in zip(ii.tolist(), jj.tolist()):
                req.append((int(of_frame[t]), *yx[tb][i], *yx[t][j]))
                where.append((lbl, i, j))
        for b0 in range(0, len(req), batch):
            r = np.asarray(req[b0:b0 + batch], np.int32).reshape(-1, 5)
            pyx, ln = ops.trace_paths(bits, shape[1], r, max_px)
            ops.raise_status(device=bits.device)
            pyx, ln = pyx.cpu().numpy(), ln.cpu().numpy()
            for k, (lbl, i, j) in enumerate(where[b0:b0 + batch]):
                L = int(ln[k])
                assert L == int(dists[lbl][i, j]), 'traced path and cost matrix disagree'
                p = pyx[k, :L]
                out[lbl][i][j] = coo_matrix((np.ones(L), (p[:, 0], p[:, 1])), shape, bool)      # utils.py:381-382
        return out

    # ---- get_frame_and_truedets: AxonDetections.py:355-376 ----------------------------------------------
    def get_frame_and_truedets(self, t, unstitched=False):
        """The stitched frame comes straight from the dense ``dataset.X`` (K6 ``axb_stitched_frame``: the frame with the
        tiles that are empty at every timepoint zeroed) instead of tiles -> ``stitch_tiles``; the true detections
        are the reference's (``stitch_tiles`` without an image).  ``unstitched=True`` and datasets without a dense
        ``X`` take the reference's path."""
        ds = self.dataset
        if unstitched or getattr(ds, 'X', None) is None or not torch.cuda.is_available() or not getattr(self, 'frames_from_dense', True):
            return super().get_frame_and_truedets(t, unstitched)
        tiler = getattr(self, '_frame_tiler', None)
        if tiler is None:
            from .tiling import FrameTiler
            dev = torch.device('cuda', torch.cuda.current_device())
            X = ds.X.to_dense() if ds.X.is_sparse else ds.X
            tiler = self._frame_tiler = FrameTiler(X.to(dev, torch.float32), ds.tilesize, ds.temporal_context,
                                                   ds.use_motion_filtered, ds.timepoints_indices)
        pd_tiled_true_det = self._yolo_Y2pandas_det(self._yolo_targets[t], conf_thr=1)
        pd_frame_true_det, _ = ds.stitch_tiles(pd_tiled_true_det)
        return tiler.stitched_frame(t).cpu(), pd_frame_true_det

    # ---- _assign_IDs_to_detections: AxonDetections.py:631-715 ---------------------------------------------
    def _assign_IDs_to_detections(self, cache=None):
        """The reference's method, unchanged -- but the ``MinCostFlowTracker`` it instantiates (a module global of the
        reference, AxonDetections.py:11,663) is libmot_bridge's subclass for the duration of the call: same nodes,
        same ``AddArc...`` calls in the same order, the transition arcs taken from arrays (K3 / K7) instead of the
        per-pair Python loop (offline_mincostflow.py:166-198).  ``arcs_from_arrays = False`` restores the loop."""
        ref_fn = super()._assign_IDs_to_detections
        if cache == 'from' or not getattr(self, 'arcs_from_arrays', True):
            return ref_fn(cache)
        from .libmot_bridge import make_arc_tracker
        g = ref_fn.__func__.__globals__
        Ref = g['MinCostFlowTracker']
        g['MinCostFlowTracker'] = make_arc_tracker(Ref)
        try:
            return ref_fn(cache)
        finally:
            g['MinCostFlowTracker'] = Ref

    # ---- _get_astar_path_distances: AxonDetections.py:717-752 ---------------------------------------
    def _get_astar_path_distances(self, astar_paths):
        if isinstance(astar_paths, AStarPaths):
            # None -> max_px_assoc_dist is already in the matrix (AxonDetections.py:736)
            return {k: v.copy() for k, v in astar_paths.dists.items()}
        return super()._get_astar_path_distances(astar_paths)


def make_drop_in(reference_cls):
    """-> subclass of the reference's AxonDetections with the hot path on the B200."""
    return type('AxonDetections', (B200AssociationMixin, reference_cls), {'__doc__': reference_cls.__doc__})
