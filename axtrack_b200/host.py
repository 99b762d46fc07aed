"""numpy-in / numpy-out binding of the HOST-buffer entry points of the C ABI
(axb_decode_nms_host, axb_cost_matrix_batch_host).  No torch: this is what a reference-side
ctypes stub would call and what bench.py times as ``e2e`` -- host<->device
copies happen inside the library call.
"""
import ctypes

import numpy as np

from . import _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def frame_pair_keys(T, max_num_misses=1, t0=0):
    """[(t, t_bef)] in the reference's loop order (AxonDetections.py:553-566)."""
    return [(t, tb) for t in range(t0, T) for tb in range(t - 1, t - (max_num_misses + 2), -1) if tb >= 0]


class HostAssociator:
    def __init__(self, device=0):
        self.ctx = _lib.Context(device)
        self.lib = self.ctx.lib

    @property
    def launches(self):
        return self.ctx.launches

    def _raise(self, rc, pairs=None, dets=None):
        if rc == _lib.OK:
            return
        if rc in (_lib.ERR_OOB_START, _lib.ERR_OOB_GOAL, _lib.ERR_BAD_WEIGHT):
            raise ValueError(self.lib.axb_last_error(self.ctx.handle).decode() or f'axb error {rc}')
        self.ctx.check(rc)

    def decode_nms(self, Y, tile_yx, tilesize=512, conf_thr=0.55, nms_dist=23, cap=None, want_pre=False, out=None):
        """Y float32 [F, ntiles, Sx, Sy, 3] (host) -> dict of numpy arrays (see axb_decode_nms)."""
        Y = np.ascontiguousarray(Y, dtype=np.float32)
        if Y.ndim == 4:
            Y = Y[None]
        tile_yx = np.ascontiguousarray(tile_yx, dtype=np.int32)
        F, ntiles, Sx, Sy, _ = Y.shape
        cap = int(cap or ntiles * Sx * Sy)
        o = out
        if o is None:
            o = dict(n_pre=np.zeros(F, np.int32), n_out=np.zeros(F, np.int32), conf=np.empty((F, cap), np.float32),
                     yx=np.empty((F, cap, 2), np.int32), keep_idx=np.empty((F, cap), np.int32))
            if want_pre:
                o.update(pre_conf=np.empty((F, cap), np.float32), pre_yx=np.empty((F, cap, 2), np.int32),
                         pre_cell=np.empty((F, cap), np.int32))
        rc = self.lib.axb_decode_nms_host(
            self.ctx.handle, _ptr(Y), _ptr(tile_yx), F, ntiles, Sx, Sy, int(tilesize),
            ctypes.c_float(float(np.float32(conf_thr))), int(nms_dist), cap,
            _ptr(o.get('pre_conf')), _ptr(o.get('pre_yx')), _ptr(o.get('pre_cell')), _ptr(o['n_pre']),
            _ptr(o['conf']), _ptr(o['yx']), _ptr(o['keep_idx']), _ptr(o['n_out']))
        self._raise(rc)
        return o

    def cost_matrices_flat(self, masks, dets_yx, pairs, total, max_px=500, want_cost=True, out=None):
        """masks uint8 [n_masks,H,W] host; dets_yx int32 [n,2] host; pairs numpy struct array."""
        masks = np.ascontiguousarray(masks, dtype=np.uint8)
        if masks.ndim == 2:
            masks = masks[None]
        dets_yx = np.ascontiguousarray(dets_yx, dtype=np.int32).reshape(-1, 2)
        n_masks, H, W = masks.shape
        if out is None:
            out = (np.empty(max(total, 1), np.int32), np.empty(max(total, 1), np.uint8),
                   np.empty(max(total, 1), np.float32) if want_cost else None)
        pl, un, wc = out
        rc = self.lib.axb_cost_matrix_batch_host(
            self.ctx.handle, _ptr(masks), n_masks, H, W, _ptr(dets_yx), dets_yx.shape[0],
            _ptr(pairs), len(pairs), int(max_px), _ptr(pl), _ptr(un), _ptr(wc))
        self._raise(rc)
        return pl, un, wc

    def cost_matrices(self, masks, frame_dets_yx, max_px=500, max_num_misses=1, mask_of_frame=None, want_cost=True):
        """frame_dets_yx: list[T] of int32 [n_t,2] (y, x).  -> {(t, t_bef): (path_len, unreachable, wcost)}."""
        from .ops_pairs import make_pairs
        T = len(frame_dets_yx)
        offs = np.concatenate([[0], np.cumsum([len(d) for d in frame_dets_yx])]).astype(np.int64)
        dets = (np.concatenate([np.asarray(d, np.int32).reshape(-1, 2) for d in frame_dets_yx])
                if T else np.zeros((0, 2), np.int32))
        keys = frame_pair_keys(T, max_num_misses)
        pairs, total, _ = make_pairs([(0 if mask_of_frame is None else int(mask_of_frame[t]), int(offs[tb]),
                                       len(frame_dets_yx[tb]), int(offs[t]), len(frame_dets_yx[t]), t - tb) for t, tb in keys])
        pl, un, wc = self.cost_matrices_flat(masks, dets, pairs, total, max_px, want_cost)
        res = {}
        for k, pd in zip(keys, pairs):
            o, ns, nd = int(pd['out_off']), int(pd['n_src']), int(pd['n_dst'])
            sl = slice(o, o + ns * nd)
            res[k] = (pl[sl].reshape(ns, nd), un[sl].reshape(ns, nd), wc[sl].reshape(ns, nd) if wc is not None else None)
        return res
