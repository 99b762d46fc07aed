"""The association hot path of a whole timelapse on one GPU, device resident.

``associate_frames`` = what AxonDetections.detect_dataset's loop body does after the model
forward (decode + stitch + NMS, /root/reference/axtrack/AxonDetections.py:122-128) for ALL frames
in one launch, followed by AxonDetections._compute_detections_astar_paths +
_get_astar_path_distances (AxonDetections.py:526-585, 717-752) for ALL (t, t_bef) keys in one
launch.  PyTorch only owns the buffers and the stream.
"""
import numpy as np
import torch

from . import _lib, ops


def frame_pair_keys(T, max_num_misses=1, t0=0):
    """[(t, t_bef)] in the reference's loop order (AxonDetections.py:553-566):
    for t: for t_bef in t-1 .. t-(max_num_misses+1), skipping t_bef < 0."""
    keys = []
    for t in range(t0, T):
        for tb in range(t - 1, t - (max_num_misses + 2), -1):
            if tb >= 0:
                keys.append((t, tb))
    return keys


def build_pairs(n_out, cap, keys, mask_of_frame=None):
    """Pair descriptors for detections stored as [F, cap, 2]: frame f's rows start at f*cap.
    Vectorised: this runs on the host between K1 and K2 while the GPU waits."""
    k = np.asarray(keys, np.int64).reshape(-1, 2)
    t, tb = k[:, 0], k[:, 1]
    n_out = np.asarray(n_out, np.int64)
    arr = np.zeros(len(k), dtype=_lib.PAIR_DTYPE)
    ns, nd = n_out[tb], n_out[t]
    arr['mask_idx'] = 0 if mask_of_frame is None else np.asarray(mask_of_frame, np.int64)[t]
    arr['src_off'], arr['n_src'] = tb * cap, ns
    arr['dst_off'], arr['n_dst'] = t * cap, nd
    arr['time_gap'] = t - tb
    sizes = ns * nd
    arr['out_off'] = np.cumsum(sizes) - sizes
    return arr, int(sizes.sum()), int(ns.sum())


class Association:
    """Device-resident result of ``associate_frames``."""

    def __init__(self, det, n_out, keys, pairs, path_len, unreachable, wcost, launches):
        self.det = det                # dict of CUDA tensors from ops.decode_nms
        self.n_out = n_out            # numpy int32 [T]
        self.keys = keys              # [(t, t_bef)]
        self.pairs = pairs            # numpy struct array (axb_pair_desc)
        self.path_len = path_len      # CUDA int32 flat
        self.unreachable = unreachable
        self.wcost = wcost
        self.launches = launches
        self._host = None

    def frame_dets(self, t):
        """numpy: conf (desc), x, y, keep_idx of frame t's NMS survivors."""
        n = int(self.n_out[t])
        yx = self.det['yx'][t, :n].cpu().numpy()
        return dict(conf=self.det['conf'][t, :n].cpu().numpy(), x=yx[:, 1].astype(np.int64),
                    y=yx[:, 0].astype(np.int64), keep_idx=self.det['keep_idx'][t, :n].cpu().numpy().astype(np.int64))

    def transition_costs(self, max_px=500, miss_rate=0.6, max_num_misses=1, cost_threshold=None):
        """K3: (cost float64 CUDA flat, arc int64 CUDA flat | None) for every pair, laid out like
        ``path_len`` (mincostflow_models.py:100-119, offline_mincostflow.py:192-198; weight 0)."""
        from .mincostflow_models import transition_tables
        ct, at = transition_tables(max_px, miss_rate, max_num_misses, 0.0, cost_threshold)
        return ops.transition_cost_batch(self.path_len, self.pairs, ct, at)

    def pair_matrices(self):
        """{(t, t_bef): (path_len [n_pred,n_cur] int32, unreachable uint8, wcost float32|None)} on the host."""
        if self._host is None:
            pl = self.path_len.cpu().numpy()
            un = self.unreachable.cpu().numpy()
            wc = self.wcost.cpu().numpy() if self.wcost is not None else None
            out = {}
            for k, pd in zip(self.keys, self.pairs):
                o, ns, nd = int(pd['out_off']), int(pd['n_src']), int(pd['n_dst'])
                sl = slice(o, o + ns * nd)
                out[k] = (pl[sl].reshape(ns, nd), un[sl].reshape(ns, nd),
                          wc[sl].reshape(ns, nd) if wc is not None else None)
            self._host = out
        return self._host


_pinned = {}


def _pinned_counts(T, device):
    key = (torch.device(device).index, int(T))
    if key not in _pinned:
        _pinned[key] = torch.empty(int(T), dtype=torch.int32).pin_memory()
    return _pinned[key]


def associate_frames(Y, tile_yx, mask, max_px=500, max_num_misses=1, tilesize=512, conf_thr=0.55,
                     nms_dist=23, cap=None, mask_of_frame=None, want_cost=True, want_pre=False,
                     mask_bits=None, t0=0, out_alloc=None):
    """Y float32 CUDA [T, ntiles, Sx, Sy, 3]; tile_yx int32 CUDA [ntiles, 2]; mask uint8/bool CUDA
    [H, W] (static, the reference stacks one 2-D mask: Timelapse.py:213-214) or [n_masks, H, W] with
    ``mask_of_frame`` [T].  ``t0`` > 0: frames < t0 are halo frames (sources only), used by the
    timepoint sharding.  ``out_alloc(total, pairs)`` -> (path_len, unreachable, wcost|None) CUDA tensors
    the kernels write into (sharding.PeerSlabs: rank 0's HBM over NVLink).  Returns an ``Association``."""
    ctx = ops.get_context(Y.device)
    l0 = ctx.launches
    det = ops.decode_nms(Y, tile_yx, tilesize, conf_thr, nms_dist, cap, want_pre=want_pre)
    T = Y.shape[0]
    cap = det['conf'].shape[1]
    if mask_bits is None:
        mask_bits = ops.pack_mask(mask)
    W = mask.shape[-1]
    # the one host sync of the path: T int32 counts, through pinned memory.  The tiles / components of the masks are
    # queued right after it, so the GPU derives them while the host builds the pair list and its device copy.
    pin = _pinned_counts(T, Y.device)
    pin.copy_(det['n_out'], non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(Y.device))
    ev.synchronize()
    n_out = pin.numpy().copy()
    ops.raise_status(device=Y.device)           # (synchronises the stream: before anything else is queued on it)
    if T >= 3:                                  # >= 2 frame pairs
        ops.prepare_masks(mask_bits, W)
    try:
        keys = frame_pair_keys(T, max_num_misses, t0)
        pairs, total, _ = build_pairs(n_out, cap, keys, mask_of_frame)
        out = out_alloc(total, pairs) if out_alloc is not None else None
        pl, un, wc, _ = ops.cost_matrix_batch(mask_bits, W, det['yx'].view(-1, 2), pairs, max_px, want_cost, out=out)
    except BaseException:
        ops.discard_prepared_masks(Y.device)
        raise
    return Association(det, n_out, keys, pairs, pl, un, wc, ctx.launches - l0)
