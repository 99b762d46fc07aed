"""torch-facing wrappers of the C ABI: tensors in, tensors out, on torch's current stream.

PyTorch is used for device memory and streams only; all arithmetic happens in
libaxtrack_b200.so.  Every function here requires CUDA tensors and fails loudly otherwise.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .ops_pairs import make_pairs  # noqa: F401  (re-exported)

_contexts = {}


def get_context(device=None):
    """The cached axb context of a CUDA device (default: torch's current device)."""
    if not torch.cuda.is_available():
        raise RuntimeError('axtrack_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback')
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    if idx is None:
        idx = torch.cuda.current_device()
    if idx not in _contexts:
        _contexts[idx] = _lib.Context(idx)
    return _contexts[idx]


def _stream(device=None):
    """torch's current stream ON THE TENSORS' DEVICE (not on torch's current device: the C ABI runs on
    the ctx's device and a stream of another device is an invalid handle there)."""
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _need_cuda(t, name, dtype=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f'{name} must be a CUDA tensor')
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f'{name} must have dtype {dtype}, got {t.dtype}')
    return t.contiguous()


def mask_words(W):
    return (int(W) + 31) // 32


def pack_mask(mask):
    """uint8/bool CUDA tensor [H,W] or [n,H,W] (value 1 = on mask) -> int32 [n,H,ceil(W/32)].

    Replaces AxonDetections._get_maskweights (AxonDetections.py:587-598)."""
    if mask.dtype == torch.bool:
        mask = mask.view(torch.uint8)
    mask = _need_cuda(mask, 'mask', torch.uint8)
    if mask.dim() == 2:
        mask = mask.unsqueeze(0)
    n, H, W = mask.shape
    ctx = get_context(mask.device)
    bits = torch.empty((n, H, mask_words(W)), dtype=torch.int32, device=mask.device)
    ctx.check(ctx.lib.axb_pack_mask(ctx.handle, mask.data_ptr(), n, H, W, bits.data_ptr(), _stream(mask.device)))
    return bits


def prepare_masks(mask_bits, W):
    """Optional: tile words and components of ``mask_bits`` (from pack_mask) ahead of the next ``cost_matrix_batch`` on
    this device, which then skips that work -- call it before a host synchronisation the pair list depends on."""
    mask_bits = _need_cuda(mask_bits, 'mask_bits', torch.int32)
    n_masks, H, _ = mask_bits.shape
    ctx = get_context(mask_bits.device)
    ctx.check(ctx.lib.axb_prepare_masks(ctx.handle, mask_bits.data_ptr(), n_masks, H, int(W), _stream(mask_bits.device)))


def discard_prepared_masks(device=None):
    """Drop what ``prepare_masks`` left for the next ``cost_matrix_batch`` (when that call is not going to happen)."""
    ctx = get_context(device)
    ctx.check(ctx.lib.axb_prepare_masks(ctx.handle, None, 0, 0, 0, None))


def pack_weights(weights):
    """float32 CUDA weight grid(s) as pyastar2d.astar_path takes them -> packed mask bits.
    Weight errors (min < 1: astar_wrapper.py:35-37) surface in ``raise_status``."""
    weights = _need_cuda(weights, 'weights', torch.float32)
    if weights.dim() == 2:
        weights = weights.unsqueeze(0)
    n, H, W = weights.shape
    ctx = get_context(weights.device)
    bits = torch.empty((n, H, mask_words(W)), dtype=torch.int32, device=weights.device)
    ctx.check(ctx.lib.axb_pack_weights(ctx.handle, weights.data_ptr(), n, H, W, bits.data_ptr(), _stream(weights.device)))
    return bits


def cost_matrix_batch(mask_bits, W, dets_yx, pairs, max_px=500, want_cost=True, out=None, pairs_dev=None):
    """Path-length cost matrices of many frame pairs in one launch.

    mask_bits int32 CUDA [n_masks, H, Wp] from pack_mask; dets_yx int32 CUDA [n, 2] rows (y, x);
    pairs = numpy struct array from make_pairs (or the entries themselves).
    Returns (path_len int32 flat, unreachable uint8 flat, wcost float32 flat | None, pairs);
    pair p's block is [out_off : out_off + n_src*n_dst] reshaped [n_src, n_dst].
    Replaces AxonDetections._compute_detections_astar_paths + _get_astar_path_distances
    (AxonDetections.py:526-585, 717-752)."""
    mask_bits = _need_cuda(mask_bits, 'mask_bits', torch.int32)
    dets_yx = _need_cuda(dets_yx, 'dets_yx', torch.int32)
    if not isinstance(pairs, np.ndarray):
        pairs, _, _ = make_pairs(pairs)
    n_pairs = len(pairs)
    total = int((pairs['out_off'] + pairs['n_src'].astype(np.int64) * pairs['n_dst']).max()) if n_pairs else 0
    total_src = int(pairs['n_src'].sum()) if n_pairs else 0
    dev = mask_bits.device
    ctx = get_context(dev)
    if out is None:
        path_len = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
        unreach = torch.empty(max(total, 1), dtype=torch.uint8, device=dev)
        wcost = torch.empty(max(total, 1), dtype=torch.float32, device=dev) if want_cost else None
    else:
        path_len, unreach, wcost = out
    if n_pairs and total:
        if pairs_dev is None:
            pairs_dev = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev, non_blocking=False)
        n_masks, H, _ = mask_bits.shape
        cap_out = min(int(path_len.numel()), int(unreach.numel()), int(wcost.numel()) if wcost is not None else 1 << 62)
        ctx.check(ctx.lib.axb_cost_matrix_batch(
            ctx.handle, mask_bits.data_ptr(), n_masks, H, int(W), dets_yx.data_ptr(), int(dets_yx.shape[0]),
            pairs_dev.data_ptr(), n_pairs, total_src, int(max_px),
            path_len.data_ptr(), unreach.data_ptr(), wcost.data_ptr() if wcost is not None else None,
            cap_out, _stream(dev)))
    else:
        discard_prepared_masks(dev)      # nothing to search: products of prepare_masks must not outlive this call
    return path_len[:total], unreach[:total], (wcost[:total] if wcost is not None else None), pairs


def cost_matrix(mask_bits, W, src_yx, dst_yx, max_px=500, want_cost=True):
    """One frame pair: -> (path_len [n_src,n_dst] int32, unreachable uint8, wcost float32|None)."""
    mask_bits = _need_cuda(mask_bits, 'mask_bits', torch.int32)
    if mask_bits.dim() == 3:
        assert mask_bits.shape[0] == 1
        mask_bits = mask_bits[0]
    src_yx = _need_cuda(src_yx, 'src_yx', torch.int32).reshape(-1, 2)
    dst_yx = _need_cuda(dst_yx, 'dst_yx', torch.int32).reshape(-1, 2)
    ns, nd = src_yx.shape[0], dst_yx.shape[0]
    dev = mask_bits.device
    ctx = get_context(dev)
    path_len = torch.empty((ns, nd), dtype=torch.int32, device=dev)
    unreach = torch.empty((ns, nd), dtype=torch.uint8, device=dev)
    wcost = torch.empty((ns, nd), dtype=torch.float32, device=dev) if want_cost else None
    if ns and nd:
        ctx.check(ctx.lib.axb_cost_matrix(
            ctx.handle, mask_bits.data_ptr(), mask_bits.shape[0], int(W), src_yx.data_ptr(), ns,
            dst_yx.data_ptr(), nd, int(max_px), path_len.data_ptr(), unreach.data_ptr(),
            wcost.data_ptr() if wcost is not None else None, _stream(dev)))
    return path_len, unreach, wcost


def decode_nms(Y, tile_yx, tilesize=512, conf_thr=0.55, nms_dist=23, cap=None, want_pre=True):
    """Fused decode + threshold + stitch + NMS of F frames.

    Y float32 CUDA [F, ntiles, Sx, Sy, 3]; tile_yx int32 CUDA [ntiles, 2].
    Returns dict of CUDA tensors: n_pre [F], pre_conf [F,cap], pre_yx [F,cap,2] (y, x),
    pre_cell [F,cap] (tile*Sx*Sy + ix*Sy + iy), n_out [F], conf [F,cap], yx [F,cap,2] (y, x), keep_idx [F,cap].
    Replaces _yolo_Y2pandas_det + stitch_tiles + _non_max_supression
    (AxonDetections.py:178-278, Timelapse.py:166-197)."""
    Y = _need_cuda(Y, 'Y', torch.float32)
    if Y.dim() == 4:
        Y = Y.unsqueeze(0)
    tile_yx = _need_cuda(tile_yx, 'tile_yx', torch.int32)
    F, ntiles, Sx, Sy, C = Y.shape
    assert C == 3 and tile_yx.shape == (ntiles, 2)
    if cap is None:
        cap = ntiles * Sx * Sy
    dev = Y.device
    ctx = get_context(dev)
    o = dict(n_pre=torch.zeros(F, dtype=torch.int32, device=dev),
             n_out=torch.zeros(F, dtype=torch.int32, device=dev),
             conf=torch.empty((F, cap), dtype=torch.float32, device=dev),
             yx=torch.empty((F, cap, 2), dtype=torch.int32, device=dev),
             keep_idx=torch.empty((F, cap), dtype=torch.int32, device=dev))
    if want_pre:
        o.update(pre_conf=torch.empty((F, cap), dtype=torch.float32, device=dev),
                 pre_yx=torch.empty((F, cap, 2), dtype=torch.int32, device=dev),
                 pre_cell=torch.empty((F, cap), dtype=torch.int32, device=dev))
    ptr = lambda k: o[k].data_ptr() if k in o else None
    ctx.check(ctx.lib.axb_decode_nms(
        ctx.handle, Y.data_ptr(), tile_yx.data_ptr(), F, ntiles, Sx, Sy, int(tilesize),
        ctypes.c_float(float(np.float32(conf_thr))), int(nms_dist), int(cap),
        ptr('pre_conf'), ptr('pre_yx'), ptr('pre_cell'), o['n_pre'].data_ptr(),
        o['conf'].data_ptr(), o['yx'].data_ptr(), o['keep_idx'].data_ptr(), o['n_out'].data_ptr(),
        _stream(dev)))
    return o


def raise_status(pairs=None, dets_yx_host=None, device=None):
    """Synchronise and raise the reference's exception for the first input error, if any
    (astar_wrapper.py:35-45: ValueError for weights < 1, out-of-grid start / goal)."""
    ctx = get_context(device)
    rc, (p, i, j) = ctx.status(_stream(torch.device('cuda', ctx.device)))
    if rc == _lib.OK:
        return
    if rc == _lib.ERR_BAD_WEIGHT:
        raise ValueError('Minimum cost to move must be 1, but got a smaller weight')
    if rc in (_lib.ERR_OOB_START, _lib.ERR_OOB_GOAL) and pairs is not None and dets_yx_host is not None and p >= 0:
        d = np.asarray(dets_yx_host).reshape(-1, 2)
        if rc == _lib.ERR_OOB_START:
            y, x = d[int(pairs['src_off'][p]) + i]
            raise ValueError(f'Start of {(int(y), int(x))} lies outside grid.')
        y, x = d[int(pairs['dst_off'][p]) + j]
        raise ValueError(f'Goal of {(int(y), int(x))} lies outside grid.')
    if rc == _lib.ERR_OOB_START:
        raise ValueError('Start lies outside grid.')
    if rc == _lib.ERR_OOB_GOAL:
        raise ValueError('Goal lies outside grid.')
    msg = ctx.lib.axb_last_error(ctx.handle).decode() if rc == _lib.ERR_BAD_ARG else ''
    raise _lib.AxbError(rc, msg or 'device-side status')


def transition_cost_batch(path_len, pairs, cost_tab, arc_tab=None, pairs_dev=None):
    """Transition costs of all frame pairs in one launch (K3).

    path_len int32 CUDA flat (output of ``cost_matrix_batch`` for the same ``pairs``); pairs numpy
    struct array with ``time_gap`` filled; cost_tab float64 [n_gaps, max_px+1] / arc_tab int64 (host
    numpy or CUDA) from ``mincostflow_models.transition_tables``.  -> (cost float64 CUDA flat,
    arc int64 CUDA flat | None).  Replaces mincostflow_models.py:100-119 and
    offline_mincostflow.py:192-198 (vis_sim_weight = 0)."""
    path_len = _need_cuda(path_len, 'path_len', torch.int32)
    dev = path_len.device
    ctx = get_context(dev)
    cost_tab = torch.as_tensor(np.ascontiguousarray(cost_tab) if isinstance(cost_tab, np.ndarray) else cost_tab,
                               dtype=torch.float64).to(dev).contiguous()
    n_gaps, n_d = cost_tab.shape
    if arc_tab is not None:
        arc_tab = torch.as_tensor(np.ascontiguousarray(arc_tab) if isinstance(arc_tab, np.ndarray) else arc_tab,
                                  dtype=torch.int64).to(dev).contiguous()
        assert arc_tab.shape == cost_tab.shape
    n_pairs = len(pairs)
    total = int(path_len.numel())
    cost = torch.empty(total, dtype=torch.float64, device=dev)
    arc = torch.empty(total, dtype=torch.int64, device=dev) if arc_tab is not None else None
    if n_pairs and total:
        if pairs_dev is None:
            pairs_dev = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev)
        max_elems = int((pairs['n_src'].astype(np.int64) * pairs['n_dst']).max())
        ctx.check(ctx.lib.axb_transition_cost_batch(
            ctx.handle, path_len.data_ptr(), pairs_dev.data_ptr(), n_pairs, max_elems,
            cost_tab.data_ptr(), arc_tab.data_ptr() if arc_tab is not None else None, int(n_gaps), int(n_d - 1),
            cost.data_ptr(), arc.data_ptr() if arc is not None else None, _stream(dev)))
    return cost, arc


def trace_paths(mask_bits, W, requests, max_px=500):
    """Path PIXELS of selected (source, target) pairs (K4).

    mask_bits int32 CUDA [n_masks, H, Wp]; requests int32 [n, 5] rows (mask_idx, sy, sx, ty, tx)
    (numpy).  -> (yx int32 CUDA [n, max_px, 2] (row, col) start -> goal, length int32 CUDA [n]; 0 = the
    reference's None).  Replaces utils._compute_astar_path -> pyastar2d.astar_path
    (utils.py:351-390) for the pairs whose pixels are wanted."""
    mask_bits = _need_cuda(mask_bits, 'mask_bits', torch.int32)
    if mask_bits.dim() == 2:
        mask_bits = mask_bits.unsqueeze(0)
    req = np.ascontiguousarray(np.asarray(requests, np.int32).reshape(-1, 5))
    n = len(req)
    dev = mask_bits.device
    ctx = get_context(dev)
    yx = torch.zeros((n, int(max_px), 2), dtype=torch.int32, device=dev)
    ln = torch.zeros(n, dtype=torch.int32, device=dev)
    if n:
        n_masks, H, _ = mask_bits.shape
        ctx.check(ctx.lib.axb_trace_paths(ctx.handle, mask_bits.data_ptr(), n_masks, H, int(W), req.ctypes.data, n,
                                          int(max_px), yx.data_ptr(), ln.data_ptr(), _stream(dev)))
    return yx, ln


def feature_hist(image, boxes, bins=180, lo=0.0, hi=1.0):
    """Appearance features of the boxes of one frame (K5).

    image float32 CUDA [H, W]; boxes int32 CUDA [n, 4] rows (x, y, w, h).  -> float32 CUDA [n, bins]:
    cv2.calcHist([crop], [0], None, [bins], [lo, hi]) min-max normalised, exactly as
    mincostflow_models.feature_model computes it (mincostflow_models.py:30-64)."""
    image = _need_cuda(image, 'image', torch.float32)
    assert image.dim() == 2, 'image must be one [H, W] frame'
    boxes = _need_cuda(boxes.reshape(-1, 4), 'boxes', torch.int32)
    n = boxes.shape[0]
    dev = image.device
    ctx = get_context(dev)
    out = torch.empty((n, int(bins)), dtype=torch.float32, device=dev)
    if n:
        H, W = image.shape
        ctx.check(ctx.lib.axb_feature_hist(ctx.handle, image.data_ptr(), int(H), int(W), boxes.data_ptr(), n,
                                           int(bins), float(lo), float(hi), out.data_ptr(), _stream(dev)))
    return out


def feature_hist_frames(images, boxes, box_frame, bins=180, lo=0.0, hi=1.0):
    """K5 for the boxes of a BLOCK of frames in one launch: images float32 CUDA [F, H, W]; boxes int32
    CUDA [n, 4] rows (x, y, w, h); box_frame int32 CUDA [n] = frame of each box.  -> float32 CUDA [n, bins]."""
    images = _need_cuda(images, 'images', torch.float32)
    assert images.dim() == 3, 'images must be [F, H, W]'
    boxes = _need_cuda(boxes.reshape(-1, 4), 'boxes', torch.int32)
    box_frame = _need_cuda(box_frame.reshape(-1), 'box_frame', torch.int32)
    n = boxes.shape[0]
    assert box_frame.shape[0] == n
    dev = images.device
    ctx = get_context(dev)
    out = torch.empty((n, int(bins)), dtype=torch.float32, device=dev)
    if n:
        F, H, W = images.shape
        ctx.check(ctx.lib.axb_feature_hist_frames(ctx.handle, images.data_ptr(), int(F), int(H), int(W), boxes.data_ptr(),
                                                  box_frame.data_ptr(), n, int(bins), float(lo), float(hi), out.data_ptr(), _stream(dev)))
    return out


def vis_sim(feats_src, feats_dst, path_len=None, max_px=500):
    """K7: ``nan_to_num(1 - cv2.compareHist(f1, f2, HISTCMP_BHATTACHARYYA))`` for every (predecessor, current)
    pair of feature histograms (mincostflow_models.py:106-113), bit-identical to cv2 4.13.

    feats_* float32 CUDA [n, bins] (or [n, bins, 1] as feature_model returns them); path_len optional int32 CUDA
    [n_src, n_dst]: entries equal to max_px are skipped (their cost is inf anyway).  -> float64 CUDA [n_src, n_dst]."""
    fs = _need_cuda(feats_src.flatten(1) if feats_src.dim() > 2 else feats_src, 'feats_src', torch.float32)
    fd = _need_cuda(feats_dst.flatten(1) if feats_dst.dim() > 2 else feats_dst, 'feats_dst', torch.float32)
    assert fs.shape[1] == fd.shape[1]
    ns, nd, bins = fs.shape[0], fd.shape[0], fs.shape[1]
    dev = fs.device
    ctx = get_context(dev)
    vis = torch.zeros((ns, nd), dtype=torch.float64, device=dev)
    if path_len is not None:
        path_len = _need_cuda(path_len.reshape(ns, nd), 'path_len', torch.int32)
    if ns and nd:
        ctx.check(ctx.lib.axb_vis_sim(ctx.handle, fs.data_ptr(), ns, fd.data_ptr(), nd, int(bins),
                                      path_len.data_ptr() if path_len is not None else None, int(max_px), vis.data_ptr(), _stream(dev)))
    return vis


#: USE_MOTION_DATA -> (first channel, number of channels) of X (Timelapse.py:116-124, 415-424)
MOTION_CHANNELS = {'exclude': (0, 1), 'temp_context': (0, 1), 'only': (1, 2), 'include': (0, 3)}


def tile_nonempty(X, tilesize=512):
    """Non-empty tile flags of a dense timelapse (K6).

    X float32 CUDA [T, C, H, W] -> bool CUDA [ytiles, xtiles, T] = ``tile_info[..., 0]`` of
    Timelapse.construct_tiles (Timelapse.py:551)."""
    X = _need_cuda(X, 'X', torch.float32)
    assert X.dim() == 4, 'X must be [T, C, H, W]'
    T, C, H, W = X.shape
    yt, xt = -(-H // tilesize), -(-W // tilesize)
    ctx = get_context(X.device)
    out = torch.empty((yt * xt, T), dtype=torch.uint8, device=X.device)
    ctx.check(ctx.lib.axb_tile_nonempty(ctx.handle, X.data_ptr(), int(T), int(C), int(H), int(W), int(tilesize),
                                        out.data_ptr(), _stream(X.device)))
    return out.view(yt, xt, T).bool()


def compact_tiles(nonempty):
    """int32 CUDA [ntiles, 2] rows (tile_y, tile_x) of the tiles that are non-empty at any t, row-major:
    the order of X_tiled's tile axis (Timelapse.py:558) and of flat_tile_idx2yx_tile_idx (:141-148)."""
    return torch.nonzero(nonempty.any(-1)).to(torch.int32).contiguous()


def gather_frametiles(X, tile_yx, frames, tilesize=512, temporal_context=2, use_motion_filtered='exclude'):
    """Model input of a batch of frames (K6): Timelapse.get_frametiles_stack (Timelapse.py:150-157) for
    every raw frame index in ``frames`` at once.

    X float32 CUDA [T, C, H, W]; tile_yx int32 CUDA [ntiles, 2]; frames int32 [B] (CUDA tensor, or any
    sequence of raw frame indices = ``timepoints_indices[t]``, Timelapse.py:113).
    -> float32 CUDA [B, ntiles, (2 * temporal_context + 1) * nch, tilesize, tilesize]."""
    X = _need_cuda(X, 'X', torch.float32)
    assert X.dim() == 4, 'X must be [T, C, H, W]'
    tile_yx = _need_cuda(tile_yx.reshape(-1, 2), 'tile_yx', torch.int32)
    if not isinstance(frames, torch.Tensor):
        frames = torch.as_tensor(np.asarray(frames, np.int32).reshape(-1)).to(X.device)
    frames = _need_cuda(frames.reshape(-1), 'frames', torch.int32)
    lo, nch = MOTION_CHANNELS[use_motion_filtered]
    T, C, H, W = X.shape
    B, nt = frames.shape[0], tile_yx.shape[0]
    ctx = get_context(X.device)
    out = torch.empty((B, nt, (2 * int(temporal_context) + 1) * nch, int(tilesize), int(tilesize)),
                      dtype=torch.float32, device=X.device)
    ctx.check(ctx.lib.axb_gather_frametiles(ctx.handle, X.data_ptr(), int(T), int(C), int(H), int(W), int(tilesize),
                                            tile_yx.data_ptr(), int(nt), frames.data_ptr(), int(B),
                                            int(temporal_context), int(lo), int(nch), out.data_ptr(), _stream(X.device)))
    return out


def stitched_frame(X, nonempty, frame, tilesize=512, use_motion_filtered='exclude'):
    """The frame of AxonDetections.get_frame_and_truedets(t)[0] (AxonDetections.py:355-376) straight from the dense
    timelapse: X float32 CUDA [T, C, H, W]; nonempty bool CUDA [ytiles, xtiles, T] (``tile_nonempty``); frame = raw
    frame index.  -> float32 CUDA [nch, H, W]: the frame with the tiles that are empty at every t zeroed, which is
    what Timelapse.stitch_tiles (Timelapse.py:173-188) rebuilds from the kept tiles."""
    X = _need_cuda(X, 'X', torch.float32)
    T, C, H, W = X.shape
    keep = nonempty.any(-1).to(torch.uint8).contiguous()
    lo, nch = MOTION_CHANNELS[use_motion_filtered]
    if use_motion_filtered == 'temp_context':
        raise ValueError("Timelapse._get_channelsizes has no 'temp_context' branch (Timelapse.py:415-424)")
    ctx = get_context(X.device)
    out = torch.empty((nch, H, W), dtype=torch.float32, device=X.device)
    ctx.check(ctx.lib.axb_stitched_frame(ctx.handle, X.data_ptr(), int(T), int(C), int(H), int(W), int(tilesize), keep.data_ptr(),
                                         int(frame), int(lo), int(nch), out.data_ptr(), _stream(X.device)))
    return out
