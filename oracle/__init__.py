"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the AxTrack association hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker or as the timed
CPU baseline.  The product (``axtrack_b200``) never imports it.

Contents
--------
* ctypes bindings of ``oracle/liboracle.so`` (``astar_oracle.cpp``): the forked pyastar2d
  search restated (``astar_weights``, ``astar_mask``), the reference's pair loop
  (``cost_matrix_astar``) and the closed form the CUDA kernel implements (``cost_matrix_rule``).
* ``ref_astar_path`` / ``ref_cost_matrix``: the reference's OWN compiled pyastar2d fork
  (``oracle/_ref/astar*.so``, built by ``oracle/Makefile`` from
  /root/reference/pyastar2d/src/cpp/astar.cpp) driven the way the reference drives it
  (astar_wrapper.py:25-54, AxonDetections.py:553-579,600-629).
* ``decode_stitch_nms``: numpy restatement of ``_yolo_Y2pandas_det`` + ``stitch_tiles`` +
  ``_non_max_supression`` (AxonDetections.py:178-278, Timelapse.py:141-148,166-197).
* ``transition_cost``: ``transition_model`` with ``vis_sim_weight`` handled as in the
  reference (mincostflow_models.py:67-119) for the deployed ``MCF_VIS_SIM_WEIGHT = 0``.

Parity status: pinned (tests/test_oracle_golden.py, tests/golden/).
"""
import ctypes
import importlib.machinery
import importlib.util
import glob
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from itertools import repeat

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

ERR_OOB_START = -2
ERR_OOB_GOAL = -3


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(_HERE, 'liboracle.so')
    src = os.path.join(_HERE, 'astar_oracle.cpp')
    need = force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src)
    have_ref = bool(glob.glob(os.path.join(_HERE, '_ref', 'astar*.so')))
    if need or (not have_ref and os.path.exists('/root/reference/pyastar2d/src/cpp/astar.cpp')):
        subprocess.run(['make', '-C', _HERE, '-B' if force else '-s'], check=True,
                       stdout=subprocess.DEVNULL)


def lib():
    global _LIB
    if _LIB is None:
        build()
        L = ctypes.CDLL(os.path.join(_HERE, 'liboracle.so'))
        i32p = np.ctypeslib.ndpointer(np.int32, flags='C_CONTIGUOUS')
        u8p = np.ctypeslib.ndpointer(np.uint8, flags='C_CONTIGUOUS')
        f32p = np.ctypeslib.ndpointer(np.float32, flags='C_CONTIGUOUS')
        c_int = ctypes.c_int
        L.axo_astar_weights.restype = c_int
        L.axo_astar_weights.argtypes = [f32p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                        ctypes.POINTER(ctypes.c_float), ctypes.c_void_p]
        L.axo_astar_mask.restype = c_int
        L.axo_astar_mask.argtypes = [u8p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                     ctypes.POINTER(ctypes.c_float)]
        for fn in (L.axo_cost_matrix_astar, L.axo_cost_matrix_rule):
            fn.restype = c_int
            fn.argtypes = [u8p, c_int, c_int, i32p, c_int, i32p, c_int, c_int,
                           i32p, u8p, f32p, ctypes.POINTER(ctypes.c_int64)]
        L.axo_cost_matrix_rule_ex.restype = c_int
        L.axo_cost_matrix_rule_ex.argtypes = [u8p, c_int, c_int, i32p, c_int, i32p, c_int, c_int,
                                              i32p, u8p, f32p, ctypes.POINTER(ctypes.c_int64), u8p]
        _LIB = L
    return _LIB


# ----------------------------------------------------------------------------------------
# restated search
# ----------------------------------------------------------------------------------------
def astar_weights(weights, start, goal, max_path_length=10000, return_cost=False):
    """Forked pyastar2d.astar_path restated (astar_wrapper.py:25-54 + astar.cpp:46-172).

    Returns int32 [L,2] (row, col) or None, raising the wrapper's errors."""
    assert weights.dtype == np.float32, (
        f"weights must have np.float32 data type, but has {weights.dtype}")
    if weights.min(axis=None) < 1.:
        raise ValueError("Minimum cost to move must be 1, but got %f" % (weights.min(axis=None)))
    H, W = weights.shape
    if start[0] < 0 or start[0] >= H or start[1] < 0 or start[1] >= W:
        raise ValueError(f"Start of {start} lies outside grid.")
    if goal[0] < 0 or goal[0] >= H or goal[1] < 0 or goal[1] >= W:
        raise ValueError(f"Goal of {goal} lies outside grid.")
    w = np.ascontiguousarray(weights)
    cap = int(min(max_path_length, H * W))
    path = np.empty((max(cap, 1), 2), np.int32)
    cost = ctypes.c_float(0)
    L = lib().axo_astar_weights(w.reshape(-1), H, W, int(start[0]), int(start[1]), int(goal[0]),
                                int(goal[1]), int(max_path_length), ctypes.byref(cost),
                                path.ctypes.data_as(ctypes.c_void_p))
    if L < 0:
        return (None, None) if return_cost else None
    return (path[:L].copy(), cost.value) if return_cost else path[:L].copy()


def _prep(mask, src_yx, dst_yx):
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    src = np.ascontiguousarray(np.asarray(src_yx, dtype=np.int32).reshape(-1, 2))
    dst = np.ascontiguousarray(np.asarray(dst_yx, dtype=np.int32).reshape(-1, 2))
    n = max(src.shape[0] * dst.shape[0], 1)
    return mask, src, dst, np.empty(n, np.int32), np.empty(n, np.uint8), np.empty(n, np.float32)


def _raise(code, pair, src, dst):
    i, j = divmod(int(pair), max(dst.shape[0], 1))
    if code == ERR_OOB_START:
        raise ValueError(f"Start of {(int(src[i, 0]), int(src[i, 1]))} lies outside grid.")
    if code == ERR_OOB_GOAL:
        raise ValueError(f"Goal of {(int(dst[j, 0]), int(dst[j, 1]))} lies outside grid.")
    raise RuntimeError(f'oracle error {code}')


def _cost_matrix(fn, mask, src_yx, dst_yx, max_px):
    mask, src, dst, pl, un, wc = _prep(mask, src_yx, dst_yx)
    err = ctypes.c_int64(-1)
    rc = fn(mask.reshape(-1), mask.shape[0], mask.shape[1], src.reshape(-1), src.shape[0],
            dst.reshape(-1), dst.shape[0], int(max_px), pl, un, wc, ctypes.byref(err))
    if rc != 0:
        _raise(rc, err.value, src, dst)
    shp = (src.shape[0], dst.shape[0])
    k = shp[0] * shp[1]
    return pl[:k].reshape(shp), un[:k].reshape(shp), wc[:k].reshape(shp)


def cost_matrix_astar(mask, src_yx, dst_yx, max_px=500):
    """Per-pair restated A* (the reference's loop).  -> (path_len i32, unreachable u8, wcost f32)."""
    return _cost_matrix(lib().axo_cost_matrix_astar, mask, src_yx, dst_yx, max_px)


def cost_matrix_rule(mask, src_yx, dst_yx, max_px=500):
    """Closed form (one bounded search per source).  Same outputs as cost_matrix_astar."""
    return _cost_matrix(lib().axo_cost_matrix_rule, mask, src_yx, dst_yx, max_px)


def cost_matrix_rule_ex(mask, src_yx, dst_yx, max_px=500):
    """cost_matrix_rule plus ``tie_regime`` uint8 [n_src, n_dst]: 1 where the reference's answer depends
    on the pop order of equal-priority heap entries (see astar_oracle.cpp); the rule is exact elsewhere."""
    mask, src, dst, pl, un, wc = _prep(mask, src_yx, dst_yx)
    tie = np.zeros(pl.shape[0], np.uint8)
    err = ctypes.c_int64(-1)
    rc = lib().axo_cost_matrix_rule_ex(mask.reshape(-1), mask.shape[0], mask.shape[1], src.reshape(-1), src.shape[0],
                                       dst.reshape(-1), dst.shape[0], int(max_px), pl, un, wc, ctypes.byref(err), tie)
    if rc != 0:
        _raise(rc, err.value, src, dst)
    shp = (src.shape[0], dst.shape[0])
    k = shp[0] * shp[1]
    return pl[:k].reshape(shp), un[:k].reshape(shp), wc[:k].reshape(shp), tie[:k].reshape(shp)


# ----------------------------------------------------------------------------------------
# the reference's own compiled search (oracle/_ref)
# ----------------------------------------------------------------------------------------
def ref_available():
    return bool(glob.glob(os.path.join(_HERE, '_ref', 'astar*.so')))


def ref_module():
    """Import oracle/_ref/astar*.so -- the reference's CPython extension, unmodified."""
    global _REF
    if _REF is None:
        build()
        cands = glob.glob(os.path.join(_HERE, '_ref', 'astar*.so'))
        if not cands:
            raise ImportError('oracle/_ref/astar*.so not built (needs /root/reference at build time)')
        loader = importlib.machinery.ExtensionFileLoader('astar', cands[0])
        spec = importlib.util.spec_from_loader('astar', loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
        _REF = mod
    return _REF


def ref_astar_path(weights, start, goal, max_path_length=10000):
    """What pyastar2d.astar_path does around the compiled call (astar_wrapper.py:31-53):
    dtype assert, O(H*W) min scan, bounds checks, ravel, O(H*W) flatten copy, native call."""
    assert weights.dtype == np.float32, (
        f"weights must have np.float32 data type, but has {weights.dtype}")
    if weights.min(axis=None) < 1.:
        raise ValueError("Minimum cost to move must be 1, but got %f" % (weights.min(axis=None)))
    if (start[0] < 0 or start[0] >= weights.shape[0] or
            start[1] < 0 or start[1] >= weights.shape[1]):
        raise ValueError(f"Start of {start} lies outside grid.")
    if (goal[0] < 0 or goal[0] >= weights.shape[0] or
            goal[1] < 0 or goal[1] >= weights.shape[1]):
        raise ValueError(f"Goal of {goal} lies outside grid.")
    height, width = weights.shape
    s = int(np.ravel_multi_index(start, (height, width)))
    g = int(np.ravel_multi_index(goal, (height, width)))
    return ref_module().astar(weights.flatten(), height, width, s, g, int(max_path_length))


def mask_weights(mask):
    """AxonDetections._get_maskweights (AxonDetections.py:598)."""
    return np.where(np.asarray(mask) == 1, 1, 2 ** 16).astype(np.float32)


def _ref_pair(src, dst, weights, max_px):
    # _get_path_between_dets (AxonDetections.py:600-629): dets are (conf, x, y)
    sy, sx = src
    ty, tx = dst
    dist = np.sqrt((sy - ty) ** 2 + (sx - tx) ** 2)
    if dist < max_px:
        p = ref_astar_path(weights, (int(sy), int(sx)), (int(ty), int(tx)), max_px)
        return None if p is None else p.shape[0]
    return None


def ref_cost_matrix(mask, src_yx, dst_yx, max_px=500, threads=True):
    """The reference's source loop with its ThreadPoolExecutor over targets
    (AxonDetections.py:570-579) on the compiled reference search; leaves mapped as in
    _get_astar_path_distances (AxonDetections.py:736).  -> (path_len i64, unreachable u8)."""
    weights = mask_weights(mask)
    src = np.asarray(src_yx, dtype=np.int64).reshape(-1, 2)
    dst = np.asarray(dst_yx, dtype=np.int64).reshape(-1, 2)
    rows = []
    for s in src:
        if threads:
            with ThreadPoolExecutor() as ex:
                row = list(ex.map(_ref_pair, repeat(s), dst, repeat(weights), repeat(max_px)))
        else:
            row = [_ref_pair(s, d, weights, max_px) for d in dst]
        rows.append(row)
    pl = np.array([[max_px if v is None else v for v in r] for r in rows], dtype=np.int64).reshape(len(src), len(dst))
    un = np.array([[v is None for v in r] for r in rows], dtype=np.uint8).reshape(len(src), len(dst))
    return pl, un


# ----------------------------------------------------------------------------------------
# decode + stitch + NMS
# ----------------------------------------------------------------------------------------
def decode_stitch_nms(Y, tile_yx, tilesize=512, conf_thr=0.55, nms_dist=23):
    """numpy restatement of the per-frame detection path.

    Y       float32 [ntiles, Sx, Sy, 3] = (conf, x, y) per grid cell (model.py:119-125)
    tile_yx int     [ntiles, 2]         = (tile_y, tile_x) of each compact tile
                                          (Timelapse.flat_tile_idx2yx_tile_idx, Timelapse.py:141-148)
    Returns dict with
      pre_conf/pre_x/pre_y  the stitched pre-NMS frame detections in reference order
                            (tile-major; inside a tile conf ascending; AxonDetections.py:235,
                            Timelapse.py:191-196)
      conf/x/y              NMS survivors, conf descending (AxonDetections.py:261-277)
      keep_idx              their pre-NMS indices ("survivor indices")
    Ties in conf: cell order inside a tile (stable), and (conf desc, pre-NMS index asc) for
    NMS -- the reference's pandas quicksort leaves tie order implementation-defined
    (SURVEY Appendix B.6); generators avoid ties.
    """
    Y = np.asarray(Y, dtype=np.float32)
    ntiles, Sx, Sy, _ = Y.shape
    ts = np.float32(tilesize)
    ix = np.arange(Sx, dtype=np.float32).reshape(1, Sx, 1)
    iy = np.arange(Sy, dtype=np.float32).reshape(1, 1, Sy)
    noanchor = (Y == 0).all(-1)
    # separate f32 add, mul, div, round-half-even (AxonDetections.py:203-207)
    x = np.rint(((Y[..., 1] + ix).astype(np.float32) * ts).astype(np.float32) / np.float32(Sx))
    y = np.rint(((Y[..., 2] + iy).astype(np.float32) * ts).astype(np.float32) / np.float32(Sy))
    conf = Y[..., 0].copy()
    x[noanchor] = 0
    y[noanchor] = 0
    conf[noanchor] = 0
    thr = np.float32(conf_thr)
    pre_c, pre_x, pre_y = [], [], []
    for t in range(ntiles):
        c = conf[t].reshape(-1)
        keep = np.nonzero(c >= thr)[0]                       # cell order ix*Sy+iy
        order = keep[np.argsort(c[keep], kind='stable')]      # conf ascending
        pre_c.append(c[order])
        pre_x.append(x[t].reshape(-1)[order].astype(np.int64) + int(tile_yx[t][1]) * tilesize)
        pre_y.append(y[t].reshape(-1)[order].astype(np.int64) + int(tile_yx[t][0]) * tilesize)
    pre_c = np.concatenate(pre_c) if pre_c else np.zeros(0, np.float32)
    pre_x = np.concatenate(pre_x) if pre_x else np.zeros(0, np.int64)
    pre_y = np.concatenate(pre_y) if pre_y else np.zeros(0, np.int64)
    n = pre_c.shape[0]
    order = np.argsort(-pre_c.astype(np.float64), kind='stable')   # conf desc, idx asc on ties
    alive = np.ones(n, bool)
    keep_idx = []
    d2max = int(nms_dist) * int(nms_dist)
    for i in order:
        if not alive[i]:
            continue
        keep_idx.append(i)
        d2 = (pre_x - pre_x[i]) ** 2 + (pre_y - pre_y[i]) ** 2
        alive &= ~(d2 < d2max)
    keep_idx = np.array(keep_idx, dtype=np.int64)
    return dict(pre_conf=pre_c, pre_x=pre_x, pre_y=pre_y, keep_idx=keep_idx,
                conf=pre_c[keep_idx], x=pre_x[keep_idx], y=pre_y[keep_idx])


# ----------------------------------------------------------------------------------------
# transition_model (consumer; the "final cost matrix")
# ----------------------------------------------------------------------------------------
def transition_cost(path_len, max_px=500, miss_rate=0.6, time_gap=1, vis_sim_weight=0.0, vis_sim=None):
    """mincostflow_models.py:100-119 for given A* distances (float64)."""
    distances = ((np.asarray(path_len) / max_px) - 1) * -1
    inf_dist = distances == 0
    if vis_sim is None:
        vis_sim = np.zeros_like(distances)
    costs = -np.log((1 - vis_sim_weight) * distances * (miss_rate ** (time_gap - 1))
                    + vis_sim_weight * vis_sim + 1e-6)
    costs[inf_dist] = np.inf
    return costs


def bhattacharyya(h1, h2):
    """cv2.compareHist(h1, h2, cv2.HISTCMP_BHATTACHARYYA) for float32 histograms, restated: OpenCV 4.x sums
    s1 = sum h1, s2 = sum h2, r = sum sqrt(h1*h2) in double in the order of its baseline SSE build (four floats
    per step: the low pair then the high pair added to a two-lane accumulator, lanes added 0 + lane0 + lane1, the
    len % 4 tail one by one) and returns sqrt(max(1 - r / sqrt(s1*s2), 0)), with 1 / sqrt(s1*s2) := 1 when
    |s1*s2| <= FLT_EPSILON.  Pinned bit for bit against cv2 4.13 (tests/test_vis_sim.py)."""
    a = np.asarray(h1, np.float32).reshape(-1).astype(np.float64)
    b = np.asarray(h2, np.float32).reshape(-1).astype(np.float64)
    n = len(a)
    body = n & ~3

    def lanes(v):
        l0 = l1 = 0.0
        for k in range(0, body, 2):
            l0 += float(v[k]); l1 += float(v[k + 1])
        t = (0.0 + l0) + l1
        for k in range(body, n):
            t += float(v[k])
        return t
    with np.errstate(invalid='ignore'):
        r, s1, s2 = lanes(np.sqrt(a * b)), lanes(a), lanes(b)     # a normalised bin can be -1e-8: sqrt -> NaN, as in cv2
    s = s1 * s2
    s = 1.0 / np.sqrt(s) if abs(s) > float(np.finfo(np.float32).eps) else 1.0
    x = 1.0 - r * s
    with np.errstate(invalid='ignore'):
        return float(np.sqrt(0.0 if x < 0.0 else x))          # std::max(x, 0.): a NaN stays a NaN (-> nan_to_num: 0)


def vis_sim(feats_src, feats_dst):
    """mincostflow_models.py:106-113: nan_to_num(1 - Bhattacharyya distance) of every (predecessor, current) pair."""
    out = np.array([[1 - bhattacharyya(f1, f2) for f2 in feats_dst] for f1 in feats_src], dtype=np.float64)
    return np.nan_to_num(out.reshape(len(feats_src), len(feats_dst)))


def feature_hist(image, boxes, bins=180, lo=0.0, hi=1.0):
    """mincostflow_models.py:30-64 (feature_model) without cv2: the crop of each box (x, y, w, h),
    cv2.calcHist over `bins` uniform bins of [lo, hi) -- idx = floor(double(v)*a + b), a = bins/(hi-lo),
    b = -lo*a -- and cv2.normalize(NORM_MINMAX, 0, 1): scale = float(1/(max-min)) (0 when max-min <=
    DBL_EPSILON), shift = 0.f - float(min*scale), feature = fma(count, scale, shift) rounded once to
    float32 (OpenCV 4.x; pinned against cv2 4.13 by tests/golden/feature_hist_*.npz).  -> float32 [n, bins]."""
    image = np.asarray(image, np.float32)
    boxes = np.atleast_2d(np.asarray(boxes)).astype(np.int64)
    a = bins / (float(hi) - float(lo))
    b = -float(lo) * a
    out = np.zeros((len(boxes), bins), np.float32)
    for i, roi in enumerate(boxes):
        x1, x2 = max(int(roi[1]), 0), max(int(roi[0]), 0)
        x3, x4 = max(x1 + 1, x1 + int(roi[3])), max(x2 + 1, x2 + int(roi[2]))
        v = image[x1:x3, x2:x4].astype(np.float64).ravel()
        with np.errstate(invalid='ignore'):
            idx = np.floor(v * a + b)
            idx[np.isnan(v)] = 0          # observed: cv2.calcHist counts a NaN pixel in bin 0 (any range)
            ok = (idx >= 0) & (idx < bins)
        cnt = np.bincount(idx[ok].astype(np.int64), minlength=bins).astype(np.float32)
        smin, smax = float(cnt.min()), float(cnt.max())
        scale = float(np.float32(1.0 / (smax - smin) if smax - smin > np.finfo(np.float64).eps else 0.0))
        shift = float(np.float32(0.0) - np.float32(smin * scale))
        out[i] = (cnt.astype(np.float64) * scale + shift).astype(np.float32)   # exact product + one rounding = fma
    return out


MOTION_CHANNELS = {'exclude': (0, 1), 'temp_context': (0, 1), 'only': (1, 2), 'include': (0, 3)}


def tile_frames(X, tilesize=512):
    """The image half of Timelapse.construct_tiles (Timelapse.py:508-525, 539-558), numpy.
    X float32 [T, C, H, W] -> (X_tiled [T, ntiles, C, ts, ts] of the tiles non-empty at any t, row-major;
    nonempty bool [ytiles, xtiles, T] = tile_info[..., 0]; tile_yx int32 [ntiles, 2])."""
    X = np.asarray(X, np.float32)
    T, C, H, W = X.shape
    ts = int(tilesize)
    yt, xt = -(-H // ts), -(-W // ts)
    pad = np.zeros((T, C, yt * ts, xt * ts), np.float32)          # ZeroPad2d right / bottom (:521-525)
    pad[:, :, :H, :W] = X
    tiles = pad.reshape(T, C, yt, ts, xt, ts).transpose(2, 4, 0, 1, 3, 5)   # ytile, xtile, t, c, y, x (:543)
    nonempty = (tiles[:, :, :, 0] > 0).any(-1).any(-1)            # :550
    keep = nonempty.any(-1)                                       # :558
    X_tiled = tiles[keep].swapaxes(0, 1)
    tile_yx = np.argwhere(keep).astype(np.int32)
    return np.ascontiguousarray(X_tiled), nonempty, tile_yx


def frametiles_stack(X_tiled, frame, temporal_context=2, use_motion_filtered='exclude'):
    """Timelapse.get_frametiles_stack (Timelapse.py:150-157) over __getitem__ (:111-125) for the raw
    frame index ``frame`` = timepoints_indices[t]: float32 [ntiles, (2tc+1)*nch, ts, ts]."""
    lo, n = MOTION_CHANNELS[use_motion_filtered]
    a, b = frame - temporal_context, frame + temporal_context + 1
    assert a >= 0 and b <= X_tiled.shape[0], "a frame's temporal context leaves the timelapse"
    st = X_tiled[a:b, :, lo:lo + n]                               # [2tc+1, ntiles, nch, ts, ts]
    st = st.transpose(1, 0, 2, 3, 4)
    return np.ascontiguousarray(st.reshape(st.shape[0], -1, *st.shape[3:]))


def stitched_frame(X, nonempty, frame, tilesize=512, use_motion_filtered='exclude'):
    """AxonDetections.get_frame_and_truedets(t)[0] (AxonDetections.py:355-376 -> Timelapse.stitch_tiles image half,
    Timelapse.py:173-188) restated on the dense X: the colour channels of raw frame ``frame`` with every tile that is
    empty at ALL timepoints (dropped from X_tiled, Timelapse.py:558) zeroed.  -> float32 [ncol, H, W]."""
    X = np.asarray(X, np.float32)
    lo, n = MOTION_CHANNELS[use_motion_filtered]
    keep = np.asarray(nonempty).any(-1)
    ts = int(tilesize)
    H, W = X.shape[2:]
    up = np.repeat(np.repeat(keep, ts, 0), ts, 1)[:H, :W]
    return np.where(up[None], X[frame, lo:lo + n], np.float32(0)).astype(np.float32)
