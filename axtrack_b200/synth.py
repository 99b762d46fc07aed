"""Seeded synthetic inputs for the association hot path.

Pure numpy; shared by bench.py and tests/.  Nothing here touches the GPU or the oracle.

* ``make_mask``       one connected channel network (the reference's masks are the flood fill of
                      a segmented micro-channel structure, data_prep_nbs/00_segment_bg.ipynb).
* ``make_tracks``     growth cones random-walking ON the mask, a few jittered <= 3 px off it.
* ``make_head_tensor`` the detector-head tensor [ntiles, Sx, Sy, 3] whose decode reproduces the
                      given detections (inverse of AxonDetections._yolo_Y2pandas_det,
                      /root/reference/axtrack/AxonDetections.py:192-210) plus sub-threshold
                      background cells and near-duplicate distractors for NMS.
* ``tile_table``      (tile_y, tile_x) of the compact non-empty tiles in row-major order
                      (Timelapse.flat_tile_idx2yx_tile_idx, Timelapse.py:141-148).
"""
import numpy as np


def _largest_component_from(mask, seed_yx):
    """4-connected component of ``mask`` containing ``seed_yx`` (the flood fill of the
    reference's mask segmentation notebook)."""
    from scipy import ndimage
    lab, _ = ndimage.label(mask, structure=[[0, 1, 0], [1, 1, 1], [0, 1, 0]])
    if lab[seed_yx] == 0:
        return np.zeros_like(mask)
    return lab == lab[seed_yx]


def make_mask(H, W, seed=0, pitch=(150, 300), width=(8, 20), n_diag=4):
    """bool [H,W]: axis-aligned corridors on a jittered pitch plus a few diagonals, reduced to
    the connected component of one seed pixel.  Foreground fraction lands near 5-15 %."""
    rng = np.random.default_rng(seed)
    m = np.zeros((H, W), bool)
    ys, y = [], int(rng.integers(20, max(21, min(pitch[0], H - 1))))
    while y < H - 2:
        w = int(rng.integers(width[0], width[1] + 1))
        m[y:min(H, y + w), :] = True
        ys.append(y + w // 2)
        y += int(rng.integers(pitch[0], pitch[1] + 1))
    xs, x = [], int(rng.integers(20, max(21, min(pitch[0], W - 1))))
    while x < W - 2:
        w = int(rng.integers(width[0], width[1] + 1))
        m[:, x:min(W, x + w)] = True
        xs.append(x + w // 2)
        x += int(rng.integers(pitch[0], pitch[1] + 1))
    yy, xx = np.mgrid[0:H, 0:W]
    for _ in range(n_diag):
        y0, x0 = int(rng.integers(0, H)), int(rng.integers(0, W))
        sgn = 1 if rng.random() < 0.5 else -1
        w = int(rng.integers(width[0], width[1] + 1))
        m |= np.abs((yy - y0) - sgn * (xx - x0)) < w
    # punch a few gaps so that some corridors dead-end
    for _ in range(max(2, (H * W) // (512 * 512))):
        y0, x0 = int(rng.integers(0, H)), int(rng.integers(0, W))
        m[max(0, y0 - 15):y0 + 15, max(0, x0 - 15):x0 + 15] = False
    if not ys or not xs:
        return m
    sy, sx = ys[len(ys) // 2], xs[len(xs) // 2]
    if not m[sy, sx]:
        cand = np.argwhere(m)
        sy, sx = map(int, cand[len(cand) // 2])
    return _largest_component_from(m, (sy, sx))


def make_tracks(mask, n, T, seed=0, step=40, min_sep=23, off_mask_frac=0.03, birth_death=0.02):
    """list[T] of int32 [n_t, 2] (y, x): growth cones random-walking on the mask.

    Per frame the detections are pairwise >= ``min_sep`` px apart (they are NMS survivors),
    in-grid, and a fraction is jittered <= 3 px off the mask."""
    rng = np.random.default_rng(seed + 1000)
    H, W = mask.shape
    fg = np.argwhere(mask)
    assert len(fg) > 0
    pos = fg[rng.integers(0, len(fg), n)].astype(np.int64)
    frames = []
    for t in range(T):
        if t > 0:
            for i in range(len(pos)):
                for _ in range(8):
                    cand = pos[i] + rng.integers(-step, step + 1, 2)
                    if 0 <= cand[0] < H and 0 <= cand[1] < W and mask[cand[0], cand[1]]:
                        pos[i] = cand
                        break
            kill = rng.random(len(pos)) < birth_death
            pos[kill] = fg[rng.integers(0, len(fg), int(kill.sum()))]
        p = pos.copy()
        jit = rng.random(len(p)) < off_mask_frac
        p[jit] += rng.integers(-3, 4, (int(jit.sum()), 2))
        p[:, 0] = np.clip(p[:, 0], 0, H - 1)
        p[:, 1] = np.clip(p[:, 1], 0, W - 1)
        # enforce the NMS separation greedily (grid hash)
        keep, cell = [], {}
        for i, (y, x) in enumerate(p):
            cy, cx = int(y) // min_sep, int(x) // min_sep
            ok = True
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    for k in cell.get((cy + dy, cx + dx), ()):
                        if (p[k, 0] - y) ** 2 + (p[k, 1] - x) ** 2 < min_sep * min_sep:
                            ok = False
            if ok:
                keep.append(i)
                cell.setdefault((cy, cx), []).append(i)
        frames.append(p[keep].astype(np.int32))
    return frames


def tile_table(H, W, tilesize=512):
    """int32 [ntiles,2] (tile_y, tile_x), all tiles non-empty, row-major."""
    yt, xt = -(-H // tilesize), -(-W // tilesize)
    ty, tx = np.divmod(np.arange(yt * xt), xt)
    return np.stack([ty, tx], 1).astype(np.int32)


def make_head_tensor(dets_yx, H, W, seed=0, tilesize=512, Sx=12, Sy=12, conf_thr=0.55,
                     dup_frac=0.2, nms_dist=23, bg_hi=0.5):
    """float32 [ntiles, Sx, Sy, 3] detector-head output for one frame.

    Every detection lands in its grid cell (ix, iy) = (floor(x_t*Sx/ts), floor(y_t*Sy/ts)) with
    in-cell offsets chosen so the reference's f32 decode (add, mul, div, rint) reproduces the
    integer pixel; when two detections share a cell only the first is kept (one anchor per cell
    is a model limit).  ``dup_frac`` of the detections get a lower-confidence distractor in a
    neighbouring cell less than ``nms_dist`` px away.  Background cells carry conf ~ U(0, bg_hi).
    Confidences are tie-free.  Returns (Y, placed_yx, placed_conf)."""
    rng = np.random.default_rng(seed + 2000)
    tt = tile_table(H, W, tilesize)
    ntiles = len(tt)
    xt = -(-W // tilesize)
    Y = np.zeros((ntiles, Sx, Sy, 3), np.float32)
    Y[..., 0] = rng.uniform(0.01, bg_hi, (ntiles, Sx, Sy)).astype(np.float32)
    Y[..., 1:] = rng.uniform(0.05, 0.95, (ntiles, Sx, Sy, 2)).astype(np.float32)
    used = np.zeros((ntiles, Sx, Sy), bool)
    n = len(dets_yx)
    # tie-free confidences: a shuffled strictly increasing ladder in [thr+0.01, 1.15]
    ladder = np.linspace(conf_thr + 0.01, 1.15, 2 * n + 2, dtype=np.float64)[1:-1]
    ladder = rng.permutation(ladder).astype(np.float32)
    assert len(np.unique(ladder)) == len(ladder)
    ts = np.float32(tilesize)

    def place(y, x, conf):
        ty, tx = int(y) // tilesize, int(x) // tilesize
        if ty * xt + tx >= ntiles:
            return False
        t = ty * xt + tx
        yl, xl = int(y) - ty * tilesize, int(x) - tx * tilesize
        ix, iy = (xl * Sx) // tilesize, (yl * Sy) // tilesize
        if used[t, ix, iy]:
            return False
        # offset so that rint(((off+i)*ts)/S) == pixel, checked in f32 like the reference
        ox = np.float32(xl * Sx / tilesize - ix)
        oy = np.float32(yl * Sy / tilesize - iy)
        rx = np.rint(((ox + np.float32(ix)) * ts) / np.float32(Sx))
        ry = np.rint(((oy + np.float32(iy)) * ts) / np.float32(Sy))
        if int(rx) != xl or int(ry) != yl:
            return False
        used[t, ix, iy] = True
        Y[t, ix, iy] = (conf, ox, oy)
        return True

    placed, pconf = [], []
    for i, (y, x) in enumerate(np.asarray(dets_yx)):
        if place(y, x, ladder[i]):
            placed.append((int(y), int(x)))
            pconf.append(ladder[i])
    # distractors: lower conf, < nms_dist px away, in another cell
    k = n
    for (y, x), c in list(zip(placed, pconf)):
        if rng.random() >= dup_frac:
            continue
        for _ in range(6):
            dy, dx = rng.integers(-nms_dist + 6, nms_dist - 5, 2)
            if dy * dy + dx * dx >= (nms_dist - 1) ** 2 or (dy == 0 and dx == 0):
                continue
            yy, xx = y + int(dy), x + int(dx)
            if not (0 <= yy < H and 0 <= xx < W):
                continue
            lower = np.float32(conf_thr + (float(c) - conf_thr) * rng.uniform(0.05, 0.9))
            if lower < conf_thr or lower in pconf:
                continue
            if place(yy, xx, lower):
                pconf.append(lower)
                k += 1
                break
    return Y, np.array(placed, np.int32).reshape(-1, 2), np.array(pconf[:len(placed)], np.float32)


def make_timelapse(H, W, T, n_dets, seed=0):
    """Convenience bundle: (mask bool[H,W], list[T] of (y,x) int32 arrays)."""
    mask = make_mask(H, W, seed)
    return mask, make_tracks(mask, n_dets, T, seed)
