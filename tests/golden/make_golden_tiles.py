"""Golden vectors for the tiling front end (K6), made by the UNMODIFIED reference:
Timelapse.construct_tiles (Timelapse.py:492-561) and Timelapse.get_frametiles_stack (:150-157) run on a
Timelapse that skips __init__ (no tif on disk) and carries a small synthetic X.  Build container only:

    python tests/golden/make_golden_tiles.py

Stored: the inputs (X, sizes, mode) and the reference's tile_info[..., 0], X_tiled shape, the compact
tile coordinates from flat_tile_idx2yx_tile_idx and a CRC32 + a strided sample of every requested
get_frametiles_stack output (the full stacks would be tens of MB)."""
import os
import sys
import zlib

import numpy as np
import pandas as pd
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

CASES = [  # seed, T, H, W, tilesize, temporal_context, use_motion_filtered
    (0, 7, 100, 150, 64, 2, 'exclude'),
    (1, 6, 128, 96, 32, 1, 'include'),
    (2, 5, 70, 130, 48, 0, 'only'),
    (3, 9, 90, 200, 64, 2, 'temp_context'),
]


def make_X(seed, T, H, W, ts):
    """Sparse-ish image stack: some tiles empty at all t, some only at some t; channel 1/2 = motion."""
    rng = np.random.default_rng(seed)
    X = rng.random((T, 3, H, W), dtype=np.float32)
    X[:, 0][rng.random((T, H, W)) < 0.7] = 0
    yt, xt = -(-H // ts), -(-W // ts)
    for ty in range(yt):
        for tx in range(xt):
            u = rng.random()
            sl = (slice(ty * ts, (ty + 1) * ts), slice(tx * ts, (tx + 1) * ts))
            if u < 0.25:
                X[:, 0][(slice(None),) + sl] = 0                      # empty at every t
            elif u < 0.5:
                X[rng.integers(0, T), 0][sl] = 0                      # empty at one t
            elif u < 0.6:
                X[:, 0][(slice(None),) + sl] *= -1                    # negative values are "empty" too
    return X


def run_reference(X, ts, tc, mode):
    ax = ref_harness.load()
    TL = ax.axtrack.Timelapse.Timelapse
    T, _, H, W = X.shape
    tl = TL.__new__(TL)
    tl.name = 'golden'
    tl.X = torch.from_numpy(X).to_sparse()
    tl.transform_configs = None
    tl.tilesize, tl.sizey, tl.sizex, tl.sizet = ts, H, W, T - 2 * tc
    tl.ytiles, tl.xtiles = -(-H // ts), -(-W // ts)
    tl.Sy = tl.Sx = 12
    tl.temporal_context, tl.use_motion_filtered = tc, mode
    tl.timepoints_indices = list(range(tc, T - tc))
    cols = pd.MultiIndex.from_product([['Axon_000'], ['anchor_x', 'anchor_y']])
    tl.target = pd.DataFrame(np.full((T, 2), np.nan), columns=cols)
    tl.target.iloc[tc] = [3, 5]
    tl.construct_tiles('cpu')
    ntiles = tl.X_tiled.shape[1]
    tile_yx = np.array([tl.flat_tile_idx2yx_tile_idx(i) for i in range(ntiles)], np.int32).reshape(-1, 2)
    stacks = [tl.get_frametiles_stack(t)[0].numpy() for t in range(tl.sizet)]
    # the frame AxonDetections.get_frame_and_truedets hands to the tracker: tiles -> Timelapse.stitch_tiles
    # (Timelapse.py:166-197; AxonDetections.py:370-372).  _get_channelsizes has no 'temp_context' branch.
    frames = None
    if mode != 'temp_context':
        empty = [pd.DataFrame({'conf': [], 'anchor_x': [], 'anchor_y': []}) for _ in range(ntiles)]
        frames = [tl.stitch_tiles(empty, torch.from_numpy(st))[1].numpy() for st in stacks]
    return tl.tile_info[..., 0].numpy().astype(bool), tuple(tl.X_tiled.shape), tile_yx, stacks, frames


def main():
    for seed, T, H, W, ts, tc, mode in CASES:
        X = make_X(seed, T, H, W, ts)
        nonempty, xt_shape, tile_yx, stacks, frames = run_reference(X, ts, tc, mode)
        st = np.stack(stacks)
        extra = {}
        if frames is not None:
            fr = np.stack(frames)
            extra = dict(frame_shape=np.array(fr.shape), frame_crc=np.array([zlib.crc32(np.ascontiguousarray(f).tobytes()) for f in frames], np.uint32),
                         frame_sample=fr.reshape(-1)[::89].copy())
        np.savez_compressed(os.path.join(HERE, f'tiles_s{seed}.npz'), seed=seed, T=T, H=H, W=W, tilesize=ts,
                            temporal_context=tc, mode=mode, nonempty=nonempty, x_tiled_shape=np.array(xt_shape),
                            tile_yx=tile_yx, stack_shape=np.array(st.shape),
                            stack_crc=np.array([zlib.crc32(s.tobytes()) for s in stacks], np.uint32),
                            stack_sample=st.reshape(-1)[::97].copy(), **extra)
        print(seed, mode, 'tiles', tile_yx.shape[0], 'of', nonempty.shape[:2], 'stack', st.shape)


if __name__ == '__main__':
    main()
