"""K6: the tiling front end = the image half of Timelapse.construct_tiles
(Timelapse.py:508-525, 539-558) + Timelapse.get_frametiles_stack (:150-157).  A copy and a comparison
with zero: the bar is BITWISE equality with the unmodified reference (tests/golden/tiles_s*.npz,
tests/golden/make_golden_tiles.py) and with the numpy oracle."""
import glob
import os
import sys
import zlib

import numpy as np
import pytest

import oracle


def _fixtures(golden_dir):
    files = sorted(glob.glob(os.path.join(golden_dir, 'tiles_s*.npz')))
    assert len(files) >= 4
    return [np.load(f) for f in files]


def _make_X(golden_dir, g):
    sys.path.insert(0, golden_dir)
    from make_golden_tiles import make_X
    return make_X(int(g['seed']), int(g['T']), int(g['H']), int(g['W']), int(g['tilesize']))


def _check_against_fixture(g, nonempty, tile_yx, stacks):
    assert np.array_equal(nonempty, g['nonempty'])
    assert np.array_equal(tile_yx, g['tile_yx'])
    assert tuple(stacks.shape) == tuple(g['stack_shape'])
    assert np.array_equal(np.array([zlib.crc32(np.ascontiguousarray(s).tobytes()) for s in stacks], np.uint32), g['stack_crc'])
    assert np.array_equal(stacks.reshape(-1)[::97], g['stack_sample'])


def test_oracle_matches_reference_fixtures(golden_dir):
    for g in _fixtures(golden_dir):
        X = _make_X(golden_dir, g)
        tc, mode, ts = int(g['temporal_context']), str(g['mode']), int(g['tilesize'])
        X_tiled, nonempty, tile_yx = oracle.tile_frames(X, ts)
        assert tuple(X_tiled.shape) == tuple(g['x_tiled_shape'])
        stacks = np.stack([oracle.frametiles_stack(X_tiled, t, tc, mode) for t in range(tc, X.shape[0] - tc)])
        _check_against_fixture(g, nonempty, tile_yx, stacks)


def _check_frames(g, frames):
    if 'frame_crc' not in g.files:
        return 0
    frames = np.asarray(frames)
    assert tuple(frames.shape) == tuple(g['frame_shape'])
    assert np.array_equal(np.array([zlib.crc32(np.ascontiguousarray(f).tobytes()) for f in frames], np.uint32), g['frame_crc'])
    assert np.array_equal(frames.reshape(-1)[::89], g['frame_sample'])
    return len(frames)


def test_oracle_stitched_frame_matches_reference_fixtures(golden_dir):
    """get_frame_and_truedets(t)[0] = tiles -> Timelapse.stitch_tiles in the reference; = the dense frame with the
    dropped tiles zeroed (oracle.stitched_frame)."""
    n = 0
    for g in _fixtures(golden_dir):
        X = _make_X(golden_dir, g)
        tc, mode, ts = int(g['temporal_context']), str(g['mode']), int(g['tilesize'])
        if mode == 'temp_context':
            continue
        _, nonempty, _ = oracle.tile_frames(X, ts)
        n += _check_frames(g, [oracle.stitched_frame(X, nonempty, t, ts, mode) for t in range(tc, X.shape[0] - tc)])
    assert n >= 10


@pytest.mark.gpu
def test_cuda_stitched_frame_matches_reference_fixtures(golden_dir):
    import torch
    from axtrack_b200.tiling import FrameTiler
    n = 0
    for g in _fixtures(golden_dir):
        X = _make_X(golden_dir, g)
        tc, mode, ts = int(g['temporal_context']), str(g['mode']), int(g['tilesize'])
        if mode == 'temp_context':
            continue
        tiler = FrameTiler(torch.from_numpy(X).cuda(), ts, tc, mode)
        n += _check_frames(g, [tiler.stitched_frame(t).cpu().numpy() for t in range(len(tiler))])
    assert n >= 10


def test_oracle_context_outside_timelapse():
    X = np.ones((3, 3, 8, 8), np.float32)
    X_tiled, _, _ = oracle.tile_frames(X, 4)
    with pytest.raises(AssertionError):
        oracle.frametiles_stack(X_tiled, 0, 1)


@pytest.mark.gpu
def test_cuda_matches_reference_fixtures_and_oracle(golden_dir):
    import torch
    from axtrack_b200 import ops
    for g in _fixtures(golden_dir):
        X = _make_X(golden_dir, g)
        tc, mode, ts = int(g['temporal_context']), str(g['mode']), int(g['tilesize'])
        Xd = torch.from_numpy(X).cuda()
        ne = ops.tile_nonempty(Xd, ts)
        tile_yx = ops.compact_tiles(ne)
        frames = list(range(tc, X.shape[0] - tc))
        st = ops.gather_frametiles(Xd, tile_yx, frames, ts, tc, mode)
        ops.raise_status(device=Xd.device)
        _check_against_fixture(g, ne.cpu().numpy(), tile_yx.cpu().numpy(), st.cpu().numpy())


@pytest.mark.gpu
@pytest.mark.parametrize('H,W,ts', [(1000, 1030, 512), (1024, 1024, 512), (130, 67, 50), (37, 41, 512), (600, 520, 512), (200, 24, 128),
                                    (1000, 1028, 256), (300, 516, 128), (1536, 2048, 512)])
def test_cuda_matches_oracle_ragged(H, W, ts):
    """Vector path (W % 4 == 0) and scalar path, images smaller than a tile, ragged borders, NaN pixels."""
    import torch
    from axtrack_b200 import ops
    rng = np.random.default_rng(H + W)
    T, tc = 5, 1
    X = rng.standard_normal((T, 3, H, W)).astype(np.float32)
    X[:, 0][rng.random((T, H, W)) < 0.9] = 0
    X[1, 0, :min(H, 60), :min(W, 60)] = np.nan          # NaN > 0 is False in the reference too
    if W > 600:
        X[:, 0, :512, 512:] = 0                          # an all-empty tile
    Xd = torch.from_numpy(X).cuda()
    X_tiled, nonempty, tile_yx = oracle.tile_frames(X, ts)
    ne = ops.tile_nonempty(Xd, ts)
    assert np.array_equal(ne.cpu().numpy(), nonempty)
    tyx = ops.compact_tiles(ne)
    assert np.array_equal(tyx.cpu().numpy(), tile_yx)
    for mode in ('exclude', 'only', 'include'):
        frames = [1, 3, 2]
        st = ops.gather_frametiles(Xd, tyx, frames, ts, tc, mode).cpu().numpy()
        want = np.stack([oracle.frametiles_stack(X_tiled, t, tc, mode) for t in frames])
        assert st.shape == want.shape
        assert np.array_equal(st.view(np.uint32), want.view(np.uint32))      # bitwise, NaNs included
    ops.raise_status(device=Xd.device)


@pytest.mark.gpu
def test_cuda_context_outside_timelapse_is_an_error():
    import torch
    from axtrack_b200 import ops, _lib
    Xd = torch.ones((3, 3, 8, 8), device='cuda')
    tyx = ops.compact_tiles(ops.tile_nonempty(Xd, 4))
    ops.gather_frametiles(Xd, tyx, [0], 4, 1)
    with pytest.raises(Exception) as e:
        ops.raise_status(device=Xd.device)
    assert 'temporal context' in str(e.value)
    code, _ = ops.get_context(Xd.device).status()
    assert code == _lib.OK                               # sticky status cleared


@pytest.mark.gpu
def test_frame_tiler_batches_the_forward():
    """FrameTiler = construct_tiles + get_frametiles_stack on the device; detect_frames feeds
    `frames_per_forward` frames per model call and returns [B, ntiles, Sx, Sy, 3]."""
    import torch
    from axtrack_b200.tiling import FrameTiler
    rng = np.random.default_rng(11)
    T, H, W, ts, tc = 9, 100, 150, 64, 2
    X = rng.random((T, 3, H, W), dtype=np.float32)
    X[:, 0, :64, 64:128] = 0
    X_tiled, nonempty, tile_yx = oracle.tile_frames(X, ts)
    tiler = FrameTiler(torch.from_numpy(X).cuda(), ts, tc, 'exclude')
    assert len(tiler) == T - 2 * tc and tiler.ntiles == len(tile_yx) == 5
    assert np.array_equal(tiler.tile_yx.cpu().numpy(), tile_yx)
    for t in range(len(tiler)):
        want = oracle.frametiles_stack(X_tiled, t + tc, tc, 'exclude')
        assert np.array_equal(tiler.get_frametiles_stack(t).cpu().numpy(), want)

    class Head:                       # stands in for YOLO_AXTrack.detect_axons (model.py:119-125)
        Sx = Sy = 2
        calls = []

        def detect_axons(self, Xb):
            self.calls.append(Xb.shape[0])
            return Xb.mean((-1, -2))[:, :1].repeat(1, 12).reshape(Xb.shape[0], 2, 2, 3)

    head = Head()
    Y = tiler.detect_frames(head, range(len(tiler)), frames_per_forward=2)
    assert head.calls == [10, 10, 5] and tuple(Y.shape) == (5, 5, 2, 2, 3)
    one = torch.stack([head.detect_axons(tiler.get_frametiles_stack(t)) for t in range(len(tiler))])
    assert torch.equal(Y, one)


@pytest.mark.gpu
def test_drop_in_detect_dataset_with_frames_per_forward(tmp_path):
    """detect_dataset() with the opt-in K6 front end (`frames_per_forward`) == the default path that tiles
    through dataset.get_frametiles_stack, for a head that is a per-tile function of the pixels."""
    import torch
    from axtrack_b200.axon_detections import make_drop_in
    from tests.test_gpu_dropin import _Base
    rng = np.random.default_rng(21)
    T, H, W, ts, tc = 8, 1000, 1400, 512, 2
    X = np.zeros((T, 3, H, W), np.float32)
    for t in range(T):                                     # sparse bright spots -> a few cells above threshold
        yy, xx = rng.integers(0, H, 300), rng.integers(0, W, 300)
        X[t, 0, yy, xx] = rng.random(300, dtype=np.float32) * 2000
    X[:, 0, 512:, :512] = 0                                # an empty tile
    X_tiled, nonempty, tile_yx = oracle.tile_frames(X, ts)

    class DS:
        name, sizet, sizey, sizex, tilesize, temporal_context = 'k6', T - 2 * tc, H, W, ts, tc
        use_motion_filtered, timepoints_indices = 'exclude', list(range(tc, T - tc))
        ytiles, xtiles = nonempty.shape[:2]
        tile_info = torch.from_numpy(np.stack([nonempty, nonempty], -1))
        target_tiled = torch.zeros((T, len(tile_yx), 12, 12, 4))

        def __init__(self):
            self.X = torch.from_numpy(X).to_sparse()

        def construct_tiles(self, *a, **k):
            pass

        def get_frametiles_stack(self, t, device=None):
            x = torch.from_numpy(oracle.frametiles_stack(X_tiled, self.timepoints_indices[t], tc, 'exclude'))
            return x.to(device), self.target_tiled[self.timepoints_indices[t]].to(device)

    class Model:
        def detect_axons(self, Xb):                        # per-tile, batch-size independent
            m = torch.nn.functional.adaptive_max_pool2d(Xb[:, tc:tc + 1], 12)[:, 0]          # [n, 12(y), 12(x)]
            conf = (m / 2000).transpose(1, 2)
            return torch.stack([conf, torch.full_like(conf, .5), torch.full_like(conf, .25)], -1)

    def run(fpf):
        ad = make_drop_in(_Base)(Model(), DS(), str(tmp_path))
        ad.device = 'cuda'
        ad.timepoint_subset = range(T - 2 * tc)
        if fpf:
            ad.frames_per_forward = fpf
        ad.detect_dataset()
        return ad

    a, b = run(None), run(3)
    assert sum(len(d) for d in a._detections) > 20
    for da, db in zip(a._detections, b._detections):
        assert da.equals(db)
    for ta, tb in zip(a._yolo_targets, b._yolo_targets):
        assert torch.equal(ta.cpu(), tb.cpu())
