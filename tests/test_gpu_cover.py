"""K2 cover (axb_search.cu k2_prep_kernel, axb_components.cu): frame pairs answered from their TARGET side.
Every case is compared pair by pair with the oracle (the closed form pinned to the compiled reference in
tests/test_oracle_golden.py), bit-exact.  The masks are chosen to stress what the cover relies on:
components smaller and larger than max_px, detections off the mask on either side, empty frames, chains of
three pairs, errors raised inside reverse pairs, the 256-column kernels."""
import numpy as np
import pytest

import oracle
from axtrack_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('-m gpu tests need a CUDA device (no CPU fallback exists)')
    torch.cuda.set_device(0)
    return torch


def _timelapse_pairs(frames, misses):
    """pairs in the reference's loop order (AxonDetections.py:553-566): (mask 0, source frame t_bef, target frame t)"""
    spec = []
    for t in range(len(frames)):
        for tb in range(t - 1, t - (misses + 2), -1):
            if tb >= 0:
                spec.append((0, tb, t))
    return spec


def _run_and_check(torch, mask, frames, spec, max_px, tag):
    from axtrack_b200 import ops
    offs = np.concatenate([[0], np.cumsum([len(f) for f in frames])])
    dets = np.concatenate(frames).astype(np.int32).reshape(-1, 2)
    pairs, total, _ = ops.make_pairs([(mk, int(offs[a]), len(frames[a]), int(offs[b]), len(frames[b])) for mk, a, b in spec])
    masks = mask if mask.ndim == 3 else mask[None]
    bits = ops.pack_mask(torch.from_numpy(np.ascontiguousarray(masks, dtype=np.uint8)).cuda())
    pl, un, wc, _ = ops.cost_matrix_batch(bits, masks.shape[2], torch.from_numpy(dets).cuda(), pairs, max_px)
    ops.raise_status()
    pl, un, wc = pl.cpu().numpy(), un.cpu().numpy(), wc.cpu().numpy()
    n_reached = 0
    for (mk, a, b), pd in zip(spec, pairs):
        o, ns, nd = int(pd['out_off']), int(pd['n_src']), int(pd['n_dst'])
        if ns * nd == 0:
            continue
        ref = oracle.cost_matrix_rule(masks[mk].astype(bool), frames[a], frames[b], max_px)
        got_pl, got_un = pl[o:o + ns * nd].reshape(ns, nd), un[o:o + ns * nd].reshape(ns, nd)
        bad = np.argwhere((got_pl != ref[0].reshape(ns, nd)) | (got_un != ref[1].reshape(ns, nd)))
        assert len(bad) == 0, (tag, (a, b), bad[:5].tolist(), got_pl[tuple(bad[0])], ref[0].reshape(ns, nd)[tuple(bad[0])])
        ok = got_un == 0
        got_wc = wc[o:o + ns * nd].reshape(ns, nd)
        assert np.all(np.isinf(got_wc[~ok])), tag
        assert np.allclose(got_wc[ok], np.asarray(ref[2]).reshape(ns, nd)[ok], rtol=1e-5, atol=0), tag
        n_reached += int(ok.sum())
    return n_reached


def _islands_mask(H, W, seed):
    """Corridors cut into pieces + blobs of every size + speckle: components well below and above max_px."""
    rng = np.random.default_rng(seed)
    m = synth.make_mask(H, W, seed, pitch=(30, 70), width=(3, 8)).astype(bool)
    for _ in range(40):                                     # cuts
        y, x = rng.integers(0, H), rng.integers(0, W)
        if rng.random() < 0.5:
            m[y:y + 3, :] = False
        else:
            m[:, x:x + 3] = False
    for _ in range(30):                                     # blobs
        y, x, h, w = rng.integers(0, H - 12), rng.integers(0, W - 12), rng.integers(1, 12), rng.integers(1, 12)
        m[y:y + h, x:x + w] = True
    m |= rng.random((H, W)) < 0.02                          # speckle
    return m


def _frames_on(mask, n_frames, n, seed, off_frac):
    rng = np.random.default_rng(seed)
    fg = np.argwhere(mask)
    H, W = mask.shape
    out = []
    for t in range(n_frames):
        d = fg[rng.integers(0, len(fg), n)]
        jit = rng.integers(-2, 3, (n, 2)) * (rng.random((n, 1)) < off_frac)
        d = d + jit
        d[:, 0] = np.clip(d[:, 0], 0, H - 1); d[:, 1] = np.clip(d[:, 1], 0, W - 1)
        out.append(d.astype(np.int32))
    return out


@pytest.mark.parametrize('seed,max_px,misses', [(0, 60, 1), (1, 35, 1), (2, 90, 2), (3, 60, 2), (4, 20, 1)])
def test_cover_on_island_masks(cuda, seed, max_px, misses):
    H, W = 260, 300
    mask = _islands_mask(H, W, seed)
    frames = _frames_on(mask, 7, 40, seed + 10, off_frac=0.3)
    frames[4] = frames[4][:0]                               # an empty frame in the middle
    frames[2] = np.vstack([frames[2], frames[1][:5]])       # detections that coincide with those of a neighbour frame
    n = _run_and_check(cuda, mask, frames, _timelapse_pairs(frames, misses), max_px, ('islands', seed))
    assert n > 50


def test_cover_on_a_connected_corridor_mask(cuda):
    """The regime of the real data: one component that reaches far beyond max_px; no directed search is needed
    except for off-mask detections."""
    H, W = 700, 900
    mask = synth.make_mask(H, W, 5, pitch=(60, 130), width=(4, 10)).astype(bool)
    tracks = synth.make_tracks(mask, 60, 8, 5, off_mask_frac=0.1)
    n = _run_and_check(cuda, mask, [t.astype(np.int32) for t in tracks], _timelapse_pairs(tracks, 1), 300, 'corridor')
    assert n > 500


def test_cover_with_pairs_in_any_order_and_duplicates(cuda):
    """The cover is decided from the pair list alone: shuffled pairs, a pair given twice, both directions of a
    pair, a frame paired with itself."""
    H, W = 200, 240
    mask = _islands_mask(H, W, 11)
    frames = _frames_on(mask, 6, 30, 12, off_frac=0.25)
    spec = _timelapse_pairs(frames, 1)
    rng = np.random.default_rng(0)
    spec = [spec[i] for i in rng.permutation(len(spec))]
    spec += [(0, 1, 2), (0, 2, 1), (0, 3, 3), (0, 5, 0)]
    n = _run_and_check(cuda, mask, frames, spec, 70, 'shuffled')
    assert n > 50


def test_cover_wide_radius(cuda):
    """max_px > 511: the 256-column kernels take the same chains."""
    H, W = 900, 1500
    mask = synth.make_mask(H, W, 7, pitch=(90, 200), width=(4, 9)).astype(bool)
    mask[:, 700:704] = False                                 # two halves: components that do not reach max_px everywhere
    tracks = synth.make_tracks(mask, 16, 5, 7, off_mask_frac=0.15)
    n = _run_and_check(cuda, mask, [t.astype(np.int32) for t in tracks], _timelapse_pairs(tracks, 1), 700, 'wide')
    assert n > 30


def test_errors_inside_reverse_pairs(cuda):
    """An out-of-grid detection in a frame that is only a TARGET side of the search still raises the reference's error,
    first offender in the reference's order, with start / goal the right way round."""
    from axtrack_b200 import ops
    m = np.ones((50, 60), np.uint8)
    bits = ops.pack_mask(cuda.from_numpy(m).cuda())
    # frames 0..3 of two detections each; pairs (1,0) (2,0) (2,1) (3,1) (3,2): frame 0 is not searched
    spec = [(0, 0, 2, 2, 2), (0, 0, 2, 4, 2), (0, 2, 2, 4, 2), (0, 2, 2, 6, 2), (0, 4, 2, 6, 2)]
    pairs, total, nsrc = ops.make_pairs(spec)
    base = np.array([[5, 5], [8, 3], [6, 6], [7, 7], [9, 9], [10, 4], [11, 11], [12, 5]], np.int32)
    for idx, bad, msg in [(1, (50, 3), r'Start of \(50, 3\) lies outside grid'),       # a source of the unsearched frame 0
                          (0, (-1, 5), r'Start of \(-1, 5\) lies outside grid'),
                          (3, (7, 60), r'Goal of \(7, 60\) lies outside grid'),          # frame 1: goal of pair (1,0) comes first
                          (7, (12, -2), r'Goal of \(12, -2\) lies outside grid')]:
        dets = base.copy()
        dets[idx] = bad
        ops.cost_matrix_batch(bits, 60, cuda.from_numpy(dets).cuda(), pairs, 500)
        with pytest.raises(ValueError, match=msg):
            ops.raise_status(pairs, dets)
        ops.raise_status()


def test_components_kernel_vs_scipy(cuda):
    """axb_components.cu through its effect: on a mask of islands, a pair answered in reverse equals the oracle for
    every max_px around the islands' sizes (the extents decide None vs directed search)."""
    from scipy import ndimage
    H, W = 120, 150
    rng = np.random.default_rng(21)
    mask = np.zeros((H, W), bool)
    for k in range(14):                                      # diagonal, L-shaped and straight islands of growing size
        y, x = 6 + 8 * k, 5 + rng.integers(0, 20)
        L = 6 + 7 * k
        mask[y:y + 2, x:min(W - 1, x + L)] = True
        if k % 3 == 0:
            mask[max(0, y - 4):y, x:x + 2] = True
    lab, ncomp = ndimage.label(mask)
    assert ncomp >= 10
    frames = _frames_on(mask, 4, 45, 22, off_frac=0.2)
    for max_px in (8, 15, 30, 55, 100):
        _run_and_check(cuda, mask, frames, _timelapse_pairs(frames, 1), max_px, ('extents', max_px))


_FULL_SIZE_SNIPPET = r'''
import sys, zlib, numpy as np, torch
sys.path.insert(0, '.')
import bench
from axtrack_b200 import ops, pipeline
wl = dict(bench.WORKLOADS['c2'], T=14)
mask, Y, tt, t0 = bench.make_block(wl)[:4]
dev = torch.device('cuda', 0)
bits = ops.pack_mask(torch.from_numpy(mask.astype(np.uint8)).to(dev))
det = ops.decode_nms(torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev))
n_out = det['n_out'].cpu().numpy(); cap = det['conf'].shape[1]
for max_px in (500, 800):
    pairs, total, _ = pipeline.build_pairs(n_out, cap, pipeline.frame_pair_keys(14))
    pl, un, wc, _ = ops.cost_matrix_batch(bits, wl['W'], det['yx'].view(-1, 2), pairs, max_px)
    ops.raise_status()
    print('CRC', max_px, zlib.crc32(pl.cpu().numpy().tobytes()), zlib.crc32(un.cpu().numpy().tobytes()), zlib.crc32(wc.cpu().numpy().tobytes()), int((un == 0).sum()))
'''


def test_cover_equals_directed_searches_at_full_size(cuda):
    """configs[1] shape (2048 x 2048, ~200 detections per frame, 14 frames), max_px 500 and 800: the cover with the
    early all-levels launch gives bit for bit what one directed search per source gives (AXB_K2_NO_REVERSE,
    AXB_K2_NO_EARLY are read once per process, hence the subprocesses)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for extra in ({}, {'AXB_K2_NO_REVERSE': '1', 'AXB_K2_NO_EARLY': '1'}):
        env = dict(os.environ, **extra)
        r = subprocess.run([sys.executable, '-c', _FULL_SIZE_SNIPPET], cwd=root, env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append([l for l in r.stdout.splitlines() if l.startswith('CRC')])
    assert len(outs[0]) == 2 and outs[0] == outs[1], outs
    assert int(outs[0][0].split()[-1]) > 10000


def test_cover_with_more_pairs_than_the_shared_memory_path_holds(cuda):
    """> 11 000 pairs: k2_prep_kernel decides the cover in global memory instead of shared memory; 5 800 frames of
    two or three detections on a small island mask, every pair against the oracle."""
    H, W = 48, 64
    mask = _islands_mask(H, W, 31)
    rng = np.random.default_rng(32)
    fg = np.argwhere(mask)
    frames = []
    for t in range(5800):
        n = 2 + (t % 3 == 0)
        d = fg[rng.integers(0, len(fg), n)] + rng.integers(-1, 2, (n, 2)) * (rng.random((n, 1)) < 0.2)
        d[:, 0] = np.clip(d[:, 0], 0, H - 1); d[:, 1] = np.clip(d[:, 1], 0, W - 1)
        frames.append(d.astype(np.int32))
    spec = _timelapse_pairs(frames, 1)
    assert len(spec) > 11000
    n = _run_and_check(cuda, mask, frames, spec, 25, 'many pairs')
    assert n > 5000


def test_cost_matrix_batch_under_graph_capture(cuda):
    """axb_cost_matrix_batch forks the all-levels searches of off-mask detections onto a ctx-owned side stream and
    joins it back before its last launch: to the caller it is ONE call ordered on `stream`, so it can be captured
    into a CUDA graph (after one warm-up call that sizes the scratch) and replayed on new detections."""
    from axtrack_b200 import ops
    torch = cuda
    H, W = 400, 520
    mask = synth.make_mask(H, W, 3, pitch=(50, 110), width=(4, 9)).astype(bool)
    tracks = [t.astype(np.int32) for t in synth.make_tracks(mask, 40, 6, 3, off_mask_frac=0.15)]
    n = min(len(t) for t in tracks)
    tracks = [t[:n] for t in tracks]                       # fixed shapes: the graph is replayed on other coordinates
    offs = np.arange(7) * n
    spec = _timelapse_pairs(tracks, 1)
    pairs, total, _ = ops.make_pairs([(mk, int(offs[a]), n, int(offs[b]), n) for mk, a, b in spec])
    dev = torch.device('cuda', 0)
    bits = ops.pack_mask(torch.from_numpy(mask.astype(np.uint8)).to(dev))
    dets = torch.from_numpy(np.concatenate(tracks)).to(dev)
    pd = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev)
    out = (torch.empty(total, dtype=torch.int32, device=dev), torch.empty(total, dtype=torch.uint8, device=dev),
           torch.empty(total, dtype=torch.float32, device=dev))
    ops.cost_matrix_batch(bits, W, dets, pairs, 150, out=out, pairs_dev=pd)          # warm-up: allocations happen here
    torch.cuda.synchronize()
    want = [o.clone() for o in out]
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        ops.cost_matrix_batch(bits, W, dets, pairs, 150, out=out, pairs_dev=pd)
    for o in out:
        o.zero_()
    g.replay()
    torch.cuda.synchronize()
    for a, b in zip(out, want):
        assert torch.equal(a, b)
    # other coordinates, same graph: frames rolled by one
    rolled = np.concatenate(tracks[1:] + tracks[:1])
    dets.copy_(torch.from_numpy(rolled).to(dev))
    g.replay()
    torch.cuda.synchronize()
    ops.raise_status()
    pl, un = out[0].cpu().numpy(), out[1].cpu().numpy()
    fr = tracks[1:] + tracks[:1]
    for (mk, a, b), p in zip(spec, pairs):
        o = int(p['out_off'])
        ref = oracle.cost_matrix_rule(mask, fr[a], fr[b], 150)
        assert np.array_equal(pl[o:o + n * n].reshape(n, n), ref[0].reshape(n, n)), (a, b)
        assert np.array_equal(un[o:o + n * n].reshape(n, n), ref[1].reshape(n, n)), (a, b)


def test_prepare_masks_is_one_shot_and_changes_nothing(cuda):
    """axb_prepare_masks ahead of axb_cost_matrix_batch: same results; the products are used once, for the same
    pointer and sizes only, and never for another mask."""
    from axtrack_b200 import ops
    torch = cuda
    H, W = 260, 300
    frames = None
    dev = torch.device('cuda', 0)
    ctx = ops.get_context(dev)
    mask_a, mask_b = _islands_mask(H, W, 41), _islands_mask(H, W, 42)
    frames = _frames_on(mask_a | mask_b, 5, 30, 43, off_frac=0.2)
    offs = np.concatenate([[0], np.cumsum([len(f) for f in frames])])
    spec = _timelapse_pairs(frames, 1)
    pairs, total, _ = ops.make_pairs([(mk, int(offs[a]), len(frames[a]), int(offs[b]), len(frames[b])) for mk, a, b in spec])
    dets = torch.from_numpy(np.concatenate(frames).astype(np.int32)).to(dev)
    bits = ops.pack_mask(torch.from_numpy(mask_a.astype(np.uint8)).to(dev))

    def run():
        pl, un, wc, _ = ops.cost_matrix_batch(bits, W, dets, pairs, 60)
        ops.raise_status()
        return pl.clone(), un.clone(), wc.clone()

    plain = run()
    l0 = ctx.launches
    run()
    n_plain = ctx.launches - l0
    ops.prepare_masks(bits, W)
    l0 = ctx.launches
    prepared = run()
    n_prepared = ctx.launches - l0
    assert n_prepared == n_plain - 5                       # tile words + the four component kernels ran ahead
    for a, b in zip(plain, prepared):
        assert torch.equal(a, b)
    l0 = ctx.launches
    run()                                                  # one shot: the next call derives them again
    assert ctx.launches - l0 == n_plain
    # prepared for mask A, then the SAME buffer is given mask B's bits via another pointer of the same size: not used
    bits_b = ops.pack_mask(torch.from_numpy(mask_b.astype(np.uint8)).to(dev))
    ops.prepare_masks(bits, W)
    pl_b, un_b, _, _ = ops.cost_matrix_batch(bits_b, W, dets, pairs, 60)
    ops.raise_status()
    for (mk, a, b), p in zip(spec, pairs):
        o, ns, nd = int(p['out_off']), int(p['n_src']), int(p['n_dst'])
        ref = oracle.cost_matrix_rule(mask_b, frames[a], frames[b], 60)
        assert np.array_equal(pl_b[o:o + ns * nd].cpu().numpy().reshape(ns, nd), ref[0].reshape(ns, nd))
    # discarded explicitly
    ops.prepare_masks(bits, W)
    ops.discard_prepared_masks(dev)
    l0 = ctx.launches
    run()
    assert ctx.launches - l0 == n_plain


def test_cover_with_more_candidates_than_one_pass_holds(cuda):
    """~200 detections per frame inside one radius: a searched detection has up to 800 in-range candidates over its
    forward and reverse members, more than the target hash of one pass holds (320), so the search repeats in passes;
    islands and off-mask detections keep the directed and all-levels fallbacks busy in the later passes too."""
    H, W = 110, 150
    mask = _islands_mask(H, W, 51)
    mask[40:70, 20:130] = True                              # a large blob: most detections see each other
    frames = _frames_on(mask, 5, 200, 52, off_frac=0.15)
    n = _run_and_check(cuda, mask, frames, _timelapse_pairs(frames, 1), 90, 'multi-pass')
    assert n > 100000
