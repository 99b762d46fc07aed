"""Parity of the CUDA path (through the C ABI) against the oracle and the committed golden
fixtures produced by the unmodified reference.  Bit-exact: path lengths, unreachable flags,
NMS survivor indices, decoded coordinates; weighted path cost within 1e-5 relative."""
import ctypes
import glob
import os

import numpy as np
import pytest

import oracle
from axtrack_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5      # north_star: float-weighted path costs agree within 1e-5 relative


@pytest.fixture(scope='module')
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('-m gpu tests need a CUDA device (no CPU fallback exists)')
    torch.cuda.set_device(0)
    return torch


def _gpu_cost(torch, mask, src, dst, max_px):
    from axtrack_b200 import ops
    bits = ops.pack_mask(torch.from_numpy(np.ascontiguousarray(mask, dtype=np.uint8)).cuda())
    pl, un, wc = ops.cost_matrix(bits, mask.shape[1], torch.from_numpy(np.ascontiguousarray(src, dtype=np.int32)).cuda(),
                                 torch.from_numpy(np.ascontiguousarray(dst, dtype=np.int32)).cuda(), max_px)
    ops.raise_status()
    return pl.cpu().numpy(), un.cpu().numpy(), wc.cpu().numpy()


def _check(got, ref, tag=''):
    pl, un, wc = got
    assert np.array_equal(pl, ref[0]), tag
    assert np.array_equal(un, ref[1]), tag
    ok = ref[1] == 0
    assert np.all(np.isinf(wc[~ok])), tag
    assert np.allclose(wc[ok], np.asarray(ref[2])[ok], rtol=RTOL, atol=0), tag


def test_pack_mask(cuda):
    from axtrack_b200 import ops
    rng = np.random.default_rng(0)
    for H, W in [(5, 7), (33, 64), (40, 100), (64, 2048), (3, 16), (17, 48)]:
        m = (rng.random((2, H, W)) < 0.4).astype(np.uint8)
        m[0, 0, 0] = 2          # only == 1 is on the mask (AxonDetections.py:598)
        bits = ops.pack_mask(cuda.from_numpy(m).cuda()).cpu().numpy().view(np.uint32)
        exp = np.zeros((2, H, (W + 31) // 32), np.uint32)
        for x in range(W):
            exp[:, :, x >> 5] |= (m[:, :, x] == 1).astype(np.uint32) << np.uint32(x & 31)
        assert np.array_equal(bits, exp), (H, W)
        w = np.where(m == 1, 1, 2 ** 16).astype(np.float32)
        bits2 = ops.pack_weights(cuda.from_numpy(w).cuda()).cpu().numpy().view(np.uint32)
        assert np.array_equal(bits2, exp), (H, W)
        ops.raise_status()


def test_golden_astar(cuda, golden_dir):
    files = sorted(glob.glob(os.path.join(golden_dir, 'astar_s*.npz')))
    assert len(files) >= 4
    for f in files:
        g = np.load(f)
        H, W, max_px = int(g['H']), int(g['W']), int(g['max_px'])
        mask = np.unpackbits(g['mask_bits'])[:H * W].reshape(H, W)
        for t, tb in g['keys']:
            got = _gpu_cost(cuda, mask, g[f'dets_yx_{tb}'], g[f'dets_yx_{t}'], max_px)
            _check(got, (g[f'dist_{t}_{tb}'], g[f'none_{t}_{tb}'], g[f'wcost_{t}_{tb}']), (f, t, tb))
            # downstream "final cost matrix" (transition_model, mincostflow_models.py:100-119)
            tc = oracle.transition_cost(got[0], max_px, 0.6, int(t - tb))
            ref_t = g[f'tcost_{t}_{tb}']
            assert np.array_equal(np.isinf(tc), np.isinf(ref_t))
            fin = ~np.isinf(tc)
            assert np.allclose(tc[fin], ref_t[fin], rtol=RTOL, atol=0)


def test_golden_decode_nms(cuda, golden_dir):
    from axtrack_b200 import ops
    for f in sorted(glob.glob(os.path.join(golden_dir, 'decode_nms_s*.npz'))):
        g = np.load(f)
        o = ops.decode_nms(cuda.from_numpy(g['Y']).cuda(), cuda.from_numpy(g['tile_yx'].astype(np.int32)).cuda(),
                           int(g['tilesize']), float(g['conf_thr']), int(g['nms_dist']))
        ops.raise_status()
        npre, nout = int(o['n_pre'][0]), int(o['n_out'][0])
        assert npre == len(g['pre_conf']) and nout == len(g['conf']), f
        assert np.array_equal(o['pre_conf'][0, :npre].cpu().numpy(), g['pre_conf'])
        pyx = o['pre_yx'][0, :npre].cpu().numpy()
        assert np.array_equal(pyx[:, 1], g['pre_x']) and np.array_equal(pyx[:, 0], g['pre_y'])
        assert np.array_equal(o['keep_idx'][0, :nout].cpu().numpy(), g['keep_idx'])
        assert np.array_equal(o['conf'][0, :nout].cpu().numpy(), g['conf'])
        yx = o['yx'][0, :nout].cpu().numpy()
        assert np.array_equal(yx[:, 1], g['x']) and np.array_equal(yx[:, 0], g['y'])


@pytest.mark.parametrize('seed', [0, 1, 2, 3])
def test_cost_matrix_vs_oracle_random(cuda, seed):
    """Small adversarial grids: gaps, off-mask detections, short cut-offs; vs the closed form AND
    the per-pair restated A*."""
    rng = np.random.default_rng(100 + seed)
    for it in range(25):
        H, W = int(rng.integers(20, 160)), int(rng.integers(20, 160))
        m = np.zeros((H, W), bool)
        for _ in range(int(rng.integers(2, 7))):
            y = int(rng.integers(0, H)); w_ = int(rng.integers(1, 6)); m[y:y + w_, :] = True
            x = int(rng.integers(0, W)); w_ = int(rng.integers(1, 6)); m[:, x:x + w_] = True
        for _ in range(int(rng.integers(0, 8))):
            y, x, g = int(rng.integers(0, H)), int(rng.integers(0, W)), int(rng.integers(1, 5))
            m[y:y + g, x:x + g] = False
        max_px = int(rng.choice([1, 2, 12, 25, 40, 80, 150, 500, 511, 512, 1023]))
        fg = np.argwhere(m)

        def pick(n):
            p = fg[rng.integers(0, len(fg), n)] + rng.integers(-2, 3, (n, 2)) * (rng.random((n, 1)) < 0.3)
            p[:, 0] = np.clip(p[:, 0], 0, H - 1)
            p[:, 1] = np.clip(p[:, 1], 0, W - 1)
            return p.astype(np.int32)
        src, dst = pick(int(rng.integers(1, 9))), pick(int(rng.integers(1, 14)))
        got = _gpu_cost(cuda, m, src, dst, max_px)
        _check(got, oracle.cost_matrix_rule(m, src, dst, max_px), (seed, it, 'rule'))


def test_cost_matrix_edge_cases(cuda):
    from axtrack_b200 import ops
    m = np.ones((1, 20), np.uint8)
    # SURVEY Appendix A known answer: 1x20 row, max=19 -> None, max=20 -> 20
    pl, un, _ = _gpu_cost(cuda, m, [[0, 0]], [[0, 19]], 19)
    assert pl[0, 0] == 19 and un[0, 0] == 1
    pl, un, wc = _gpu_cost(cuda, m, [[0, 0]], [[0, 19]], 20)
    assert pl[0, 0] == 20 and un[0, 0] == 0 and wc[0, 0] == 19.0
    # source == target -> path of one pixel (astar.cpp:67,99-102)
    pl, un, wc = _gpu_cost(cuda, m, [[0, 5]], [[0, 5]], 500)
    assert pl[0, 0] == 1 and un[0, 0] == 0 and wc[0, 0] == 0.0
    # pyastar2d/tests/test_astar.py:63-78 narrow grids
    for shp, a, b in [((5, 1), (0, 0), (4, 0)), ((1, 5), (0, 0), (0, 4))]:
        pl, un, _ = _gpu_cost(cuda, np.ones(shp, np.uint8), [a], [b], 500)
        assert pl[0, 0] == 5 and un[0, 0] == 0
    # empty inputs
    bits = ops.pack_mask(cuda.from_numpy(m).cuda())
    e = cuda.zeros((0, 2), dtype=cuda.int32, device='cuda')
    one = cuda.zeros((1, 2), dtype=cuda.int32, device='cuda')
    assert ops.cost_matrix(bits, 20, e, one)[0].shape == (0, 1)
    assert ops.cost_matrix(bits, 20, one, e)[0].shape == (1, 0)
    # all-background mask: every step costs 65536 (levels only)
    z = np.zeros((30, 30), np.uint8)
    src, dst = np.array([[3, 3], [15, 15]], np.int32), np.array([[3, 9], [20, 11], [29, 29]], np.int32)
    _check(_gpu_cost(cuda, z, src, dst, 40), oracle.cost_matrix_rule(z, src, dst, 40))
    _check(_gpu_cost(cuda, z, src, dst, 40), oracle.cost_matrix_astar(z, src, dst, 40))


def test_out_of_grid_errors(cuda):
    """astar_wrapper.py:39-45 -> the same ValueError text, first offender in reference order."""
    from axtrack_b200 import ops
    m = np.ones((50, 60), np.uint8)
    bits = ops.pack_mask(cuda.from_numpy(m).cuda())
    pairs, total, nsrc = ops.make_pairs([(0, 0, 2, 2, 2)])
    for dets, msg in [(np.array([[5, 5], [50, 3], [6, 6], [7, 7]], np.int32), r'Start of \(50, 3\) lies outside grid'),
                      (np.array([[5, 5], [8, 3], [6, 6], [7, -1]], np.int32), r'Goal of \(7, -1\) lies outside grid')]:
        ops.cost_matrix_batch(bits, 60, cuda.from_numpy(dets).cuda(), pairs, 500)
        with pytest.raises(ValueError, match=msg):
            ops.raise_status(pairs, dets)
        ops.raise_status()      # cleared
    # out-of-grid but out of Euclid range: the reference never calls pyastar2d -> no error
    dets = np.array([[5, 5], [6, 6], [7, 7], [900, 900]], np.int32)
    pl, un, _, _ = ops.cost_matrix_batch(bits, 60, cuda.from_numpy(dets).cuda(), pairs, 500)
    ops.raise_status(pairs, dets)
    assert un.cpu().numpy().reshape(2, 2)[:, 1].all()
    # weights < 1 (astar_wrapper.py:35-37)
    w = np.ones((8, 8), np.float32); w[2, 2] = 0.25
    ops.pack_weights(cuda.from_numpy(w).cuda())
    with pytest.raises(ValueError, match='Minimum cost to move must be 1'):
        ops.raise_status()


def test_bad_pair_descriptors_get_no_work(cuda):
    """the DEVICE entry point validates its descriptors -- a pair that points outside the
    detections / masks / outputs gets no work and sets AXB_ERR_BAD_ARG; the valid pairs of the same batch
    are still answered; a total_src smaller than the pairs' sources runs nothing."""
    from axtrack_b200 import ops, _lib
    mask = synth.make_mask(300, 320, 2, pitch=(50, 90), width=(5, 10))
    a, b = synth.make_tracks(mask, 20, 2, 2)
    dets = np.concatenate([a, b]).astype(np.int32)
    bits = ops.pack_mask(cuda.from_numpy(mask.astype(np.uint8)).cuda())
    dd = cuda.from_numpy(dets).cuda()
    good = (0, 0, len(a), len(a), len(b))
    ref = oracle.cost_matrix_rule(mask, a, b, 200)
    n = len(a) * len(b)
    for bad in [(0, 0, len(a), len(a), len(b) + 1),        # targets run past the detections
                (0, -1, 3, 0, 3), (1, 0, 3, 0, 3)]:         # negative offset, mask index out of range
        pairs, total, _ = ops.make_pairs([good, bad])
        out = (cuda.full((total,), -7, dtype=cuda.int32).cuda(), cuda.full((total,), 9, dtype=cuda.uint8).cuda(), None)
        pl, un, _, _ = ops.cost_matrix_batch(bits, 320, dd, pairs, 200, want_cost=False, out=out)
        with pytest.raises(_lib.AxbError, match='pair descriptor'):
            ops.raise_status()
        ops.raise_status()
        assert np.array_equal(pl[:n].cpu().numpy().reshape(ref[0].shape), ref[0]), bad
        assert (out[0][n:] == -7).all() and (out[1][n:] == 9).all(), bad
    # output block outside the buffers: out_capacity is what the caller's tensors hold
    pairs, total, _ = ops.make_pairs([good])
    small = (cuda.full((n - 1,), -7, dtype=cuda.int32).cuda(), cuda.full((n - 1,), 9, dtype=cuda.uint8).cuda(), None)
    ops.cost_matrix_batch(bits, 320, dd, pairs, 200, want_cost=False, out=small)
    with pytest.raises(_lib.AxbError, match='pair descriptor'):
        ops.raise_status()
    assert (small[0] == -7).all()
    # total_src too small for the pairs' sources: nothing runs
    ctx = ops.get_context()
    pd = cuda.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).cuda()
    out = (cuda.full((n,), -7, dtype=cuda.int32).cuda(), cuda.full((n,), 9, dtype=cuda.uint8).cuda())
    ctx.check(ctx.lib.axb_cost_matrix_batch(ctx.handle, bits.data_ptr(), 1, 300, 320, dd.data_ptr(), len(dets), pd.data_ptr(), 1,
                                            len(a) - 1, 200, out[0].data_ptr(), out[1].data_ptr(), None, n, None))
    cuda.cuda.synchronize()
    with pytest.raises(_lib.AxbError, match='total_src'):
        ops.raise_status()
    assert (out[0] == -7).all()


def test_entry_points_keep_the_callers_device(cuda):
    """tensors on another device than torch's current one -- the call runs on the tensors'
    device with that device's stream and leaves the thread's current device alone."""
    if cuda.cuda.device_count() < 2:
        pytest.skip('needs two visible GPUs')
    from axtrack_b200 import ops
    mask = synth.make_mask(300, 320, 2, pitch=(50, 90), width=(5, 10))
    a, b = synth.make_tracks(mask, 20, 2, 2)
    cuda.cuda.set_device(0)
    d1 = cuda.device('cuda', 1)
    bits = ops.pack_mask(cuda.from_numpy(mask.astype(np.uint8)).to(d1))
    pl, un, wc = ops.cost_matrix(bits, 320, cuda.from_numpy(a).to(d1), cuda.from_numpy(b).to(d1), 200)
    ops.raise_status(device=d1)
    assert cuda.cuda.current_device() == 0
    _check((pl.cpu().numpy(), un.cpu().numpy(), wc.cpu().numpy()), oracle.cost_matrix_rule(mask, a, b, 200))


@pytest.mark.parametrize('seed', [0, 1])
def test_timelapse_pipeline_vs_oracle(cuda, seed):
    """decode+NMS of all frames then all (t, t_bef) cost matrices, vs the oracle; corridor mask."""
    from axtrack_b200.pipeline import associate_frames
    H, W, T = 1024, 1536, 4
    mask = synth.make_mask(H, W, seed, pitch=(90, 200), width=(6, 16))
    tracks = synth.make_tracks(mask, 70, T, seed, off_mask_frac=0.08)
    Y = np.stack([synth.make_head_tensor(tracks[t], H, W, seed * 10 + t)[0] for t in range(T)])
    tt = synth.tile_table(H, W)
    res = associate_frames(cuda.from_numpy(Y).cuda(), cuda.from_numpy(tt).cuda(),
                           cuda.from_numpy(mask.astype(np.uint8)).cuda())
    dets = []
    for t in range(T):
        o = oracle.decode_stitch_nms(Y[t], tt)
        d = res.frame_dets(t)
        for k in ('conf', 'x', 'y', 'keep_idx'):
            assert np.array_equal(d[k], o[k]), (t, k)
        dets.append(np.stack([o['y'], o['x']], 1).astype(np.int32))
    mats = res.pair_matrices()
    assert list(mats) == [(1, 0), (2, 1), (2, 0), (3, 2), (3, 1)]
    for (t, tb), got in mats.items():
        _check(got, oracle.cost_matrix_rule(mask, dets[tb], dets[t], 500), (t, tb))


def test_host_entry_points(cuda):
    """axb_decode_nms_host / axb_cost_matrix_batch_host: numpy in, numpy out (what e2e times)."""
    from axtrack_b200 import host
    H, W, T = 640, 900, 3
    mask = synth.make_mask(H, W, 5, pitch=(70, 160), width=(6, 14))
    tracks = synth.make_tracks(mask, 40, T, 5, off_mask_frac=0.1)
    Y = np.stack([synth.make_head_tensor(tracks[t], H, W, 50 + t)[0] for t in range(T)])
    tt = synth.tile_table(H, W)
    ha = host.HostAssociator(0)
    det = ha.decode_nms(Y, tt)
    dets = []
    for t in range(T):
        o = oracle.decode_stitch_nms(Y[t], tt)
        n = int(det['n_out'][t])
        assert np.array_equal(det['keep_idx'][t, :n], o['keep_idx'])
        assert np.array_equal(det['yx'][t, :n, 0], o['y']) and np.array_equal(det['yx'][t, :n, 1], o['x'])
        dets.append(det['yx'][t, :n])
    mats = ha.cost_matrices(mask.astype(np.uint8), dets, max_px=500, max_num_misses=1)
    for (t, tb), got in mats.items():
        _check(got, oracle.cost_matrix_rule(mask, dets[tb], dets[t], 500), (t, tb))


def test_full_size_properties(cuda):
    """BASELINE config-2 sizes (2048^2, ~200 dets): size-independent properties -- symmetry of
    on-mask/on-mask path lengths where both directions are reachable, self distance 1, the
    Euclid pre-filter, L >= L1+1, and agreement with the closed-form oracle on a row sample."""
    H = W = 2048
    mask = synth.make_mask(H, W, 0)
    a, b = synth.make_tracks(mask, 200, 2, 0, off_mask_frac=0.0)
    pl, un, wc = _gpu_cost(cuda, mask, a, b, 500)
    pl2, un2, _ = _gpu_cost(cuda, mask, b, a, 500)
    both = (un == 0) & (un2.T == 0)
    assert both.sum() > 500
    assert np.array_equal(pl[both], pl2.T[both])
    d2 = ((a[:, None, :].astype(np.int64) - b[None, :, :]) ** 2).sum(-1)
    assert un[d2 >= 500 * 500].all() and (pl[un == 1] == 500).all()
    l1 = np.abs(a[:, None, :].astype(np.int64) - b[None, :, :]).sum(-1)
    assert (pl[un == 0] >= l1[un == 0] + 1).all() and (pl[un == 0] <= 500).all()
    # level-0 paths cost exactly L-1
    lvl0 = (un == 0) & (wc < 65536)
    assert np.array_equal(wc[lvl0], (pl[lvl0] - 1).astype(np.float32))
    pls, uns, _ = _gpu_cost(cuda, mask, a, a, 500)
    assert (np.diag(pls) == 1).all() and (np.diag(uns) == 0).all()
    rows = np.arange(0, len(a), 9)
    _check((pl[rows], un[rows], wc[rows]), oracle.cost_matrix_rule(mask, a[rows], b, 500))


def test_many_targets_multi_pass(cuda):
    """> kHashMaxLoad (640) in-range targets per source (several search passes), duplicate target
    cells, targets equal to the source, open field (frontier spills to the global queue)."""
    rng = np.random.default_rng(7)
    H, W = 420, 380
    m = np.ones((H, W), np.uint8)
    m[100:300, 190] = 0                      # a wall with a detour
    m[40, 20:200] = 0
    src = np.stack([rng.integers(0, H, 5), rng.integers(0, W, 5)], 1).astype(np.int32)
    dst = np.stack([rng.integers(0, H, 1500), rng.integers(0, W, 1500)], 1).astype(np.int32)
    dst[10] = dst[11] = dst[12]              # duplicates
    dst[20] = src[0]; dst[21] = src[3]       # start == goal
    for max_px in (500, 90):
        _check(_gpu_cost(cuda, m, src, dst, max_px), oracle.cost_matrix_rule(m, src, dst, max_px), max_px)


def test_chain_sharing_and_batches(cuda):
    """Pairs that share (mask, sources) are answered by one search; pairs that do not, are not.
    Compared per pair with the oracle; includes per-frame masks and an empty frame."""
    from axtrack_b200 import ops
    rng = np.random.default_rng(3)
    H, W = 300, 340
    masks = np.stack([synth.make_mask(H, W, s, pitch=(40, 90), width=(4, 10)) for s in (1, 2)]).astype(np.uint8)
    frames = [synth.make_tracks(masks[0].astype(bool), 25, 1, s, off_mask_frac=0.2)[0] for s in range(5)]
    frames[3] = np.zeros((0, 2), np.int32)
    offs = np.concatenate([[0], np.cumsum([len(f) for f in frames])])
    dets = np.concatenate(frames).astype(np.int32)
    # (mask, src frame, dst frame): 0-2 share a search (same mask + sources), 3 differs by mask,
    # 4 has an empty target frame, 5 an empty source frame, 6-10 a chain longer than kMaxChain
    spec = [(0, 0, 1), (0, 0, 2), (0, 0, 4), (1, 0, 1), (0, 1, 3), (0, 3, 1)] + [(1, 2, 4)] * 5
    pairs, total, _ = ops.make_pairs([(mk, int(offs[a]), len(frames[a]), int(offs[b]), len(frames[b])) for mk, a, b in spec])
    bits = ops.pack_mask(cuda.from_numpy(masks).cuda())
    pl, un, wc, _ = ops.cost_matrix_batch(bits, W, cuda.from_numpy(dets).cuda(), pairs, 120)
    ops.raise_status()
    pl, un, wc = pl.cpu().numpy(), un.cpu().numpy(), wc.cpu().numpy()
    for (mk, a, b), pd in zip(spec, pairs):
        o, ns, nd = int(pd['out_off']), int(pd['n_src']), int(pd['n_dst'])
        sl = slice(o, o + ns * nd)
        ref = oracle.cost_matrix_rule(masks[mk], frames[a], frames[b], 120)
        _check((pl[sl].reshape(ns, nd), un[sl].reshape(ns, nd), wc[sl].reshape(ns, nd)),
               (ref[0].reshape(ns, nd), ref[1].reshape(ns, nd), ref[2].reshape(ns, nd)), (mk, a, b))


def test_general_path_sources(cuda):
    """Sources off the mask / in small components (deferred to the all-levels kernel) at
    max_px = 500 on a corridor mask; vs the closed form and, on a sample, the per-pair A*."""
    H, W = 700, 900
    mask = synth.make_mask(H, W, 9, pitch=(80, 170), width=(5, 12))
    mask[300:306, 400:420] = True            # an island
    mask[296:300, :] &= False
    rng = np.random.default_rng(9)
    fg = np.argwhere(mask)
    src = fg[rng.integers(0, len(fg), 24)] + rng.integers(-6, 7, (24, 2))
    src = np.vstack([src, [[302, 405], [303, 418]]])
    dst = fg[rng.integers(0, len(fg), 60)] + rng.integers(-3, 4, (60, 2)) * (rng.random((60, 1)) < 0.4)
    for a in (src, dst):
        a[:, 0] = np.clip(a[:, 0], 0, H - 1); a[:, 1] = np.clip(a[:, 1], 0, W - 1)
    src, dst = src.astype(np.int32), dst.astype(np.int32)
    got = _gpu_cost(cuda, mask, src, dst, 500)
    _check(got, oracle.cost_matrix_rule(mask, src, dst, 500))
    sub = np.arange(0, len(src), 5)
    a = oracle.cost_matrix_astar(mask, src[sub], dst, 500)
    assert np.array_equal(got[0][sub], a[0]) and np.array_equal(got[1][sub], a[1])


@pytest.mark.gpu
def test_sparse_window_multi_pass_and_dense_fallback(cuda):
    """(a) a corridor mask whose windows fit the sparse layout with more in-range targets per source
    than one pass of the sparse kernel's target hash holds (several passes, k2_sparse_kernel);
    (b) an open field at full radius: the window has more non-zero tile words than the sparse
    layout holds -> k2_fast_kernel (dense window), diamond-shaped frontiers of ~2000 cells that
    spill into the global frontier segments."""
    rng = np.random.default_rng(11)
    H, W = 1100, 1200
    mask = synth.make_mask(H, W, seed=5, pitch=(90, 200), width=(6, 16))
    fg = np.argwhere(mask)
    src = fg[rng.integers(0, len(fg), 6)].astype(np.int32)
    dst = fg[rng.integers(0, len(fg), 2500)].astype(np.int32)
    _check(_gpu_cost(cuda, mask, src, dst, 500), oracle.cost_matrix_rule(mask, src, dst, 500), 'sparse multi-pass')
    # the same with enough sources to fill the CTA slots (small windows, target hash in global scratch)
    src = fg[rng.integers(0, len(fg), 1500)].astype(np.int32)
    _check(_gpu_cost(cuda, mask, src, dst[:1200], 500), oracle.cost_matrix_rule(mask, src, dst[:1200], 500), 'sparse multi-pass, full launch')
    ones = np.ones((H, W), np.uint8)
    ones[300:800, 600] = 0
    src = np.array([[550, 560], [10, 10], [1099, 1199], [550, 640]], np.int32)
    dst = np.stack([rng.integers(0, H, 300), rng.integers(0, W, 300)], 1).astype(np.int32)
    for max_px in (500, 511, 130):
        _check(_gpu_cost(cuda, ones, src, dst, max_px), oracle.cost_matrix_rule(ones, src, dst, max_px), ('dense', max_px))
    # (c) a launch that fills the CTA slots keeps the small sparse windows (a small launch grows them
    # to what its residency leaves): 1500 open-field sources at radius 300 overflow them -> dense kernel
    src = np.stack([rng.integers(0, H, 1500), rng.integers(0, W, 1500)], 1).astype(np.int32)
    dst = np.stack([rng.integers(0, H, 40), rng.integers(0, W, 40)], 1).astype(np.int32)
    _check(_gpu_cost(cuda, ones, src, dst, 300), oracle.cost_matrix_rule(ones, src, dst, 300), 'dense, full launch')


@pytest.mark.gpu
def test_window_rim_at_largest_radii(cuda):
    """max_px > 504: the two halves of a folded window row come closer than one tile, so one folded
    tile word holds rim pixels of an upper and of a lower image row.  Targets on and just inside the
    rim of the diamond (hop max_px, max_px-1, ...) on a comb mask (sparse windows) and on an open
    field (dense windows), sources at every (x & 31, y & 3) alignment class that matters."""
    H, W = 1100, 1200
    for max_px in (505, 508, 511):
        for field in ('comb', 'open'):
            srcs = [(548 + k % 4, 590 + 5 * k) for k in range(8)] if field == 'comb' else [(549, 601), (550, 622)]
            for sy, sx in srcs:
                if field == 'comb':
                    mask = np.zeros((H, W), np.uint8)
                    mask[sy % 12::12, :] = 1          # rows every 12 through the source row
                    mask[:, sx] = 1
                else:
                    mask = np.ones((H, W), np.uint8)
                rows = np.arange(sy % 12, H, 12) if field == 'comb' else np.arange(max(sy - max_px, 0), min(sy + max_px + 1, H), 7)
                dst = []
                for y in rows:
                    hw = max_px - abs(int(y) - sy)
                    if hw < 0:
                        continue
                    for side in (-1, 1):
                        for inward in range(0, 10, 1 if field == 'comb' else 3):
                            x = sx + side * (hw - inward)
                            if hw - inward >= 0 and 0 <= x < W:
                                dst.append((y, x))
                src = np.array([[sy, sx]], np.int32)
                dst = np.array(dst, np.int32)
                _check(_gpu_cost(cuda, mask, src, dst, max_px), oracle.cost_matrix_rule(mask, src, dst, max_px),
                       (field, max_px, sy, sx))


@pytest.mark.gpu
def test_wide_radius(cuda):
    """max_px_assoc_dist is an attribute the reference honours at any value (AxonDetections.py:623-625,
    736; SURVEY 8d sweeps {500, 1000}): 512 <= max_px <= 1023 runs on sparse tile-word windows folded on a
    256-column circle (k2_sparse_wide_kernel); sources level 0 cannot answer and windows too dense for the
    layout go to the wide all-levels kernel (2048-bit folded rows, bitmaps in global scratch).  Corridor mask
    with sources on and off it, an open field with targets on the rim (dense: all-levels kernel), comb masks
    with targets on the rim (sparse), and a frame smaller than the window."""
    rng = np.random.default_rng(21)
    H, W = 1500, 1700
    mask = synth.make_mask(H, W, seed=9, pitch=(90, 200), width=(6, 16))
    fg = np.argwhere(mask)
    src = fg[rng.integers(0, len(fg), 5)].astype(np.int32)
    src = np.concatenate([src, np.array([[3, 4], [H - 1, W - 1], [700, 811]], np.int32)])
    dst = np.concatenate([fg[rng.integers(0, len(fg), 700)], np.stack([rng.integers(0, H, 40), rng.integers(0, W, 40)], 1)]).astype(np.int32)
    for max_px in (512, 777, 1000, 1023):
        _check(_gpu_cost(cuda, mask, src, dst, max_px), oracle.cost_matrix_rule(mask, src, dst, max_px), ('corridors', max_px))
    ones = np.ones((2100, 2100), np.uint8)
    for max_px, (sy, sx) in ((1023, (1040, 1051)), (1000, (1030, 1040))):
        dst = []
        for y in range(sy - max_px, sy + max_px + 1, 37):
            hw = max_px - abs(y - sy)
            for side in (-1, 1):
                for inward in (0, 1, 2, 9):
                    if hw - inward >= 0:
                        dst.append((y, sx + side * (hw - inward)))
        dst = np.array(dst, np.int32)
        src = np.array([[sy, sx]], np.int32)
        _check(_gpu_cost(cuda, ones, src, dst, max_px), oracle.cost_matrix_rule(ones, src, dst, max_px), ('rim', max_px))
    # sparse wide windows at the rim of the diamond (256-column circle): comb mask, every source alignment class
    Hc, Wc = 2100, 2200
    for max_px in (1000, 1016, 1023):
        for k in range(8):
            sy, sx = 1048 + k % 4, 1090 + 5 * k
            comb = np.zeros((Hc, Wc), np.uint8)
            comb[sy % 24::24, :] = 1
            comb[:, sx] = 1
            dst = []
            for y in np.arange(sy % 24, Hc, 24):
                hw = max_px - abs(int(y) - sy)
                for side in (-1, 1):
                    for inward in range(0, 10):
                        x = sx + side * (hw - inward)
                        if hw - inward >= 0 and 0 <= x < Wc:
                            dst.append((y, x))
            src = np.array([[sy, sx]], np.int32)
            dst = np.array(dst, np.int32)
            _check(_gpu_cost(cuda, comb, src, dst, max_px), oracle.cost_matrix_rule(comb, src, dst, max_px), ('comb', max_px, sy, sx))
    small = (rng.random((90, 130)) < 0.7).astype(np.uint8)
    pts = np.stack([rng.integers(0, 90, 30), rng.integers(0, 130, 30)], 1).astype(np.int32)
    _check(_gpu_cost(cuda, small, pts[:10], pts, 900), oracle.cost_matrix_rule(small, pts[:10], pts, 900), 'small frame')


@pytest.mark.gpu
def test_decode_nms_wide_kernel_equals_one_cta_per_frame(cuda):
    """Few frames per launch run on the cooperative wide kernel (a frame spread over many CTAs), many frames on one CTA
    per frame: every output identical, also with conf ties inside a tile and across tiles, an over-full frame
    (AXB_ERR_CAPACITY) and an empty frame."""
    import os
    from axtrack_b200 import ops, _lib
    rng = np.random.default_rng(4)
    H, W = 3072, 2048
    tt = synth.tile_table(H, W)
    frames = []
    for k, n in enumerate((1500, 0, 400, 3000)):
        dets = np.stack([rng.integers(0, H, n), rng.integers(0, W, n)], 1)
        Y = synth.make_head_tensor(dets, H, W, seed=k)[0]
        if k == 2:
            Y[..., 0] = np.round(Y[..., 0], 1)                    # many exact conf ties
        frames.append(Y)
    Yd = cuda.from_numpy(np.stack(frames)).cuda()
    td = cuda.from_numpy(tt).cuda()
    outs = []
    for env in (None, '1'):
        if env is None:
            os.environ.pop('AXB_K1_NO_WIDE', None)
        else:
            os.environ['AXB_K1_NO_WIDE'] = env
        o = ops.decode_nms(Yd, td)
        ops.raise_status()
        outs.append({k: v.cpu().numpy() for k, v in o.items()})
        small = ops.decode_nms(Yd, td, cap=500, want_pre=False)    # frames 0 and 3 hold more candidates than cap
        with pytest.raises(_lib.AxbError):
            ops.raise_status()
        assert small['n_out'].cpu().numpy().tolist()[1] == 0 and small['n_out'].cpu().numpy()[0] == 0
    os.environ.pop('AXB_K1_NO_WIDE', None)
    a, b = outs
    assert np.array_equal(a['n_pre'], b['n_pre']) and np.array_equal(a['n_out'], b['n_out']) and a['n_out'][1] == 0
    for f in range(4):
        npre, nout = int(a['n_pre'][f]), int(a['n_out'][f])
        for k in ('pre_conf', 'pre_yx', 'pre_cell'):
            assert np.array_equal(a[k][f, :npre], b[k][f, :npre]), (f, k)
        for k in ('conf', 'yx', 'keep_idx'):
            assert np.array_equal(a[k][f, :nout], b[k][f, :nout]), (f, k)
        want = oracle.decode_stitch_nms(frames[f], tt)
        assert np.array_equal(a['keep_idx'][f, :nout], want['keep_idx']), f


@pytest.mark.gpu
def test_decode_nms_8192_frames(cuda):
    """BASELINE configs[4] shape: 8192x8192 frames = 256 tiles x 12x12 cells, thousands of candidates
    per frame (the sort leaves shared memory), several frames per launch; survivor indices,
    coordinates and confidences bit-exact vs the numpy port of the reference's pandas code."""
    from axtrack_b200 import ops
    H = W = 8192
    tt = synth.tile_table(H, W)
    rng = np.random.default_rng(3)
    Ys, wants = [], []
    for n_det in (50, 2500, 9000):
        dets = np.stack([rng.integers(0, H, n_det), rng.integers(0, W, n_det)], 1)
        Y = synth.make_head_tensor(dets, H, W, seed=n_det)[0]
        Ys.append(Y); wants.append(oracle.decode_stitch_nms(Y, tt))
    o = ops.decode_nms(cuda.from_numpy(np.stack(Ys)).cuda(), cuda.from_numpy(tt).cuda())
    ops.raise_status()
    for f, w in enumerate(wants):
        n = int(o['n_out'][f])
        assert n == len(w['keep_idx']) and int(o['n_pre'][f]) == len(w['pre_conf'])
        assert np.array_equal(o['keep_idx'][f, :n].cpu().numpy(), w['keep_idx'])
        yx = o['yx'][f, :n].cpu().numpy()
        assert np.array_equal(yx[:, 0], w['y']) and np.array_equal(yx[:, 1], w['x'])
        assert np.array_equal(o['conf'][f, :n].cpu().numpy(), w['conf'])
