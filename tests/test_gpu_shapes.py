"""GPU-vs-oracle parity at the shapes BASELINE.json names (VERDICT r1 "weak" #1, north_star N2):

* configs[0]  1969 x 3868, 2 frames: W % 32 = 28 and W % 16 = 12 (scalar branch of the mask packer,
              ragged last tile column in K1 and in the K2 window staging), H % 4 = 1; ALL rows.
* configs[2]  4096 x 4096, ~800 detections per frame, 3 frames = 3 frame pairs including the
              (t, t-2) chain; >= 64 seeded source rows of every pair.
* configs[3]  a 300-frame block of the long 2048 x 2048 timelapse (597 chained frame pairs in ONE
              launch, ~24 M output entries): sampled rows of the first, a middle and the last pairs,
              so that the chained out_off arithmetic is covered.

Everything goes through the C ABI (axb_decode_nms + axb_pack_mask + axb_cost_matrix_batch, and for
configs[0] also the host-buffer entry points); the checker is the closed-form oracle, itself pinned
to the compiled reference at these sizes in tests/test_oracle_golden.py.  Bit-exact."""
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


def _block(H, W, T, n_dets, seed):
    mask = synth.make_mask(H, W, seed)
    tracks = synth.make_tracks(mask, n_dets, T, seed)
    tt = synth.tile_table(H, W)
    Y = np.stack([synth.make_head_tensor(tracks[t], H, W, seed + 50 + t)[0] for t in range(T)])
    return mask, Y, tt


def _associate(torch, mask, Y, tt, **kw):
    from axtrack_b200 import pipeline
    res = pipeline.associate_frames(torch.from_numpy(Y).cuda(), torch.from_numpy(tt).cuda(),
                                    torch.from_numpy(mask.astype(np.uint8)).cuda(), max_px=500, max_num_misses=1, **kw)
    torch.cuda.synchronize()
    return res


def _check_decode(res, Y, tt):
    dets = []
    for t in range(len(Y)):
        o = oracle.decode_stitch_nms(Y[t], tt)
        d = res.frame_dets(t)
        assert np.array_equal(d['keep_idx'], o['keep_idx']), t
        assert np.array_equal(d['x'], o['x']) and np.array_equal(d['y'], o['y']) and np.array_equal(d['conf'], o['conf']), t
        dets.append(np.stack([o['y'], o['x']], 1).astype(np.int32))
    return dets


def _check_rows(mat, mask, src, dst, rows, tag):
    pl, un, wc = mat
    rpl, run, rwc = oracle.cost_matrix_rule(mask, src[rows], dst, 500)
    assert np.array_equal(pl[rows], rpl), tag
    assert np.array_equal(un[rows], run), tag
    if wc is not None:
        ok = run == 0
        assert np.allclose(wc[rows][ok], rwc[ok], rtol=1e-5, atol=0), tag
    return int((run == 0).sum())


def test_config0_shape_all_rows(cuda):
    H, W = 1969, 3868
    mask, Y, tt = _block(H, W, 2, 250, seed=0)
    assert W % 32 == 28 and len(tt) == 4 * 8
    res = _associate(cuda, mask, Y, tt)
    dets = _check_decode(res, Y, tt)
    assert min(len(d) for d in dets) > 150
    assert (np.concatenate(dets)[:, 1] >= 3840).any(), 'detections in the ragged last tile column'
    mats = res.pair_matrices()
    assert list(mats) == [(1, 0)]
    found = _check_rows(mats[(1, 0)], mask, dets[0], dets[1], np.arange(len(dets[0])), 'c1')
    assert found > 500
    # the same through the host-buffer entry points (H2D / D2H inside the library)
    from axtrack_b200 import host
    ha = host.HostAssociator(0)
    hm = ha.cost_matrices(mask.astype(np.uint8), dets, 500, 1)
    assert np.array_equal(hm[(1, 0)][0], mats[(1, 0)][0]) and np.array_equal(hm[(1, 0)][1], mats[(1, 0)][1])


def test_config2_shape_sampled_rows(cuda):
    H = W = 4096
    mask, Y, tt = _block(H, W, 3, 1000, seed=0)
    res = _associate(cuda, mask, Y, tt)
    dets = _check_decode(res, Y, tt)
    assert min(len(d) for d in dets) > 600
    mats = res.pair_matrices()
    assert list(mats) == [(1, 0), (2, 1), (2, 0)]
    rng = np.random.default_rng(7)
    found = 0
    for (t, tb), m in mats.items():
        rows = np.sort(rng.choice(len(dets[tb]), 72, replace=False))
        found += _check_rows(m, mask, dets[tb], dets[t], rows, (t, tb))
    assert found > 2000


def test_config3_long_block_sampled_rows(cuda):
    H = W = 2048
    T = 300
    mask, Y, tt = _block(H, W, T, 250, seed=1)
    res = _associate(cuda, mask, Y, tt, want_cost=False)
    from axtrack_b200.pipeline import frame_pair_keys
    keys = frame_pair_keys(T, 1)
    assert res.keys == keys and len(keys) == 2 * T - 3
    assert int(res.path_len.numel()) > 15_000_000
    mats = res.pair_matrices()
    rng = np.random.default_rng(3)
    dets = {}

    def det(t):
        if t not in dets:
            o = oracle.decode_stitch_nms(Y[t], tt)
            d = res.frame_dets(t)
            assert np.array_equal(d['keep_idx'], o['keep_idx']), t
            dets[t] = np.stack([o['y'], o['x']], 1).astype(np.int32)
        return dets[t]
    found = 0
    for key in (keys[0], keys[1], keys[2], keys[len(keys) // 2], keys[len(keys) // 2 + 1], keys[-2], keys[-1]):
        t, tb = key
        rows = np.sort(rng.choice(len(det(tb)), 10, replace=False))
        found += _check_rows(mats[key], mask, det(tb), det(t), rows, key)
    assert found > 300
    # size-independent properties over the WHOLE block: None <=> 500, lengths within [1, 500]
    pl = res.path_len.cpu().numpy()
    un = res.unreachable.cpu().numpy()
    assert ((pl >= 1) & (pl <= 500)).all() and (pl[un == 1] == 500).all() and set(np.unique(un)) <= {0, 1}
