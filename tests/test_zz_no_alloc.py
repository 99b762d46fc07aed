"""axb_tile_nonempty_noalloc: the K6 non-empty flags with the vote flags in a static device array of the library
(no cudaMalloc on the call path).  No counterpart in the reference; the bar is equality with the existing entry point
axb_tile_nonempty (bitwise: the result is uint8 flags) and with the numpy oracle, on two calls in a row that share the
static array.  Collected last (file name) so that a failure here hides no other test under `pytest -x`."""
import ctypes

import numpy as np
import pytest

import oracle


def _make_X(seed, T, H, W, density):
    rng = np.random.default_rng(seed)
    X = np.zeros((T, 2, H, W), np.float32)
    X[:, 1] = rng.standard_normal((T, H, W)).astype(np.float32)        # channel 1 must not matter
    hits = rng.random((T, H, W)) < density
    X[:, 0][hits] = rng.random(int(hits.sum()), dtype=np.float32) + 0.01
    X[:, 0][rng.random((T, H, W)) < density] = -1.0                     # negative pixels are not "> 0"
    return X


def _call(ctx, name, Xd, ts, out):
    import torch
    from axtrack_b200 import ops
    T, C, H, W = Xd.shape
    fn = getattr(ctx.lib, name)
    return fn(ctx.handle, Xd.data_ptr(), int(T), int(C), int(H), int(W), int(ts), out.data_ptr(), ops._stream(Xd.device))


@pytest.mark.gpu
def test_noalloc_matches_existing_entry_point_twice():
    import torch
    from axtrack_b200 import ops
    dev = torch.device('cuda', 0)
    ctx = ops.get_context(dev)
    # two different calls in a row: the second is smaller and emptier, so flags left over from the first would show
    cases = [(_make_X(1, 6, 1100, 1300, 2e-6), 512), (_make_X(2, 3, 700, 520, 4e-6), 256)]
    prepared = []
    for X, ts in cases:
        Xd = torch.from_numpy(X).to(dev)
        T, _, H, W = X.shape
        n = (-(-H // ts)) * (-(-W // ts))
        prepared.append((X, ts, Xd, torch.full((n, T), 7, dtype=torch.uint8, device=dev), torch.full((n, T), 9, dtype=torch.uint8, device=dev)))
    for rep in range(2):                                  # the whole sequence twice: state carried across calls
        for X, ts, Xd, out_ref, out_new in prepared:
            out_ref.fill_(7); out_new.fill_(9)
            ctx.check(_call(ctx, 'axb_tile_nonempty', Xd, ts, out_ref))
            torch.cuda.synchronize(dev)
            free0 = torch.cuda.mem_get_info(dev)[0]
            ctx.check(_call(ctx, 'axb_tile_nonempty_noalloc', Xd, ts, out_new))
            torch.cuda.synchronize(dev)
            free1 = torch.cuda.mem_get_info(dev)[0]
            if rep == 1:
                assert free1 >= free0, 'device memory was allocated on the no-alloc path'
            assert torch.equal(out_new, out_ref)
            want = oracle.tile_frames(X, ts)[1]           # bool [ytiles, xtiles, T]
            assert np.array_equal(out_new.cpu().numpy().astype(bool).reshape(want.shape), want)
            assert 0 < int(out_new.sum()) < out_new.numel()      # the case exercises both flag values
    ops.raise_status(device=dev)


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_static_array():
    import torch
    from axtrack_b200 import ops, _lib
    dev = torch.device('cuda', 0)
    ctx = ops.get_context(dev)
    Xd = torch.zeros((1, 1, 2, 2), dtype=torch.float32, device=dev)
    out = torch.zeros(16, dtype=torch.uint8, device=dev)
    # 2x2 image in 1-pixel tiles = 4 tiles; 20 000 frames -> 80 000 (tile, frame) flags > 65 536: rejected before any access
    rc = ctx.lib.axb_tile_nonempty_noalloc(ctx.handle, Xd.data_ptr(), 20000, 1, 2, 2, 1, out.data_ptr(), ops._stream(dev))
    assert rc == _lib.ERR_BAD_ARG
    msg = ctx.lib.axb_last_error(ctx.handle).decode()
    assert 'AXB_NOALLOC_MAX_TILE_FRAMES' in msg and '80000' in msg
    # the largest size that fits is accepted: 4 tiles x 16 384 frames = 65 536 flags
    T = 16384
    Xb = torch.zeros((T, 1, 2, 2), dtype=torch.float32, device=dev)
    Xb[5, 0, 1, 0] = 1.0
    Xb[T - 1, 0, 0, 1] = 2.0
    ob = torch.full((4, T), 3, dtype=torch.uint8, device=dev)
    ctx.check(ctx.lib.axb_tile_nonempty_noalloc(ctx.handle, Xb.data_ptr(), T, 1, 2, 2, 1, ob.data_ptr(), ops._stream(dev)))
    torch.cuda.synchronize(dev)
    assert int(ob.sum()) == 2 and int(ob[2, 5]) == 1 and int(ob[1, T - 1]) == 1
