"""Host logic of the timepoint sharding (N > 1 path) on CPU: partition, key order, and the
gather of per-pair matrices to rank 0 over gloo with world_size 2 and 3.  The per-rank "compute"
here is the oracle (checker) so that the assembled result can be compared with a single-rank
run; the CUDA path is covered by the -m gpu tests and bench.py --gpus N."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from axtrack_b200 import sharding, synth
from axtrack_b200.host import frame_pair_keys


def test_partition_covers_all_pairs():
    for T in (1, 2, 3, 7, 100, 2000):
        for world in (1, 2, 3, 4, 8):
            for mm in (0, 1, 3):
                blocks = sharding.partition(T, world, mm)
                keys = [k for b in blocks for k in sharding.global_keys(b, mm)]
                assert keys == frame_pair_keys(T, mm), (T, world, mm)
                for first, own, stop in blocks:
                    assert first == max(0, own - (mm + 1)) and own <= stop


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, T, out_path):
    import oracle
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        mask = synth.make_mask(200, 240, 4, pitch=(40, 80), width=(4, 9))
        masks = [mask, synth.make_mask(200, 240, 9, pitch=(40, 80), width=(4, 9))]     # per-frame masks
        mask_of_frame = np.array([(t // 2) % 2 for t in range(T)])
        tracks = synth.make_tracks(mask, 12, T, 4, off_mask_frac=0.2)
        tracks[2] = np.zeros((0, 2), np.int32)          # an empty frame
        blocks = sharding.partition(T, world, 1)
        first = blocks[rank][0]
        _, t0, mof = sharding.block_inputs(blocks[rank], np.arange(T), mask_of_frame)    # what the rank's kernels get
        assert t0 == blocks[rank][1] - first
        pls, uns, shapes = [], [], []
        for t, tb in sharding.global_keys(blocks[rank], 1):
            pl, un, _ = oracle.cost_matrix_rule(masks[mof[t - first]], tracks[tb], tracks[t], 60)   # block-local index, as build_pairs
            pls.append(pl.reshape(-1)); uns.append(un.reshape(-1)); shapes.append((len(tracks[tb]), len(tracks[t])))
        cat = lambda xs, dt: torch.from_numpy(np.concatenate(xs).astype(dt)) if xs else torch.zeros(0, dtype=torch.from_numpy(np.zeros(0, dt)).dtype)
        g = sharding.gather_to_rank0(cat(pls, np.int32), cat(uns, np.uint8),
                                     torch.tensor(shapes, dtype=torch.int64).reshape(-1, 2))
        if rank == 0:
            res = sharding.assemble(g, blocks, 1)
            assert list(res) == frame_pair_keys(T, 1)
            for (t, tb), (pl, un) in res.items():
                rpl, run, _ = oracle.cost_matrix_rule(masks[mask_of_frame[t]], tracks[tb], tracks[t], 60)   # the frame's own mask (AxonDetections.py:557)
                assert np.array_equal(pl, rpl) and np.array_equal(un, run), (t, tb)
            open(out_path, 'w').write('ok %d' % len(res))
        else:
            assert g is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_gather_gloo(world, tmp_path):
    out = str(tmp_path / 'ok.txt')
    T = 7
    mp.spawn(_worker, args=(world, _free_port(), T, out), nprocs=world, join=True)
    assert open(out).read() == 'ok %d' % len(frame_pair_keys(T, 1))


def test_block_inputs_cut_the_mask_table_like_the_frames():
    T = 11
    Y = np.arange(T)
    mof = np.arange(T) % 3
    for world in (1, 2, 3, 4):
        for block in sharding.partition(T, world, 1):
            Yb, t0, m = sharding.block_inputs(block, Y, mof)
            assert len(Yb) == len(m) == block[2] - block[0] and t0 == block[1] - block[0]
            assert np.array_equal(m, mof[Yb])              # local frame l is global frame Yb[l]
            assert sharding.block_inputs(block, Y, None)[2] is None


@pytest.mark.gpu
def test_sharded_blocks_with_per_frame_masks():
    """Every rank's block of a timelapse with TWO masks (per-frame mask table) on one GPU, assembled as
    rank 0 does, against the unsharded result (the table must be cut like the frames)."""
    if not torch.cuda.is_available():
        pytest.fail('-m gpu tests need a CUDA device')
    from axtrack_b200 import pipeline
    dev = torch.device('cuda', 0)
    H, W, T = 600, 700, 9
    masks = np.stack([synth.make_mask(H, W, seed=3), synth.make_mask(H, W, seed=8)]).astype(np.uint8)
    mof = np.array([0, 0, 1, 1, 1, 0, 1, 0, 0], np.int32)
    tt = synth.tile_table(H, W)
    rng = np.random.default_rng(5)
    fg = np.argwhere(masks[0] & masks[1])
    Y = np.stack([synth.make_head_tensor(fg[rng.integers(0, len(fg), 50)], H, W, seed=20 + t)[0] for t in range(T)])
    Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(masks).to(dev)
    ref = pipeline.associate_frames(Yd, td, md, want_cost=False, mask_of_frame=mof).pair_matrices()
    one = pipeline.associate_frames(Yd, td, md[:1], want_cost=False).pair_matrices()
    assert any(not np.array_equal(ref[k][0], one[k][0]) for k in ref), 'the two masks must matter'
    for world in (2, 3):
        blocks = sharding.partition(T, world, 1)
        g = []
        for b in blocks:
            r = sharding.associate_block(b, Yd, td, md, mask_of_frame=mof)
            g.append((r.path_len, r.unreachable,
                      torch.from_numpy(np.stack([r.pairs['n_src'], r.pairs['n_dst']], 1).astype(np.int64).reshape(-1, 2))))
        got = sharding.assemble(g, blocks, 1)
        assert list(got) == list(ref)
        for k in ref:
            assert np.array_equal(got[k][0], ref[k][0]) and np.array_equal(got[k][1], ref[k][1]), (world, k)


def _peer_worker(rank, world, port, out_path):
    """One NCCL rank on cuda:0: PeerSlabs as the kernels' output (slab layout, collective growth,
    publish / collect) against the plain single-process result."""
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(0)
    dev = torch.device('cuda', 0)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        from axtrack_b200 import pipeline
        H, W, T = 700, 800, 6
        mask = synth.make_mask(H, W, seed=3)
        tt = synth.tile_table(H, W)
        rng = np.random.default_rng(2)
        fg = np.argwhere(mask)
        Y = np.stack([synth.make_head_tensor(fg[rng.integers(0, len(fg), 60)], H, W, seed=10 + t)[0] for t in range(T)])
        Yd, td = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev)
        md = torch.from_numpy(mask.astype(np.uint8)).to(dev)
        ref = pipeline.associate_frames(Yd, td, md, want_cost=False).pair_matrices()
        slabs = sharding.PeerSlabs(dev, cap_elems=64, cap_pairs=2)            # far too small: must grow
        for _ in range(2):
            got, _ = sharding.associate_sharded(Yd, td, md, slabs=slabs)
            assert set(got) == set(ref) and len(got) == 2 * T - 3
            for k in ref:
                assert np.array_equal(got[k][0], ref[k][0]) and np.array_equal(got[k][1], ref[k][1]), k
        assert slabs.cap_elems >= sum(v[0].size for v in ref.values())
        open(out_path, 'w').write('ok %d' % len(got))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_peer_slabs_single_rank(tmp_path):
    """The N > 1 product path (kernels store into rank 0's symmetric-memory slab) with world_size 1
    on one GPU; the 2- and 8-GPU runs are tools/peer_check.py and bench.py --gpus N."""
    if not torch.cuda.is_available():
        pytest.fail('-m gpu tests need a CUDA device')
    out = str(tmp_path / 'ok.txt')
    mp.spawn(_peer_worker, args=(1, _free_port(), out), nprocs=1, join=True)
    assert open(out).read().startswith('ok ')
