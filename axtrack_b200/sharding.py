"""Timepoint sharding of the association path across the GPUs of one box (one process per GPU).

Frame pairs are independent (AxonDetections.py:553-566 only reads mask[t] and the detections of
t and t_bef), so rank r owns a contiguous block of timepoints plus a halo of
``max_num_misses + 1`` preceding frames (sources only).  There is no data-path collective; the
only exchange is the gather of the per-pair matrices to rank 0, where the reference's unchanged
host min-cost-flow runs (AxonDetections.py:663-690).  Works on NCCL (CUDA tensors) and on gloo
(CPU tensors; used by the world_size-2 CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist


def partition(T, world, max_num_misses=1):
    """[(first_frame_with_halo, first_owned_frame, stop)] per rank: contiguous, near-equal blocks.
    Rank r computes the pairs (t, t_bef) with first_owned <= t < stop."""
    halo = max_num_misses + 1
    bounds = [(T * r) // world for r in range(world + 1)]
    return [(max(0, bounds[r] - halo), bounds[r], bounds[r + 1]) for r in range(world)]


def global_keys(block, max_num_misses=1):
    """(t, t_bef) keys rank-block ``block`` computes, in the reference's loop order."""
    _, own, stop = block
    return [(t, tb) for t in range(own, stop) for tb in range(t - 1, t - (max_num_misses + 2), -1) if tb >= 0]


def gather_to_rank0(path_len, unreachable, pair_shapes, group=None):
    """Gather flat per-rank results to rank 0.

    path_len int32 / unreachable uint8: flat tensors (this rank's pairs back to back);
    pair_shapes: int64 tensor [n_pairs, 2] (n_pred, n_cur) of this rank's pairs.
    Returns on rank 0 a list over ranks of (path_len, unreachable, pair_shapes) tensors, else None.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = path_len.device
    pair_shapes = pair_shapes.to(dev, torch.int64).reshape(-1, 2)
    meta = torch.tensor([path_len.numel(), pair_shapes.shape[0]], dtype=torch.int64, device=dev)
    metas = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(metas, meta, group=group)
    if world == 1:
        return [(path_len, unreachable, pair_shapes)]
    ops, out = [], None
    if rank == 0:
        out = [(path_len, unreachable, pair_shapes)]
        for r in range(1, world):
            n, k = int(metas[r][0]), int(metas[r][1])
            bufs = (torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.uint8, device=dev),
                    torch.empty((k, 2), dtype=torch.int64, device=dev))
            out.append(bufs)
            for b in bufs:
                if b.numel():
                    ops.append(dist.P2POp(dist.irecv, b, r, group))
    else:
        for b in (path_len.contiguous(), unreachable.contiguous(), pair_shapes.contiguous()):
            if b.numel():
                ops.append(dist.P2POp(dist.isend, b, 0, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return out


def assemble(gathered, blocks, max_num_misses=1):
    """rank 0: {(t, t_bef): (path_len [n_pred,n_cur] int32 ndarray, unreachable uint8 ndarray)} over the
    whole timelapse, from ``gather_to_rank0`` output and the partition."""
    res = {}
    for (pl, un, shp), block in zip(gathered, blocks):
        pl, un, shp = pl.cpu().numpy(), un.cpu().numpy(), shp.cpu().numpy()
        off = 0
        for key, (ns, nd) in zip(global_keys(block, max_num_misses), shp):
            n = int(ns) * int(nd)
            res[key] = (pl[off:off + n].reshape(int(ns), int(nd)), un[off:off + n].reshape(int(ns), int(nd)))
            off += n
    return res


def associate_sharded(Y, tile_yx, mask, max_px=500, max_num_misses=1, group=None, **kw):
    """Each rank passes the WHOLE head tensor Y [T, ...] (or at least its own block + halo rows at
    their global positions) and processes its block on its own GPU; rank 0 gets the assembled dict."""
    from . import pipeline
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    T = Y.shape[0]
    blocks = partition(T, world, max_num_misses)
    first, own, stop = blocks[rank]
    res = pipeline.associate_frames(Y[first:stop], tile_yx, mask, max_px=max_px, max_num_misses=max_num_misses,
                                    t0=own - first, want_cost=False, **kw)
    shapes = torch.from_numpy(np.stack([res.pairs['n_src'], res.pairs['n_dst']], 1).astype(np.int64).reshape(-1, 2))
    g = gather_to_rank0(res.path_len, res.unreachable, shapes, group)
    return (assemble(g, blocks, max_num_misses) if rank == 0 else None), res
