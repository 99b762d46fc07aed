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


class PeerSlabs:
    """Gather without a payload transfer: rank 0's HBM holds one result slab per rank (symmetric
    memory, peer-mapped over NVLink / NVSwitch) and every rank's K2 kernels store their path-length
    and flag matrices straight into their slab, so the gather is the kernels' own epilogue.  A step
    ends with one 16-byte all-reduce (which also orders rank 0's reads after every peer's stores)
    instead of world-1 receives of tens of MB after the kernels.  CUDA + NCCL only.

    Per step and rank: ``out_alloc`` (views of the rank's slab for the kernels), ``publish`` (pair
    shapes + counts from pinned staging, then the all-reduce); rank 0 then ``collect``s
    [(path_len, unreachable, pair_shapes)] per rank from its own memory.

    Sizing.  ``collective_sizing`` = True (default; a one-shot ``associate_sharded``): ``out_alloc``
    first agrees on the capacity with one more 16-byte all-reduce + host read (``ensure``) and grows
    the slabs collectively when some rank needs more.  A caller that repeats steps of a known size
    (bench.py after its warm-up) sets it to False: ``out_alloc`` then only checks the rank's own need
    against the capacity -- no collective and no host sync before the kernels -- and raises if the
    slab is too small."""

    def __init__(self, device, group=None, cap_elems=1 << 20, cap_pairs=1024):
        self.device, self.group = device, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.cap_elems = self.cap_pairs = 0
        self.collective_sizing = True
        self._need = torch.zeros(2, dtype=torch.int64, device=device)
        self._need_host = torch.zeros(2, dtype=torch.int64).pin_memory()
        self._step = 0
        self._alloc(cap_elems, cap_pairs)

    def _alloc(self, cap_elems, cap_pairs):
        import torch.distributed._symmetric_memory as symm
        al = lambda n: (int(n) + 255) & ~255
        self.cap_elems, self.cap_pairs = int(cap_elems), int(cap_pairs)
        self.o_un = al(4 * self.cap_elems)
        self.o_shp = self.o_un + al(self.cap_elems)
        self.o_meta = self.o_shp + al(16 * self.cap_pairs)
        self.slab = self.o_meta + 256
        nbytes = self.slab * self.world
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=self.device)      # rank 0's copy is the one in use
        grp = self.group if self.group is not None else dist.group.WORLD
        self.hdl = symm.rendezvous(self.buf, grp.group_name)
        self.rank0 = self.buf if self.rank == 0 else self.hdl.get_buffer(0, (nbytes,), torch.uint8)
        # pinned staging of (pair shapes, meta), double buffered: a step's copy may still be in flight
        # while the host prepares the next one
        self._stage = [torch.zeros(2 * self.cap_pairs + 2, dtype=torch.int64).pin_memory() for _ in range(2)]

    def _views(self, base, r):
        s = base[r * self.slab:(r + 1) * self.slab]
        return (s[:4 * self.cap_elems].view(torch.int32), s[self.o_un:self.o_un + self.cap_elems],
                s[self.o_shp:self.o_shp + 16 * self.cap_pairs].view(torch.int64).view(-1, 2),
                s[self.o_meta:self.o_meta + 16].view(torch.int64))

    def ensure(self, total, n_pairs, headroom=0.25):
        """COLLECTIVE.  Makes every slab hold max-over-ranks (total, n_pairs) (+ headroom when it has to
        grow); one 16-byte all-reduce and a host read of its result."""
        self._need_host[0], self._need_host[1] = int(total), int(n_pairs)
        self._need.copy_(self._need_host, non_blocking=True)
        dist.all_reduce(self._need, op=dist.ReduceOp.MAX, group=self.group)
        need_e, need_p = (int(v) for v in self._need.cpu())
        if need_e > self.cap_elems or need_p > self.cap_pairs:
            torch.cuda.synchronize(self.device)
            dist.barrier(self.group)                  # nobody still writes into the old slabs
            self._alloc(max(int(need_e * (1 + headroom)), self.cap_elems), max(2 * need_p, self.cap_pairs))

    reserve = ensure          # round-1 name

    def out_alloc(self, total, pairs):
        if self.collective_sizing:
            self.ensure(total, len(pairs))
        elif int(total) > self.cap_elems or len(pairs) > self.cap_pairs:
            raise RuntimeError(f'PeerSlabs: this rank needs {int(total)} elements / {len(pairs)} pairs but the slabs hold '
                               f'{self.cap_elems} / {self.cap_pairs}; size them with ensure() (collective) first')
        pl, un, _, _ = self._views(self.rank0, self.rank)
        return pl, un, None

    def publish(self, total, pair_shapes):
        """pair_shapes int64 [n_pairs, 2] (host).  Ends the step on this rank: shapes and counts go to
        the rank's slab from pinned memory (asynchronous, stream ordered after the kernels), then the
        all-reduce that orders rank 0's reads after every rank's stores."""
        _, _, shp, meta = self._views(self.rank0, self.rank)
        k = int(pair_shapes.shape[0])
        st = self._stage[self._step & 1]
        self._step += 1
        if k:
            st[:2 * k].view(-1, 2).copy_(torch.as_tensor(pair_shapes, dtype=torch.int64).reshape(-1, 2))
            shp[:k].copy_(st[:2 * k].view(-1, 2), non_blocking=True)
        st[2 * self.cap_pairs], st[2 * self.cap_pairs + 1] = int(total), k
        meta.copy_(st[2 * self.cap_pairs:2 * self.cap_pairs + 2], non_blocking=True)
        dist.all_reduce(self._need, op=dist.ReduceOp.MAX, group=self.group)   # stream-ordered after every rank's stores

    def collect(self):
        """rank 0, after ``publish``: [(path_len, unreachable, pair_shapes)] per rank (views of the slabs)."""
        assert self.rank == 0
        out = []
        metas = torch.stack([self._views(self.buf, r)[3] for r in range(self.world)]).cpu()
        for r in range(self.world):
            pl, un, shp, _ = self._views(self.buf, r)
            n, k = int(metas[r, 0]), int(metas[r, 1])
            out.append((pl[:n], un[:n], shp[:k]))
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


def block_inputs(block, Y, mask_of_frame=None):
    """What one rank feeds to ``associate_frames`` for its ``partition`` block: the head tensors of its
    frames (halo included) and -- with per-frame masks -- the slice of ``mask_of_frame`` that belongs to
    them.  ``build_pairs`` indexes ``mask_of_frame`` with the block-LOCAL frame index, so the global
    table must be cut exactly like Y."""
    first, own, stop = block
    mof = None if mask_of_frame is None else np.asarray(mask_of_frame)[first:stop]
    return Y[first:stop], own - first, mof


def associate_block(block, Y, tile_yx, mask, max_px=500, max_num_misses=1, mask_of_frame=None, out_alloc=None, **kw):
    """The association path of one timepoint block (``partition``) on the current GPU."""
    from . import pipeline
    Yb, t0, mof = block_inputs(block, Y, mask_of_frame)
    return pipeline.associate_frames(Yb, tile_yx, mask, max_px=max_px, max_num_misses=max_num_misses, t0=t0,
                                     want_cost=False, mask_of_frame=mof, out_alloc=out_alloc, **kw)


def associate_sharded(Y, tile_yx, mask, max_px=500, max_num_misses=1, group=None, slabs=None, mask_of_frame=None,
                      to_host=True, **kw):
    """Each rank passes the WHOLE head tensor Y [T, ...] (or at least its own block + halo rows at
    their global positions) and processes its block on its own GPU; rank 0 gets the result.
    ``mask_of_frame`` [T] (global frame -> mask index) goes with per-frame masks [n_masks, H, W].
    ``slabs`` (a PeerSlabs, CUDA + NCCL): the kernels write into rank 0's memory directly; without
    it the matrices are gathered with NCCL / gloo sends after the kernels.
    ``to_host``: True -> rank 0 gets the assembled {(t, t_bef): (path_len, unreachable)} dict of host
    arrays; False -> the gathered per-rank tensors as they lie in rank 0's memory (``assemble`` or
    ``HostGather`` bring them to the host); a ``HostGather`` -> copied into its pinned buffers.
    -> (rank 0's result | None, this rank's Association)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    T = Y.shape[0]
    blocks = partition(T, world, max_num_misses)
    res = associate_block(blocks[rank], Y, tile_yx, mask, max_px, max_num_misses, mask_of_frame,
                          out_alloc=slabs.out_alloc if slabs is not None else None, **kw)
    shapes = torch.from_numpy(np.stack([res.pairs['n_src'], res.pairs['n_dst']], 1).astype(np.int64).reshape(-1, 2))
    if slabs is not None:
        slabs.publish(res.path_len.numel(), shapes)
        g = slabs.collect() if rank == 0 else None
    else:
        g = gather_to_rank0(res.path_len, res.unreachable, shapes, group)
    if rank != 0:
        return None, res
    if isinstance(to_host, HostGather):
        return to_host.fetch(g, blocks, max_num_misses), res
    return (assemble(g, blocks, max_num_misses) if to_host else g), res


class HostGather:
    """rank 0: the gathered matrices device -> pinned host memory with asynchronous copies and ONE
    synchronisation (what a tracker that consumes the dict on the host needs; bench.py's N > 1 ``e2e``)."""

    def __init__(self):
        self.pl = self.un = None

    def fetch(self, gathered, blocks, max_num_misses=1):
        n = sum(int(pl.numel()) for pl, _, _ in gathered)
        if self.pl is None or self.pl.numel() < n:
            self.pl = torch.empty(max(n, 1), dtype=torch.int32).pin_memory()
            self.un = torch.empty(max(n, 1), dtype=torch.uint8).pin_memory()
        off, parts = 0, []
        for pl, un, shp in gathered:
            k = int(pl.numel())
            self.pl[off:off + k].copy_(pl, non_blocking=True)
            self.un[off:off + k].copy_(un, non_blocking=True)
            parts.append((off, k, shp.cpu() if shp.is_cuda else shp))
            off += k
        if gathered and gathered[0][0].is_cuda:
            torch.cuda.current_stream(gathered[0][0].device).synchronize()
        hp, hu = self.pl.numpy(), self.un.numpy()
        res = {}
        for (o, k, shp), block in zip(parts, blocks):
            shp = shp.numpy()
            for key, (ns, nd) in zip(global_keys(block, max_num_misses), shp):
                m = int(ns) * int(nd)
                res[key] = (hp[o:o + m].reshape(int(ns), int(nd)), hu[o:o + m].reshape(int(ns), int(nd)))
                o += m
        self.bytes = 5 * n
        return res
