#!/usr/bin/env python
"""Benchmark of the AxTrack association hot path (BASELINE.json metric: frame-pairs/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2|c3|c1]

One STEP = one pass of the hot path over one timelapse block: fused decode+threshold+NMS of all
frames of the block (K1) and the A* path-length cost matrices of all its (t, t_bef) frame pairs
(K2).  `value` = frame pairs per second with inputs resident in HBM; `e2e` = the same through the
host-buffer C-ABI calls (numpy in / numpy out, H2D + D2H inside the timed region).

N > 1 (torchrun): the timelapse is partitioned by timepoint, T frames per rank (weak scaling)
plus a 2-frame halo that each rank decodes itself; the only collective is the gather of the
per-pair matrices to rank 0 (NCCL send/recv), inside the timed step.

--impl reference: the reference's own CPU implementation (its compiled pyastar2d fork from
oracle/_ref driven by the reference's pair loop + ThreadPoolExecutor, and the numpy decode/NMS
port) on the host cores, each step a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]: synthetic 2048x2048 timelapse, 100 frames, ~200 detections/frame
    'c2': dict(H=2048, W=2048, T=100, n_dets=250, desc='configs[1]: synthetic 2048x2048 timelapse, 100 frames, ~200 dets/frame, deployed params.pkl tiling (512 px tiles, 12x12 cells)'),
    # configs[2]: dense 4096x4096, 200 frames, ~800 dets/frame
    'c3': dict(H=4096, W=4096, T=200, n_dets=1000, desc='configs[2]: dense synthetic 4096x4096 timelapse, 200 frames, ~800 dets/frame'),
    # configs[3]: long 2048x2048 timelapse, 2000 frames (per rank under torchrun: weak scaling)
    'c4': dict(H=2048, W=2048, T=2000, n_dets=250, desc='configs[3]: long synthetic 2048x2048 timelapse, 2000 frames, ~200 dets/frame, sharded by timepoint'),
    # configs[0] shape (examples/test.py), 2 frames
    'c1': dict(H=1969, W=3868, T=2, n_dets=250, desc='configs[0]: example timelapse shape 1969x3868, 2 frames'),
}
MAX_PX = 500          # AxonDetections.py:77
MAX_MISSES = 1        # deployed MCF_MAX_NUM_MISSES


# --------------------------------------------------------------------------------------------
# synthetic workload
# --------------------------------------------------------------------------------------------
def make_block(wl, rank=0, world=1, seed=0):
    """(mask bool[H,W], Y f32 [Tb,ntiles,12,12,3], tile_yx, t0, block): rank's block of ONE synthetic
    timelapse of T * world frames (every rank derives the same timelapse from the seed and cuts its
    block out of it with sharding.partition, so a halo frame IS the neighbour rank's frame);
    ranks > 0 carry the 2-frame halo in front (t0 = 2).  block = (first, own, stop) in global frames."""
    from axtrack_b200 import synth, sharding
    H, W, T = wl['H'], wl['W'], wl['T'] * world
    mask = synth.make_mask(H, W, seed)
    tracks = synth.make_tracks(mask, wl['n_dets'], T, seed)
    first, own, stop = sharding.partition(T, world, MAX_MISSES)[rank]
    Y = np.stack([synth.make_head_tensor(tracks[t], H, W, seed + t)[0] for t in range(first, stop)])
    return mask, Y, synth.tile_table(H, W), own - first, (first, own, stop)


def ball_cells(dets_yx, H, W, r):
    """|L1 ball of radius r around each detection, clipped to the image| (cells)."""
    out = np.zeros(len(dets_yx), np.int64)
    dy = np.arange(-r, r + 1)
    for i, (y, x) in enumerate(np.asarray(dets_yx, np.int64)):
        rows = y + dy
        ok = (rows >= 0) & (rows < H)
        hw = r - np.abs(dy)
        lo = np.maximum(0, x - hw)
        hi = np.minimum(W - 1, x + hw)
        out[i] = np.maximum(0, hi - lo + 1)[ok].sum()
    return out


def algorithmic_bytes(frame_dets, keys, H, W, max_px):
    """SURVEY.md 8(d) B_pair summed over the pairs: the L1 ball of radius max_px-1 of every
    source at 1 B/cell (the uint8 mask the reference starts from) + 5 B per matrix entry
    (int32 length + flag) + 8 B per detection."""
    balls = [ball_cells(d, H, W, max_px - 1) for d in frame_dets]
    tot = 0
    for t, tb in keys:
        tot += int(balls[tb].sum()) + len(frame_dets[tb]) * len(frame_dets[t]) * 5 + (len(frame_dets[tb]) + len(frame_dets[t])) * 8
    return tot


# --------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        time.sleep(0.15)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# --------------------------------------------------------------------------------------------
# CPU reference arm
# --------------------------------------------------------------------------------------------
def cpu_sample(mask, frame_dets, Y, tile_yx, target_calls):
    """One bounded sample of the reference CPU path: decode+NMS of one frame (numpy port of the
    pandas code) + the reference pair loop over the first sources of frame pair (1, 0) until
    ~target_calls in-range A* calls.  -> (seconds, in-range calls, kind)."""
    import oracle
    a, b = frame_dets[0], frame_dets[1]
    d2 = ((a[:, None, :].astype(np.int64) - b[None, :, :]) ** 2).sum(-1)
    inr = (d2 < MAX_PX * MAX_PX).sum(1)
    ns = int(np.searchsorted(np.cumsum(inr), target_calls) + 1)
    ns = max(1, min(ns, len(a)))
    use_ref = oracle.ref_available()
    t0 = time.perf_counter()
    oracle.decode_stitch_nms(Y[0], tile_yx)
    t_dec = time.perf_counter() - t0
    t0 = time.perf_counter()
    if use_ref:
        oracle.ref_cost_matrix(mask, a[:ns], b, MAX_PX, threads=True)
    else:
        oracle.cost_matrix_astar(mask, a[:ns], b, MAX_PX)
    t_astar = time.perf_counter() - t0
    return t_astar, int(inr[:ns].sum()), t_dec, ('reference' if use_ref else 'port'), ns


def cpu_pairs_per_s(t_astar, calls, t_dec, calls_per_pair, frames_per_pair):
    per_pair = t_astar / max(calls, 1) * calls_per_pair + t_dec * frames_per_pair
    return 1.0 / per_pair


def inrange_calls_per_pair(frame_dets, keys):
    tot = 0
    for t, tb in keys:
        a, b = frame_dets[tb].astype(np.int64), frame_dets[t].astype(np.int64)
        tot += int((((a[:, None, :] - b[None, :, :]) ** 2).sum(-1) < MAX_PX * MAX_PX).sum())
    return tot / max(len(keys), 1)


def oracle_frame_dets(Y, tile_yx):
    import oracle
    out = []
    for t in range(len(Y)):
        o = oracle.decode_stitch_nms(Y[t], tile_yx)
        out.append(np.stack([o['y'], o['x']], 1).astype(np.int32))
    return out


def run_reference(args, wl):
    from axtrack_b200.host import frame_pair_keys
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    sub = dict(wl, T=3)
    mask, Y, tile_yx, _, _ = make_block(sub)
    frame_dets = oracle_frame_dets(Y, tile_yx)
    keys = frame_pair_keys(3, MAX_MISSES)
    cpp = inrange_calls_per_pair(frame_dets, keys)
    T = wl['T'] * args.gpus
    frames_per_pair = T / max(2 * T - 3, 1)
    target = args.ref_calls
    times = []
    for i in range(args.warmup + args.steps):
        t_astar, calls, t_dec, kind, ns = cpu_sample(mask, frame_dets, Y, tile_yx, target)
        if i >= args.warmup:
            times.append((t_astar, calls, t_dec))
    ta = sum(x[0] for x in times); ca = sum(x[1] for x in times); td = sum(x[2] for x in times) / len(times)
    v = cpu_pairs_per_s(ta, ca, td, cpp, frames_per_pair)
    cores = len(os.sched_getaffinity(0))
    sample = (f'per step: decode+NMS of 1 frame + the reference pair loop (ThreadPoolExecutor over targets, GIL-bound) '
              f'for the first {ns} sources x {len(frame_dets[1])} targets of frame pair (1,0) = {calls} in-range A* calls; '
              f'extrapolated linearly to {cpp:.0f} in-range calls per frame pair')
    line = dict(metric='frame-pairs/sec (A* cost matrix + decode/NMS)', value=v, unit='frame-pairs/s', impl='reference',
                n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * (ta + td * len(times)) / len(times), higher_is_better=True, scaling='weak',
                vs_baseline=None, dtype='f32', data='synthetic',
                config=dict(workload=wl['desc'], max_px=MAX_PX, max_num_misses=MAX_MISSES),
                cpu_baseline=dict(value=v, unit='frame-pairs/s', cores=cores, kind=kind, sample=sample),
                e2e=dict(value=v, unit='frame-pairs/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    emit(line)


# --------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------
class Job:
    """One workload on this rank: its block of the synthetic timelapse resident in HBM, and the step
    that is timed.  N = 1: pipeline.associate_frames.  N > 1: sharding.associate_sharded -- the product
    entry -- on ONE timelapse of T * N frames cut by sharding.partition, every rank's K2 kernels
    storing straight into rank 0's HBM (PeerSlabs) or NCCL sends after the kernels (--gather nccl)."""

    def __init__(self, wl, dev, rank, world, gather):
        import torch
        import torch.distributed as dist
        from axtrack_b200 import sharding
        self.torch, self.dist, self.wl, self.dev, self.rank, self.world = torch, dist, wl, dev, rank, world
        self.mask, self.Y, self.tile_yx, self.t0, self.block = make_block(wl, rank, world)
        self.T_total = wl['T'] * world
        first, own, stop = self.block
        H, W = self.mask.shape
        ntiles = len(self.tile_yx)
        self.td = torch.from_numpy(self.tile_yx).to(dev)
        self.md = torch.from_numpy(self.mask.astype(np.uint8)).to(dev)
        if world == 1:
            self.Yd = torch.from_numpy(self.Y).to(dev)
        else:       # the whole timelapse's head tensor as associate_sharded takes it; only this rank's rows are filled
            self.Yd = torch.zeros((self.T_total, ntiles) + self.Y.shape[2:], dtype=torch.float32, device=dev)
            self.Yd[first:stop].copy_(torch.from_numpy(self.Y))
        self.slabs = None
        if world > 1 and gather == 'peer':
            try:
                self.slabs = sharding.PeerSlabs(dev)
            except Exception as e:                   # no peer mapping on this box: every rank falls back together
                print(f'[bench] rank {rank}: peer slabs unavailable ({e!r}); using the NCCL gather', file=sys.stderr)
            ok = torch.tensor([1 if self.slabs is not None else 0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 0:
                self.slabs = None

    def step(self, to_host=False):
        from axtrack_b200 import pipeline, sharding
        if self.world == 1:
            return None, pipeline.associate_frames(self.Yd, self.td, self.md, max_px=MAX_PX, max_num_misses=MAX_MISSES,
                                                   want_cost=False)
        return sharding.associate_sharded(self.Yd, self.td, self.md, max_px=MAX_PX, max_num_misses=MAX_MISSES,
                                          slabs=self.slabs, to_host=to_host)

    def sync_all(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def timed(self, steps, warmup, flush, ctx=None, sampler=None):
        """-> (ms summed over the steps [max over ranks], frame pairs of all ranks per step, launches, last res)"""
        torch, dist = self.torch, self.dist
        for _ in range(max(warmup, 3)):
            _, res = self.step()
        if self.slabs is not None:           # sized by the warm-up: the timed steps need no sizing collective / host sync
            self.slabs.collective_sizing = False
        self.sync_all()
        if sampler is not None:
            sampler.start()
        l0 = ctx.launches if ctx is not None else 0
        evs = []
        self.sync_all()
        for _ in range(steps):
            flush.fill_(1)                               # L2 flush between timed iterations (untimed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _, res = self.step()
            e1.record()
            evs.append((e0, e1))
        self.sync_all()
        launches = (ctx.launches - l0) if ctx is not None else 0
        ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([ms, float(len(res.keys))], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ts = t.clone(); dist.all_reduce(ts, op=dist.ReduceOp.SUM)
            every = [torch.zeros_like(t) for _ in range(self.world)]
            dist.all_gather(every, t)
            self.rank_ms = [float(e[0]) / steps for e in every]      # each rank's own device time per step (the step is their max)
            return float(tm[0]), float(ts[1]), launches, res
        return ms, float(len(res.keys)), launches, res

    def verify_gather(self):
        """Outside the timed region: what rank 0 assembled from the gather equals what every rank computes
        for its own block into LOCAL buffers (per-pair sums of path_len and of the flags)."""
        from axtrack_b200 import sharding
        if self.world == 1:
            return None
        if self.slabs is not None:
            self.slabs.collective_sizing = True
        got, _ = self.step(to_host=True)
        local = sharding.associate_block(self.block, self.Yd, self.td, self.md, MAX_PX, MAX_MISSES).pair_matrices()
        mine = {(k[0] + self.block[0], k[1] + self.block[0]): (int(v[0].sum()), int(v[1].sum()), v[0].shape) for k, v in local.items()}
        allsums = [None] * self.world
        self.dist.all_gather_object(allsums, mine)
        if self.rank != 0:
            return None
        want = {}
        for d in allsums:
            want.update(d)
        from axtrack_b200.host import frame_pair_keys
        assert list(got) == frame_pair_keys(self.T_total, MAX_MISSES), 'assembled keys are not the reference loop order'
        for k, (spl, sun, shp) in want.items():
            assert got[k][0].shape == shp and int(got[k][0].sum()) == spl and int(got[k][1].sum()) == sun, k
        return len(got)


def run_b200(args, wl):
    import torch
    import torch.distributed as dist
    from axtrack_b200 import ops, pipeline, host, sharding

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: axtrack_b200 has no CPU fallback')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    job = Job(wl, dev, rank, world, args.gather)
    mask, Y, tile_yx = job.mask, job.Y, job.tile_yx
    H, W = mask.shape
    first, own, stop = job.block
    Yd_block = job.Yd if world == 1 else job.Yd[first:stop]
    td, md, slabs = job.td, job.md, job.slabs
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    ctx = ops.get_context(dev)
    sampler = ClockSampler(local) if rank == 0 else None
    ms, total_pairs, launches, res = job.timed(args.steps, args.warmup, flush, ctx, sampler)
    rank_ms = getattr(job, 'rank_ms', None)
    clocks = sampler.stop() if rank == 0 else None
    value = total_pairs * args.steps / (ms * 1e-3)
    gathered_pairs = job.verify_gather()

    # ---- dominant kernel alone (K2) for the roofline: CUDA events around the launch ------------
    bits = ops.pack_mask(md)
    det = ops.decode_nms(Yd_block, td)
    n_out = det['n_out'].cpu().numpy()
    cap = det['conf'].shape[1]
    pairs, total, _ = pipeline.build_pairs(n_out, cap, res.keys)
    pairs_dev = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev)
    outb = (torch.empty(max(total, 1), dtype=torch.int32, device=dev), torch.empty(max(total, 1), dtype=torch.uint8, device=dev), None)
    yx = det['yx'].view(-1, 2)
    k2ms = []
    for i in range(3 + args.steps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.cost_matrix_batch(bits, W, yx, pairs, MAX_PX, want_cost=False, out=outb, pairs_dev=pairs_dev)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            k2ms.append(e0.elapsed_time(e1))
    k1ms = []
    for i in range(3 + args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.decode_nms(Yd_block, td, want_pre=False)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            k1ms.append(e0.elapsed_time(e1))
    k3ms = []          # K3 (transition costs + arc costs of every pair), reported beside the metric, not inside it
    for i in range(3 + args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res.transition_costs(MAX_PX, 0.6, MAX_MISSES, cost_threshold=0.7)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            k3ms.append(e0.elapsed_time(e1))
    yx_h = det['yx'].cpu().numpy()
    frame_dets = [yx_h[t, :int(n_out[t])] for t in range(len(n_out))]
    # K5 (feature_model histograms): the 70x70 boxes of every frame of the block on a
    # synthetic float32 frame, one launch per frame as libmot calls it; beside the metric, not inside it
    k5ms = []
    frame_img = torch.rand((H, W), dtype=torch.float32, device=dev)
    box_t = [torch.from_numpy(np.concatenate([fd[:, ::-1] - 35, np.full((len(fd), 2), 70)], 1).astype(np.int32)).to(dev)
             for fd in frame_dets]
    for i in range(3 + args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for b in box_t:
            ops.feature_hist(frame_img, b)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            k5ms.append(e0.elapsed_time(e1))
    # the same boxes in ONE launch (axb_feature_hist_frames; the frames of a timelapse are known up front)
    k5b = []
    nfr = min(len(box_t), 16)
    imgs = torch.rand((nfr, H, W), dtype=torch.float32, device=dev)
    all_boxes = torch.cat(box_t) if box_t else torch.zeros((0, 4), dtype=torch.int32, device=dev)
    all_fr = torch.cat([torch.full((len(b),), t % nfr, dtype=torch.int32, device=dev) for t, b in enumerate(box_t)]) if box_t else torch.zeros(0, dtype=torch.int32, device=dev)
    for i in range(3 + args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.feature_hist_frames(imgs, all_boxes, all_fr)
        e1.record()
        torch.cuda.synchronize()
        if i >= 3:
            k5b.append(e0.elapsed_time(e1))
    del imgs
    # K6 (tiling front end): non-empty tile flags of a 20-frame dense image block and the
    # model input of 16 frames per launch (temporal context 2, USE_MOTION_DATA exclude); beside the metric
    k6 = None
    if H * W <= 4096 * 4096:
        Tb, tcx = 20, 2
        Xb = torch.rand((Tb, 1, H, W), dtype=torch.float32, device=dev)
        fr = torch.arange(tcx, Tb - tcx, dtype=torch.int32, device=dev)
        ne_ms, ga_ms = [], []
        for i in range(3 + args.steps):
            e0, e1, e2, e3 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
            e0.record()
            ne = ops.tile_nonempty(Xb, 512)
            e1.record()
            tyx = ops.compact_tiles(ne)                       # torch.nonzero of the [ytiles, xtiles] table (host sync), not timed
            e2.record()
            st = ops.gather_frametiles(Xb, tyx, fr, 512, tcx, 'exclude')
            e3.record()
            torch.cuda.synchronize()
            if i >= 3:
                ne_ms.append(e0.elapsed_time(e1)); ga_ms.append(e2.elapsed_time(e3))
        k6_bytes = st.numel() * 4 + Tb * H * W * 4          # every output element written once, every input frame read once
        k6 = dict(nonempty_ms=float(np.mean(ne_ms)), gather_ms=float(np.mean(ga_ms)), frames_per_launch=int(fr.numel()),
                  gather_gbs=k6_bytes / (float(np.mean(ga_ms)) * 1e-3) / 1e9, algorithmic_bytes=int(k6_bytes),
                  note='random image: every tile non-empty, the flags kernel exits after its first round')
        del Xb, st
    alg = algorithmic_bytes(frame_dets, res.keys, H, W, MAX_PX)
    k2 = float(np.mean(k2ms)) * 1e-3
    peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.exists(os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {}
    peak = float(peaks.get('hbm_gbs', 6650.0))
    traffic, traffic_src = None, None
    tf = os.path.join(ROOT, 'profiles', 'k2_dram_traffic.json')
    if os.path.exists(tf) and MAX_PX == 500:      # the committed ncu captures are of max_px = 500 runs
        try:
            tj = json.load(open(tf))
            traffic, traffic_src = tj.get(args.workload), tj.get('source')
        except Exception:
            traffic = None
    klabel = ('K2 launch group = the level-0 search kernel (dominant, > 90 % by the ncu launch list) + tile/prep/dense/general; one launch = all frame pairs of the block'
              if MAX_PX <= 511 else 'K2 launch group at max_px > 511 (wide windows); one launch = all frame pairs of the block')
    roofline = dict(bound='hbm', kernel=klabel,
                    achieved=alg / k2 / 1e9, peak=peak, unit='GB/s', frac=alg / k2 / 1e9 / peak,
                    peak_source='MEASURED_PEAKS.json hbm_gbs (measured copy)' if peaks else 'fallback 6650 GB/s',
                    algorithmic_bytes_per_launch=alg, kernel_ms=k2 * 1e3, k1_decode_nms_ms=float(np.mean(k1ms)), k3_transition_ms=float(np.mean(k3ms)), k5_feature_hist_ms=float(np.mean(k5ms)), k5_feature_hist_one_launch_ms=float(np.mean(k5b)), k6_tiling=k6,
                    traffic=traffic, traffic_source=traffic_src)

    # ---- e2e: host buffers in, host results out ------------------------------------------------------
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    if world == 1:
        # the reference-facing C-ABI calls with HOST buffers (numpy in / numpy out): axb_decode_nms_host +
        # axb_cost_matrix_batch_host; H2D and D2H happen inside the library
        ha = host.HostAssociator(local)
        Yh, th, mh = pin(Y), pin(tile_yx), pin(mask.astype(np.uint8))
        F = Y.shape[0]
        dout = dict(n_pre=pin(np.zeros(F, np.int32)), n_out=pin(np.zeros(F, np.int32)), conf=pin(np.empty((F, cap), np.float32)),
                    yx=pin(np.empty((F, cap, 2), np.int32)), keep_idx=pin(np.empty((F, cap), np.int32)))
        mout = (pin(np.empty(max(total, 1), np.int32)), pin(np.empty(max(total, 1), np.uint8)), None)

        def e2e_step():
            d = ha.decode_nms(Yh, th, out=dout)
            prs, tot, _ = pipeline.build_pairs(d['n_out'], cap, res.keys)
            ha.cost_matrices_flat(mh, d['yx'].reshape(-1, 2), prs, tot, MAX_PX, want_cost=False, out=mout)
            return tot
        for _ in range(3):
            tot = e2e_step()
        torch.cuda.synchronize()
        t_0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t_0
        h2d = Yh.nbytes + th.nbytes + mh.nbytes + F * cap * 8 + len(res.keys) * 32
        d2h = sum(v.nbytes for v in dout.values()) + tot * 5
        e2e = dict(value=total_pairs * args.steps / dt, unit='frame-pairs/s', h2d_bytes_per_step=int(h2d),
                   d2h_bytes_per_step=int(d2h), ms_per_step=1e3 * dt / args.steps,
                   api='axb_decode_nms_host + axb_cost_matrix_batch_host (numpy in/out, pinned)')
        # what was timed equals the device path's result
        assert np.array_equal(mout[0][:tot], res.path_len.cpu().numpy()) and np.array_equal(mout[1][:tot], res.unreachable.cpu().numpy())
    else:
        # N > 1: every rank uploads its block's inputs from pinned host memory, runs the sharded product
        # entry, and the step ENDS ON RANK 0'S HOST: rank 0 copies every rank's gathered matrices out of its
        # slabs into pinned memory (sharding.HostGather) inside the timed region; wall clock, max over ranks
        Yh, th, mh = pin(Y), pin(tile_yx), pin(mask.astype(np.uint8))
        hg = sharding.HostGather()
        if slabs is not None:
            slabs.collective_sizing = False

        def e2e_step():
            job.Yd[first:stop].copy_(torch.from_numpy(Yh), non_blocking=True)
            td.copy_(torch.from_numpy(th), non_blocking=True)
            md.copy_(torch.from_numpy(mh), non_blocking=True)
            got, _ = sharding.associate_sharded(job.Yd, td, md, max_px=MAX_PX, max_num_misses=MAX_MISSES, slabs=slabs, to_host=hg)
            dist.barrier()                 # the step is over when rank 0 holds the dict on its host
            return got
        for _ in range(3):
            got = e2e_step()
        job.sync_all()
        t_0 = time.perf_counter()
        for _ in range(args.steps):
            got = e2e_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t_0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = None
        if rank == 0:
            assert len(got) == int(total_pairs)
            e2e = dict(value=total_pairs * args.steps / float(tt[0]), unit='frame-pairs/s',
                       h2d_bytes_per_step=int(Yh.nbytes + th.nbytes + mh.nbytes), d2h_bytes_per_step=int(hg.bytes),
                       ms_per_step=1e3 * float(tt[0]) / args.steps,
                       api='sharding.associate_sharded(to_host=HostGather): per-rank H2D of its block from pinned memory, gather into rank 0 HBM, '
                           'rank 0 D2H of ALL ranks matrices into pinned memory; h2d is per rank, d2h is rank 0')

    cpu = None
    if rank == 0 and world == 1:
        t_astar, calls, t_dec, kind, ns = cpu_sample(mask, frame_dets, Y, tile_yx, args.cpu_calls)
        cpp = inrange_calls_per_pair(frame_dets, res.keys)
        v = cpu_pairs_per_s(t_astar, calls, t_dec, cpp, len(frame_dets) / max(len(res.keys), 1))
        cpu = dict(value=v, unit='frame-pairs/s', cores=len(os.sched_getaffinity(0)), kind=kind,
                   sample=(f'{t_astar:.1f} s: reference pair loop (compiled pyastar2d fork + ThreadPoolExecutor over targets, GIL-bound) '
                           f'for the first {ns} sources of frame pair (1,0) = {calls} in-range A* calls, + numpy decode/NMS of 1 frame '
                           f'({t_dec * 1e3:.1f} ms); extrapolated to {cpp:.0f} in-range calls per frame pair'))

    # ---- configs[3] as BASELINE states it: ONE 2000-frame timelapse sharded by timepoint (strong scaling) ----
    strong = None
    if args.strong_c4 and args.workload == 'c2' and args.scaling == 'weak' and MAX_PX == 500:
        del job, res
        torch.cuda.empty_cache()
        c4 = WORKLOADS['c4']
        wl4 = dict(c4, T=-(-c4['T'] // world))
        j4 = Job(wl4, dev, rank, world, args.gather)
        n4 = max(2, min(args.steps, 5))
        ms4, pairs4, _, _ = j4.timed(n4, 3, flush)
        g4 = j4.verify_gather()
        strong = dict(workload=c4['desc'], frames_total=wl4['T'] * world, frames_per_gpu=wl4['T'], frame_pairs_total=int(pairs4),
                      value=pairs4 * n4 / (ms4 * 1e-3), unit='frame-pairs/s', ms_per_step=ms4 / n4, steps=n4, scaling='strong',
                      gathered_pairs_verified=g4)

    if rank == 0:
        par = ('single GPU' if world == 1 else 'K2 stores into rank 0 HBM over NVLink (peer slabs)' if slabs is not None else 'NCCL gather to rank 0')
        line = dict(metric='frame-pairs/sec (A* cost matrix + decode/NMS)', value=value, unit='frame-pairs/s', n_gpus=world,
                    steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=ms / args.steps, higher_is_better=True,
                    scaling=args.scaling, vs_baseline=None, dtype='int32', data='synthetic',
                    config=dict(workload=wl['desc'], frames_per_gpu=wl['T'], frames_total=wl['T'] * world, frame_pairs_total=int(total_pairs),
                                dets_per_frame=float(np.mean(n_out)), max_px=MAX_PX, max_num_misses=MAX_MISSES,
                                parallelism=f'one timelapse, timepoint shards x{world} (sharding.associate_sharded), ' + par,
                                gathered_pairs_verified=gathered_pairs, l2='flushed between timed steps (256 MiB write)'),
                    e2e=e2e, gpu_launches=int(launches), clocks=clocks, roofline=roofline, cpu_baseline=cpu, strong_c4=strong)
        if rank_ms:
            line['rank_ms_per_step'] = [round(v, 4) for v in rank_ms]     # ms_per_step is their maximum
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_c5(args):
    """BASELINE configs[4] / SURVEY 8(d) C5: detector-output stress, 8192x8192 frames (256 tiles of
    12x12 cells) through the fused decode+threshold+stitch+NMS kernel K1; sweep of frames per launch
    and of candidate density.  One JSON line; value = frames/s at 64 frames per launch, ~3000
    detections per frame.  Not the headline metric (that is frame-pairs/s on configs[1])."""
    import torch
    from axtrack_b200 import ops, synth
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: axtrack_b200 has no CPU fallback')
    import oracle
    dev = torch.device('cuda', 0)
    H = W = 8192
    tt = synth.tile_table(H, W)
    td = torch.from_numpy(tt).to(dev)
    rng = np.random.default_rng(0)
    sweep = []
    peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.exists(os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {}
    peak = float(peaks.get('hbm_gbs', 6650.0))
    headline = None
    for n_det in (800, 3000, 12000):
        dets = np.stack([rng.integers(0, H, n_det), rng.integers(0, W, n_det)], 1)
        Y1 = synth.make_head_tensor(dets, H, W, seed=n_det)[0]
        want = oracle.decode_stitch_nms(Y1, tt)                       # checker: the timed kernel's result is right
        for F in (1, 8, 64, 512):
            Yd = torch.from_numpy(Y1).to(dev).unsqueeze(0).repeat(F, 1, 1, 1, 1).contiguous()
            ms = []
            for i in range(3 + args.steps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                o = ops.decode_nms(Yd, td, want_pre=False)
                e1.record()
                torch.cuda.synchronize()
                if i >= 3:
                    ms.append(e0.elapsed_time(e1))
            n_out = int(o['n_out'][F - 1])
            assert n_out == len(want['keep_idx']) and np.array_equal(o['keep_idx'][F - 1, :n_out].cpu().numpy(), want['keep_idx'])
            t = float(np.mean(ms)) * 1e-3
            alg = F * (Y1.nbytes + n_out * 16)
            row = dict(dets_per_frame=n_out, frames_per_launch=F, ms=t * 1e3, frames_per_s=F / t, gbs=alg / t / 1e9, frac=alg / t / 1e9 / peak)
            sweep.append(row)
            if n_det == 3000 and F == 64:
                headline = row
            del Yd
    line = dict(metric='frames/sec (fused decode+threshold+stitch+NMS, K1)', value=headline['frames_per_s'], unit='frames/s',
                n_gpus=1, steps=args.steps, warmup=3, ms_per_step=headline['ms'], higher_is_better=True, scaling='weak',
                vs_baseline=None, dtype='f32', data='synthetic',
                config=dict(workload='configs[4]: detector-output stress, 8192x8192 frames tiled through the YOLO_AXTrack head (256 tiles x 12x12 cells), fused decode+NMS sweep',
                            frames_per_launch=64, dets_per_frame=headline['dets_per_frame'], l2='inputs are re-read from HBM/L2 each launch; 64 frames = 28 MB'),
                roofline=dict(bound='hbm', kernel='k1_decode_nms_kernel', achieved=headline['gbs'], peak=peak, unit='GB/s', frac=headline['frac'],
                              note='one CTA per frame: latency bound below ~148 frames per launch'),
                sweep=sweep)
    emit(line)


_JSON_OUT = None


def emit(line):
    """The one JSON line, on the process's original stdout."""
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + '\n')
    out.flush()


def main():
    global MAX_PX, _JSON_OUT
    # stdout carries the JSON line only: whatever libraries print on fd 1 (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='c2', choices=sorted(WORKLOADS) + ['c5'])
    ap.add_argument('--cpu-calls', type=int, default=1200, help='in-range A* calls of the cpu_baseline sample (~15 ms each at 2048^2)')
    ap.add_argument('--ref-calls', type=int, default=250, help='in-range A* calls per --impl reference step')
    ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                    help="weak: the workload's T frames per GPU (default); strong: T frames in total, T/N per GPU")
    ap.add_argument('--gather', default='peer', choices=['peer', 'nccl'],
                    help='N > 1: K2 writes into rank 0 memory over NVLink (peer) or NCCL sends after the kernels (nccl)')
    ap.add_argument('--no-strong-c4', dest='strong_c4', action='store_false',
                    help='skip the strong_c4 sub-record (configs[3]: one 2000-frame timelapse over the N GPUs)')
    ap.add_argument('--max-px', type=int, default=500,
                    help='max_px_assoc_dist (SURVEY 8d: configs[2] is also swept at 1000; > 511 runs the wide all-levels kernel)')
    args = ap.parse_args()
    MAX_PX = args.max_px
    if args.workload == 'c5':
        return run_c5(args)
    wl = WORKLOADS[args.workload]
    if args.scaling == 'strong':       # configs[3]: ONE timelapse of T frames sharded by timepoint over the ranks
        world = int(os.environ.get('WORLD_SIZE', '1'))
        wl = dict(wl, T=-(-wl['T'] // world), desc=wl['desc'] + f' -- {wl["T"]} frames in total, sharded over {world} GPU(s)')
    if args.impl == 'reference':
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == '__main__':
    main()
