"""Randomised A/B: the K2 cover (+ early all-levels launch) against one directed search per source, as CRCs of all
outputs.  Run twice (default, and with AXB_K2_NO_REVERSE=1 AXB_K2_NO_EARLY=1) and diff the output:
    python tools/cover_stress.py > a.txt; AXB_K2_NO_REVERSE=1 AXB_K2_NO_EARLY=1 python tools/cover_stress.py > b.txt; diff a.txt b.txt"""
import sys, zlib, numpy as np, torch
sys.path.insert(0, '.')
from axtrack_b200 import ops, synth
dev = torch.device('cuda', 0)
n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 12
for case in range(n_cases):
    rng = np.random.default_rng(1000 + case)
    H, W = int(rng.integers(300, 1400)), int(rng.integers(300, 1400))
    kind = case % 3
    mask = synth.make_mask(H, W, case, pitch=(int(rng.integers(40, 90)), int(rng.integers(100, 220))), width=(3, int(rng.integers(6, 14)))).astype(bool)
    if kind >= 1:                                   # cut into islands
        for _ in range(int(rng.integers(10, 60))):
            y, x = rng.integers(0, H), rng.integers(0, W)
            if rng.random() < 0.5: mask[y:y + 3, :] = False
            else: mask[:, x:x + 3] = False
    if kind == 2:
        mask |= rng.random((H, W)) < 0.03
    T = int(rng.integers(4, 12)); n = int(rng.integers(20, 120)); misses = int(rng.integers(1, 3))
    max_px = int(rng.choice([60, 150, 300, 500, 511, 512, 700, 1000]))
    fg = np.argwhere(mask)
    frames = []
    for t in range(T):
        d = fg[rng.integers(0, len(fg), n)] + rng.integers(-3, 4, (n, 2)) * (rng.random((n, 1)) < 0.2)
        d[:, 0] = np.clip(d[:, 0], 0, H - 1); d[:, 1] = np.clip(d[:, 1], 0, W - 1)
        frames.append(d.astype(np.int32))
    if case % 4 == 1:
        frames[T // 2] = frames[T // 2][:0]
    offs = np.concatenate([[0], np.cumsum([len(f) for f in frames])])
    spec = [(0, tb, t) for t in range(T) for tb in range(t - 1, t - (misses + 2), -1) if tb >= 0]
    pairs, total, _ = ops.make_pairs([(mk, int(offs[a]), len(frames[a]), int(offs[b]), len(frames[b])) for mk, a, b in spec])
    bits = ops.pack_mask(torch.from_numpy(mask.astype(np.uint8)).to(dev))
    dets = torch.from_numpy(np.concatenate(frames).astype(np.int32).reshape(-1, 2)).to(dev)
    pl, un, wc, _ = ops.cost_matrix_batch(bits, W, dets, pairs, max_px)
    ops.raise_status()
    print('case %2d %4dx%-4d T %2d n %3d misses %d max_px %4d kind %d  reached %7d  crc %08x %08x %08x' % (
        case, H, W, T, n, misses, max_px, kind, int((un == 0).sum()), zlib.crc32(pl.cpu().numpy().tobytes()),
        zlib.crc32(un.cpu().numpy().tobytes()), zlib.crc32(wc.cpu().numpy().tobytes())), flush=True)
