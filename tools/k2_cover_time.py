"""What a search from 2/3 of the frames with the targets of 4 neighbour frames costs (the vertex-cover decomposition
of the frame pairs: every pair (t, t_bef) is answered from whichever of its two frames is searched)."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import bench
from axtrack_b200 import ops, pipeline
T = int(sys.argv[1]) if len(sys.argv) > 1 else 100
wl = dict(bench.WORKLOADS[sys.argv[2] if len(sys.argv) > 2 else 'c2'], T=T)
mask, Y, tt, t0 = bench.make_block(wl)[:4]
dev = torch.device('cuda', 0)
Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)
bits = ops.pack_mask(md)
det = ops.decode_nms(Yd, td)
n_out = det['n_out'].cpu().numpy(); cap = det['conf'].shape[1]
keys = pipeline.frame_pair_keys(T)
cover = []
for (t, tb) in keys:                       # source frame tb, target frame t
    cover.append((t, tb) if tb % 3 != 0 else (tb, t))
cover.sort(key=lambda k: (k[1], k[0]))     # pairs of one source frame next to each other (chains)
for name, kk in (('reference pairs', keys), ('cover pairs', cover)):
    pairs, total, _ = pipeline.build_pairs(n_out, cap, kk)
    pd = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev)
    outb = (torch.empty(total, dtype=torch.int32, device=dev), torch.empty(total, dtype=torch.uint8, device=dev), None)
    ms = []
    for i in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.cost_matrix_batch(bits, wl['W'], det['yx'].view(-1, 2), pairs, 500, want_cost=False, out=outb, pairs_dev=pd); e1.record()
        torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    srcs = sorted(set(k[1] for k in kk))
    print(name, 'K2 ms', ['%.2f' % m for m in ms], 'pairs', len(kk), 'searched frames', len(srcs), 'sources', int(sum(n_out[s] for s in srcs)),
          'reached', int((outb[1] == 0).sum()), 'sum L', int(outb[0].sum()))
