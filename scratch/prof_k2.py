import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import bench
from axtrack_b200 import ops, pipeline
T = int(sys.argv[1]) if len(sys.argv) > 1 else 12
wl = dict(bench.WORKLOADS[sys.argv[2] if len(sys.argv) > 2 else 'c2'], T=T)
mask, Y, tt, t0 = bench.make_block(wl, 0)
dev = torch.device('cuda', 0)
Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)
for i in range(3):
    torch.cuda.synchronize(); t = time.perf_counter()
    res = pipeline.associate_frames(Yd, td, md, want_cost=False)
    torch.cuda.synchronize(); print('step', i, (time.perf_counter() - t) * 1e3, 'ms', len(res.keys), 'pairs', int(res.n_out.sum()), 'dets')
