"""Host-side timeline of one pipeline.associate_frames step on the config-2 block (where the time outside K2 goes)."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import bench
from axtrack_b200 import ops, pipeline
wl = dict(bench.WORKLOADS['c2'])
mask, Y, tt, t0 = bench.make_block(wl)[:4]
dev = torch.device('cuda', 0)
Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)
for _ in range(3):
    pipeline.associate_frames(Yd, td, md, max_px=500, max_num_misses=1, want_cost=False)
torch.cuda.synchronize()
marks = []
orig = {}
def wrap(mod, name):
    f = getattr(mod, name); orig[(mod, name)] = f
    def g(*a, **k):
        t = time.perf_counter(); r = f(*a, **k); marks.append((name, t, time.perf_counter())); return r
    setattr(mod, name, g)
for mod, name in ((ops, 'decode_nms'), (ops, 'pack_mask'), (ops, 'prepare_masks'), (ops, 'raise_status'), (pipeline, 'build_pairs'), (ops, 'cost_matrix_batch')):
    wrap(mod, name)
for rep in range(3):
    marks.clear()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    T0 = time.perf_counter(); e0.record()
    pipeline.associate_frames(Yd, td, md, max_px=500, max_num_misses=1, want_cost=False)
    T1 = time.perf_counter(); e1.record()
    torch.cuda.synchronize(); T2 = time.perf_counter()
    print('rep', rep, 'device %.3f ms; host returns after %.3f ms; drained after %.3f ms' % (e0.elapsed_time(e1), 1e3 * (T1 - T0), 1e3 * (T2 - T0)))
    for name, a, b in marks:
        print('   %-18s start %.3f  end %.3f ms' % (name, 1e3 * (a - T0), 1e3 * (b - T0)))
