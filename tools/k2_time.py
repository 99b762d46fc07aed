import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from axtrack_b200 import _lib
if len(sys.argv) > 1 and sys.argv[1] != '-':
    _lib.LIB_PATH = sys.argv[1]
import bench
from axtrack_b200 import ops, pipeline
T = int(sys.argv[2]) if len(sys.argv) > 2 else 12
wl = dict(bench.WORKLOADS[sys.argv[3] if len(sys.argv) > 3 else 'c2'], T=T)
mask, Y, tt, t0 = bench.make_block(wl)[:4]
dev = torch.device('cuda', 0)
Yd, td, md = torch.from_numpy(Y).to(dev), torch.from_numpy(tt).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)
bits = ops.pack_mask(md)
det = ops.decode_nms(Yd, td)
n_out = det['n_out'].cpu().numpy(); cap = det['conf'].shape[1]
keys = pipeline.frame_pair_keys(T)
pairs, total, _ = pipeline.build_pairs(n_out, cap, keys)
pd = torch.from_numpy(pairs.view(np.uint8).reshape(-1).copy()).to(dev)
outb = (torch.empty(total, dtype=torch.int32, device=dev), torch.empty(total, dtype=torch.uint8, device=dev), None)
ms = []
for i in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.cost_matrix_batch(bits, wl['W'], det['yx'].view(-1, 2), pairs, 500, want_cost=False, out=outb, pairs_dev=pd); e1.record()
    torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
print(sys.argv[1] if len(sys.argv) > 1 else 'default', 'K2 ms', ['%.2f' % m for m in ms], 'pairs', len(keys), 'sources', int(n_out[:-1].sum()), 'checksum', int(outb[0].sum()), int(outb[1].sum()))
if len(sys.argv) > 1 and 'timing' in sys.argv[1]:
    import ctypes
    L = _lib.load()
    buf = (ctypes.c_ulonglong * 24)()
    L.axb_dbg_fetch(buf, 1)
    ops.cost_matrix_batch(bits, wl['W'], det['yx'].view(-1, 2), pairs, 500, want_cost=False, out=outb, pairs_dev=pd); torch.cuda.synchronize()
    L.axb_dbg_fetch(buf, 1)
    b = list(buf)
    if b[4]:
      print('timing: items', b[4], 'waves/item %.1f' % (b[3] / max(b[4], 1)), 'entries/wave %.1f' % (b[5] / max(b[3], 1)),
          'cycles/item: setup %.0f staging %.0f waves %.0f' % (b[0] / b[4], b[1] / b[4], b[2] / b[4]), 'cycles/wave %.0f' % (b[2] / max(b[3], 1)), 'warp0 body %.0f barrier+lds %.0f' % (b[6] / max(b[3], 1), b[7] / max(b[3], 1)))
    if b[19]:
        names = ['entry fetch', 'keys+PB resolve', 'window loads', 'claims (atomics)', 'target probe', 'ballots+stores', 'wave end (syncwarp)']
        print('warp kernel: searches', b[19], 'waves/search %.1f' % (b[15] / b[19]), 'iterations/wave %.2f' % (b[16] / max(b[15], 1)),
              'entries/wave %.1f' % (b[17] / max(b[15], 1)), 'probes/wave %.2f' % (b[18] / max(b[15], 1)))
        tot = sum(b[8:15])
        for k, n in enumerate(names):
            print('   %-22s %8.0f cycles/wave  %5.1f %%' % (n, b[8 + k] / max(b[15], 1), 100.0 * b[8 + k] / max(tot, 1)))
        print('   total %.0f cycles/wave' % (tot / max(b[15], 1)))
