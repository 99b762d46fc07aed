import sys, numpy as np
sys.path.insert(0, '.')
import bench
from axtrack_b200 import synth
wl = dict(bench.WORKLOADS['c2'], T=3)
mask, Y, tt, t0 = bench.make_block(wl, 0)
print('fg frac', mask.mean())
import oracle
dets = bench.oracle_frame_dets(Y, tt)
H, W = mask.shape
rng = np.random.default_rng(0)
for s in dets[0][:6]:
    sy, sx = int(s[0]), int(s[1])
    vis = np.zeros_like(mask); vis[sy, sx] = True
    fr = [(sy, sx)]
    cells = []; words = []
    j = 0
    while fr and j < 500:
        nxt = []
        for (y, x) in fr:
            for dy, dx in ((1,0),(-1,0),(0,1),(0,-1)):
                yy, xx = y+dy, x+dx
                if 0 <= yy < H and 0 <= xx < W and mask[yy, xx] and not vis[yy, xx]:
                    vis[yy, xx] = True; nxt.append((yy, xx))
        fr = nxt; j += 1
        cells.append(len(fr)); words.append(len({(y>>2, x>>3) for y, x in fr}))
    print('src', sy, sx, 'waves', j, 'cells total', sum(cells), 'avg frontier', np.mean(cells), 'max', max(cells), 'avg tilewords', np.mean(words), 'max', max(words))
