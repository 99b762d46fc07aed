import sys, numpy as np
sys.path.insert(0, '.')
from axtrack_b200 import synth
from scipy import ndimage
H = W = 2048
mask = synth.make_mask(H, W, 0)
tr = synth.make_tracks(mask, 250, 1, 0)[0]
tc = tw = 0
for (sy, sx) in tr[:12]:
    if not mask[sy, sx]: continue
    y0, y1, x0, x1 = max(0, sy-500), min(H, sy+501), max(0, sx-500), min(W, sx+501)
    sub = mask[y0:y1, x0:x1]
    yy, xx = np.mgrid[y0:y1, x0:x1]
    sub = sub & (np.abs(yy-sy)+np.abs(xx-sx) <= 500)
    dist = np.full(sub.shape, -1, np.int32)
    fr = np.zeros(sub.shape, bool); fr[sy-y0, sx-x0] = True; dist[fr] = 0
    cells = words = 0
    for j in range(1, 501):
        nb = np.zeros_like(fr)
        nb[1:] |= fr[:-1]; nb[:-1] |= fr[1:]; nb[:, 1:] |= fr[:, :-1]; nb[:, :-1] |= fr[:, 1:]
        fr = nb & sub & (dist < 0)
        n = int(fr.sum())
        if n == 0: break
        dist[fr] = j
        ys, xs = np.nonzero(fr)
        wd = len(set(zip(ys.tolist(), ((xs + x0) >> 5).tolist())))
        cells += n; words += wd
    print((sy, sx), 'waves', j, 'cells', cells, 'words', words, 'ratio %.2f' % (cells / max(words, 1)), 'cells/wave %.0f' % (cells / j))
    tc += cells; tw += words
print('total ratio', tc / tw)
