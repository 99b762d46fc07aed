import sys, glob, numpy as np, torch
sys.path.insert(0, '.')
import oracle
from axtrack_b200 import ops
for f in sorted(glob.glob('tests/golden/astar_s*.npz')):
    g = np.load(f)
    H, W, max_px = int(g['H']), int(g['W']), int(g['max_px'])
    mask = np.unpackbits(g['mask_bits'])[:H * W].reshape(H, W)
    bits = ops.pack_mask(torch.from_numpy(mask.astype(np.uint8)).cuda())
    for t, tb in g['keys']:
        src, dst = g[f'dets_yx_{tb}'], g[f'dets_yx_{t}']
        for rep in range(3):
            pl, un, wc = ops.cost_matrix(bits, W, torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda(), max_px)
            pl, un, wc = pl.cpu().numpy(), un.cpu().numpy(), wc.cpu().numpy()
            ref = g[f'dist_{t}_{tb}']; refn = g[f'none_{t}_{tb}']; refc = g[f'wcost_{t}_{tb}']
            bad = np.argwhere((pl != ref) | (un != refn))
            print(f, t, tb, 'rep', rep, 'mismatch', len(bad), 'of', pl.size)
            for i, j in bad[:6]:
                print('   src', src[i], 'on' if mask[src[i][0], src[i][1]] else 'off', 'dst', dst[j], 'on' if mask[dst[j][0], dst[j][1]] else 'off',
                      'got', pl[i, j], un[i, j], wc[i, j], 'ref', ref[i, j], refn[i, j], refc[i, j])
