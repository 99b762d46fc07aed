import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, torch
import oracle
from test_gpu_parity import _gpu_cost
H, W = 1100, 1200
for max_px in (509, 510, 511):
    for sy, sx in [(548, 590), (549, 595), (550, 600), (551, 605), (548, 610)]:
        mask = np.zeros((H, W), np.uint8); mask[sy % 12::12, :] = 1; mask[:, sx] = 1
        dst = []
        for y in np.arange(sy % 12, H, 12):
            hw = max_px - abs(int(y) - sy)
            if hw < 0: continue
            for side in (-1, 1):
                for inward in range(10):
                    x = sx + side * (hw - inward)
                    if hw - inward >= 0 and 0 <= x < W: dst.append((y, x))
        src = np.array([[sy, sx]], np.int32); dst = np.array(dst, np.int32)
        got = _gpu_cost(torch, mask, src, dst, max_px); ref = oracle.cost_matrix_rule(mask, src, dst, max_px)
        bad = np.nonzero(got[0][0] != ref[0][0])[0]
        print(max_px, sy, sx, 'bad', len(bad), 'of', len(dst))
        for b in bad[:12]:
            print('   dy', dst[b, 0] - sy, 'dx', dst[b, 1] - sx, 'got', got[0][0, b], got[1][0, b], 'want', ref[0][0, b], ref[1][0, b])
