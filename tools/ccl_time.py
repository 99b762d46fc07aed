"""Time of the mask-component kernels (axb_components.cu) inside a small K2 call on adversarial masks."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from axtrack_b200 import ops
dev = torch.device('cuda', 0)
H = W = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:H, 0:W]
masks = {
    'all ones': np.ones((H, W), np.uint8),
    'checkerboard': ((yy + xx) & 1).astype(np.uint8),
    'speckle 50 %': (rng.random((H, W)) < 0.5).astype(np.uint8),
    'speckle 59 % (near the percolation threshold)': (rng.random((H, W)) < 0.59).astype(np.uint8),
    'vertical stripes': (xx & 1).astype(np.uint8),
    'serpentine': (((yy % 4 == 0) | ((yy % 8 == 1) & (xx == W - 1)) | ((yy % 8 == 2) & (xx == W - 1)) | ((yy % 8 == 3) & (xx == W - 1))
                    | ((yy % 8 == 5) & (xx == 0)) | ((yy % 8 == 6) & (xx == 0)) | ((yy % 8 == 7) & (xx == 0)))).astype(np.uint8),
}
dets = np.array([[10, 10], [20, 30], [40, 50], [60, 70], [80, 90], [100, 110]], np.int32)
pairs, total, _ = ops.make_pairs([(0, 0, 2, 2, 2), (0, 0, 2, 4, 2), (0, 2, 2, 4, 2)])
dd = torch.from_numpy(dets).to(dev)
for name, m in masks.items():
    bits = ops.pack_mask(torch.from_numpy(m).to(dev))
    ms = []
    for max_px in (30,):
        for i in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.cost_matrix_batch(bits, W, dd, pairs, max_px); e1.record()
            torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print('%-48s whole call ms %s' % (name, ['%.3f' % v for v in ms]))
