"""K6 alone on the configs[1] shape: 20 frames of 2048x2048, 16 frames gathered per launch (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from axtrack_b200 import ops
H = W = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
X = torch.rand((20, 1, H, W), device='cuda')
fr = torch.arange(2, 18, dtype=torch.int32, device='cuda')
for i in range(3):
    ne = ops.tile_nonempty(X, 512)
    tyx = ops.compact_tiles(ne)
    st = ops.gather_frametiles(X, tyx, fr, 512, 2, 'exclude')
torch.cuda.synchronize()
Z = torch.zeros_like(X)                      # all-empty image: the flags kernel reads everything
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
e[0].record(); ne = ops.tile_nonempty(Z, 512); e[1].record(); st = ops.gather_frametiles(X, tyx, fr, 512, 2, 'exclude'); e[2].record()
torch.cuda.synchronize()
nb, gb = Z.numel() * 4, st.numel() * 4 + X.numel() * 4
print('nonempty (all-empty image, full read): %.3f ms %.0f GB/s; gather: %.3f ms %.0f GB/s' % (
    e[0].elapsed_time(e[1]), nb / e[0].elapsed_time(e[1]) / 1e6, e[1].elapsed_time(e[2]), gb / e[1].elapsed_time(e[2]) / 1e6))
