"""K6 alone: 20 frames of HxW (default 2048x2048), 16 frames gathered per launch; CUDA-event times over 5 launches
(for ncu: the same launches).  AXB_K6_NO_TMA=1 selects the LDG/STG gather kernel."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from axtrack_b200 import _lib
if os.environ.get("AXB_LIB"):
    _lib.LIB_PATH = os.environ["AXB_LIB"]
from axtrack_b200 import ops
H = W = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
X = torch.rand((20, 1, H, W), device='cuda')
Z = torch.zeros_like(X)                      # all-empty image: the flags kernel reads everything
fr = torch.arange(2, 18, dtype=torch.int32, device='cuda')
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
tyx = ops.compact_tiles(ops.tile_nonempty(X, 512))
ne_ms, ga_ms = [], []
st = None
for i in range(8):
    del st                                   # the 1.3 GB output block goes back to the caching allocator
    flush.fill_(1)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    e[0].record(); ne = ops.tile_nonempty(Z, 512); e[1].record()
    flush.fill_(1)
    e[2].record(); st = ops.gather_frametiles(X, tyx, fr, 512, 2, 'exclude'); e[3].record()
    torch.cuda.synchronize()
    if i >= 3:
        ne_ms.append(e[0].elapsed_time(e[1])); ga_ms.append(e[2].elapsed_time(e[3]))
nb, gb = Z.numel() * 4, st.numel() * 4 + X.numel() * 4
n, g = min(ne_ms), min(ga_ms)
print('%dx%d nonempty (all-empty image, full read): %.3f ms %.0f GB/s; gather (%s): %.3f ms %.0f GB/s' % (
    H, W, n, nb / n / 1e6, 'LDG/STG' if os.environ.get('AXB_K6_NO_TMA') else 'TMA', g, gb / g / 1e6))
