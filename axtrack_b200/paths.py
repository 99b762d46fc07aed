"""Mirror of ``utils._compute_astar_path`` (axtrack/utils.py:351-390) for the pairs whose path
PIXELS are wanted: a bool ``coo_matrix`` of the frame shape with the pixels of a minimum-cost path,
or ``None`` where the reference returns ``(None, None)``.  The search runs on the GPU
(``axb_trace_paths``); there is no CPU path.  Which of several equal-cost paths is returned is
heap-tie dependent in the reference (``std::priority_queue`` order) and not reproduced: length and
cost are."""
import numpy as np
import torch
from scipy.sparse import coo_matrix

from . import ops


def compute_astar_paths(mask, sources_yx, targets_yx, max_path_length=500, mask_bits=None):
    """mask bool/uint8 [H, W] (numpy or CUDA tensor; weights are where(mask==1, 1, 65536),
    AxonDetections.py:598); sources_yx / targets_yx int [n, 2] rows (y, x), paired row by row.
    -> list of n (coo_matrix bool [H, W], L) or (None, None), as utils._compute_astar_path returns."""
    if not torch.cuda.is_available():
        raise RuntimeError('axtrack_b200 needs a CUDA device; there is no CPU fallback')
    m = mask if isinstance(mask, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(mask).astype(np.uint8)))
    m = m.to('cuda', torch.uint8)
    H, W = m.shape
    if mask_bits is None:
        mask_bits = ops.pack_mask(m)
    s = np.asarray(sources_yx, np.int32).reshape(-1, 2)
    t = np.asarray(targets_yx, np.int32).reshape(-1, 2)
    assert len(s) == len(t)
    # the reference's Euclid pre-filter (AxonDetections.py:620-628) is part of the search kernel
    req = np.concatenate([np.zeros((len(s), 1), np.int32), s, t], 1)
    yx, ln = ops.trace_paths(mask_bits, W, req, max_path_length)
    ops.raise_status(device=m.device)
    yx, ln = yx.cpu().numpy(), ln.cpu().numpy()
    out = []
    for i, L in enumerate(ln):
        if L <= 0:
            out.append((None, None))
            continue
        p = yx[i, :L]
        out.append((coo_matrix((np.ones(L, bool), (p[:, 0], p[:, 1])), shape=(H, W), dtype=bool), int(L)))
    return out


def compute_astar_path(source, target, mask, max_path_length=500):
    """Single pair, argument order of utils._compute_astar_path(source, target, weights, ...)."""
    return compute_astar_paths(mask, [source], [target], max_path_length)[0]
