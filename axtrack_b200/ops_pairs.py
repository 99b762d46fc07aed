"""Pair-descriptor helper (torch-free)."""
import numpy as np

from . import _lib


def make_pairs(entries):
    """entries: iterable of (mask_idx, src_off, n_src, dst_off, n_dst[, time_gap]) -> (numpy struct
    array with out_off filled back to back, total output elements, total sources).  time_gap =
    t - t_bef is only read by the transition-cost kernel (0 = not given)."""
    entries = list(entries)
    arr = np.zeros(len(entries), dtype=_lib.PAIR_DTYPE)
    off = 0
    tot_src = 0
    for i, e in enumerate(entries):
        m, so, ns, do, nd = e[:5]
        arr[i] = (m, so, ns, do, nd, int(e[5]) if len(e) > 5 else 0, off)
        off += int(ns) * int(nd)
        tot_src += int(ns)
    return arr, off, tot_src
