"""Pair-descriptor helper (torch-free)."""
import numpy as np

from . import _lib


def make_pairs(entries):
    """entries: iterable of (mask_idx, src_off, n_src, dst_off, n_dst) -> (numpy struct array with
    out_off filled back to back, total output elements, total sources)."""
    entries = list(entries)
    arr = np.zeros(len(entries), dtype=_lib.PAIR_DTYPE)
    off = 0
    tot_src = 0
    for i, (m, so, ns, do, nd) in enumerate(entries):
        arr[i] = (m, so, ns, do, nd, 0, off)
        off += int(ns) * int(nd)
        tot_src += int(ns)
    return arr, off, tot_src
