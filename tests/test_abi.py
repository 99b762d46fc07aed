"""The C-ABI library builds, loads and exports every symbol include/axtrack_b200.h declares.
No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

from axtrack_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, 'include', 'axtrack_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(axb_[a-z_0-9]+)\s*\(', src)))


def test_header_symbols_listed():
    assert _declared_symbols() == sorted(_lib.SYMBOLS)


def test_library_loads_and_exports():
    assert os.path.exists(_lib.LIB_PATH), 'run __graft_entry__.build() first'
    L = _lib.load()
    for s in _declared_symbols():
        assert hasattr(L, s), s
    assert b'sm_100a' in L.axb_version()
    assert L.axb_mask_words(2048) == 64 and L.axb_mask_words(33) == 2


def test_pair_desc_layout():
    import numpy as np
    assert ctypes.sizeof(_lib.PairDesc) == 32
    assert np.dtype(_lib.PAIR_DTYPE).itemsize == 32


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from axtrack_b200 import ops
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.get_context()
    h = ctypes.c_void_p()
    assert _lib.load().axb_create(0, ctypes.byref(h)) == _lib.ERR_CUDA


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'axtrack_b200')
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith(('.py', '.cu', '.cuh', '.h')):
                txt = open(os.path.join(dp, fn)).read()
                assert 'import oracle' not in txt and 'from oracle' not in txt and 'liboracle' not in txt, fn
