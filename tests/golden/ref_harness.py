"""Import the UNMODIFIED Python reference from /root/reference (build container only).

Used by tests/golden/make_golden.py to generate the committed fixtures and by the optional
``-m "not gpu"`` test that re-checks them when /root/reference is mounted.  Never used on the
GPU box, never by the product.

The reference imports plotting / IO / solver packages that are not installed here
(matplotlib, tifffile, skimage, torchsummary, cycler, motmetrics, ortools, ...).  None of
them is touched by the association hot path, so they are replaced by permissive stub modules
before the import.  pyastar2d is the reference's own fork compiled
into oracle/_ref/ plus a two-line module that exposes ``astar_path`` = the reference's
astar_wrapper.py loaded from where it lies.
"""
import importlib.machinery
import importlib.util
import os
import sys
import types

REF = '/root/reference'
_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))


class _Anything:
    """Callable, indexable, attribute-able dummy."""
    def __init__(self, *a, **k): pass
    def __call__(self, *a, **k): return _Anything()
    def __getattr__(self, n):
        if n.startswith('__') and n.endswith('__'):
            raise AttributeError(n)
        return _Anything()
    def __getitem__(self, k): return _Anything()
    def __setitem__(self, k, v): pass
    def __iter__(self): return iter(())
    def __mro_entries__(self, bases): return (object,)


class _StubModule(types.ModuleType):
    def __getattr__(self, n):
        if n.startswith('__') and n.endswith('__'):
            raise AttributeError(n)
        return _Anything()


_STUBS = ['torchsummary', 'matplotlib', 'matplotlib.pyplot', 'matplotlib.patches',
          'matplotlib.animation', 'matplotlib.lines', 'matplotlib.colors', 'matplotlib.cm',
          'matplotlib.gridspec', 'matplotlib.ticker', 'matplotlib.collections',
          'cycler', 'motmetrics', 'ortools', 'ortools.graph', 'ortools.graph.pywrapgraph',
          'tifffile', 'skimage', 'skimage.util', 'skimage.exposure', 'skimage.filters',
          'skimage.morphology', 'skimage.io', 'skimage.measure', 'skimage.transform',
          'skimage.draw', 'skimage.feature', 'skimage.segmentation',
          'addict', 'progressbar', 'path', 'seaborn', 'imageio', 'tensorboardX',
          'dateutil', 'dateutil.relativedelta', 'mpl_toolkits', 'mpl_toolkits.axes_grid1',
          'mpl_toolkits.axes_grid1.anchored_artists', 'pims', 'filterpy', 'lap',
          'scikit_image', 'napari']


def available():
    return os.path.isdir(os.path.join(REF, 'axtrack'))


def _install_stubs():
    for name in _STUBS:
        if name in sys.modules:
            continue
        try:
            if importlib.util.find_spec(name.split('.')[0]) is not None and '.' not in name:
                continue
        except (ImportError, ValueError):
            pass
        try:
            importlib.import_module(name)
            continue
        except Exception:
            pass
        m = _StubModule(name)
        m.__path__ = []
        sys.modules[name] = m
        if '.' in name:
            parent, _, child = name.rpartition('.')
            if parent in sys.modules:
                setattr(sys.modules[parent], child, m)


def _install_pyastar2d():
    """pyastar2d = compiled reference extension + the reference's own wrapper file."""
    if 'pyastar2d' in sys.modules and hasattr(sys.modules['pyastar2d'], 'astar_path'):
        return
    sys.path.insert(0, _ROOT)
    import oracle
    ext = oracle.ref_module()
    pkg = types.ModuleType('pyastar2d')
    pkg.__path__ = []
    pkg.astar = ext
    sys.modules['pyastar2d'] = pkg
    sys.modules['pyastar2d.astar'] = ext
    wrapper = os.path.join(REF, 'pyastar2d/src/pyastar2d/astar_wrapper.py')
    spec = importlib.util.spec_from_file_location('pyastar2d.astar_wrapper', wrapper)
    mod = importlib.util.module_from_spec(spec)
    sys.modules['pyastar2d.astar_wrapper'] = mod
    spec.loader.exec_module(mod)
    pkg.astar_path = mod.astar_path


def _install_libmot():
    """``libmot.data_association.MinCostFlowTracker`` loaded from its file; the rest of libmot
    (DAN tracker, loggers ... never called by axtrack) is not imported."""
    if 'libmot.data_association' in sys.modules:
        return
    pkg = types.ModuleType('libmot')
    pkg.__path__ = []
    sub = types.ModuleType('libmot.data_association')
    sub.__path__ = []
    f = os.path.join(REF, 'libmot/libmot/data_association/offline_mincostflow.py')
    spec = importlib.util.spec_from_file_location('libmot.data_association.offline_mincostflow', f)
    mod = importlib.util.module_from_spec(spec)
    sys.modules['libmot'] = pkg
    sys.modules['libmot.data_association'] = sub
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    sub.MinCostFlowTracker = mod.MinCostFlowTracker
    sub.offline_mincostflow = mod
    pkg.data_association = sub


def load():
    """-> the reference's root package (``axref``); ``axref.axtrack.AxonDetections`` etc."""
    if 'axref' in sys.modules:
        return sys.modules['axref']
    assert available(), '/root/reference is not mounted'
    import numpy as np
    if not hasattr(np, 'float'):
        np.float = float          # mincostflow_models.py:19 (reference pins numpy 1.20)
    _install_stubs()
    _install_pyastar2d()
    _install_libmot()
    spec = importlib.util.spec_from_file_location(
        'axref', os.path.join(REF, '__init__.py'), submodule_search_locations=[REF])
    mod = importlib.util.module_from_spec(spec)
    sys.modules['axref'] = mod
    spec.loader.exec_module(mod)
    return mod


def make_dataset_stub(name, mask, T, tilesize=512, temporal_context=2):
    """A Timelapse that skips __init__ (no tif on disk) and carries exactly what the hot path
    reads: name, sizes, tile grid, tile_info (all tiles non-empty), mask[t] as coo_matrix."""
    import numpy as np
    import pandas as pd
    import torch
    from scipy.sparse import coo_matrix
    ax = load()
    Timelapse = sys.modules['axref.axtrack.Timelapse'].Timelapse

    class _DS(Timelapse):
        def __init__(self):
            torch.utils.data.Dataset.__init__(self)

        def construct_tiles(self, *a, **k):
            pass

    ds = _DS()
    H, W = mask.shape
    ds.name = name
    ds.sizet, ds.sizey, ds.sizex = T, H, W
    ds.tilesize = tilesize
    ds.ytiles = int(np.ceil(H / tilesize))
    ds.xtiles = int(np.ceil(W / tilesize))
    ds.tile_info = torch.ones((ds.ytiles, ds.xtiles, T, 2))
    ds.use_motion_filtered = 'exclude'
    ds.temporal_context = temporal_context
    ds.target = pd.DataFrame([])
    ds.mask = [coo_matrix(mask.astype(bool))] * T
    return ds


def make_axon_detections(ds, params=None):
    import pickle
    ax = load()
    AD = sys.modules['axref.axtrack.AxonDetections'].AxonDetections
    if params is None:
        with open(os.path.join(REF, 'deployed_model/params.pkl'), 'rb') as f:
            params = pickle.load(f)
        params['DEVICE'] = 'cpu'
    return AD(None, ds, params, None)
