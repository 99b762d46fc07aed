"""ctypes binding of libaxtrack_b200.so (include/axtrack_b200.h).

The library is built in-tree by ``axtrack_b200/csrc/Makefile`` (``__graft_entry__.build()``).
There is no fallback: if the shared library is missing this module raises at import of the
first symbol, and without a CUDA device ``Context()`` raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# AXTRACK_B200_LIB: another build of the same library (csrc/Makefile `check` / `timing`), for debugging only
LIB_PATH = os.environ.get('AXTRACK_B200_LIB') or os.path.join(_HERE, 'libaxtrack_b200.so')

OK = 0
ERR_CUDA = -1
ERR_OOB_START = -2
ERR_OOB_GOAL = -3
ERR_BAD_WEIGHT = -4
ERR_BAD_ARG = -5
ERR_CAPACITY = -6
ERR_UNSUPPORTED_WEIGHT = -7
MAX_PX = 1023

#: every symbol include/axtrack_b200.h declares
SYMBOLS = ('axb_create', 'axb_destroy', 'axb_last_error', 'axb_version', 'axb_launch_count',
           'axb_mask_words', 'axb_pack_mask', 'axb_pack_weights', 'axb_prepare_masks', 'axb_cost_matrix_batch',
           'axb_cost_matrix', 'axb_cost_matrix_batch_host', 'axb_cost_matrix_host',
           'axb_decode_nms', 'axb_decode_nms_host', 'axb_transition_cost_batch', 'axb_trace_paths', 'axb_feature_hist',
           'axb_feature_hist_frames', 'axb_vis_sim',
           'axb_tile_nonempty', 'axb_tile_nonempty_noalloc', 'axb_gather_frametiles', 'axb_stitched_frame', 'axb_status')


class PairDesc(ctypes.Structure):
    """struct axb_pair_desc"""
    _fields_ = [('mask_idx', ctypes.c_int32), ('src_off', ctypes.c_int32), ('n_src', ctypes.c_int32),
                ('dst_off', ctypes.c_int32), ('n_dst', ctypes.c_int32), ('time_gap', ctypes.c_int32),
                ('out_off', ctypes.c_int64)]


#: numpy dtype with the same layout as axb_pair_desc
PAIR_DTYPE = [('mask_idx', '<i4'), ('src_off', '<i4'), ('n_src', '<i4'), ('dst_off', '<i4'),
              ('n_dst', '<i4'), ('time_gap', '<i4'), ('out_off', '<i8')]

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
            '(nvcc, sm_100a). axtrack_b200 has no CPU fallback.')
    L = ctypes.CDLL(LIB_PATH)
    vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
    L.axb_create.restype = ci
    L.axb_create.argtypes = [ci, ctypes.POINTER(vp)]
    L.axb_destroy.restype = None
    L.axb_destroy.argtypes = [vp]
    L.axb_last_error.restype = ctypes.c_char_p
    L.axb_last_error.argtypes = [vp]
    L.axb_version.restype = ctypes.c_char_p
    L.axb_version.argtypes = []
    L.axb_launch_count.restype = ctypes.c_int64
    L.axb_launch_count.argtypes = [vp]
    L.axb_mask_words.restype = ci
    L.axb_mask_words.argtypes = [ci]
    L.axb_pack_mask.restype = ci
    L.axb_pack_mask.argtypes = [vp, vp, ci, ci, ci, vp, vp]
    L.axb_pack_weights.restype = ci
    L.axb_pack_weights.argtypes = [vp, vp, ci, ci, ci, vp, vp]
    L.axb_prepare_masks.restype = ci
    L.axb_prepare_masks.argtypes = [vp, vp, ci, ci, ci, vp]
    L.axb_cost_matrix_batch.restype = ci
    L.axb_cost_matrix_batch.argtypes = [vp, vp, ci, ci, ci, vp, ci, vp, ci, ci, ci, vp, vp, vp, ctypes.c_int64, vp]
    L.axb_cost_matrix.restype = ci
    L.axb_cost_matrix.argtypes = [vp, vp, ci, ci, vp, ci, vp, ci, ci, vp, vp, vp, vp]
    L.axb_cost_matrix_batch_host.restype = ci
    L.axb_cost_matrix_batch_host.argtypes = [vp, vp, ci, ci, ci, vp, ci, vp, ci, ci, vp, vp, vp]
    L.axb_cost_matrix_host.restype = ci
    L.axb_cost_matrix_host.argtypes = [vp, vp, ci, ci, vp, ci, vp, ci, ci, vp, vp, vp]
    L.axb_decode_nms.restype = ci
    L.axb_decode_nms.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, cf, ci, ci,
                                 vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.axb_decode_nms_host.restype = ci
    L.axb_decode_nms_host.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, cf, ci, ci,
                                      vp, vp, vp, vp, vp, vp, vp, vp]
    L.axb_trace_paths.restype = ci
    L.axb_trace_paths.argtypes = [vp, vp, ci, ci, ci, vp, ci, ci, vp, vp, vp]
    L.axb_transition_cost_batch.restype = ci
    L.axb_transition_cost_batch.argtypes = [vp, vp, vp, ci, ci, vp, vp, ci, ci, vp, vp, vp]
    L.axb_feature_hist.restype = ci
    L.axb_feature_hist.argtypes = [vp, vp, ci, ci, vp, ci, ci, cf, cf, vp, vp]
    L.axb_feature_hist_frames.restype = ci
    L.axb_feature_hist_frames.argtypes = [vp, vp, ci, ci, ci, vp, vp, ci, ci, cf, cf, vp, vp]
    L.axb_vis_sim.restype = ci
    L.axb_vis_sim.argtypes = [vp, vp, ci, vp, ci, ci, vp, ci, vp, vp]
    L.axb_tile_nonempty.restype = ci
    L.axb_tile_nonempty.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp]
    L.axb_tile_nonempty_noalloc.restype = ci
    L.axb_tile_nonempty_noalloc.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp]
    L.axb_gather_frametiles.restype = ci
    L.axb_gather_frametiles.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, ci, vp, ci, ci, ci, ci, vp, vp]
    L.axb_stitched_frame.restype = ci
    L.axb_stitched_frame.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, ci, ci, ci, vp, vp]
    L.axb_status.restype = ci
    L.axb_status.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_int64 * 3)]
    _lib = L
    return L


class AxbError(RuntimeError):
    def __init__(self, code, msg=''):
        super().__init__(f'axtrack_b200 error {code}: {msg}')
        self.code = code


class Context:
    """Owns one axb_ctx (device scratch, sticky status).  One per device per thread."""

    def __init__(self, device=0):
        self.lib = load()
        h = ctypes.c_void_p()
        rc = self.lib.axb_create(int(device), ctypes.byref(h))
        if rc != OK:
            raise AxbError(rc, 'axb_create failed: no usable CUDA device (axtrack_b200 has no CPU fallback)')
        self.handle = h
        self.device = int(device)

    def close(self):
        if getattr(self, 'handle', None):
            self.lib.axb_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != OK:
            raise AxbError(rc, self.lib.axb_last_error(self.handle).decode())

    @property
    def launches(self):
        return int(self.lib.axb_launch_count(self.handle))

    def status(self, stream=None):
        """-> (code, (pair, source, target)); synchronises the stream and clears the status."""
        info = (ctypes.c_int64 * 3)()
        rc = self.lib.axb_status(self.handle, stream, ctypes.byref(info))
        return rc, (int(info[0]), int(info[1]), int(info[2]))
