"""ctypes binding of libccm_orb_cuda.so (include/ccm_orb.h). Fails loudly when the library is missing:
there is no CPU fallback anywhere in this package."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libccm_orb_cuda.so")

CCM_OK, CCM_E_BADARG, CCM_E_CAPACITY, CCM_E_CUDA = 0, -1, -2, -3


class Keypoint(ctypes.Structure):
    _fields_ = [("x", ctypes.c_float), ("y", ctypes.c_float), ("size", ctypes.c_float), ("angle", ctypes.c_float),
                ("response", ctypes.c_float), ("octave", ctypes.c_int)]


class OrbConfig(ctypes.Structure):
    _fields_ = [("nfeatures", ctypes.c_int), ("scale_factor", ctypes.c_float), ("nlevels", ctypes.c_int),
                ("ini_th_fast", ctypes.c_int), ("min_th_fast", ctypes.c_int), ("max_width", ctypes.c_int),
                ("max_height", ctypes.c_int), ("max_batch", ctypes.c_int), ("device", ctypes.c_int)]


_c = ctypes
_vp, _i, _sz, _ip, _fp = _c.c_void_p, _c.c_int, _c.c_size_t, _c.POINTER(_c.c_int), _c.POINTER(_c.c_float)

# every symbol include/ccm_orb.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "ccm_orb_create": (_i, [_c.POINTER(OrbConfig), _c.POINTER(_vp)]),
    "ccm_orb_destroy": (None, [_vp]),
    "ccm_last_error": (_c.c_char_p, [_vp]),
    "ccm_orb_get_tables": (_i, [_vp, _fp, _fp, _fp, _fp, _ip]),
    "ccm_orb_max_keypoints": (_i, [_vp]),
    "ccm_orb_extract": (_i, [_vp, _vp, _i, _i, _sz, _vp, _vp, _i, _ip]),
    "ccm_orb_extract_batch": (_i, [_vp, _c.POINTER(_vp), _i, _i, _i, _sz, _vp, _vp, _i, _ip]),
    "ccm_orb_extract_batch_device": (_i, [_vp, _vp, _i, _i, _i, _sz, _sz, _vp, _vp, _i, _vp]),
    "ccm_orb_sync": (_i, [_vp]),
    "ccm_orb_stream": (_vp, [_vp]),
    "ccm_orb_launch_count": (_c.c_longlong, [_vp]),
    "ccm_orb_set_stage_timing": (_i, [_vp, _i]),
    "ccm_orb_get_stage_ms": (_i, [_vp, _c.POINTER(_c.c_double), _c.POINTER(_c.c_longlong), _i]),
    "ccm_orb_get_pyramid_level": (_i, [_vp, _i, _i, _i, _vp, _sz, _ip, _ip]),
    "ccm_orb_debug_blurred_level": (_i, [_vp, _i, _i, _vp, _sz, _ip, _ip]),
    "ccm_orb_debug_level_corners": (_i, [_vp, _i, _i, _vp, _i, _ip]),
}

_lib = None


def load():
    """Load the CUDA library; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libccm_orb_cuda.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class CcmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("ccm error %d: %s" % (code, msg))
        self.code = code
