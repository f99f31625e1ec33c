"""ctypes binding of libccm_orb_cuda.so (include/ccm_orb.h, include/ccm_match.h). Fails loudly when the library is missing:
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
    "ccm_orb_extract_batch_submit": (_i, [_vp, _c.POINTER(_vp), _i, _i, _i, _sz, _vp, _vp, _i, _ip, _ip]),
    "ccm_orb_extract_batch_wait": (_i, [_vp, _i]),
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



class GridParams(ctypes.Structure):
    _fields_ = [("min_x", ctypes.c_float), ("min_y", ctypes.c_float), ("inv_w", ctypes.c_float), ("inv_h", ctypes.c_float),
                ("cols", ctypes.c_int), ("rows", ctypes.c_int)]


class Camera(ctypes.Structure):
    """ccm_camera: mK and mDistCoef of a Frame (Tracking.cpp:38-58)"""
    _fields_ = [("fx", ctypes.c_float), ("fy", ctypes.c_float), ("cx", ctypes.c_float), ("cy", ctypes.c_float),
                ("dist", ctypes.c_float * 5), ("n_dist", ctypes.c_int)]


class EpipolarGeom(ctypes.Structure):
    _fields_ = [("f12", ctypes.c_float * 9), ("ex", ctypes.c_float), ("ey", ctypes.c_float),
                ("scale_factors", ctypes.POINTER(ctypes.c_float)), ("level_sigma2", ctypes.POINTER(ctypes.c_float)),
                ("n_levels", ctypes.c_int)]


class WindowQuery(ctypes.Structure):
    _fields_ = [("u", ctypes.c_float), ("v", ctypes.c_float), ("r", ctypes.c_float), ("angle", ctypes.c_float),
                ("min_level", ctypes.c_int16), ("max_level", ctypes.c_int16), ("flags", ctypes.c_uint32)]


class WindowParams(ctypes.Structure):
    _fields_ = [("mode", ctypes.c_int), ("nnratio", ctypes.c_float), ("th_dist", ctypes.c_int),
                ("check_orientation", ctypes.c_int), ("inv_sigma2", ctypes.POINTER(ctypes.c_float)), ("n_levels", ctypes.c_int)]


_ll, _u32, _f = _c.c_longlong, _c.c_uint32, _c.c_float
_gp, _wp, _cp = _c.POINTER(GridParams), _c.POINTER(WindowParams), _c.POINTER(Camera)

# every symbol include/ccm_match.h declares
SIGNATURES.update({
    "ccm_matcher_create": (_i, [_i, _c.POINTER(_vp)]),
    "ccm_matcher_destroy": (None, [_vp]),
    "ccm_matcher_last_error": (_c.c_char_p, [_vp]),
    "ccm_matcher_sync": (_i, [_vp]),
    "ccm_matcher_stream": (_vp, [_vp]),
    "ccm_matcher_launch_count": (_ll, [_vp]),
    "ccm_descriptor_distance": (_i, [_vp, _vp, _vp, _i, _vp]),
    "ccm_frame_pack_create": (_i, [_vp, _i, _c.POINTER(_vp)]),
    "ccm_frame_pack_destroy": (None, [_vp]),
    "ccm_frame_pack_upload": (_i, [_vp, _gp, _i, _vp, _vp, _vp, _vp, _vp]),
    "ccm_frame_pack_from_extractor_device": (_i, [_vp, _gp, _vp, _vp, _vp, _vp]),
    "ccm_frame_pack_from_extractor_batch_device": (_i, [_c.POINTER(_vp), _i, _gp, _vp, _vp, _vp, _i, _vp]),
    "ccm_frame_pack_from_extractor_batch_device_cam": (_i, [_c.POINTER(_vp), _i, _gp, _cp, _vp, _vp, _vp, _i]),
    "ccm_undistort_points": (_i, [_vp, _cp, _i, _vp, _vp]),
    "ccm_undistort_keypoints_batch_device": (_i, [_vp, _cp, _i, _vp, _vp, _i, _vp]),
    "ccm_image_bounds": (_i, [_vp, _cp, _i, _i, _gp, _vp]),
    "ccm_wire_encode_keyframe_device": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "ccm_wire_encode_keyframe": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "ccm_frame_pack_from_wire": (_i, [_vp, _gp, _vp, _sz, _vp, _sz, _vp]),
    "ccm_frame_pack_get_grid": (_i, [_vp, _vp, _vp, _ip]),
    "ccm_frame_pack_count": (_i, [_vp, _ip]),
    "ccm_search_window": (_i, [_vp, _vp, _wp, _i, _vp, _vp, _vp, _vp, _ip]),
    "ccm_search_window_batch_device": (_i, [_vp, _c.POINTER(_vp), _i, _wp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "ccm_search_bow": (_i, [_vp, _i, _f, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ip]),
    "ccm_vocabulary_create": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _c.POINTER(_vp)]),
    "ccm_vocabulary_destroy": (None, [_vp]),
    "ccm_vocabulary_transform": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp]),
    "ccm_vocabulary_transform_device": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp]),
    "ccm_distinctive_descriptors": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "ccm_search_triangulation": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ip]),
    "ccm_exchange_create": (_i, [_vp, _i, _i, _i, _c.POINTER(_vp), _vp]),
    "ccm_exchange_connect_ipc": (_i, [_vp, _vp]),
    "ccm_exchange_connect_ptrs": (_i, [_vp, _c.POINTER(_vp)]),
    "ccm_exchange_buffer": (_vp, [_vp]),
    "ccm_exchange_destroy": (None, [_vp]),
    "ccm_exchange_check": (_i, [_vp]),
    "ccm_match_allpairs_best2_exchange_device": (_i, [_vp, _vp, _vp, _i, _vp, _ll, _u32, _f, _vp, _vp]),
    "ccm_match_allpairs_best2_device": (_i, [_vp, _vp, _i, _vp, _ll, _u32, _vp]),
    "ccm_match_allpairs_merge_device": (_i, [_vp, _vp, _i, _i, _f, _vp, _vp]),
    "ccm_match_allpairs_best2": (_i, [_vp, _vp, _i, _vp, _ll, _u32, _vp]),
    "ccm_match_allpairs_best2_noalloc_device": (_i, [_vp, _vp, _i, _vp, _ll, _u32, _vp]),
})

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
