"""Python mirror of cslam::ORBextractor (cslam/include/cslam/ORBextractor.h:97-164) over the C ABI.

Same constructor arguments, same getters, `__call__` in place of `operator()`. Used by tests and bench.py; the
C++ drop-in class with the reference's exact declaration lives in motioncheck_ccm_slam_b200/host/.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import CcmError, Keypoint, OrbConfig

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])
assert KP_DTYPE.itemsize == ctypes.sizeof(Keypoint) == 24


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — ORBextractor.h:103-105."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width=752, max_height=480,
                 max_batch=1, device=0):
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        cfg = OrbConfig(int(nfeatures), float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST),
                        int(max_width), int(max_height), int(max_batch), int(device))
        rc = self._lib.ccm_orb_create(ctypes.byref(cfg), ctypes.byref(self._h))
        if rc != 0:
            raise CcmError(rc, self._lib.ccm_last_error(None).decode())
        self.nfeatures, self.nlevels, self.max_batch = int(nfeatures), int(nlevels), int(max_batch)
        self._scale_factor = float(np.float32(scaleFactor))
        n = self.nlevels
        arr = [(ctypes.c_float * n)() for _ in range(4)]
        q = (ctypes.c_int * n)()
        self._check(self._lib.ccm_orb_get_tables(self._h, arr[0], arr[1], arr[2], arr[3], q))
        self._tables = [np.array(a[:], dtype=np.float32) for a in arr]
        self.features_per_level = np.array(q[:], dtype=np.int32)
        self.mvImagePyramid = _PyramidView(self)

    # -- reference getters (ORBextractor.h:112-136) --
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self._scale_factor

    def GetScaleFactors(self):
        return self._tables[0].copy()

    def GetInverseScaleFactors(self):
        return self._tables[1].copy()

    def GetScaleSigmaSquares(self):
        return self._tables[2].copy()

    def GetInverseScaleSigmaSquares(self):
        return self._tables[3].copy()

    def max_keypoints(self):
        return int(self._lib.ccm_orb_max_keypoints(self._h))

    def launch_count(self):
        return int(self._lib.ccm_orb_launch_count(self._h))

    def _check(self, rc):
        if rc != 0:
            raise CcmError(rc, self._lib.ccm_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.ccm_orb_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- operator() (ORBextractor.cpp:1216-1278). mask is ignored, as in the reference (ORBextractor.h:109-111) --
    def __call__(self, image, mask=None):
        """image: uint8 [H, W] (any row stride). Returns (keypoints structured array, descriptors uint8 [N, 32])."""
        im = np.asarray(image)
        if im.dtype != np.uint8 or im.ndim != 2:
            raise TypeError("images must be 2-D uint8 (CV_8UC1), as ORBextractor.cpp:1223 asserts")
        rows, cols = im.shape
        if rows == 0 or cols == 0:      # ORBextractor.cpp:1219-1220: empty image, outputs untouched
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        if im.strides[1] != 1:
            raise ValueError("unit pixel stride expected")
        cap = self.max_keypoints_for(rows, cols)
        kp = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        cnt = ctypes.c_int()
        self._check(self._lib.ccm_orb_extract(self._h, im.ctypes.data, rows, cols, im.strides[0], kp.ctypes.data, desc.ctypes.data, cap,
                                              ctypes.byref(cnt)))
        return kp[:cnt.value].copy(), desc[:cnt.value].copy()

    def extract_raw(self, img_ptr, rows, cols, step, out_kp_ptr, out_desc_ptr, capacity, out_n):
        """ccm_orb_extract (one frame, synchronous: the call Frame::ExtractORB makes) with caller-owned host buffers."""
        self._check(self._lib.ccm_orb_extract(self._h, img_ptr, rows, cols, step, out_kp_ptr, out_desc_ptr, capacity, out_n))

    def extract_batch(self, images):
        """images: sequence of uint8 [H, W] arrays of one size (or one [B, H, W] array). Host in, host out."""
        imgs = [np.asarray(im) for im in images]
        n = len(imgs)
        if n == 0:
            return [], []
        for im in imgs:
            if im.dtype != np.uint8 or im.ndim != 2:
                raise TypeError("images must be 2-D uint8 (CV_8UC1), as ORBextractor.cpp:1223 asserts")
        rows, cols = imgs[0].shape
        if rows == 0 or cols == 0:      # ORBextractor.cpp:1219-1220: empty image, outputs untouched
            return [np.zeros(0, KP_DTYPE) for _ in imgs], [np.zeros((0, 32), np.uint8) for _ in imgs]
        for im in imgs:
            if im.shape != imgs[0].shape or im.strides != imgs[0].strides or im.strides[1] != 1:
                raise ValueError("batch images must share shape and row stride, with unit pixel stride")
        cap = self.max_keypoints_for(rows, cols)
        kp = np.zeros((n, cap), dtype=KP_DTYPE)
        desc = np.zeros((n, cap, 32), dtype=np.uint8)
        cnt = (ctypes.c_int * n)()
        ptrs = (ctypes.c_void_p * n)(*[im.ctypes.data for im in imgs])
        self._check(self._lib.ccm_orb_extract_batch(self._h, ptrs, n, rows, cols, imgs[0].strides[0],
                                                    kp.ctypes.data, desc.ctypes.data, cap, cnt))
        return [kp[f, :cnt[f]].copy() for f in range(n)], [desc[f, :cnt[f]].copy() for f in range(n)]

    def max_keypoints_for(self, rows, cols):
        # generous, shape-independent bound: quota + 3 + 4*nIni per level with nIni <= 16
        return max(self.max_keypoints(), int(self.features_per_level.sum() + self.nlevels * 67))

    def extract_batch_device(self, d_imgs_ptr, n, rows, cols, pitch, frame_stride, d_kp_ptr, d_desc_ptr, capacity, d_n_ptr):
        """Device-resident path (raw device pointers, e.g. torch tensors' data_ptr()); asynchronous."""
        self._check(self._lib.ccm_orb_extract_batch_device(self._h, d_imgs_ptr, n, rows, cols, pitch, frame_stride,
                                                           d_kp_ptr, d_desc_ptr, capacity, d_n_ptr))

    def extract_batch_raw(self, img_ptrs, n, rows, cols, step, out_kp_ptr, out_desc_ptr, capacity, out_n):
        """ccm_orb_extract_batch with caller-owned host buffers (e.g. pinned): no numpy allocation per call."""
        self._check(self._lib.ccm_orb_extract_batch(self._h, img_ptrs, n, rows, cols, step, out_kp_ptr, out_desc_ptr,
                                                    capacity, out_n))

    def extract_batch_submit(self, img_ptrs, n, rows, cols, step, out_kp_ptr, out_desc_ptr, capacity, out_n):
        """ccm_orb_extract_batch_submit: enqueue one batch and return its ticket; the buffers stay owned by the caller
        until extract_batch_wait(ticket) returns. Two batches may be in flight (their frames together <= max_batch)."""
        t = ctypes.c_int(-1)
        self._check(self._lib.ccm_orb_extract_batch_submit(self._h, img_ptrs, n, rows, cols, step, out_kp_ptr, out_desc_ptr,
                                                           capacity, out_n, ctypes.byref(t)))
        return t.value

    def extract_batch_wait(self, ticket):
        self._check(self._lib.ccm_orb_extract_batch_wait(self._h, int(ticket)))

    def set_stage_timing(self, on):
        self._check(self._lib.ccm_orb_set_stage_timing(self._h, 1 if on else 0))

    def get_stage_ms(self, reset=False):
        """(per-stage summed CUDA-event ms [resize, fast, blur, quadtree, describe], number of timed batches)."""
        ms = (ctypes.c_double * 5)()
        calls = ctypes.c_longlong()
        self._check(self._lib.ccm_orb_get_stage_ms(self._h, ms, ctypes.byref(calls), 1 if reset else 0))
        return list(ms), int(calls.value)

    def sync(self):
        self._check(self._lib.ccm_orb_sync(self._h))

    def stream(self):
        return self._lib.ccm_orb_stream(self._h)

    # -- stage taps --
    def pyramid_level(self, level, frame=0, with_border=False, blurred=False):
        w = ctypes.c_int()
        h = ctypes.c_int()
        side = 4096 + 38
        buf = np.zeros((side, side), dtype=np.uint8)
        if blurred:
            rc = self._lib.ccm_orb_debug_blurred_level(self._h, frame, level, buf.ctypes.data, side, ctypes.byref(w), ctypes.byref(h))
        else:
            rc = self._lib.ccm_orb_get_pyramid_level(self._h, frame, level, 1 if with_border else 0, buf.ctypes.data, side,
                                                     ctypes.byref(w), ctypes.byref(h))
        self._check(rc)
        return buf[:h.value, :w.value].copy()

    def level_corners(self, level, frame=0):
        n = ctypes.c_int()
        cap = 1 << 19
        buf = np.zeros((cap, 3), dtype=np.int32)
        self._check(self._lib.ccm_orb_debug_level_corners(self._h, frame, level, buf.ctypes.data, cap, ctypes.byref(n)))
        return buf[:n.value].copy()


class _PyramidView:
    """mvImagePyramid (ORBextractor.h:138): level images of the last call, read back on demand."""

    def __init__(self, ex):
        self._ex = ex

    def __len__(self):
        return self._ex.nlevels

    def __getitem__(self, level):
        return self._ex.pyramid_level(level, 0, with_border=False)
