"""ctypes access to the CPU oracle (oracle/liborb_oracle.so) and, when present, the compiled reference
(oracle/_ref). TEST INFRASTRUCTURE: imported only from tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])

_c = ctypes
_vp, _i, _sz, _ip, _fp = _c.c_void_p, _c.c_int, _c.c_size_t, _c.POINTER(_c.c_int), _c.POINTER(_c.c_float)


def build_oracle():
    lib = os.path.join(ORACLE_DIR, "liborb_oracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))]
    if not os.path.exists(lib) or any(os.path.getmtime(s) > os.path.getmtime(lib) for s in srcs):
        subprocess.run(["make", "-C", ORACLE_DIR, "liborb_oracle.so"], check=True, capture_output=True)
    return lib


_oracle = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        L = ctypes.CDLL(build_oracle())
        L.orb_oracle_create.restype = _vp
        L.orb_oracle_create.argtypes = [_i, _c.c_float, _i, _i, _i]
        L.orb_oracle_destroy.argtypes = [_vp]
        L.orb_oracle_extract.argtypes = [_vp, _vp, _i, _i, _sz, _vp, _vp, _i]
        L.orb_oracle_tables.argtypes = [_vp, _fp, _fp, _fp, _fp, _ip, _ip]
        L.orb_oracle_level_size.argtypes = [_vp, _i, _ip, _ip]
        L.orb_oracle_level_image.argtypes = [_vp, _i, _i, _vp, _i]
        L.orb_oracle_level_corners.argtypes = [_vp, _i, _vp, _i]
        L.orb_oracle_level_keypoints.argtypes = [_vp, _i, _vp, _i]
        L.orb_oracle_distribute.argtypes = [_vp, _i, _i, _i, _i, _vp, _i]
        _oracle = L
    return _oracle


class OracleExtractor:
    """CPU oracle with the reference's constructor arguments (cslam/src/ORBextractor.cpp:579)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST):
        self.L = oracle_lib()
        self.nlevels = nlevels
        self.h = self.L.orb_oracle_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orb_oracle_destroy(self.h)
            self.h = None

    def __call__(self, image, cap=20000):
        image = np.ascontiguousarray(image)
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.orb_oracle_extract(self.h, image.ctypes.data, image.shape[0], image.shape[1], image.strides[0],
                                      kp.ctypes.data, desc.ctypes.data, cap)
        assert n <= cap
        return kp[:n].copy(), desc[:n].copy()

    def tables(self):
        n = self.nlevels
        a = [(ctypes.c_float * n)() for _ in range(4)]
        q = (ctypes.c_int * n)()
        um = (ctypes.c_int * 16)()
        self.L.orb_oracle_tables(self.h, a[0], a[1], a[2], a[3], q, um)
        return [np.array(x[:], np.float32) for x in a], np.array(q[:]), np.array(um[:])

    def level_image(self, level, blurred=False):
        w, h = ctypes.c_int(), ctypes.c_int()
        self.L.orb_oracle_level_size(self.h, level, ctypes.byref(w), ctypes.byref(h))
        buf = np.zeros((h.value, w.value), np.uint8)
        n = self.L.orb_oracle_level_image(self.h, level, 1 if blurred else 0, buf.ctypes.data, buf.size)
        return buf if n == buf.size else None

    def level_corners(self, level, cap=1 << 19):
        buf = np.zeros((cap, 3), np.int32)
        n = self.L.orb_oracle_level_corners(self.h, level, buf.ctypes.data, cap)
        return buf[:n].copy()

    def level_keypoints(self, level, cap=8192):
        kp = np.zeros(cap, KP_DTYPE)
        n = self.L.orb_oracle_level_keypoints(self.h, level, kp.ctypes.data, cap)
        return kp[:n].copy()


def ref_available(mono=True):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_orb_mono.so" if mono else "libref_orb.so"))


class RefExtractor:
    """The reference's own ORBextractor.cpp (compiled unmodified, oracle/_ref). mono=True: monotone operator new."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, mono=True):
        L = ctypes.CDLL(os.path.join(ORACLE_DIR, "_ref", "libref_orb_mono.so" if mono else "libref_orb.so"))
        L.ref_orb_create.restype = _vp
        L.ref_orb_create.argtypes = [_i, _c.c_float, _i, _i, _i]
        L.ref_orb_destroy.argtypes = [_vp]
        L.ref_orb_extract.argtypes = [_vp, _vp, _i, _i, _sz, _vp, _vp, _i]
        L.ref_orb_pyramid_level.argtypes = [_vp, _i, _vp, _i, _ip, _ip]
        L.ref_orb_tables.argtypes = [_i, _c.c_float, _i, _i, _i, _fp, _fp, _fp, _fp, _ip, _ip]
        self.L = L
        self.args = (nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self.h = L.ref_orb_create(*self.args)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_orb_destroy(self.h)
            self.h = None

    def __call__(self, image, cap=20000):
        image = np.ascontiguousarray(image)
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.ref_orb_extract(self.h, image.ctypes.data, image.shape[0], image.shape[1], image.strides[0],
                                   kp.ctypes.data, desc.ctypes.data, cap)
        assert n <= cap
        return kp[:n].copy(), desc[:n].copy()

    def tables(self):
        n = self.args[2]
        a = [(ctypes.c_float * n)() for _ in range(4)]
        q = (ctypes.c_int * n)()
        um = (ctypes.c_int * 16)()
        self.L.ref_orb_tables(*self.args, a[0], a[1], a[2], a[3], q, um)
        return [np.array(x[:], np.float32) for x in a], np.array(q[:]), np.array(um[:])

    def pyramid_level(self, level):
        w, h = ctypes.c_int(), ctypes.c_int()
        buf = np.zeros(4096 * 4096, np.uint8)
        rc = self.L.ref_orb_pyramid_level(self.h, level, buf.ctypes.data, buf.size, ctypes.byref(w), ctypes.byref(h))
        assert rc == 0
        return buf[:w.value * h.value].reshape(h.value, w.value).copy()
