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


def oracle_undistort(xy, K, dist):
    """oracle/cv_primitives.cpp ocv_undistort_points_f32: cv::undistortPoints(xy, K, dist, None, K) as Frame.cpp:295 calls it."""
    L = oracle_lib()
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float32).reshape(9)
    d = np.ascontiguousarray(dist, np.float32).ravel()
    out = np.empty_like(xy)
    vp = ctypes.c_void_p
    L.ocv_undistort_points_f32.argtypes = [vp, vp, ctypes.c_int, vp, vp, ctypes.c_int, vp]
    L.ocv_undistort_points_f32.restype = None
    L.ocv_undistort_points_f32(xy.ctypes.data, out.ctypes.data, len(xy), K.ctypes.data, d.ctypes.data, d.size, K.ctypes.data)
    return out


def oracle_image_bounds(K, dist, w, h, cols=75, rows=48):
    """Frame::ComputeImageBounds + Frame.cpp:87-88 -> (minX, maxX, minY, maxY, invW, invH) as float32."""
    f = np.float32
    if f(dist[0]) == 0:
        mn_x, mx_x, mn_y, mx_y = f(0), f(w), f(0), f(h)
    else:
        c = oracle_undistort(np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32), K, dist)
        mn_x, mx_x = min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0])
        mn_y, mx_y = min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])
    return np.array([mn_x, mx_x, mn_y, mx_y, f(cols) / (mx_x - mn_x), f(rows) / (mx_y - mn_y)], np.float32)


def oracle_wire_encode(kp, desc):
    """oracle/match_oracle.cpp mo_wire_encode_keyframe on KP_DTYPE-like records -> (kp_bytes, desc_bytes)."""
    L = match_lib()
    n = len(kp)
    a = {k: np.ascontiguousarray(kp[k], np.int32 if k == "octave" else np.float32) for k in ("x", "y", "size", "angle", "response", "octave")}
    d = np.ascontiguousarray(desc, np.uint8).reshape(n, 32)
    kb, db = np.zeros(4 + 15 * n, np.uint8), np.zeros(4 + 32 * n, np.uint8)
    L.mo_wire_encode_keyframe.restype = None
    L.mo_wire_encode_keyframe.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 9
    L.mo_wire_encode_keyframe(n, *[a[k].ctypes.data for k in ("x", "y", "size", "angle", "response", "octave")], d.ctypes.data, kb.ctypes.data, db.ctypes.data)
    return kb.tobytes(), db.tobytes()


def oracle_wire_decode(kp_bytes, desc_bytes, cap=1 << 16):
    """mo_wire_decode_keyframe -> (dict of x, y, size, angle, response, octave; desc [n][32]) or None when the counts disagree."""
    L = match_lib()
    kb, db = np.frombuffer(kp_bytes, np.uint8), np.frombuffer(desc_bytes, np.uint8)
    f = {k: np.zeros(cap, np.float32) for k in ("x", "y", "size", "angle", "response")}
    oc, d = np.zeros(cap, np.int32), np.zeros((cap, 32), np.uint8)
    L.mo_wire_decode_keyframe.restype = ctypes.c_int
    L.mo_wire_decode_keyframe.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 7
    n = L.mo_wire_decode_keyframe(kb.ctypes.data, db.ctypes.data, cap, f["x"].ctypes.data, f["y"].ctypes.data, f["size"].ctypes.data,
                                  f["angle"].ctypes.data, f["response"].ctypes.data, oc.ctypes.data, d.ctypes.data)
    if n < 0:
        return None
    out = {k: v[:n] for k, v in f.items()}
    out["octave"] = oc[:n]
    return out, d[:n]


def ref_available(mono=True):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_orb_mono.so" if mono else "libref_orb.so"))


def dropin_available():
    """oracle/_ref/libdropin_orb.so: the product's drop-in cslam::ORBextractor class behind the reference glue."""
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libdropin_orb.so"))


class RefExtractor:
    """The reference's own ORBextractor.cpp (compiled unmodified, oracle/_ref). mono=True: monotone operator new."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, mono=True, lib=None):
        L = ctypes.CDLL(os.path.join(ORACLE_DIR, "_ref", lib or ("libref_orb_mono.so" if mono else "libref_orb.so")))
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


# ---------------------------------------------------------------------------------------------------------------
# matcher oracle (oracle/match_oracle.cpp): restated ORBmatcher loops over plain arrays
# ---------------------------------------------------------------------------------------------------------------
RESULT_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("best_dist2", "<i4"), ("accepted", "<i4")])
_u8p = _vp
_mo = None


def match_lib():
    global _mo
    if _mo is None:
        L = oracle_lib()
        L.mo_descriptor_distance.argtypes = [_vp, _vp]
        L.mo_frame_create.restype = _vp
        L.mo_frame_create.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _i, _i]
        L.mo_frame_destroy.argtypes = [_vp]
        L.mo_frame_grid.argtypes = [_vp, _vp, _vp]
        L.mo_features_in_area.argtypes = [_vp, _c.c_float, _c.c_float, _c.c_float, _i, _i, _i, _vp, _i]
        L.mo_search_projection_mappoints.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _c.c_float, _vp, _vp, _vp]
        L.mo_search_projection_lastframe.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _vp, _i, _vp, _vp]
        L.mo_search_projection_kf_sim3.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp]
        L.mo_search_best_in_window.restype = None
        L.mo_search_best_in_window.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp]
        L.mo_search_bow_kf_frame.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _c.c_float, _i, _vp, _vp]
        L.mo_search_bow_kf_kf.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _c.c_float, _i, _vp, _vp]
        L.mo_search_for_initialization.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _i, _c.c_float, _i, _vp]
        L.mo_match_allpairs_best2.restype = None
        L.mo_match_allpairs_best2.argtypes = [_i, _vp, _c.c_longlong, _vp, _c.c_float, _vp]
        _mo = L
    return _mo


def _a(x, dt):
    return np.ascontiguousarray(x, dt)


class OracleFrame:
    """Matching data of a Frame / KeyFrame: descriptors, undistorted keypoints and the 75x48 feature grid."""

    def __init__(self, desc, x, y, octave, angle, grid):
        self.L = match_lib()
        self.desc = _a(desc, np.uint8).reshape(-1, 32)
        self.x, self.y, self.octave, self.angle = _a(x, np.float32), _a(y, np.float32), _a(octave, np.int32), _a(angle, np.float32)
        self.n = len(self.desc)
        self.grid = grid
        self.h = self.L.mo_frame_create(self.n, self.desc.ctypes.data, self.x.ctypes.data, self.y.ctypes.data,
                                        self.octave.ctypes.data, self.angle.ctypes.data, grid.min_x, grid.min_y,
                                        grid.inv_w, grid.inv_h, grid.cols, grid.rows)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.mo_frame_destroy(self.h)
            self.h = None

    def csr(self):
        cs = np.zeros(self.grid.cols * self.grid.rows + 1, np.int32)
        items = np.zeros(max(self.n, 1), np.int32)
        k = self.L.mo_frame_grid(self.h, cs.ctypes.data, items.ctypes.data)
        return cs, items[:k].copy()

    def features_in_area(self, x, y, r, min_level, max_level, kf_variant=False):
        out = np.zeros(max(self.n, 1), np.int32)
        k = self.L.mo_features_in_area(self.h, x, y, r, min_level, max_level, 1 if kf_variant else 0, out.ctypes.data, len(out))
        return out[:k].copy()


class OracleMatcher:
    """CPU oracle with the reference's ORBmatcher(nnratio, checkOri) arguments."""

    def __init__(self, nnratio=0.6, checkOri=True):
        self.L = match_lib()
        self.nnratio = float(np.float32(nnratio))
        self.ori = 1 if checkOri else 0

    def descriptor_distance(self, a, b):
        a = _a(a, np.uint8).reshape(-1, 32); b = _a(b, np.uint8).reshape(-1, 32)
        return np.array([self.L.mo_descriptor_distance(a[i].ctypes.data, b[i].ctypes.data) for i in range(len(a))], np.int32)

    def projection_mappoints(self, F, q_valid, q_u, q_v, q_level, q_viewcos, q_desc, q_obs, th, sf, holder):
        n = len(q_u)
        res = np.zeros(n, RESULT_DTYPE)
        v, u_, v_, l, c, d, o, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_level, np.int32),
                                    _a(q_viewcos, np.float32), _a(q_desc, np.uint8), _a(q_obs, np.uint8), _a(sf, np.float32))
        nm = self.L.mo_search_projection_mappoints(F.h, n, v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                   c.ctypes.data, d.ctypes.data, o.ctypes.data, th, self.nnratio,
                                                   s.ctypes.data, holder.ctypes.data, res.ctypes.data)
        return nm, res

    def projection_lastframe(self, F, q_valid, q_u, q_v, q_octave, q_angle, q_desc, q_obs, th, sf, holder):
        n = len(q_u)
        res = np.zeros(n, RESULT_DTYPE)
        v, u_, v_, l, a, d, o, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_octave, np.int32),
                                    _a(q_angle, np.float32), _a(q_desc, np.uint8), _a(q_obs, np.uint8), _a(sf, np.float32))
        nm = self.L.mo_search_projection_lastframe(F.h, n, v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                   a.ctypes.data, d.ctypes.data, o.ctypes.data, th, s.ctypes.data, self.ori,
                                                   holder.ctypes.data, res.ctypes.data)
        return nm, res

    def projection_kf_sim3(self, F, q_valid, q_u, q_v, q_level, q_desc, q_claims, th, sf, holder):
        n = len(q_u)
        res = np.zeros(n, RESULT_DTYPE)
        v, u_, v_, l, d, c, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_level, np.int32),
                                 _a(q_desc, np.uint8), _a(q_claims, np.uint8), _a(sf, np.float32))
        nm = self.L.mo_search_projection_kf_sim3(F.h, n, v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                 d.ctypes.data, c.ctypes.data, int(th), s.ctypes.data, holder.ctypes.data,
                                                 res.ctypes.data)
        return nm, res

    def best_in_window(self, F, q_valid, q_u, q_v, q_radius, q_level, q_desc, th_dist, inv_sigma2=None):
        n = len(q_u)
        res = np.zeros(n, RESULT_DTYPE)
        v, u_, v_, r, l, d = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_radius, np.float32),
                              _a(q_level, np.int32), _a(q_desc, np.uint8))
        s = _a(inv_sigma2, np.float32) if inv_sigma2 is not None else None
        self.L.mo_search_best_in_window(F.h, n, v.ctypes.data, u_.ctypes.data, v_.ctypes.data, r.ctypes.data, l.ctypes.data,
                                        d.ctypes.data, th_dist, s.ctypes.data if s is not None else None, res.ctypes.data)
        return int(res["accepted"].sum()), res

    def bow_kf_frame(self, desc_kf, node_kf, valid_kf, angle_kf, desc_f, node_f, angle_f):
        dk, nk, vk, ak = _a(desc_kf, np.uint8), _a(node_kf, np.int32), _a(valid_kf, np.uint8), _a(angle_kf, np.float32)
        df, nf, af = _a(desc_f, np.uint8), _a(node_f, np.int32), _a(angle_f, np.float32)
        mf = np.full(max(len(nf), 1), -1, np.int32)
        res = np.zeros(max(len(nk), 1), RESULT_DTYPE)
        nm = self.L.mo_search_bow_kf_frame(len(nk), dk.ctypes.data, nk.ctypes.data, vk.ctypes.data, ak.ctypes.data, len(nf),
                                           df.ctypes.data, nf.ctypes.data, af.ctypes.data, self.nnratio, self.ori,
                                           mf.ctypes.data, res.ctypes.data)
        return nm, mf[:len(nf)], res[:len(nk)]

    def bow_kf_kf(self, d1, n1, v1, a1, d2, n2, v2, a2):
        d1, n1, v1, a1 = _a(d1, np.uint8), _a(n1, np.int32), _a(v1, np.uint8), _a(a1, np.float32)
        d2, n2, v2, a2 = _a(d2, np.uint8), _a(n2, np.int32), _a(v2, np.uint8), _a(a2, np.float32)
        m12 = np.full(max(len(n1), 1), -1, np.int32)
        res = np.zeros(max(len(n1), 1), RESULT_DTYPE)
        nm = self.L.mo_search_bow_kf_kf(len(n1), d1.ctypes.data, n1.ctypes.data, v1.ctypes.data, a1.ctypes.data, len(n2),
                                        d2.ctypes.data, n2.ctypes.data, v2.ctypes.data, a2.ctypes.data, self.nnratio, self.ori,
                                        m12.ctypes.data, res.ctypes.data)
        return nm, m12[:len(n1)], res[:len(n1)]

    def search_for_initialization(self, desc1, octave1, angle1, prev_xy, F2, windowSize):
        d1, o1, a1 = _a(desc1, np.uint8), _a(octave1, np.int32), _a(angle1, np.float32)
        assert prev_xy.dtype == np.float32 and prev_xy.flags.c_contiguous
        m12 = np.full(max(len(o1), 1), -1, np.int32)
        nm = self.L.mo_search_for_initialization(len(o1), d1.ctypes.data, o1.ctypes.data, a1.ctypes.data, prev_xy.ctypes.data,
                                                 F2.h, int(windowSize), self.nnratio, self.ori, m12.ctypes.data)
        return nm, m12[:len(o1)]

    def search_for_triangulation(self, c, sf, level_sigma2, ori):
        return _triangulation(self.L, self.L.mo_search_for_triangulation, c, sf, level_sigma2, ori)

    def allpairs_best2(self, q, t):
        q = _a(q, np.uint8).reshape(-1, 32); t = _a(t, np.uint8).reshape(-1, 32)
        res = np.zeros(max(len(q), 1), RESULT_DTYPE)
        self.L.mo_match_allpairs_best2(len(q), q.ctypes.data, len(t), t.ctypes.data, self.nnratio, res.ctypes.data)
        return res[:len(q)]


# ---------------------------------------------------------------------------------------------------------------
# the reference's own ORBmatcher.cpp, compiled unmodified over mock Frame/KeyFrame/MapPoint (oracle/_ref/libref_match.so)
# ---------------------------------------------------------------------------------------------------------------
def _triangulation(L, fn, c, sf, level_sigma2, ori):
    """SearchForTriangulation through `fn` (mo_* of the oracle or rm_* of a reference/drop-in build): c is a
    tests/_fuse_cases.tri_case dict. Returns (nmatches, match12)."""
    a, b = c["kf1"], c["kf2"]
    arrs = [_a(a["desc"], np.uint8), _a(c["node1"], np.int32), _a(c["free1"], np.uint8), _a(a["x"], np.float32), _a(a["y"], np.float32),
            _a(a["angle"], np.float32)]
    brrs = [_a(b["desc"], np.uint8), _a(c["node2"], np.int32), _a(c["free2"], np.uint8), _a(b["x"], np.float32), _a(b["y"], np.float32),
            _a(b["octave"], np.int32), _a(b["angle"], np.float32)]
    f12, s, g = _a(c["f12"], np.float32), _a(sf, np.float32), _a(level_sigma2, np.float32)
    n1, n2 = len(arrs[1]), len(brrs[1])
    m12 = np.full(max(n1, 1), -7, np.int32)
    fn.restype = _i
    fn.argtypes = [_i] + [_vp] * 6 + [_i] + [_vp] * 7 + [_vp, _c.c_float, _c.c_float, _vp, _vp, _i, _vp]
    nm = fn(n1, *[x.ctypes.data for x in arrs], n2, *[x.ctypes.data for x in brrs], f12.ctypes.data, c["ex"], c["ey"],
            s.ctypes.data, g.ctypes.data, int(ori), m12.ctypes.data)
    return nm, m12[:n1]


def ref_match_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_match.so"))


def dropin_match_available():
    """oracle/_ref/libdropin_match.so: the product's drop-in ORBmatcher bodies behind the reference glue."""
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libdropin_match.so"))


class RefMatcher:
    """Same array interface as OracleMatcher, executed by the reference's ORBmatcher methods. `F` arguments are
    OracleFrame objects whose handles must come from the SAME library (pass lib=self.L to OracleFrame)."""

    def __init__(self, nnratio=0.6, checkOri=True, lib="libref_match.so"):
        L = ctypes.CDLL(os.path.join(ORACLE_DIR, "_ref", lib))
        L.mo_frame_create.restype = _vp
        L.mo_frame_create.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _i, _i]
        L.mo_frame_destroy.argtypes = [_vp]
        L.rm_descriptor_distance.argtypes = [_vp, _vp]
        L.rm_search_projection_mappoints.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _c.c_float, _vp, _vp]
        L.rm_search_projection_lastframe.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _vp, _i, _vp]
        L.rm_search_projection_kf_sim3.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp]
        L.rm_search_projection_reloc.argtypes = [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _i, _vp, _i, _vp]
        L.rm_search_bow_kf_frame.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _c.c_float, _i, _vp]
        L.rm_search_bow_kf_kf.argtypes = [_i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _c.c_float, _i, _vp]
        L.rm_search_for_initialization.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _i, _c.c_float, _i, _vp]
        L.rm_fuse.argtypes = [_vp, _i] + [_vp] * 8 + [_c.c_float] + [_vp] * 10
        L.rm_fuse_sim3.argtypes = [_vp, _i] + [_vp] * 8 + [_c.c_float] + [_vp] * 7
        L.rm_search_by_sim3.argtypes = [_vp, _vp] + [_vp] * 11 + [_c.c_float, _vp, _vp]
        self.L = L
        self.nnratio = float(np.float32(nnratio))
        self.ori = 1 if checkOri else 0

    def frame(self, fr, grid):
        """Frame handle living in libref_match.so's copy of the grid code."""
        f = OracleFrame.__new__(OracleFrame)
        f.L = self.L
        f.desc = _a(fr["desc"], np.uint8).reshape(-1, 32)
        f.x, f.y, f.octave, f.angle = _a(fr["x"], np.float32), _a(fr["y"], np.float32), _a(fr["octave"], np.int32), _a(fr["angle"], np.float32)
        f.n, f.grid = len(f.desc), grid
        f.h = self.L.mo_frame_create(f.n, f.desc.ctypes.data, f.x.ctypes.data, f.y.ctypes.data, f.octave.ctypes.data,
                                     f.angle.ctypes.data, grid.min_x, grid.min_y, grid.inv_w, grid.inv_h, grid.cols, grid.rows)
        return f

    def descriptor_distance(self, a, b):
        a = _a(a, np.uint8).reshape(-1, 32); b = _a(b, np.uint8).reshape(-1, 32)
        return np.array([self.L.rm_descriptor_distance(a[i].ctypes.data, b[i].ctypes.data) for i in range(len(a))], np.int32)

    def projection_mappoints(self, F, q_valid, q_u, q_v, q_level, q_viewcos, q_desc, q_obs, th, sf, holder):
        v, u_, v_, l, c, d, o, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_level, np.int32),
                                    _a(q_viewcos, np.float32), _a(q_desc, np.uint8), _a(q_obs, np.uint8), _a(sf, np.float32))
        return self.L.rm_search_projection_mappoints(F.h, len(u_), v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                     c.ctypes.data, d.ctypes.data, o.ctypes.data, th, self.nnratio, s.ctypes.data,
                                                     holder.ctypes.data)

    def projection_lastframe(self, F, q_valid, q_u, q_v, q_octave, q_angle, q_desc, q_obs, th, sf, holder):
        v, u_, v_, l, a, d, o, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_octave, np.int32),
                                    _a(q_angle, np.float32), _a(q_desc, np.uint8), _a(q_obs, np.uint8), _a(sf, np.float32))
        return self.L.rm_search_projection_lastframe(F.h, len(u_), v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                     a.ctypes.data, d.ctypes.data, o.ctypes.data, th, s.ctypes.data, self.ori,
                                                     holder.ctypes.data)

    def projection_reloc(self, F, q_state, q_u, q_v, q_level, q_angle, q_desc, th, orb_dist, sf, holder):
        """SearchByProjection(Frame&, kfptr, sAlreadyFound, th, ORBdist), ORBmatcher.cpp:1478-1605"""
        st, u_, v_, l, a, d, s = (_a(q_state, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_level, np.int32),
                                  _a(q_angle, np.float32), _a(q_desc, np.uint8), _a(sf, np.float32))
        return self.L.rm_search_projection_reloc(F.h, len(u_), st.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data, a.ctypes.data,
                                                 d.ctypes.data, th, int(orb_dist), s.ctypes.data, self.ori, holder.ctypes.data)

    def projection_kf_sim3(self, F, q_valid, q_u, q_v, q_level, q_desc, q_claims, th, sf, holder):
        v, u_, v_, l, d, c, s = (_a(q_valid, np.uint8), _a(q_u, np.float32), _a(q_v, np.float32), _a(q_level, np.int32),
                                 _a(q_desc, np.uint8), _a(q_claims, np.uint8), _a(sf, np.float32))
        return self.L.rm_search_projection_kf_sim3(F.h, len(u_), v.ctypes.data, u_.ctypes.data, v_.ctypes.data, l.ctypes.data,
                                                   d.ctypes.data, c.ctypes.data, int(th), s.ctypes.data, holder.ctypes.data)

    def bow_kf_frame(self, desc_kf, node_kf, valid_kf, angle_kf, desc_f, node_f, angle_f):
        dk, nk, vk, ak = _a(desc_kf, np.uint8), _a(node_kf, np.int32), _a(valid_kf, np.uint8), _a(angle_kf, np.float32)
        df, nf, af = _a(desc_f, np.uint8), _a(node_f, np.int32), _a(angle_f, np.float32)
        mf = np.full(max(len(nf), 1), -1, np.int32)
        nm = self.L.rm_search_bow_kf_frame(len(nk), dk.ctypes.data, nk.ctypes.data, vk.ctypes.data, ak.ctypes.data, len(nf),
                                           df.ctypes.data, nf.ctypes.data, af.ctypes.data, self.nnratio, self.ori, mf.ctypes.data)
        return nm, mf[:len(nf)]

    def bow_kf_kf(self, d1, n1, v1, a1, d2, n2, v2, a2):
        d1, n1, v1, a1 = _a(d1, np.uint8), _a(n1, np.int32), _a(v1, np.uint8), _a(a1, np.float32)
        d2, n2, v2, a2 = _a(d2, np.uint8), _a(n2, np.int32), _a(v2, np.uint8), _a(a2, np.float32)
        m12 = np.full(max(len(n1), 1), -1, np.int32)
        nm = self.L.rm_search_bow_kf_kf(len(n1), d1.ctypes.data, n1.ctypes.data, v1.ctypes.data, a1.ctypes.data, len(n2),
                                        d2.ctypes.data, n2.ctypes.data, v2.ctypes.data, a2.ctypes.data, self.nnratio, self.ori,
                                        m12.ctypes.data)
        return nm, m12[:len(n1)]

    def search_for_initialization(self, desc1, octave1, angle1, prev_xy, F2, windowSize):
        d1, o1, a1 = _a(desc1, np.uint8), _a(octave1, np.int32), _a(angle1, np.float32)
        m12 = np.full(max(len(o1), 1), -1, np.int32)
        nm = self.L.rm_search_for_initialization(len(o1), d1.ctypes.data, o1.ctypes.data, a1.ctypes.data, prev_xy.ctypes.data,
                                                 F2.h, int(windowSize), self.nnratio, self.ori, m12.ctypes.data)
        return nm, m12[:len(o1)]

    def fuse(self, F, c, th, sf, inv_sigma2):
        """c: a tests/_fuse_cases.fuse_case dict. Returns the state ORBmatcher::Fuse(pKF, vpMapPoints, th) leaves."""
        n, nb = len(c["q_u"]), F.n
        a = [_a(c[k], t) for k, t in (("q_valid", np.uint8), ("q_u", np.float32), ("q_v", np.float32), ("q_level", np.int32),
                                      ("q_desc", np.uint8), ("q_obs", np.int32), ("q_flags", np.uint8), ("q_alias", np.int32))]
        s, g = _a(sf, np.float32), _a(inv_sigma2, np.float32)
        oo, of = _a(c["occ_obs"], np.int32), _a(c["occ_flags"], np.uint8)
        out = {k: np.full(m, -7, np.int32) for k, m in (("replaced_by", n), ("added_idx", n), ("q_bad", n), ("slot_id", nb),
                                                         ("slot_bad", nb), ("occ_replaced_by", nb))}
        out["n"] = self.L.rm_fuse(F.h, n, *[x.ctypes.data for x in a], th, s.ctypes.data, g.ctypes.data, oo.ctypes.data,
                                  of.ctypes.data, *[out[k].ctypes.data for k in ("replaced_by", "added_idx", "q_bad", "slot_id",
                                                                                "slot_bad", "occ_replaced_by")])
        return out

    def fuse_sim3(self, F, c, th, sf):
        n, nb = len(c["q_u"]), F.n
        a = [_a(c[k], t) for k, t in (("q_valid", np.uint8), ("q_u", np.float32), ("q_v", np.float32), ("q_level", np.int32),
                                      ("q_desc", np.uint8), ("q_obs", np.int32), ("q_flags", np.uint8), ("q_alias", np.int32))]
        s = _a(sf, np.float32)
        oo, of = _a(c["occ_obs"], np.int32), _a(c["occ_flags"], np.uint8)
        out = {k: np.full(m, -7, np.int32) for k, m in (("replace_point", n), ("added_idx", n), ("slot_id", nb), ("slot_bad", nb))}
        out["n"] = self.L.rm_fuse_sim3(F.h, n, *[x.ctypes.data for x in a], th, s.ctypes.data, oo.ctypes.data, of.ctypes.data,
                                       *[out[k].ctypes.data for k in ("replace_point", "added_idx", "slot_id", "slot_bad")])
        return out

    def search_by_sim3(self, F1, F2, c, th, sf):
        a = [_a(c[k], t) for k, t in (("st1", np.uint8), ("u1", np.float32), ("v1", np.float32), ("l1", np.int32), ("d1", np.uint8),
                                      ("st2", np.uint8), ("u2", np.float32), ("v2", np.float32), ("l2", np.int32), ("d2", np.uint8),
                                      ("prior", np.int32))]
        s = _a(sf, np.float32)
        m12 = np.full(max(F1.n, 1), -7, np.int32)
        n = self.L.rm_search_by_sim3(F1.h, F2.h, *[x.ctypes.data for x in a], th, s.ctypes.data, m12.ctypes.data)
        return n, m12[:F1.n]

    def search_for_triangulation(self, c, sf, level_sigma2, ori):
        return _triangulation(self.L, self.L.rm_search_for_triangulation, c, sf, level_sigma2, ori)


# ---- vocabulary transform (Frame::ComputeBoW) ------------------------------------------------------------------------
def ref_vocab_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_vocab.so"))


class RefVocabulary:
    """The vendored DBoW2 (oracle/_ref/libref_vocab.so) loaded from a vocabulary text file."""

    def __init__(self, path):
        L = ctypes.CDLL(os.path.join(ORACLE_DIR, "_ref", "libref_vocab.so"))
        L.rv_load.restype = _vp; L.rv_load.argtypes = [_c.c_char_p]
        L.rv_free.argtypes = [_vp]; L.rv_words.argtypes = [_vp]
        L.rv_transform.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]
        L.rv_transform_features.argtypes = [_vp, _i, _vp, _i, _vp, _vp, _vp]
        self.L = L
        self.h = L.rv_load(path.encode())
        if not self.h:
            raise RuntimeError("reference vocabulary failed to load " + path)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.rv_free(self.h); self.h = None

    def words(self):
        return self.L.rv_words(self.h)

    def transform(self, desc, levelsup=4):
        """-> (bow_words, bow_values, fv_node per feature, fv_order)"""
        d = _a(desc, np.uint8).reshape(-1, 32); n = len(d)
        nb, nf = _c.c_int(), _c.c_int()
        bw, bv = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        node, order = np.full(max(n, 1), -7, np.int32), np.zeros(max(n, 1), np.int32)
        self.L.rv_transform(self.h, n, d.ctypes.data, levelsup, _c.byref(nb), bw.ctypes.data, bv.ctypes.data, node.ctypes.data,
                            _c.byref(nf), order.ctypes.data)
        return bw[:nb.value], bv[:nb.value], node[:n], order[:nf.value]

    def transform_features(self, desc, levelsup=4):
        d = _a(desc, np.uint8).reshape(-1, 32); n = len(d)
        w, wt, nd = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64), np.zeros(max(n, 1), np.uint32)
        self.L.rv_transform_features(self.h, n, d.ctypes.data, levelsup, w.ctypes.data, wt.ctypes.data, nd.ctypes.data)
        return w[:n].astype(np.int32), wt[:n], nd[:n].astype(np.int32)


def oracle_vocab_transform(voc, desc, levelsup=4):
    """oracle/match_oracle.cpp mo_vocab_transform on a synth.make_vocabulary dict -> (word, weight, node) per feature."""
    L = match_lib()
    d = _a(desc, np.uint8).reshape(-1, 32); n = len(d)
    cs, ci, nd, nw, wd = (_a(voc["child_start"], np.int32), _a(voc["child_ids"], np.int32), _a(voc["desc"], np.uint8),
                          _a(voc["weight"], np.float64), _a(voc["word"], np.int32))
    word, weight, node = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64), np.zeros(max(n, 1), np.int32)
    L.mo_vocab_transform.argtypes = [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _vp, _vp]
    L.mo_vocab_transform.restype = None
    L.mo_vocab_transform(cs.ctypes.data, ci.ctypes.data, nd.ctypes.data, nw.ctypes.data, wd.ctypes.data, int(voc["L"]), n, d.ctypes.data,
                         int(levelsup), word.ctypes.data, weight.ctypes.data, node.ctypes.data)
    return word[:n], weight[:n], node[:n]


def assemble_bow(word, weight, node):
    """The container half of TemplatedVocabulary::transform(features, BowVector, FeatureVector, levelsup) for L1 / TF-IDF
    (TemplatedVocabulary.h:1143-1168, BowVector.cpp:28-41, 60-84): features with weight > 0 add their weight to their word
    in feature order (repeated double additions), then the vector is divided by its L1 norm accumulated in word order;
    the FeatureVector lists those features per node."""
    acc = {}
    for i in range(len(word)):
        if weight[i] > 0:
            w = int(word[i])
            acc[w] = acc[w] + float(weight[i]) if w in acc else float(weight[i])
    words = sorted(acc)
    norm = 0.0
    for w in words:
        norm += abs(acc[w])
    vals = [acc[w] / norm if norm > 0.0 else acc[w] for w in words]
    fv_node = np.where(np.asarray(weight) > 0, node, -1).astype(np.int32)
    order = [i for nd in sorted(set(fv_node[fv_node >= 0].tolist())) for i in np.nonzero(fv_node == nd)[0].tolist()]
    return np.array(words, np.uint32), np.array(vals, np.float64), fv_node, np.array(order, np.int32)


def oracle_distinctive_descriptors(start, desc):
    L = match_lib()
    st, d = _a(start, np.int32), _a(desc, np.uint8).reshape(-1, 32)
    n = len(st) - 1
    best, med = np.full(max(n, 1), -1, np.int32), np.zeros(max(n, 1), np.int32)
    L.mo_distinctive_descriptors.argtypes = [_i, _vp, _vp, _vp, _vp]
    L.mo_distinctive_descriptors.restype = None
    L.mo_distinctive_descriptors(n, st.ctypes.data, d.ctypes.data, best.ctypes.data, med.ctypes.data)
    return best[:n], med[:n]


def distinctive_case(seed, n_points=400, max_obs=40):
    """Map points with 0..max_obs observations: each point's rows are one random descriptor seen again with 0-12 % of
    its bits flipped (so medians tie often), plus a few outliers."""
    from motioncheck_ccm_slam_b200.synth import perturb_descriptors
    rng = np.random.default_rng(seed)
    counts = rng.integers(0, max_obs + 1, n_points)
    counts[:8] = [0, 1, 2, 3, 33, 64, 65, max_obs]
    start = np.zeros(n_points + 1, np.int32); start[1:] = np.cumsum(counts)
    desc = np.zeros((int(start[-1]), 32), np.uint8)
    for p in range(n_points):
        if counts[p] == 0:
            continue
        base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
        rows = perturb_descriptors(rng, np.repeat(base, counts[p], 0), rng.choice([0.0, 0.02, 0.06, 0.12]))
        out = rng.random(counts[p]) < 0.1
        rows[out] = rng.integers(0, 256, (int(out.sum()), 32), dtype=np.uint8)
        desc[start[p]:start[p + 1]] = rows
    return start, desc


# line ranges of MapPoint.cpp / Converter.cc / KeyFrame.cpp compiled as they are (oracle/_ref/libref_pin.so, oracle/Makefile)
def ref_pin_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref_pin.so"))


_pin = None


def ref_pin_lib():
    global _pin
    if _pin is None:
        L = ctypes.CDLL(os.path.join(ORACLE_DIR, "_ref", "libref_pin.so"))
        L.pin_distinctive_descriptors.restype = None
        L.pin_distinctive_descriptors.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 5
        L.pin_wire_encode.restype = None
        L.pin_wire_encode.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 9
        L.pin_wire_decode.restype = ctypes.c_int
        L.pin_wire_decode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 7
        _pin = L
    return _pin


def ref_distinctive_descriptors(start, desc, kf_bad=None):
    """The reference's MapPoint::ComputeDistinctiveDescriptors per map point -> (chosen descriptor [n][32], assigned [n])."""
    L = ref_pin_lib()
    st, d = _a(start, np.int32), _a(desc, np.uint8).reshape(-1, 32)
    n = len(st) - 1
    out, has = np.zeros((max(n, 1), 32), np.uint8), np.zeros(max(n, 1), np.int32)
    bad = None if kf_bad is None else _a(kf_bad, np.uint8)
    L.pin_distinctive_descriptors(n, st.ctypes.data, d.ctypes.data, None if bad is None else bad.ctypes.data, out.ctypes.data, has.ctypes.data)
    return out[:n], has[:n]


def ref_wire_encode(kp, desc):
    L = ref_pin_lib()
    n = len(kp)
    a = {k: np.ascontiguousarray(kp[k], np.int32 if k == "octave" else np.float32) for k in ("x", "y", "size", "angle", "response", "octave")}
    d = _a(desc, np.uint8).reshape(-1, 32)
    kb, db = np.zeros(4 + 15 * n + 4, np.uint8), np.zeros(4 + 32 * n + 4, np.uint8)
    L.pin_wire_encode(n, *[a[k].ctypes.data for k in ("x", "y", "size", "angle", "response", "octave")], d.ctypes.data, kb.ctypes.data, db.ctypes.data)
    return kb[:4 + 15 * n].tobytes(), db[:4 + 32 * n].tobytes()


def ref_wire_decode(kp_bytes, desc_bytes, cap=1 << 16):
    L = ref_pin_lib()
    kb, db = np.frombuffer(kp_bytes, np.uint8).copy(), np.frombuffer(desc_bytes, np.uint8).copy()
    n0 = int(np.frombuffer(kp_bytes[:4], np.uint32)[0])
    f = {k: np.zeros(max(n0, 1), np.int32 if k == "octave" else np.float32) for k in ("x", "y", "size", "angle", "response", "octave")}
    d = np.zeros((max(n0, 1), 32), np.uint8)
    n = L.pin_wire_decode(kb.ctypes.data, db.ctypes.data, cap, *[f[k].ctypes.data for k in ("x", "y", "size", "angle", "response", "octave")], d.ctypes.data)
    if n < 0:
        return None
    return {k: v[:n] for k, v in f.items()}, d[:n]
