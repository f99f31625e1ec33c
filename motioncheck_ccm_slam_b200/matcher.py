"""Python mirror of cslam::ORBmatcher's Hamming searches (cslam/include/cslam/ORBmatcher.h:77-132) over the C ABI
(include/ccm_match.h). The reference methods take Frame / KeyFrame / MapPoint objects; here, as in the C ABI, a
"map point" is a query index and everything the reference computes on the host before its candidate loop
(projection u, v, predicted level, viewing cosine, validity) is passed in as arrays. Used by tests and bench.py.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import Camera, CcmError, GridParams, WindowParams, WindowQuery

TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30          # ORBmatcher.cpp:63-65
FRAME_GRID_COLS, FRAME_GRID_ROWS = 75, 48            # Frame.h:51-52

WIN_PROJ_MAPPOINTS, WIN_PROJ_LASTFRAME, WIN_PROJ_KF_SIM3, WIN_BEST, WIN_INIT = 0, 1, 2, 3, 4
BOW_KF_FRAME, BOW_KF_KF = 0, 1
Q_VALID, Q_OBS = 1, 2

QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("angle", "<f4"), ("min_level", "<i2"),
                        ("max_level", "<i2"), ("flags", "<u4")])
RESULT_DTYPE = np.dtype([("best_idx", "<i4"), ("best_dist", "<i4"), ("best_dist2", "<i4"), ("accepted", "<i4")])
BEST2_DTYPE = np.dtype([("idx", "<u4"), ("d1", "<u2"), ("d2", "<u2")])
assert QUERY_DTYPE.itemsize == ctypes.sizeof(WindowQuery) == 24


def frame_grid(width, height, min_x=0.0, min_y=0.0, cols=FRAME_GRID_COLS, rows=FRAME_GRID_ROWS):
    """mfGridElementWidthInv / HeightInv as Frame.cpp:87-88 computes them (float division of the image bounds)."""
    max_x, max_y = np.float32(min_x + width), np.float32(min_y + height)
    inv_w = np.float32(cols) / (max_x - np.float32(min_x))
    inv_h = np.float32(rows) / (max_y - np.float32(min_y))
    return GridParams(float(min_x), float(min_y), float(inv_w), float(inv_h), cols, rows)


def make_camera(fx, fy, cx, cy, dist):
    """mK + mDistCoef (k1 k2 p1 p2 [k3]) as floats, the way Tracking.cpp:38-58 reads them from the settings file."""
    dist = [float(np.float32(d)) for d in dist]
    assert len(dist) in (4, 5)
    return Camera(float(np.float32(fx)), float(np.float32(fy)), float(np.float32(cx)), float(np.float32(cy)),
                  (ctypes.c_float * 5)(*(dist + [0.0] * (5 - len(dist)))), len(dist))


class Matcher:
    """One device, one CUDA stream, a grow-only scratch arena (ccm_matcher)."""

    def __init__(self, device=0):
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        rc = self._lib.ccm_matcher_create(int(device), ctypes.byref(self._h))
        if rc != 0:
            raise CcmError(rc, self._lib.ccm_matcher_last_error(None).decode())

    def _check(self, rc):
        if rc != 0:
            raise CcmError(rc, self._lib.ccm_matcher_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.ccm_matcher_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        self._check(self._lib.ccm_matcher_sync(self._h))

    def stream(self):
        return self._lib.ccm_matcher_stream(self._h)

    def launch_count(self):
        return int(self._lib.ccm_matcher_launch_count(self._h))

    # static ORBmatcher::DescriptorDistance, ORBmatcher.cpp:1653-1669
    def DescriptorDistance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        assert a.shape == b.shape
        out = np.zeros(len(a), np.int32)
        self._check(self._lib.ccm_descriptor_distance(self._h, a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data))
        return out

    # ---- server keyframe-database search (BASELINE.json configs[4]) ----
    def UndistortKeyPoints(self, cam, xy):
        """Frame::UndistortKeyPoints (Frame.cpp:277-307) on an [n][2] float array of keypoint positions."""
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        out = np.empty_like(xy)
        self._check(self._lib.ccm_undistort_points(self._h, ctypes.byref(cam), len(xy), xy.ctypes.data, out.ctypes.data))
        return out

    def undistort_keypoints_batch_device(self, cam, n_frames, d_kp, d_n, kp_stride, d_xy_un):
        self._check(self._lib.ccm_undistort_keypoints_batch_device(self._h, ctypes.byref(cam), int(n_frames), d_kp, d_n,
                                                                   int(kp_stride), d_xy_un))

    def ComputeImageBounds(self, cam, width, height, cols=FRAME_GRID_COLS, rows=FRAME_GRID_ROWS):
        """Frame::ComputeImageBounds + the grid constants (Frame.cpp:85-88, 309-343) -> (GridParams, [minX, maxX, minY, maxY])."""
        g = GridParams(0.0, 0.0, 0.0, 0.0, cols, rows)
        b = np.zeros(4, np.float32)
        self._check(self._lib.ccm_image_bounds(self._h, ctypes.byref(cam), int(width), int(height), ctypes.byref(g), b.ctypes.data))
        return g, b

    def wire_encode_keyframe(self, kp, desc, xy_un=None):
        """The CvKeyPoint[] / Descriptor[] arrays of KeyFrame::ConvertToMessage (KeyFrame.cpp:1462-1472) as ROS 1 serialises
        them -> (kp_bytes, desc_bytes). kp: extractor.KP_DTYPE records; xy_un [n][2] overrides the positions (mvKeysUn)."""
        kp = np.ascontiguousarray(kp)
        n = len(kp)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(n, 32)
        xy = None if xy_un is None else np.ascontiguousarray(xy_un, np.float32).reshape(n, 2)
        kb = np.zeros((4 + 15 * n + 3) // 4 * 4, np.uint8)
        db = np.zeros(4 + 32 * n, np.uint8)
        self._check(self._lib.ccm_wire_encode_keyframe(self._h, kp.ctypes.data, None if xy is None else xy.ctypes.data, desc.ctypes.data, n,
                                                       kb.ctypes.data, db.ctypes.data))
        return kb[:4 + 15 * n].tobytes(), db.tobytes()

    def wire_encode_keyframe_device(self, d_kp, d_xy_un, d_desc, n, d_kp_bytes, d_desc_bytes):
        self._check(self._lib.ccm_wire_encode_keyframe_device(self._h, d_kp, d_xy_un, d_desc, int(n), d_kp_bytes, d_desc_bytes))

    def match_allpairs_best2(self, q, targets, index_base=0):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(targets, np.uint8).reshape(-1, 32)
        out = np.zeros(len(q), BEST2_DTYPE)
        self._check(self._lib.ccm_match_allpairs_best2(self._h, q.ctypes.data, len(q), t.ctypes.data, len(t), index_base,
                                                       out.ctypes.data))
        return out

    def match_allpairs_best2_device(self, d_q, nq, d_targets, nt, index_base, d_out):
        self._check(self._lib.ccm_match_allpairs_best2_device(self._h, d_q, nq, d_targets, nt, index_base, d_out))

    def match_allpairs_merge_device(self, d_parts, n_parts, nq, nnratio, d_out, d_accept=None):
        self._check(self._lib.ccm_match_allpairs_merge_device(self._h, d_parts, n_parts, nq, nnratio, d_out, d_accept))


class Exchange:
    """Peer-memory exchange context of the sharded keyframe-database search (ccm_exchange_*): one per rank."""

    def __init__(self, matcher, rank, world, nq_cap):
        self.m, self.rank, self.world = matcher, rank, world
        self._h = ctypes.c_void_p()
        self.handle = np.zeros(64, np.uint8)             # cudaIpcMemHandle_t of this rank's buffer
        matcher._check(matcher._lib.ccm_exchange_create(matcher._h, rank, world, int(nq_cap), ctypes.byref(self._h), self.handle.ctypes.data))

    def buffer(self):
        return self.m._lib.ccm_exchange_buffer(self._h)

    def connect_ipc(self, handles):
        """handles: [world][64] uint8, all ranks' `handle` arrays in rank order (other processes of the same box)."""
        hs = np.ascontiguousarray(handles, np.uint8).reshape(self.world, 64)
        self.m._check(self.m._lib.ccm_exchange_connect_ipc(self._h, hs.ctypes.data))

    def connect_ptrs(self, buffers):
        """buffers: every rank's Exchange.buffer() in rank order (contexts of the same process on different devices)."""
        arr = (ctypes.c_void_p * self.world)(*buffers)
        self.m._check(self.m._lib.ccm_exchange_connect_ptrs(self._h, arr))

    def allpairs_best2_device(self, d_q, nq, d_targets, nt, index_base, nnratio, d_out, d_accept=None):
        self.m._check(self.m._lib.ccm_match_allpairs_best2_exchange_device(self.m._h, self._h, d_q, int(nq), d_targets, int(nt), int(index_base),
                                                                           float(nnratio), d_out, d_accept))

    def check(self):
        """Synchronises the stream; raises CcmError(CCM_E_CUDA) if a wait for a peer timed out (CCM_EXCHANGE_TIMEOUT_MS)."""
        self.m._check(self.m._lib.ccm_exchange_check(self._h))

    def close(self):
        if self._h:
            self.m._lib.ccm_exchange_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class FramePack:
    """Device copy of what the matcher reads of a Frame / KeyFrame: mDescriptors, mvKeysUn, mGrid (Frame.h:128-149)."""

    def __init__(self, matcher, capacity):
        self.m = matcher
        self.capacity = int(capacity)
        self._h = ctypes.c_void_p()
        matcher._check(matcher._lib.ccm_frame_pack_create(matcher._h, self.capacity, ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.m._lib.ccm_frame_pack_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, grid, desc, x, y, octave, angle):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        x = np.ascontiguousarray(x, np.float32); y = np.ascontiguousarray(y, np.float32)
        octave = np.ascontiguousarray(octave, np.int32); angle = np.ascontiguousarray(angle, np.float32)
        n = len(desc)
        assert len(x) == len(y) == len(octave) == len(angle) == n
        self.grid = grid
        self.m._check(self.m._lib.ccm_frame_pack_upload(self._h, ctypes.byref(grid), n, desc.ctypes.data, x.ctypes.data,
                                                        y.ctypes.data, octave.ctypes.data, angle.ctypes.data))
        return self

    def from_wire(self, grid, kp_bytes, desc_bytes, want_keypoints=True):
        """KeyFrame::WriteMembersFromMessage (KeyFrame.cpp:1689-1702): the CvKeyPoint[] / Descriptor[] byte strings of a
        received ccmslam_msgs/KF -> this pack; returns the decoded mvKeysUn (extractor.KP_DTYPE) or None."""
        from .extractor import KP_DTYPE
        kb = np.frombuffer(bytes(kp_bytes), np.uint8)
        db = np.frombuffer(bytes(desc_bytes), np.uint8)
        n = int(kb[:4].view("<u4")[0]) if len(kb) >= 4 else 0
        out = np.zeros(max(n, 1), KP_DTYPE) if want_keypoints else None
        self.grid = grid
        self.m._check(self.m._lib.ccm_frame_pack_from_wire(self._h, ctypes.byref(grid), kb.ctypes.data, len(kb), db.ctypes.data, len(db),
                                                           out.ctypes.data if want_keypoints else None))
        return out[:n] if want_keypoints else None

    def from_extractor_device(self, grid, d_kp, d_desc, d_n, d_xy_un=None):
        self.grid = grid
        self.m._check(self.m._lib.ccm_frame_pack_from_extractor_device(self._h, ctypes.byref(grid), d_kp, d_desc, d_n, d_xy_un))
        return self

    def count(self):
        n = ctypes.c_int()
        self.m._check(self.m._lib.ccm_frame_pack_count(self._h, ctypes.byref(n)))
        return n.value

    def get_grid(self):
        ncell = self.grid.cols * self.grid.rows
        cs = np.zeros(ncell + 1, np.int32)
        items = np.zeros(self.capacity, np.int32)
        n = ctypes.c_int()
        self.m._check(self.m._lib.ccm_frame_pack_get_grid(self._h, cs.ctypes.data, items.ctypes.data, ctypes.byref(n)))
        return cs, items[:n.value].copy()


def _queries(n):
    return np.zeros(n, QUERY_DTYPE)


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=true) — ORBmatcher.h:86; methods named after the reference's."""

    def __init__(self, matcher, nnratio=0.6, checkOri=True):
        self.m = matcher
        self.mfNNratio = float(np.float32(nnratio))
        self.mbCheckOrientation = bool(checkOri)

    def _search_window(self, pack, mode, th_dist, q, q_desc, holder, want_res=True, inv_sigma2=None):
        q_desc = np.ascontiguousarray(q_desc, np.uint8).reshape(-1, 32)
        nq = len(q)
        assert len(q_desc) == nq
        prm = WindowParams(mode, self.mfNNratio, th_dist, 1 if self.mbCheckOrientation else 0, None, 0)
        if inv_sigma2 is not None:
            arr = np.ascontiguousarray(inv_sigma2, np.float32)
            prm.inv_sigma2 = arr.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
            prm.n_levels = len(arr)
        res = np.zeros(nq, RESULT_DTYPE) if want_res else None
        nm = ctypes.c_int()
        self.m._check(self.m._lib.ccm_search_window(self.m._h, pack._h, ctypes.byref(prm), nq, q.ctypes.data, q_desc.ctypes.data,
                                                    holder.ctypes.data if holder is not None else None,
                                                    res.ctypes.data if want_res else None, ctypes.byref(nm)))
        return nm.value, res

    # SearchByProjection(Frame &F, const vector<mpptr> &vpMapPoints, const float th), ORBmatcher.cpp:71-148
    def SearchByProjection_MapPoints(self, pack, q_valid, q_u, q_v, q_level, q_viewcos, q_desc, q_obs, th, scale_factors, holder):
        n = len(q_u)
        q = _queries(n)
        lvl = np.asarray(q_level, np.int64)
        r = np.where(np.asarray(q_viewcos, np.float32).astype(np.float64) > 0.998, np.float32(2.5), np.float32(4.0)).astype(np.float32)
        if float(th) != 1.0:                       # bFactor, :75
            r = r * np.float32(th)
        sf = np.asarray(scale_factors, np.float32)
        q["u"], q["v"] = q_u, q_v
        q["r"] = r * sf[np.clip(lvl, 0, len(sf) - 1)]
        q["min_level"], q["max_level"] = lvl - 1, lvl
        q["flags"] = np.asarray(q_valid, np.uint32) * Q_VALID + np.asarray(q_obs, np.uint32) * Q_OBS
        return self._search_window(pack, WIN_PROJ_MAPPOINTS, TH_HIGH, q, q_desc, holder)

    # SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th), ORBmatcher.cpp:1350-1476
    def SearchByProjection_LastFrame(self, pack, q_valid, q_u, q_v, q_octave, q_angle, q_desc, q_obs, th, scale_factors, holder):
        n = len(q_u)
        q = _queries(n)
        octv = np.asarray(q_octave, np.int64)
        sf = np.asarray(scale_factors, np.float32)
        q["u"], q["v"], q["angle"] = q_u, q_v, q_angle
        q["r"] = np.float32(th) * sf[np.clip(octv, 0, len(sf) - 1)]
        q["min_level"], q["max_level"] = octv - 1, octv + 1
        q["flags"] = np.asarray(q_valid, np.uint32) * Q_VALID + np.asarray(q_obs, np.uint32) * Q_OBS
        return self._search_window(pack, WIN_PROJ_LASTFRAME, TH_HIGH, q, q_desc, holder)

    # SearchByProjection(kfptr pKF, cv::Mat Scw, vpPoints, vpMatched, int th), ORBmatcher.cpp:308-446 (search part)
    def SearchByProjection_KFSim3(self, pack, q_valid, q_u, q_v, q_level, q_desc, q_claims, th, scale_factors, holder):
        n = len(q_u)
        q = _queries(n)
        lvl = np.asarray(q_level, np.int64)
        sf = np.asarray(scale_factors, np.float32)
        q["u"], q["v"] = q_u, q_v
        q["r"] = np.float32(int(th)) * sf[np.clip(lvl, 0, len(sf) - 1)]
        q["min_level"], q["max_level"] = lvl - 1, lvl
        q["flags"] = np.asarray(q_valid, np.uint32) * Q_VALID + np.asarray(q_claims, np.uint32) * Q_OBS
        return self._search_window(pack, WIN_PROJ_KF_SIM3, TH_LOW, q, q_desc, holder)

    # the candidate search of SearchBySim3's passes (:1216-1252, :1296-1327) and Fuse's distance phase (:929-962, :1082-1103)
    def SearchBestInWindow(self, pack, q_valid, q_u, q_v, q_radius, q_level, q_desc, th_dist, inv_sigma2=None):
        n = len(q_u)
        q = _queries(n)
        lvl = np.asarray(q_level, np.int64)
        q["u"], q["v"], q["r"] = q_u, q_v, q_radius
        q["min_level"], q["max_level"] = lvl - 1, lvl
        q["flags"] = np.asarray(q_valid, np.uint32) * Q_VALID
        return self._search_window(pack, WIN_BEST, th_dist, q, q_desc, None, inv_sigma2=inv_sigma2)

    # SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), ORBmatcher.cpp:448-563
    def SearchForInitialization(self, desc1, octave1, angle1, prev_xy, pack2, xy2, windowSize=10):
        """prev_xy [n1, 2] float32 is updated in place (vbPrevMatched); xy2 = F2's mvKeysUn positions. Returns (nmatches, vnMatches12)."""
        n1 = len(octave1)
        q = _queries(n1)
        octv = np.asarray(octave1, np.int64)
        q["u"], q["v"], q["angle"] = prev_xy[:, 0], prev_xy[:, 1], angle1
        q["r"] = np.float32(windowSize)
        q["min_level"], q["max_level"] = octv, octv
        q["flags"] = (octv <= 0).astype(np.uint32) * Q_VALID          # `if(level1>0) continue;`, :463-465
        n2 = pack2.count()
        holder = np.full(max(n2, 1), -1, np.int32)
        nm, _ = self._search_window(pack2, WIN_INIT, TH_LOW, q, desc1, holder, want_res=False)
        matches12 = np.full(n1, -1, np.int32)
        t = np.nonzero(holder[:n2] >= 0)[0]
        matches12[holder[t]] = t
        m = matches12 >= 0
        prev_xy[m] = np.asarray(xy2, np.float32)[matches12[m]]         # :558-560
        return nm, matches12

    def _search_bow(self, mode, desc_a, node_a, valid_a, angle_a, desc_b, node_b, valid_b, angle_b):
        desc_a = np.ascontiguousarray(desc_a, np.uint8).reshape(-1, 32)
        desc_b = np.ascontiguousarray(desc_b, np.uint8).reshape(-1, 32)
        na, nb = len(desc_a), len(desc_b)
        node_a = np.ascontiguousarray(node_a, np.int32); node_b = np.ascontiguousarray(node_b, np.int32)
        valid_a = np.ascontiguousarray(valid_a, np.uint8)
        valid_b = np.ascontiguousarray(valid_b if valid_b is not None else np.ones(nb), np.uint8)
        angle_a = np.ascontiguousarray(angle_a, np.float32); angle_b = np.ascontiguousarray(angle_b, np.float32)
        ma, mb = np.full(na, -1, np.int32), np.full(nb, -1, np.int32)
        res = np.zeros(na, RESULT_DTYPE)
        nm = ctypes.c_int()
        self.m._check(self.m._lib.ccm_search_bow(
            self.m._h, mode, self.mfNNratio, 1 if self.mbCheckOrientation else 0,
            na, desc_a.ctypes.data, node_a.ctypes.data, valid_a.ctypes.data, angle_a.ctypes.data,
            nb, desc_b.ctypes.data, node_b.ctypes.data, valid_b.ctypes.data, angle_b.ctypes.data,
            ma.ctypes.data, mb.ctypes.data, res.ctypes.data, ctypes.byref(nm)))
        return nm.value, ma, mb, res

    # SearchByBoW(kfptr pKF, Frame &F, vpMapPointMatches), ORBmatcher.cpp:178-306 -> (nmatches, match_of_f, res)
    def SearchByBoW_KF_Frame(self, desc_kf, node_kf, valid_kf, angle_kf, desc_f, node_f, angle_f):
        nm, _, mb, res = self._search_bow(BOW_KF_FRAME, desc_kf, node_kf, valid_kf, angle_kf, desc_f, node_f, None, angle_f)
        return nm, mb, res

    # SearchByBoW(kfptr pKF1, kfptr pKF2, vpMatches12), ORBmatcher.cpp:565-698 -> (nmatches, match12, res)
    def SearchByBoW_KF_KF(self, desc1, node1, valid1, angle1, desc2, node2, valid2, angle2):
        nm, ma, _, res = self._search_bow(BOW_KF_KF, desc1, node1, valid1, angle1, desc2, node2, valid2, angle2)
        return nm, ma, res

    # SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs), ORBmatcher.cpp:700-852 -> (nmatches, match12);
    # vMatchedPairs = [(i, match12[i]) for i where match12[i] >= 0]. free* = the feature holds no map point.
    def SearchForTriangulation(self, desc1, node1, free1, x1, y1, angle1, desc2, node2, free2, x2, y2, octave2, angle2,
                               F12, ex, ey, scale_factors, level_sigma2):
        from ._lib import EpipolarGeom
        c = np.ascontiguousarray
        desc1 = c(desc1, np.uint8).reshape(-1, 32); desc2 = c(desc2, np.uint8).reshape(-1, 32)
        n1, n2 = len(desc1), len(desc2)
        a = [c(node1, np.int32), c(free1, np.uint8), c(x1, np.float32), c(y1, np.float32), c(angle1, np.float32)]
        b = [c(node2, np.int32), c(free2, np.uint8), c(x2, np.float32), c(y2, np.float32), c(octave2, np.int32), c(angle2, np.float32)]
        sf, ls = c(scale_factors, np.float32), c(level_sigma2, np.float32)
        g = EpipolarGeom()
        g.f12[:] = [float(v) for v in c(F12, np.float32).reshape(-1)]
        g.ex, g.ey = float(ex), float(ey)
        g.scale_factors = sf.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        g.level_sigma2 = ls.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        g.n_levels = len(sf)
        m12 = np.full(max(n1, 1), -1, np.int32)
        nm = ctypes.c_int()
        self.m._check(self.m._lib.ccm_search_triangulation(
            self.m._h, 1 if self.mbCheckOrientation else 0, ctypes.byref(g),
            n1, desc1.ctypes.data, *[v.ctypes.data for v in a], n2, desc2.ctypes.data, *[v.ctypes.data for v in b],
            m12.ctypes.data, ctypes.byref(nm)))
        return nm.value, m12[:n1]


def distinctive_descriptors(matcher, start, desc):
    """MapPoint::ComputeDistinctiveDescriptors (cslam/src/MapPoint.cpp:929-994) for a batch of map points: point p owns
    rows start[p]:start[p+1] of desc. -> (best row index within each point, -1 if it has none; its median distance)."""
    st = np.ascontiguousarray(start, np.int32)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(st) - 1
    best, med = np.full(max(n, 1), -1, np.int32), np.zeros(max(n, 1), np.int32)
    matcher._check(matcher._lib.ccm_distinctive_descriptors(matcher._h, n, st.ctypes.data, d.ctypes.data if len(d) else None,
                                                            best.ctypes.data, med.ctypes.data))
    return best[:n], med[:n]


class Vocabulary:
    """Device copy of a DBoW2 vocabulary tree for the per-feature half of Frame::ComputeBoW (cslam/src/Frame.cpp:268-275).
    `voc` = dict(child_start, child_ids, desc, word, weight, L) in loadFromTextFile's node numbering (synth.make_vocabulary
    builds a seeded one; the ORBvoc.txt blob is absent from the reference checkout)."""

    def __init__(self, matcher, voc):
        c = np.ascontiguousarray
        self.m = matcher
        self.weight = c(voc["weight"], np.float64)
        cs, ci, nd, wd = c(voc["child_start"], np.int32), c(voc["child_ids"], np.int32), c(voc["desc"], np.uint8), c(voc["word"], np.int32)
        self.n_nodes = len(wd)
        self._h = ctypes.c_void_p()
        matcher._check(matcher._lib.ccm_vocabulary_create(matcher._h, self.n_nodes, cs.ctypes.data, ci.ctypes.data, nd.ctypes.data,
                                                          wd.ctypes.data, int(voc["L"]), ctypes.byref(self._h)))
        self.word_weight = np.zeros(int(wd.max()) + 1, np.float64)
        self.word_weight[wd[wd >= 0]] = self.weight[wd >= 0]

    def close(self):
        if self._h:
            self.m._lib.ccm_vocabulary_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform_features(self, desc, levelsup=4):
        """-> (word [n], node [n]) per descriptor (TemplatedVocabulary.h:1218-1261)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        word, node = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
        self.m._check(self.m._lib.ccm_vocabulary_transform(self.m._h, self._h, n, d.ctypes.data, int(levelsup), word.ctypes.data, node.ctypes.data))
        return word[:n], node[:n]

    def transform_device(self, n, d_desc, d_word, d_node, levelsup=4):
        """Raw device pointers; asynchronous on the matcher's stream."""
        self.m._check(self.m._lib.ccm_vocabulary_transform_device(self.m._h, self._h, int(n), d_desc, int(levelsup), d_word, d_node))

    def ComputeBoW(self, desc, levelsup=4):
        """Frame::ComputeBoW: -> (bow_words, bow_values, fv_node per feature (-1: stopped word), fv_order). The descent runs
        on the device; the two containers are filled on the host exactly as TemplatedVocabulary.h:1143-1168 does for the ORB
        vocabulary's L1 / TF-IDF setting: weights added per word in feature order, then divided by the L1 norm summed in
        word order (BowVector.cpp:28-41, 62-84)."""
        word, node = self.transform_features(desc, levelsup)
        wt = self.word_weight[word] if len(word) else np.zeros(0)
        acc = {}
        for i in range(len(word)):
            if wt[i] > 0:
                w = int(word[i])
                acc[w] = acc[w] + float(wt[i]) if w in acc else float(wt[i])
        words = sorted(acc)
        norm = 0.0
        for w in words:
            norm += abs(acc[w])
        vals = [acc[w] / norm if norm > 0.0 else acc[w] for w in words]
        fv_node = np.where(wt > 0, node, -1).astype(np.int32)
        order = [i for nd in sorted(set(fv_node[fv_node >= 0].tolist())) for i in np.nonzero(fv_node == nd)[0].tolist()]
        return np.array(words, np.uint32), np.array(vals, np.float64), fv_node, np.array(order, np.int32)


# ---- device-resident batch path (extractor output -> packs -> searches, no host round trip) ----------------------
def packs_from_extractor_batch_device(packs, grid, d_kp, d_desc, d_n, kp_stride, d_xy_un=None, cam=None):
    """packs[j] takes frame j of ORBextractor.extract_batch_device's output (raw device pointers); asynchronous.
    cam: undistort the keypoints inside the pack fill (mvKeysUn of a distorted camera)."""
    m = packs[0].m
    arr = (ctypes.c_void_p * len(packs))(*[p._h.value for p in packs])
    for p in packs:
        p.grid = grid
    if cam is not None:
        assert d_xy_un is None
        m._check(m._lib.ccm_frame_pack_from_extractor_batch_device_cam(arr, len(packs), ctypes.byref(grid), ctypes.byref(cam), d_kp,
                                                                       d_desc, d_n, int(kp_stride)))
        return
    m._check(m._lib.ccm_frame_pack_from_extractor_batch_device(arr, len(packs), ctypes.byref(grid), d_kp, d_desc, d_n,
                                                               int(kp_stride), d_xy_un))


def search_window_batch_device(matcher, packs, mode, nnratio, th_dist, check_orientation, d_nq, q_stride, d_q, d_q_desc,
                               d_holder, holder_stride, d_res, d_nmatches):
    """ccm_search_window_batch_device: job j searches packs[j] with the q_stride-strided query block j; asynchronous."""
    arr = (ctypes.c_void_p * len(packs))(*[p._h.value for p in packs])
    prm = WindowParams(mode, float(nnratio), int(th_dist), 1 if check_orientation else 0, None, 0)
    matcher._check(matcher._lib.ccm_search_window_batch_device(matcher._h, arr, len(packs), ctypes.byref(prm), d_nq, int(q_stride),
                                                               d_q, d_q_desc, d_holder, int(holder_stride), d_res, d_nmatches))
