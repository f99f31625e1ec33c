"""us per host entry point of one agent's frame loop (configs[3]): extract, pack upload (+ grid), vocabulary transform,
SearchByBoW, two SearchByProjection calls. python tools/time_frame_calls.py"""
import ctypes
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import bench_legs as bl
from motioncheck_ccm_slam_b200.synth import make_stream, make_vocabulary_complete

T = 40
ag = bl._Agent(0, 0, make_stream(1000, T, 752, 480), make_vocabulary_complete(77, 10, 6), 1000, 8)
for t in range(T):
    ag.frame(t, record=True)
L, m, cap = ag.lib, ag.m._h, ag.cap
n = ctypes.c_int(); nm = ctypes.c_int()
from motioncheck_ccm_slam_b200._lib import WindowParams
from motioncheck_ccm_slam_b200.matcher import WIN_PROJ_LASTFRAME, WIN_PROJ_MAPPOINTS


def timeit(name, f, reps=200):
    for _ in range(10):
        f()
    t0 = time.perf_counter()
    for _ in range(reps):
        f()
    print("%-28s %.1f us" % (name, 1e6 * (time.perf_counter() - t0) / reps))


t = 20; p = t - 1; k = int(ag.n[t]); pack = ag.packs[t & 1]
timeit("ccm_orb_extract", lambda: ag.ex.extract_raw(ag.frames[t].ctypes.data, ag.h, ag.w, ag.w, ag.kp[t].ctypes.data, ag.desc[t].ctypes.data, cap, ctypes.byref(n)))
timeit("ccm_frame_pack_upload", lambda: L.ccm_frame_pack_upload(pack._h, ctypes.byref(ag.grid), k, ag.desc[t].ctypes.data, ag.xs[t].ctypes.data, ag.ys[t].ctypes.data, ag.oc[t].ctypes.data, ag.an[t].ctypes.data))
timeit("ccm_vocabulary_transform", lambda: L.ccm_vocabulary_transform(m, ag.V._h, k, ag.desc[t].ctypes.data, 4, ag.word[t].ctypes.data, ag.node[t].ctypes.data))
timeit("ccm_search_bow", lambda: L.ccm_search_bow(m, 0, 0.7, 1, int(ag.n[p]), ag.desc[p].ctypes.data, ag.node[p].ctypes.data, ag.valid.ctypes.data, ag.an[p].ctypes.data, k, ag.desc[t].ctypes.data, ag.node[t].ctypes.data, ag.valid.ctypes.data, ag.an[t].ctypes.data, ag.match_a.ctypes.data, ag.match_b.ctypes.data, ag.res.ctypes.data, ctypes.byref(nm)))
prm = WindowParams(WIN_PROJ_LASTFRAME, 0.9, 100, 1, None, 0)
def lastf():
    ag.hold[0, :k] = -1
    L.ccm_search_window(m, pack._h, ctypes.byref(prm), int(ag.n[p]), ag.q1[t].ctypes.data, ag.desc[p].ctypes.data, ag.hold[0].ctypes.data, None, ctypes.byref(nm))
timeit("ccm_search_window LastFrame", lastf)
prm2 = WindowParams(WIN_PROJ_MAPPOINTS, 0.8, 100, 0, None, 0)
def mapp():
    ag.hold[1, :k] = -1
    L.ccm_search_window(m, pack._h, ctypes.byref(prm2), int(ag.n[t - 2]), ag.q2[t].ctypes.data, ag.desc[t - 2].ctypes.data, ag.hold[1].ctypes.data, None, ctypes.byref(nm))
timeit("ccm_search_window MapPoints", mapp)
timeit("whole frame()", lambda: ag.frame(t), 100)
