"""Extra legs of bench.py (kept apart so that bench.py stays readable):
  link_ceiling      the e2e copy pattern alone (23.1 MB in + 3.7 MB out per step at C1), all ranks at once
  single_frame      the per-image call Frame::ExtractORB makes (ccm_orb_extract), one frame at a time
  searches_leg      every converted ORBmatcher method, drop-in class on the GPU vs the reference class on one host thread
  c4_leg            BASELINE.json configs[3]: four agent streams, frame by frame, with the tracking-side searches
The CPU side of searches_leg is the reference's own ORBmatcher.cpp (oracle/_ref/libref_match.so) and c4_leg checks a
sample of its outputs against the oracle: both import tests/_oracle.py, as a baseline and as a checker only."""
import ctypes
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
SCALE, INI_TH, MIN_TH = 1.2, 20, 7


def _tests_path():
    p = os.path.join(ROOT, "tests")
    if p not in sys.path:
        sys.path.insert(0, p)


# ----------------------------------------------------------------------------------------------------------
def link_ceiling(torch, dev, pin_batches, out_kp, out_desc, steps, barrier):
    """The host<->device traffic of one e2e step with no kernel in between, through the SAME pinned buffers the e2e leg uses
    (host-to-device bandwidth differs between pinned allocations on this VM: 33-45 GB/s measured for 23 MB copies): one batch
    up on one stream, keypoints + descriptors down on another, every rank at the same time. ms per step on this rank."""
    R = pin_batches.shape[0]
    d_in = torch.zeros(pin_batches.shape[1:], dtype=torch.uint8, device=dev)
    d_kp = torch.zeros(out_kp[0].shape, dtype=torch.float32, device=dev)
    d_desc = torch.zeros(out_desc[0].shape, dtype=torch.uint8, device=dev)
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

    def run(k):
        for i in range(k):
            with torch.cuda.stream(s_in):
                d_in.copy_(pin_batches[i % R], non_blocking=True)
            with torch.cuda.stream(s_out):
                out_kp[i & 1].copy_(d_kp, non_blocking=True); out_desc[i & 1].copy_(d_desc, non_blocking=True)
        torch.cuda.synchronize()

    run(3)
    barrier()
    t0 = time.perf_counter()
    run(steps)
    return 1e3 * (time.perf_counter() - t0) / steps


# ----------------------------------------------------------------------------------------------------------
def single_frame(torch, local_rank, w, h, nf, nl, frames, calls=300):
    """us per ccm_orb_extract call (host image in, host keypoints + descriptors out, synchronous), pinned buffers."""
    from motioncheck_ccm_slam_b200 import ORBextractor
    ex = ORBextractor(nf, SCALE, nl, INI_TH, MIN_TH, max_width=w, max_height=h, max_batch=1, device=local_rank)
    cap = ex.max_keypoints()
    pin = torch.from_numpy(np.ascontiguousarray(frames[:16])).pin_memory()
    pn = pin.numpy()
    okp = torch.zeros((cap, 6), dtype=torch.float32).pin_memory()
    ode = torch.zeros((cap, 32), dtype=torch.uint8).pin_memory()
    on = ctypes.c_int()
    ptrs = [pn[f].ctypes.data for f in range(len(pn))]
    for i in range(20):
        ex.extract_raw(ptrs[i % len(ptrs)], h, w, w, okp.data_ptr(), ode.data_ptr(), cap, ctypes.byref(on))
    l0 = ex.launch_count()
    per = np.empty(calls)
    t0 = time.perf_counter()
    for i in range(calls):
        ta = time.perf_counter()
        ex.extract_raw(ptrs[i % len(ptrs)], h, w, w, okp.data_ptr(), ode.data_ptr(), cap, ctypes.byref(on))
        per[i] = time.perf_counter() - ta
    dt = time.perf_counter() - t0
    launches = ex.launch_count() - l0
    ex.close()
    q = np.percentile(per, [10, 50, 90]) * 1e6
    return {"single_frame_latency_us": 1e6 * dt / calls, "single_frame_latency_us_p10_p50_p90": [float(q[0]), float(q[1]), float(q[2])],
            "single_frame_frames_per_s": calls / dt,
            "single_frame_api": "ccm_orb_extract: copy-in, one CUDA graph (a branch per pyramid level: FAST + quadtree of level l behind its resize; blur "
                                "branch; k_orient_describe writes into the pinned host mirrors), wait; %d calls back to back, mean and percentiles "
                                "of the per-call host time, %d kernel launches replayed" % (calls, launches)}, launches


# ----------------------------------------------------------------------------------------------------------
def searches_leg(calls=64, cpu_calls=6):
    """Every converted ORBmatcher method on the shapes the parity tests use: the drop-in C++ class (GPU, through the same
    object glue as the reference; frame packs cached per Frame / KeyFrame id) against the reference's ORBmatcher.cpp on
    one host thread. `calls` consecutive calls cycling over 4 input sets; us per call."""
    _tests_path()
    import _fuse_cases as fc
    import _oracle
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    from motioncheck_ccm_slam_b200.synth import bow_node_ids, make_feature_frame, make_next_frame
    if not (_oracle.dropin_match_available() and _oracle.ref_match_available()):
        return {"unavailable": "oracle/_ref/libdropin_match.so / libref_match.so not built"}, 0
    W, H = 752, 480
    SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    grid = frame_grid(W, H)

    def frames(seed, n=1000, **kw):
        a = make_feature_frame(seed, n, W, H)
        b, src = make_next_frame(seed + 1, a, W, H, **kw)
        return a, b, src

    def proj_inputs(a, b, src, rng):
        n = len(a["x"])
        inv = np.full(n, -1); inv[src[src >= 0]] = np.nonzero(src >= 0)[0]
        u = np.where(inv >= 0, b["x"][np.maximum(inv, 0)], a["x"] + 2).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
        v = np.where(inv >= 0, b["y"][np.maximum(inv, 0)], a["y"] + 1).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
        level = np.where(inv >= 0, b["octave"][np.maximum(inv, 0)], a["octave"]).astype(np.int32)
        level = np.clip(level + rng.choice([0, 0, 0, 1], n), 0, 7).astype(np.int32)
        valid = (rng.random(n) < 0.9).astype(np.uint8) & ((u >= 0) & (u <= W) & (v >= 0) & (v <= H)).astype(np.uint8)
        obs = (rng.random(n) < 0.8).astype(np.uint8)
        return valid, u, v, level, obs

    def pair(nnratio, ori):
        return _oracle.RefMatcher(nnratio, ori, lib="libdropin_match.so"), _oracle.RefMatcher(nnratio, ori)

    # mode -> (nnratio, ori, make(seed) -> inputs, run(M, inputs, frame handles of M) -> comparable tuple, frame names)
    def mk_proj(n=1000):
        def make(seed):
            rng = np.random.default_rng(seed)
            a, b, src = frames(seed, n)
            valid, u, v, level, obs = proj_inputs(a, b, src, rng)
            return dict(a=a, b=b, valid=valid, u=u, v=v, level=level, obs=obs, viewcos=rng.choice([0.9, 0.9985], len(u)).astype(np.float32),
                        holder0=rng.choice([-1, -1, -1, -1, -2, -3], len(b["x"])).astype(np.int32),
                        state=np.where(valid > 0, rng.choice([1, 1, 1, 1, 1, 1, 2, 3], len(u)), 0).astype(np.uint8), F=["b"])
        return make

    def mk_bow(seed):
        rng = np.random.default_rng(seed)
        a, b, src = frames(seed, flip_prob=0.03)
        na, nb = bow_node_ids(a["desc"], 1000), bow_node_ids(b["desc"], 1000)
        return dict(a=a, b=b, na=na, nb=nb, va=(rng.random(len(na)) < 0.8).astype(np.uint8), vb=(rng.random(len(nb)) < 0.8).astype(np.uint8), F=[])

    def mk_init(seed):
        a, b, src = frames(seed, n=2000, motion=(6.0, -3.0))
        return dict(a=a, b=b, F=["b"])

    def run_mp(M, c, f):
        h = c["holder0"].copy()
        return M.projection_mappoints(f["b"], c["valid"], c["u"], c["v"], c["level"], c["viewcos"], c["a"]["desc"], c["obs"], 1.0, SF, h), h

    def run_last(M, c, f):
        h = c["holder0"].copy()
        return M.projection_lastframe(f["b"], c["valid"], c["u"], c["v"], c["a"]["octave"], c["a"]["angle"], c["a"]["desc"], c["obs"], 7.0, SF, h), h

    def run_kfsim3(M, c, f):
        h = np.where(c["holder0"] == -3, -1, c["holder0"]).astype(np.int32)
        return M.projection_kf_sim3(f["b"], c["valid"], c["u"], c["v"], c["level"], c["a"]["desc"], c["obs"], 10, SF, h), h

    def run_reloc(M, c, f):
        h = c["holder0"].copy()
        return M.projection_reloc(f["b"], c["state"], c["u"], c["v"], c["level"], c["a"]["angle"], c["a"]["desc"], 10.0, 100, SF, h), h

    def run_bow_kf_f(M, c, f):
        return M.bow_kf_frame(c["a"]["desc"], c["na"], c["va"], c["a"]["angle"], c["b"]["desc"], c["nb"], c["b"]["angle"])

    def run_bow_kf_kf(M, c, f):
        return M.bow_kf_kf(c["a"]["desc"], c["na"], c["va"], c["a"]["angle"], c["b"]["desc"], c["nb"], c["vb"], c["b"]["angle"])

    def run_init(M, c, f):
        prev = np.stack([c["a"]["x"], c["a"]["y"]], 1).astype(np.float32).copy()
        return M.search_for_initialization(c["a"]["desc"], c["a"]["octave"], c["a"]["angle"], prev, f["b"], 100) + (prev,)

    def mk_fuse(seed):
        c = fc.fuse_case(seed); c["F"] = ["kf"]; return c

    def mk_sim3(seed):
        c = fc.sim3_case(seed); c["F"] = ["kf1", "kf2"]; return c

    def mk_tri(seed):
        c = fc.tri_case(seed, n_nodes=120); c["F"] = []; return c

    def run_fuse(M, c, f):
        r = M.fuse(f["kf"], c, 3.0, fc.SF, fc.INV_SIGMA2)
        return r["n"], r["replaced_by"], r["added_idx"], r["slot_id"]

    def run_fuse_sim3(M, c, f):
        r = M.fuse_sim3(f["kf"], c, 4.0, fc.SF)
        return r["n"], r["replace_point"], r["added_idx"], r["slot_id"]

    def run_sim3(M, c, f):
        return M.search_by_sim3(f["kf1"], f["kf2"], c, 7.5, fc.SF)

    def run_tri(M, c, f):
        return M.search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, False)

    # the same inputs with every query switched off: the reference method then only walks its (empty) loops, so its time is
    # the object set-up of the glue (mock Frame / KeyFrame / MapPoint construction), which both arms pay
    def null_proj(c):
        z = np.zeros_like(c["valid"])
        return dict(c, valid=z, state=z)

    def null_bow(c):
        return dict(c, va=np.zeros_like(c["va"]), vb=np.zeros_like(c["vb"]))

    def null_init(c):
        a = dict(c["a"]); a["octave"] = np.ones_like(a["octave"])          # `if(level1>0) continue`, :464-466
        return dict(c, a=a)

    def null_fuse(c):
        return dict(c, q_valid=np.zeros_like(c["q_valid"]))

    def null_sim3(c):
        return dict(c, st1=np.zeros_like(c["st1"]), st2=np.zeros_like(c["st2"]))

    def null_tri(c):
        return dict(c, free1=np.zeros_like(c["free1"]))

    # the C ABI called directly (ctypes wrappers of motioncheck_ccm_slam_b200/matcher.py): what the copies and kernels of one
    # call cost without any Frame / MapPoint object around it
    from motioncheck_ccm_slam_b200.matcher import FramePack, Matcher, ORBmatcher
    gm = Matcher(0)

    def gpack(fr):
        pk = FramePack(gm, max(len(fr["x"]), 1))
        pk.upload(grid, fr["desc"], fr["x"], fr["y"], fr["octave"], fr["angle"])
        return pk

    def d_mp(om, c, pk):
        return om.SearchByProjection_MapPoints(pk["b"], c["valid"], c["u"], c["v"], c["level"], c["viewcos"], c["a"]["desc"], c["obs"], 1.0, SF, c["holder0"].copy())

    def d_last(om, c, pk):
        return om.SearchByProjection_LastFrame(pk["b"], c["valid"], c["u"], c["v"], c["a"]["octave"], c["a"]["angle"], c["a"]["desc"], c["obs"], 7.0, SF, c["holder0"].copy())

    def d_kfsim3(om, c, pk):
        return om.SearchByProjection_KFSim3(pk["b"], c["valid"], c["u"], c["v"], c["level"], c["a"]["desc"], c["obs"], 10, SF,
                                            np.where(c["holder0"] == -3, -1, c["holder0"]).astype(np.int32))

    def d_bow_f(om, c, pk):
        return om.SearchByBoW_KF_Frame(c["a"]["desc"], c["na"], c["va"], c["a"]["angle"], c["b"]["desc"], c["nb"], c["b"]["angle"])

    def d_bow_kf(om, c, pk):
        return om.SearchByBoW_KF_KF(c["a"]["desc"], c["na"], c["va"], c["a"]["angle"], c["b"]["desc"], c["nb"], c["vb"], c["b"]["angle"])

    def d_init(om, c, pk):
        prev = np.stack([c["a"]["x"], c["a"]["y"]], 1).astype(np.float32).copy()
        return om.SearchForInitialization(c["a"]["desc"], c["a"]["octave"], c["a"]["angle"], prev, pk["b"], np.stack([c["b"]["x"], c["b"]["y"]], 1), 100)

    def d_tri(om, c, pk):
        a, b = c["kf1"], c["kf2"]
        return om.SearchForTriangulation(a["desc"], c["node1"], c["free1"], a["x"], a["y"], a["angle"], b["desc"], c["node2"], c["free2"], b["x"], b["y"],
                                         b["octave"], b["angle"], c["f12"], c["ex"], c["ey"], fc.SF, fc.LEVEL_SIGMA2)

    def d_sim3(om, c, pk):
        r1 = om.SearchBestInWindow(pk["kf2"], (c["st1"] > 0).astype(np.uint8), c["u1"], c["v1"], (np.float32(7.5) * fc.SF[c["l1"]]).astype(np.float32), c["l1"], c["d1"], 100)
        r2 = om.SearchBestInWindow(pk["kf1"], (c["st2"] > 0).astype(np.uint8), c["u2"], c["v2"], (np.float32(7.5) * fc.SF[c["l2"]]).astype(np.float32), c["l2"], c["d2"], 100)
        return r1, r2

    def d_fuse(om, c, pk):
        return om.SearchBestInWindow(pk["kf"], c["q_valid"], c["q_u"], c["q_v"], (np.float32(3.0) * fc.SF[c["q_level"]]).astype(np.float32), c["q_level"], c["q_desc"], 50, fc.INV_SIGMA2)

    def d_fuse_sim3(om, c, pk):
        return om.SearchBestInWindow(pk["kf"], c["q_valid"], c["q_u"], c["q_v"], (np.float32(4.0) * fc.SF[c["q_level"]]).astype(np.float32), c["q_level"], c["q_desc"], 50)

    modes = [
        ("SearchByProjection(F,MapPoints,th=1)", "ORBmatcher.cpp:71-148", 0.8, True, mk_proj(), run_mp, null_proj, d_mp, "1000 map points into a 1000-feature frame"),
        ("SearchByProjection(F,LastFrame,th=7)", ":1350-1476", 0.9, True, mk_proj(), run_last, null_proj, d_last, "1000 points of the last frame"),
        ("SearchByProjection(KF,Scw,th=10)", ":308-446", 0.75, True, mk_proj(3000), run_kfsim3, null_proj, d_kfsim3, "3000 loop points into a 3000-feature key frame"),
        ("SearchByProjection(F,KF,reloc,th=10)", ":1478-1605", 0.9, True, mk_proj(), run_reloc, null_proj, None, "1000 key-frame points, ORBdist 100"),
        ("SearchByBoW(KF,F)", ":178-306", 0.7, True, mk_bow, run_bow_kf_f, null_bow, d_bow_f, "1000 x 1000 features, 1000 vocabulary nodes"),
        ("SearchByBoW(KF,KF)", ":565-698", 0.75, True, mk_bow, run_bow_kf_kf, null_bow, d_bow_kf, "1000 x 1000 features, 1000 vocabulary nodes"),
        ("SearchForInitialization", ":448-563", 0.9, True, mk_init, run_init, null_init, d_init, "2000 x 2000 features, window 100"),
        ("SearchForTriangulation", ":700-852", 0.6, False, mk_tri, run_tri, null_tri, d_tri, "1000 x 1000 features, 120 nodes, epipolar gate"),
        ("SearchBySim3(th=7.5)", ":1124-1348", 0.75, True, mk_sim3, run_sim3, null_sim3, d_sim3, "two 1000-feature key frames, both directions"),
        ("Fuse(KF,MapPoints,th=3)", ":854-993", 0.6, True, mk_fuse, run_fuse, null_fuse, d_fuse, "1000 map points into a 1000-feature key frame"),
        ("Fuse(KF,Scw,th=4)", ":995-1122", 0.8, True, mk_fuse, run_fuse_sim3, null_fuse, d_fuse_sim3, "1000 loop points into a 1000-feature key frame"),
    ]

    def same(x, y):
        if isinstance(x, (tuple, list)):
            return len(x) == len(y) and all(same(p, q) for p, q in zip(x, y))
        if isinstance(x, np.ndarray):
            return np.array_equal(x, y)
        return x == y

    out, n_gpu_calls = {}, 0
    for name, cite, nnratio, ori, make, run, null, direct, shape in modes:
        D, R = pair(nnratio, ori)
        sets = [make(1000 + 37 * k) for k in range(4)]
        fd = [{k: D.frame(c[k], grid) for k in c["F"]} for c in sets]
        fr = [{k: R.frame(c[k], grid) for k in c["F"]} for c in sets]
        t0 = time.perf_counter()
        first = [run(D, sets[k], fd[k]) for k in range(4)]           # first use of every frame: pack upload + grid build
        cold = (time.perf_counter() - t0) / 4
        t0 = time.perf_counter()
        for i in range(calls):
            run(D, sets[i & 3], fd[i & 3])
        warm = (time.perf_counter() - t0) / calls
        t0 = time.perf_counter()
        ref = [run(R, sets[k % 4], fr[k % 4]) for k in range(cpu_calls)]
        cpu = (time.perf_counter() - t0) / cpu_calls
        nulls = [null(c) for c in sets]
        t0 = time.perf_counter()
        for k in range(cpu_calls):
            run(R, nulls[k % 4], fr[k % 4])
        glue = (time.perf_counter() - t0) / cpu_calls
        n_gpu_calls += calls + 4
        nm = first[0][0] if isinstance(first[0], tuple) else first[0]
        row = {"ref": cite, "shape": shape, "dropin_us_per_call": 1e6 * warm, "dropin_us_first_call": 1e6 * cold,
               "reference_cpu_us_per_call": 1e6 * cpu, "glue_us_per_call": 1e6 * glue, "reference_cpu_us_per_call_net": 1e6 * max(cpu - glue, 0.0),
               "calls": calls, "cpu_calls": cpu_calls, "matches_per_call": int(nm if np.isscalar(nm) else nm[0]),
               "identical_to_reference": bool(all(same(first[k], ref[k]) for k in range(min(4, cpu_calls))))}
        if direct is not None:
            om = ORBmatcher(gm, nnratio, ori)
            pks = [{k: gpack(c[k]) for k in c["F"]} for c in sets]
            for k in range(4):
                direct(om, sets[k], pks[k])
            t0 = time.perf_counter()
            for i in range(calls):
                direct(om, sets[i & 3], pks[i & 3])
            row["gpu_c_abi_us_per_call"] = 1e6 * (time.perf_counter() - t0) / calls
            n_gpu_calls += calls + 4
            for pk in pks:
                for v in pk.values():
                    v.close()
        out[name] = row
    gm.close()
    note = ("per call, one host thread. gpu_c_abi = the C ABI of include/ccm_match.h called directly (host arrays in, host arrays out: copies + "
            "kernels + synchronisation; for SearchBySim3 / Fuse the CCM_WIN_BEST searches without the host replay of the map mutations). "
            "dropin = the cslam::ORBmatcher bodies of motioncheck_ccm_slam_b200/host/ORBmatcher.cpp, reference = cslam/src/ORBmatcher.cpp compiled "
            "unmodified (oracle/_ref), both through the same object glue; glue = the reference arm with every query switched off, i.e. the mock "
            "Frame / KeyFrame / MapPoint set-up both arms pay; reference_cpu_us_per_call_net = reference - glue = the candidate loops the GPU "
            "replaces. first call = frame pack not cached yet")
    return {"what": note, "cpu": {"kind": "reference", "cores": 1}, "modes": out}, n_gpu_calls


# ----------------------------------------------------------------------------------------------------------
class _Agent:
    """One client of configs[3]: its own extractor handle, matcher, vocabulary copy and frame stream, driven frame by frame
    through the host entry points a drop-in Frame / Tracking would call."""

    def __init__(self, agent, device, frames, voc, nf, nl):
        from motioncheck_ccm_slam_b200 import ORBextractor, _lib
        from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
        from motioncheck_ccm_slam_b200.matcher import FramePack, Matcher, QUERY_DTYPE, RESULT_DTYPE, Vocabulary, frame_grid
        self.agent, self.frames = agent, frames
        self.T, self.h, self.w = frames.shape
        self.ex = ORBextractor(nf, SCALE, nl, INI_TH, MIN_TH, max_width=self.w, max_height=self.h, max_batch=1, device=device)
        self.m = Matcher(device)
        self.V = Vocabulary(self.m, voc)
        self.lib = _lib.load()
        self.cap = self.ex.max_keypoints()
        self.packs = [FramePack(self.m, self.cap), FramePack(self.m, self.cap)]
        self.grid = frame_grid(self.w, self.h)
        self.sf = self.ex.GetScaleFactors()
        cap, T = self.cap, self.T
        self.kp = np.zeros((T, cap), KP_DTYPE); self.desc = np.zeros((T, cap, 32), np.uint8); self.n = np.zeros(T, np.int32)
        self.word = np.zeros((T, cap), np.int32); self.node = np.zeros((T, cap), np.int32)
        self.xs = np.zeros((T, cap), np.float32); self.ys = np.zeros((T, cap), np.float32)
        self.oc = np.zeros((T, cap), np.int32); self.an = np.zeros((T, cap), np.float32)
        self.q1 = np.zeros((T, cap), QUERY_DTYPE); self.q2 = np.zeros((T, cap), QUERY_DTYPE)
        self.valid = np.ones(cap, np.uint8)
        self.match_a, self.match_b = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
        self.res = np.zeros(cap, RESULT_DTYPE)
        self.hold = np.zeros((2, cap), np.int32)
        self.nm = np.zeros((T, 3), np.int32)
        self.keep = {}
        self.done = 0

    def frame(self, t, record=False, sample=()):
        """extract -> pack (+grid) -> vocabulary transform -> SearchByBoW(KF = frame t-1, F) -> SearchByProjection(F, LastF, 7)
        -> SearchByProjection(F, MapPoints = frame t-2's points, 1); record=True also builds the queries of the next frames."""
        from motioncheck_ccm_slam_b200._lib import WindowParams
        from motioncheck_ccm_slam_b200.matcher import Q_OBS, Q_VALID, WIN_PROJ_LASTFRAME, WIN_PROJ_MAPPOINTS
        L, m, cap = self.lib, self.m._h, self.cap
        n = ctypes.c_int()
        img = self.frames[t]
        self.ex.extract_raw(img.ctypes.data, self.h, self.w, self.w, self.kp[t].ctypes.data, self.desc[t].ctypes.data, cap, ctypes.byref(n))
        k = n.value
        self.n[t] = k
        kp = self.kp[t]
        self.xs[t, :k] = kp["x"][:k]; self.ys[t, :k] = kp["y"][:k]; self.oc[t, :k] = kp["octave"][:k]; self.an[t, :k] = kp["angle"][:k]
        pack = self.packs[t & 1]
        rc = L.ccm_frame_pack_upload(pack._h, ctypes.byref(self.grid), k, self.desc[t].ctypes.data, self.xs[t].ctypes.data, self.ys[t].ctypes.data,
                                     self.oc[t].ctypes.data, self.an[t].ctypes.data)
        rc |= L.ccm_vocabulary_transform(m, self.V._h, k, self.desc[t].ctypes.data, 4, self.word[t].ctypes.data, self.node[t].ctypes.data)
        if record and t + 1 < self.T:
            kk = kp[:k]
            q1, q2 = self.q1[t + 1], self.q2[min(t + 2, self.T - 1)]
            q1["u"][:k] = kk["x"] - 2; q1["v"][:k] = kk["y"] - 1; q1["angle"][:k] = kk["angle"]; q1["r"][:k] = np.float32(7.0) * self.sf[kk["octave"]]
            q1["min_level"][:k] = kk["octave"] - 1; q1["max_level"][:k] = kk["octave"] + 1; q1["flags"][:k] = Q_VALID | Q_OBS
            if t + 2 < self.T:
                q2["u"][:k] = kk["x"] - 4; q2["v"][:k] = kk["y"] - 2; q2["r"][:k] = np.float32(2.5) * self.sf[kk["octave"]]
                q2["min_level"][:k] = kk["octave"] - 1; q2["max_level"][:k] = kk["octave"]; q2["flags"][:k] = Q_VALID | Q_OBS
        nm = ctypes.c_int()
        if t >= 1:
            p = t - 1                                       # the reference key frame = the previous frame
            rc |= L.ccm_search_bow(m, 0, 0.7, 1, int(self.n[p]), self.desc[p].ctypes.data, self.node[p].ctypes.data, self.valid.ctypes.data,
                                   self.an[p].ctypes.data, k, self.desc[t].ctypes.data, self.node[t].ctypes.data, self.valid.ctypes.data,
                                   self.an[t].ctypes.data, self.match_a.ctypes.data, self.match_b.ctypes.data, self.res.ctypes.data, ctypes.byref(nm))
            self.nm[t, 0] = nm.value
            if t in sample:
                self.keep[(t, "bow")] = (nm.value, self.match_b[:k].copy())
            prm = WindowParams(WIN_PROJ_LASTFRAME, 0.9, 100, 1, None, 0)
            self.hold[0, :k] = -1
            rc |= L.ccm_search_window(m, pack._h, ctypes.byref(prm), int(self.n[p]), self.q1[t].ctypes.data, self.desc[p].ctypes.data,
                                      self.hold[0].ctypes.data, None, ctypes.byref(nm))
            self.nm[t, 1] = nm.value
            if t in sample:
                self.keep[(t, "last")] = (nm.value, self.hold[0, :k].copy())
        if t >= 2:
            p = t - 2
            prm = WindowParams(WIN_PROJ_MAPPOINTS, 0.8, 100, 1, None, 0)
            self.hold[1, :k] = -1
            rc |= L.ccm_search_window(m, pack._h, ctypes.byref(prm), int(self.n[p]), self.q2[t].ctypes.data, self.desc[p].ctypes.data,
                                      self.hold[1].ctypes.data, None, ctypes.byref(nm))
            self.nm[t, 2] = nm.value
            if t in sample:
                self.keep[(t, "mp")] = (nm.value, self.hold[1, :k].copy())
        if rc != 0:
            raise RuntimeError("c4 agent %d frame %d: C ABI error %d" % (self.agent, t, rc))
        self.done += 1

    def run(self, sample=()):
        for t in range(self.T):
            self.frame(t, sample=sample)

    def check_against_oracle(self, voc, sample):
        """The recorded outputs of the sampled frames against the CPU oracle (checker only, outside every timed region)."""
        _tests_path()
        import _oracle
        ok = 0
        for t in sample:
            k, p, p2 = int(self.n[t]), t - 1, t - 2
            F = _oracle.OracleFrame(self.desc[t, :k], self.xs[t, :k], self.ys[t, :k], self.oc[t, :k], self.an[t, :k], self.grid)
            node_t = _oracle.oracle_vocab_transform(voc, self.desc[t, :k], 4)[2]
            node_p = _oracle.oracle_vocab_transform(voc, self.desc[p, :self.n[p]], 4)[2]
            assert np.array_equal(node_t, self.node[t, :k]), "vocabulary transform differs from the oracle"
            nm, mf, _ = _oracle.OracleMatcher(0.7, True).bow_kf_frame(self.desc[p, :self.n[p]], node_p, np.ones(self.n[p], np.uint8), self.an[p, :self.n[p]],
                                                                      self.desc[t, :k], node_t, self.an[t, :k])
            g = self.keep[(t, "bow")]
            assert nm == g[0] and np.array_equal(mf, g[1]), "SearchByBoW differs from the oracle (agent %d frame %d)" % (self.agent, t)
            one_p, one_p2 = np.ones(int(self.n[p]), np.uint8), np.ones(int(self.n[p2]), np.uint8)
            hold = np.full(k, -1, np.int32)
            nm = _oracle.OracleMatcher(0.9, True).projection_lastframe(F, one_p, self.xs[p, :self.n[p]] - 2, self.ys[p, :self.n[p]] - 1, self.oc[p, :self.n[p]],
                                                                       self.an[p, :self.n[p]], self.desc[p, :self.n[p]], one_p, 7.0, self.sf, hold)[0]
            g = self.keep[(t, "last")]
            assert nm == g[0] and np.array_equal(hold, g[1]), "SearchByProjection(LastFrame) differs from the oracle (agent %d frame %d)" % (self.agent, t)
            hold = np.full(k, -1, np.int32)
            nm = _oracle.OracleMatcher(0.8, True).projection_mappoints(F, one_p2, self.xs[p2, :self.n[p2]] - 4, self.ys[p2, :self.n[p2]] - 2, self.oc[p2, :self.n[p2]],
                                                                       one_p2.astype(np.float32), self.desc[p2, :self.n[p2]], one_p2, 1.0, self.sf, hold)[0]
            g = self.keep[(t, "mp")]
            assert nm == g[0] and np.array_equal(hold, g[1]), "SearchByProjection(MapPoints) differs from the oracle (agent %d frame %d)" % (self.agent, t)
            ok += 1
        return ok

    def close(self):
        for p in self.packs:
            p.close()
        self.V.close(); self.m.close(); self.ex.close()


def c4_leg(torch, dist, rank, world, local_rank, barrier, voc, nf=1000, nl=8, w=752, h=480, T=48):
    """configs[3]: 4 agent streams (seeds 1000..4000), agent a on GPU a mod G, one host thread per agent, frame by frame."""
    from motioncheck_ccm_slam_b200.synth import make_stream
    n_agents = 4
    mine = [a for a in range(n_agents) if a % world == rank]
    agents = [_Agent(a, local_rank, make_stream(1000 * (a + 1), T, w, h), voc, nf, nl) for a in mine]
    sample = (2, T // 2, T - 1)
    for ag in agents:                                       # untimed pass: warms every handle and records the next frames' queries
        for t in range(T):
            ag.frame(t, record=True)
    alone_fps = None
    if agents and rank == 0:                                # one agent alone on its GPU
        t0 = time.perf_counter()
        agents[0].run()
        alone_fps = T / (time.perf_counter() - t0)
    barrier()
    ths = [threading.Thread(target=ag.run, kwargs={"sample": sample}) for ag in agents]
    t0 = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    dt = time.perf_counter() - t0
    barrier()
    checked = sum(ag.check_against_oracle(voc, sample) for ag in agents[:1])
    per_agent = {str(ag.agent): T / dt for ag in agents}
    nm = np.concatenate([ag.nm[2:] for ag in agents]) if agents else np.zeros((1, 3))
    launches = sum(ag.ex.launch_count() + ag.m.launch_count() for ag in agents)
    for ag in agents:
        ag.close()
    stat = torch.tensor([dt if agents else 0.0], dtype=torch.float64, device=torch.device("cuda", local_rank))
    if world > 1:
        dist.all_reduce(stat, op=dist.ReduceOp.MAX)
        gathered = [None] * world
        dist.all_gather_object(gathered, per_agent)
        per_agent = {k: v for d in gathered for k, v in d.items()}
    dt_max = float(stat[0])
    out = {"what": "4 agent streams (seeds 1000-4000), agent a on GPU a mod %d, one host thread per agent, frame by frame through the host entry "
                   "points: ccm_orb_extract -> frame pack + 75x48 grid -> vocabulary transform (k=10, L=6) -> SearchByBoW(KF = previous frame, F, 0.7) "
                   "-> SearchByProjection(F, LastFrame, 7) -> SearchByProjection(F, MapPoints = frame t-2, 1); queries precomputed" % world,
           "frames_per_agent": T, "frames_per_s_total": n_agents * T / dt_max, "frames_per_s_per_agent": per_agent,
           "matches_per_frame": {"bow": float(nm[:, 0].mean()), "lastframe": float(nm[:, 1].mean()), "mappoints": float(nm[:, 2].mean())},
           "oracle_checked_frames": int(checked), "host": "python threads + ctypes (GIL released inside every call)"}
    if alone_fps is not None:
        out["one_agent_alone_frames_per_s"] = alone_fps
        mine_fps = np.mean([v for k, v in per_agent.items() if int(k) % world == 0])
        out["contention_factor_rank0_gpu"] = alone_fps / mine_fps
        out["agents_on_rank0_gpu"] = len(mine)
    return out, int(launches)
