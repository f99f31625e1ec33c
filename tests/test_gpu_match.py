"""GPU parity: the CUDA Hamming searches (through the C ABI, include/ccm_match.h) against the CPU oracle
(oracle/match_oracle.cpp, the restated loops of cslam/src/ORBmatcher.cpp) on the same seeded inputs. Everything here
is integer / index work: the bar is bit-exact (distances, indices, accepted sets, holder state, return values)."""
import numpy as np
import pytest

from _oracle import OracleFrame, OracleMatcher

pytestmark = pytest.mark.gpu

W, H = 752, 480
SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)   # stand-in for mvScaleFactor; both sides get the same table


@pytest.fixture(scope="module")
def gm(cuda_lib):
    from motioncheck_ccm_slam_b200.matcher import Matcher
    m = Matcher(0)
    yield m
    m.close()


def _frames(seed, n=1000, **kw):
    from motioncheck_ccm_slam_b200.synth import make_feature_frame, make_next_frame
    a = make_feature_frame(seed, n, W, H)
    b, src = make_next_frame(seed + 1, a, W, H, **kw)
    return a, b, src


def _pack(gm, fr, grid=None, cap=None):
    from motioncheck_ccm_slam_b200.matcher import FramePack, frame_grid
    grid = grid or frame_grid(W, H)
    p = FramePack(gm, cap or max(len(fr["x"]), 1))
    p.upload(grid, fr["desc"], fr["x"], fr["y"], fr["octave"], fr["angle"])
    return p, OracleFrame(fr["desc"], fr["x"], fr["y"], fr["octave"], fr["angle"], grid)


def _same_res(r_gpu, r_cpu, fields=("best_idx", "best_dist", "best_dist2", "accepted")):
    for f in fields:
        bad = np.nonzero(r_gpu[f] != r_cpu[f])[0]
        assert len(bad) == 0, (f, bad[:10], r_gpu[bad[:10]], r_cpu[bad[:10]])


def test_descriptor_distance(gm):
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (4097, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (4097, 32), dtype=np.uint8)
    b[:5] = a[:5]
    b[5] = ~a[5]
    d = gm.DescriptorDistance(a, b)
    assert np.array_equal(d, np.unpackbits(a ^ b, axis=1).sum(1))
    assert np.array_equal(d[:64], OracleMatcher().descriptor_distance(a[:64], b[:64]))
    assert d[5] == 256 and d[0] == 0


@pytest.mark.parametrize("seed,n,cap", [(11, 1000, None), (12, 2000, None), (13, 1, None), (14, 0, None), (15, 5000, None),
                                        (16, 3000, 30000)])      # a pack too large for the shared-memory working set: the grid is built in global memory
def test_grid_matches_frame_assign(gm, seed, n, cap):
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    from motioncheck_ccm_slam_b200.synth import make_feature_frame
    fr = make_feature_frame(seed, max(n, 1), W, H)
    fr = {k: v[:n] for k, v in fr.items()}
    if n > 10:   # features outside the image bounds (undistortion can push them out) and on cell boundaries
        fr["x"][:4] = [-30.0, W + 20.0, 10.0, 375.99]
        fr["y"][:4] = [10.0, 10.0, H + 9.0, -0.49]
        fr["x"][4:8] = [5.0133333, 10.026666, 15.04, 20.053333]
    grid = frame_grid(W, H, min_x=-3.5, min_y=-2.25) if seed % 2 else frame_grid(W, H)
    p, of = _pack(gm, fr, grid, cap=cap or max(n, 1) + 7)
    cs, items = p.get_grid()
    ocs, oitems = of.csr()
    assert np.array_equal(cs, ocs) and np.array_equal(items, oitems)
    assert p.count() == n


def _projection_inputs(a, b, src, rng, level_from="b"):
    """Queries = map points seen in frame a, projected into frame b at (their true position + noise)."""
    n = len(a["x"])
    inv = np.full(n, -1)
    inv[src[src >= 0]] = np.nonzero(src >= 0)[0]
    u = np.where(inv >= 0, b["x"][np.maximum(inv, 0)], a["x"] + 2).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
    v = np.where(inv >= 0, b["y"][np.maximum(inv, 0)], a["y"] + 1).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
    level = np.where(inv >= 0, b["octave"][np.maximum(inv, 0)], a["octave"]).astype(np.int32)
    level = np.clip(level + rng.choice([0, 0, 0, 1], n), 0, 7).astype(np.int32)
    valid = (rng.random(n) < 0.9).astype(np.uint8)
    obs = (rng.random(n) < 0.8).astype(np.uint8)
    return valid, u, v, level, obs


@pytest.mark.parametrize("seed,th,nnratio", [(21, 1.0, 0.8), (22, 5.0, 0.8), (23, 3.0, 0.6), (24, 12.0, 0.9)])
def test_search_by_projection_mappoints(gm, seed, th, nnratio):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed)
    valid, u, v, level, obs = _projection_inputs(a, b, src, rng)
    viewcos = rng.choice([0.9, 0.9985, 0.998, 0.99801], len(u)).astype(np.float32)
    holder0 = rng.choice([-1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
    p, of = _pack(gm, b)
    hg, hc = holder0.copy(), holder0.copy()
    nm_g, res_g = ORBmatcher(gm, nnratio, True).SearchByProjection_MapPoints(p, valid, u, v, level, viewcos, a["desc"], obs, th, SF, hg)
    nm_c, res_c = OracleMatcher(nnratio, True).projection_mappoints(of, valid, u, v, level, viewcos, a["desc"], obs, th, SF, hc)
    _same_res(res_g, res_c)
    assert nm_g == nm_c and np.array_equal(hg, hc)
    assert nm_c > 200


@pytest.mark.parametrize("seed,th,ori", [(31, 7.0, True), (32, 14.0, True), (33, 7.0, False), (34, 30.0, True)])
def test_search_by_projection_lastframe(gm, seed, th, ori):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed)
    valid, u, v, _, obs = _projection_inputs(a, b, src, rng)
    holder0 = rng.choice([-1, -1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
    p, of = _pack(gm, b)
    hg, hc = holder0.copy(), holder0.copy()
    nm_g, res_g = ORBmatcher(gm, 0.9, ori).SearchByProjection_LastFrame(p, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hg)
    nm_c, res_c = OracleMatcher(0.9, ori).projection_lastframe(of, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hc)
    _same_res(res_g, res_c)
    assert nm_g == nm_c and np.array_equal(hg, hc)
    assert nm_c > 200


@pytest.mark.parametrize("seed,th", [(41, 10), (42, 4)])
def test_search_by_projection_kf_sim3(gm, seed, th):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed, n=3000)
    valid, u, v, level, claims = _projection_inputs(a, b, src, rng)
    holder0 = rng.choice([-1, -1, -1, -2], len(b["x"])).astype(np.int32)
    p, of = _pack(gm, b)
    hg, hc = holder0.copy(), holder0.copy()
    nm_g, res_g = ORBmatcher(gm, 0.75, True).SearchByProjection_KFSim3(p, valid, u, v, level, a["desc"], claims, th, SF, hg)
    nm_c, res_c = OracleMatcher(0.75, True).projection_kf_sim3(of, valid, u, v, level, a["desc"], claims, th, SF, hc)
    _same_res(res_g, res_c)
    assert nm_g == nm_c and np.array_equal(hg, hc)
    assert nm_c > 100


@pytest.mark.parametrize("seed,gate,th_dist", [(51, False, 100), (52, True, 50), (53, False, 50)])
def test_search_best_in_window(gm, seed, gate, th_dist):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed)
    valid, u, v, level, _ = _projection_inputs(a, b, src, rng)
    radius = (np.float32(7.5) * SF[level]).astype(np.float32)
    inv_sigma2 = (1.0 / (SF * SF)).astype(np.float32) if gate else None
    p, of = _pack(gm, b)
    nm_g, res_g = ORBmatcher(gm).SearchBestInWindow(p, valid, u, v, radius, level, a["desc"], th_dist, inv_sigma2)
    nm_c, res_c = OracleMatcher().best_in_window(of, valid, u, v, radius, level, a["desc"], th_dist, inv_sigma2)
    _same_res(res_g, res_c)
    assert nm_g == nm_c and nm_c > 100


def test_greedy_chain_conflicts(gm):
    """Many identical queries fighting for few targets: the reference's sequential claim order must be reproduced."""
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    rng = np.random.default_rng(61)
    nt, nq = 40, 300
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    from motioncheck_ccm_slam_b200.synth import perturb_descriptors
    b = dict(desc=perturb_descriptors(rng, np.repeat(base, nt, 0), 0.05), x=rng.uniform(100, 130, nt).astype(np.float32),
             y=rng.uniform(100, 130, nt).astype(np.float32), octave=np.zeros(nt, np.int32), angle=rng.uniform(0, 360, nt).astype(np.float32))
    qd = perturb_descriptors(rng, np.repeat(base, nq, 0), 0.02)
    u = np.full(nq, 115, np.float32); v = np.full(nq, 115, np.float32)
    valid = np.ones(nq, np.uint8); obs = (rng.random(nq) < 0.7).astype(np.uint8)
    p, of = _pack(gm, b)
    for mode in ("last", "mp", "sim3"):
        hg, hc = np.full(nt, -1, np.int32), np.full(nt, -1, np.int32)
        if mode == "last":
            nm_g, res_g = ORBmatcher(gm, 0.9, False).SearchByProjection_LastFrame(p, valid, u, v, np.zeros(nq), np.zeros(nq), qd, obs, 30.0, SF, hg)
            nm_c, res_c = OracleMatcher(0.9, False).projection_lastframe(of, valid, u, v, np.zeros(nq), np.zeros(nq), qd, obs, 30.0, SF, hc)
        elif mode == "mp":
            nm_g, res_g = ORBmatcher(gm, 0.95, False).SearchByProjection_MapPoints(p, valid, u, v, np.zeros(nq), np.ones(nq), qd, obs, 8.0, SF, hg)
            nm_c, res_c = OracleMatcher(0.95, False).projection_mappoints(of, valid, u, v, np.zeros(nq), np.ones(nq), qd, obs, 8.0, SF, hc)
        else:
            nm_g, res_g = ORBmatcher(gm).SearchByProjection_KFSim3(p, valid, u, v, np.zeros(nq), qd, obs, 30, SF, hg)
            nm_c, res_c = OracleMatcher().projection_kf_sim3(of, valid, u, v, np.zeros(nq), qd, obs, 30, SF, hc)
        _same_res(res_g, res_c)
        assert nm_g == nm_c and np.array_equal(hg, hc), mode
        assert nm_c >= 20


@pytest.mark.parametrize("seed,window,ori", [(71, 100, True), (72, 30, True), (73, 100, False)])
def test_search_for_initialization(gm, seed, window, ori):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    a, b, src = _frames(seed, n=2000, motion=(6.0, -3.0))
    p, of = _pack(gm, b)
    prev_g = np.stack([a["x"], a["y"]], 1).astype(np.float32).copy()
    prev_c = prev_g.copy()
    nm_g, m_g = ORBmatcher(gm, 0.9, ori).SearchForInitialization(a["desc"], a["octave"], a["angle"], prev_g, p,
                                                                np.stack([b["x"], b["y"]], 1), window)
    nm_c, m_c = OracleMatcher(0.9, ori).search_for_initialization(a["desc"], a["octave"], a["angle"], prev_c, of, window)
    assert nm_g == nm_c and np.array_equal(m_g, m_c) and np.array_equal(prev_g, prev_c)
    assert nm_c > 50


@pytest.mark.parametrize("seed,nnratio,ori,n_nodes", [(81, 0.7, True, 1000), (82, 0.75, False, 100), (83, 0.9, True, 7)])
def test_search_by_bow(gm, seed, nnratio, ori, n_nodes):
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    from motioncheck_ccm_slam_b200.synth import bow_node_ids
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed, flip_prob=0.03)
    na, nb = bow_node_ids(a["desc"], n_nodes), bow_node_ids(b["desc"], n_nodes)
    na[rng.random(len(na)) < 0.05] = -1
    nb[rng.random(len(nb)) < 0.05] = -1
    va = (rng.random(len(na)) < 0.8).astype(np.uint8)
    vb = (rng.random(len(nb)) < 0.8).astype(np.uint8)
    g, c = ORBmatcher(gm, nnratio, ori), OracleMatcher(nnratio, ori)
    nm_g, mf_g, res_g = g.SearchByBoW_KF_Frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
    nm_c, mf_c, res_c = c.bow_kf_frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
    _same_res(res_g, res_c)
    assert nm_g == nm_c and np.array_equal(mf_g, mf_c) and nm_c > 20
    nm_g, m12_g, res_g = g.SearchByBoW_KF_KF(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
    nm_c, m12_c, res_c = c.bow_kf_kf(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
    _same_res(res_g, res_c)
    assert nm_g == nm_c and np.array_equal(m12_g, m12_c) and nm_c > 10


@pytest.mark.parametrize("nq,nt", [(1000, 20000), (1, 1), (255, 257), (257, 255), (1025, 4099), (500, 70001), (3, 0), (2048, 513)])
def test_allpairs_best2(gm, nq, nt):
    from motioncheck_ccm_slam_b200.synth import perturb_descriptors
    rng = np.random.default_rng(nq * 7 + nt)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    if nt > 10:
        k = min(nq, nt) // 2
        q[:k] = perturb_descriptors(rng, t[rng.integers(0, nt, k)], 0.05)
        t[nt // 2] = t[3]; t[nt - 1] = t[3]                 # exact duplicates: first index wins, second distance == best
        q[0] = t[3]
    out = gm.match_allpairs_best2(q, t, index_base=0)
    ref = OracleMatcher(0.75).allpairs_best2(q, t)
    if nt == 0:
        assert np.all(out["idx"] == 0xFFFFFFFF) and np.all(out["d1"] == 256) and np.all(out["d2"] == 256)
        return
    assert np.array_equal(out["idx"].astype(np.int64), ref["best_idx"].astype(np.int64))
    assert np.array_equal(out["d1"], ref["best_dist"]) and np.array_equal(out["d2"], ref["best_dist2"])
    if nt > 10:
        assert out["idx"][0] == 3 and out["d1"][0] == 0 and out["d2"][0] == 0


def test_allpairs_sharded_merge_equals_single(gm):
    """Shard the target database, search each shard with its index base, merge: identical to one search (the NCCL
    path of BASELINE.json configs[4] does exactly this with the allgather in between)."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE
    rng = np.random.default_rng(5)
    nq, nt, shards = 1000, 60000, 4
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t[50000] = t[100]; q[7] = t[100]
    whole = gm.match_allpairs_best2(q, t)
    dq = torch.from_numpy(q).cuda(); dt = torch.from_numpy(t).cuda()
    parts = torch.zeros((shards, nq, 8), dtype=torch.uint8, device="cuda")
    out = torch.zeros((nq, 8), dtype=torch.uint8, device="cuda")
    acc = torch.zeros(nq, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    bounds = [0, 9000, 9000 + 256, 41000, nt]
    for s in range(shards):
        lo, hi = bounds[s], bounds[s + 1]
        gm.match_allpairs_best2_device(dt.data_ptr() * 0 + dq.data_ptr(), nq, dt.data_ptr() + lo * 32, hi - lo, lo, parts[s].data_ptr())
    gm.match_allpairs_merge_device(parts.data_ptr(), shards, nq, 0.75, out.data_ptr(), acc.data_ptr())
    gm.sync()
    merged = out.cpu().numpy().view(BEST2_DTYPE).reshape(-1)
    assert np.array_equal(merged, whole)
    ref = OracleMatcher(0.75).allpairs_best2(q, t)
    assert np.array_equal(acc.cpu().numpy().astype(np.int32), ref["accepted"])
    assert merged["idx"][7] == 100 and merged["d2"][7] == 0


@pytest.mark.parametrize("B", [4, 10, 24])      # queries in flight ~3 k / ~9 k / ~23 k: k_window_eval at 32 / 16 / 8 lanes per query
def test_extractor_to_matcher_zero_copy_batch(gm, B):
    """Extractor output stays in HBM: packs are built from it on the device, frame t is searched with frame t-1's
    keypoints as queries (SearchByProjection(Frame, LastFrame)), all frames of the batch in one launch. Checked against
    the oracle fed with the host copies of the same extractor output."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200 import ORBextractor
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    from motioncheck_ccm_slam_b200.matcher import (FramePack, QUERY_DTYPE, RESULT_DTYPE, WIN_PROJ_LASTFRAME, Q_VALID, Q_OBS,
                                                   frame_grid, packs_from_extractor_batch_device, search_window_batch_device)
    from motioncheck_ccm_slam_b200.synth import make_stream
    w, h = 640, 400
    frames = make_stream(4242, B, w, h)
    ex = ORBextractor(800, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B, device=0)
    cap = ex.max_keypoints()
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(frames).to(dev)
    d_kp = torch.zeros((B, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros((B,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    ex.extract_batch_device(d_img.data_ptr(), B, h, w, w, w * h, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    ex.sync()
    kp = d_kp.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
    desc = d_desc.cpu().numpy()
    n = d_n.cpu().numpy()
    assert n.min() > 500
    grid = frame_grid(w, h)
    packs = [FramePack(gm, cap) for _ in range(B)]
    packs_from_extractor_batch_device(packs, grid, d_kp.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap)
    sf = ex.GetScaleFactors()
    # queries for job j (pack j+1): frame j's keypoints moved by the stream's motion (-2, -1)
    q = np.zeros((B - 1, cap), QUERY_DTYPE)
    for j in range(B - 1):
        k = kp[j, :n[j]]
        q["u"][j, :n[j]] = k["x"] - 2; q["v"][j, :n[j]] = k["y"] - 1; q["angle"][j, :n[j]] = k["angle"]
        q["r"][j, :n[j]] = np.float32(7.0) * sf[k["octave"]]
        q["min_level"][j, :n[j]] = k["octave"] - 1; q["max_level"][j, :n[j]] = k["octave"] + 1
        q["flags"][j, :n[j]] = Q_VALID | Q_OBS
    d_q = torch.from_numpy(q.view(np.uint8).reshape(B - 1, cap, 24)).to(dev)
    d_holder = torch.full((B - 1, cap), -1, dtype=torch.int32, device=dev)
    d_res = torch.zeros((B - 1, cap, 4), dtype=torch.int32, device=dev)
    d_nm = torch.zeros((B - 1,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    search_window_batch_device(gm, packs[1:], WIN_PROJ_LASTFRAME, 0.9, 100, True, d_n.data_ptr(), cap, d_q.data_ptr(),
                               d_desc.data_ptr(), d_holder.data_ptr(), cap, d_res.data_ptr(), d_nm.data_ptr())
    gm.sync()
    holder, res, nm = d_holder.cpu().numpy(), d_res.cpu().numpy().view(RESULT_DTYPE).reshape(B - 1, cap), d_nm.cpu().numpy()
    for j in range(B - 1):
        t = kp[j + 1, :n[j + 1]]
        of = OracleFrame(desc[j + 1, :n[j + 1]], t["x"], t["y"], t["octave"], t["angle"], grid)
        cs, items = packs[j + 1].get_grid()
        ocs, oitems = of.csr()
        assert np.array_equal(cs, ocs) and np.array_equal(items, oitems)
        k = kp[j, :n[j]]
        hc = np.full(n[j + 1], -1, np.int32)
        nm_c, res_c = OracleMatcher(0.9, True).projection_lastframe(of, np.ones(n[j]), k["x"] - 2, k["y"] - 1, k["octave"], k["angle"],
                                                                    desc[j, :n[j]], np.ones(n[j]), 7.0, sf, hc)
        _same_res(res[j, :n[j]], res_c)
        assert nm[j] == nm_c and np.array_equal(holder[j, :n[j + 1]], hc)
        assert nm_c > 300, nm_c        # consecutive frames of the stream really match
    ex.close()


@pytest.mark.parametrize("n_jobs", [1, 6, 20])      # 1 k / 6 k / 20 k queries in flight: 32 / 16 / 8 lanes per query in k_window_eval
def test_dense_windows_batch_all_lane_widths(gm, n_jobs):
    """Windows of 60-150 candidates (radius 120 px x scale): every lane of a query's group sees more than its four best, so the
    stored candidate lists end early and the resolver has to know it; hundreds of queries compete for the same features."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200.matcher import (QUERY_DTYPE, RESULT_DTYPE, WIN_PROJ_LASTFRAME, Q_VALID, Q_OBS, frame_grid,
                                                   search_window_batch_device)
    dev = torch.device("cuda", 0)
    n, th = 1000, 120.0
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    grid = frame_grid(W, H)
    packs, ofs, q, qd, rng = [], [], np.zeros((n_jobs, n), QUERY_DTYPE), np.zeros((n_jobs, n, 32), np.uint8), np.random.default_rng(77)
    inputs = []
    for j in range(n_jobs):
        a, b, _ = _frames(500 + 3 * j, n)
        p, of = _pack(gm, b, grid, cap=n)
        packs.append(p); ofs.append(of)
        u = (a["x"] + rng.normal(0, 3, n)).astype(np.float32); v = (a["y"] + rng.normal(0, 3, n)).astype(np.float32)
        q["u"][j] = u; q["v"][j] = v; q["angle"][j] = a["angle"]
        q["r"][j] = np.float32(th) * sf[a["octave"]]
        q["min_level"][j] = a["octave"] - 1; q["max_level"][j] = a["octave"] + 1
        q["flags"][j] = Q_VALID | Q_OBS
        qd[j] = a["desc"]
        inputs.append((a, u, v))
    d_q = torch.from_numpy(q.view(np.uint8).reshape(n_jobs, n, 24)).to(dev)
    d_qd = torch.from_numpy(qd).to(dev)
    d_holder = torch.full((n_jobs, n), -1, dtype=torch.int32, device=dev)
    d_res = torch.zeros((n_jobs, n, 4), dtype=torch.int32, device=dev)
    d_nm = torch.zeros((n_jobs,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    search_window_batch_device(gm, packs, WIN_PROJ_LASTFRAME, 0.9, 100, True, None, n, d_q.data_ptr(), d_qd.data_ptr(),
                               d_holder.data_ptr(), n, d_res.data_ptr(), d_nm.data_ptr())
    gm.sync()
    holder, res, nm = d_holder.cpu().numpy(), d_res.cpu().numpy().view(RESULT_DTYPE).reshape(n_jobs, n), d_nm.cpu().numpy()
    for j in range(n_jobs):
        a, u, v = inputs[j]
        hc = np.full(n, -1, np.int32)
        nm_c, res_c = OracleMatcher(0.9, True).projection_lastframe(ofs[j], np.ones(n), u, v, a["octave"], a["angle"], a["desc"], np.ones(n), th, sf, hc)
        _same_res(res[j], res_c)
        assert nm[j] == nm_c and np.array_equal(holder[j], hc)
    for p in packs:
        p.close()


def test_dropin_cpp_matcher_matches_reference_class(cuda_lib):
    """The drop-in C++ ORBmatcher bodies (motioncheck_ccm_slam_b200/host/ORBmatcher.cpp, compiled against the
    reference's own <cslam/ORBmatcher.h>) and the reference's ORBmatcher.cpp, both driven through the same glue with
    Frame / KeyFrame / MapPoint objects: identical object state after every converted method. Built only where
    /root/reference exists; the prebuilt libraries travel to the GPU box."""
    import _oracle
    if not (_oracle.dropin_match_available() and _oracle.ref_match_available()):
        pytest.skip("oracle/_ref/libdropin_match.so not built (needs the reference headers at build time)")
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    from motioncheck_ccm_slam_b200.synth import bow_node_ids
    grid = frame_grid(W, H)

    def pair(nnratio, ori):
        return _oracle.RefMatcher(nnratio, ori, lib="libdropin_match.so"), _oracle.RefMatcher(nnratio, ori)

    def inb(u, v, valid):
        return valid & ((u >= 0) & (u <= W) & (v >= 0) & (v <= H)).astype(np.uint8)

    for seed, th, nnratio in [(21, 1.0, 0.8), (22, 5.0, 0.8), (24, 12.0, 0.9)]:
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed)
        valid, u, v, level, obs = _projection_inputs(a, b, src, rng)
        viewcos = rng.choice([0.9, 0.9985, 0.998, 0.99801], len(u)).astype(np.float32)
        holder0 = rng.choice([-1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
        D, R = pair(nnratio, True)
        hd, hr = holder0.copy(), holder0.copy()
        nd = D.projection_mappoints(D.frame(b, grid), valid, u, v, level, viewcos, a["desc"], obs, th, SF, hd)
        nr = R.projection_mappoints(R.frame(b, grid), valid, u, v, level, viewcos, a["desc"], obs, th, SF, hr)
        assert nd == nr > 200 and np.array_equal(hd, hr)
    for seed, th, ori in [(31, 7.0, True), (32, 14.0, True), (33, 7.0, False)]:
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed)
        valid, u, v, _, obs = _projection_inputs(a, b, src, rng)
        valid = inb(u, v, valid)
        holder0 = rng.choice([-1, -1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
        D, R = pair(0.9, ori)
        hd, hr = holder0.copy(), holder0.copy()
        nd = D.projection_lastframe(D.frame(b, grid), valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hd)
        nr = R.projection_lastframe(R.frame(b, grid), valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hr)
        assert nd == nr > 200 and np.array_equal(hd, hr)
    # relocalisation overload (:1478-1605): every occupied feature blocks, bound = ORBdist, levels [l-1, l+1]
    for seed, th, orb_dist, ori in [(51, 10.0, 100, True), (52, 3.0, 64, True), (53, 10.0, 100, False)]:
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed)
        valid, u, v, level, _ = _projection_inputs(a, b, src, rng)
        valid = inb(u, v, valid)
        state = np.where(valid > 0, rng.choice([1, 1, 1, 1, 1, 1, 2, 3], len(u)), 0).astype(np.uint8)
        holder0 = rng.choice([-1, -1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
        D, R = pair(0.9, ori)
        hd, hr = holder0.copy(), holder0.copy()
        nd = D.projection_reloc(D.frame(b, grid), state, u, v, level, a["angle"], a["desc"], th, orb_dist, SF, hd)
        nr = R.projection_reloc(R.frame(b, grid), state, u, v, level, a["angle"], a["desc"], th, orb_dist, SF, hr)
        assert nd == nr > 150 and np.array_equal(hd, hr), (seed, nd, nr)
    for seed, th in [(41, 10), (42, 4)]:
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed, n=3000)
        valid, u, v, level, claims = _projection_inputs(a, b, src, rng)
        holder0 = rng.choice([-1, -1, -1, -2], len(b["x"])).astype(np.int32)
        D, R = pair(0.75, True)
        hd, hr = holder0.copy(), holder0.copy()
        nd = D.projection_kf_sim3(D.frame(b, grid), valid, u, v, level, a["desc"], claims, th, SF, hd)
        nr = R.projection_kf_sim3(R.frame(b, grid), valid, u, v, level, a["desc"], claims, th, SF, hr)
        assert nd == nr > 100 and np.array_equal(hd, hr)
    for seed, window, ori in [(71, 100, True), (73, 100, False)]:
        a, b, src = _frames(seed, n=2000, motion=(6.0, -3.0))
        D, R = pair(0.9, ori)
        pd_, pr = np.stack([a["x"], a["y"]], 1).astype(np.float32).copy(), np.stack([a["x"], a["y"]], 1).astype(np.float32).copy()
        nd, md = D.search_for_initialization(a["desc"], a["octave"], a["angle"], pd_, D.frame(b, grid), window)
        nr, mr = R.search_for_initialization(a["desc"], a["octave"], a["angle"], pr, R.frame(b, grid), window)
        assert nd == nr > 50 and np.array_equal(md, mr) and np.array_equal(pd_, pr)
    for seed, nnratio, ori, n_nodes in [(81, 0.7, True, 1000), (82, 0.75, False, 100)]:
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed, flip_prob=0.03)
        na, nb = bow_node_ids(a["desc"], n_nodes), bow_node_ids(b["desc"], n_nodes)
        na[rng.random(len(na)) < 0.05] = -1; nb[rng.random(len(nb)) < 0.05] = -1
        va = (rng.random(len(na)) < 0.8).astype(np.uint8); vb = (rng.random(len(nb)) < 0.8).astype(np.uint8)
        D, R = pair(nnratio, ori)
        nd, md = D.bow_kf_frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
        nr, mr = R.bow_kf_frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
        assert nd == nr > 20 and np.array_equal(md, mr)
        nd, md = D.bow_kf_kf(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
        nr, mr = R.bow_kf_kf(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
        assert nd == nr > 10 and np.array_equal(md, mr)
    rng = np.random.default_rng(5)
    x = rng.integers(0, 256, (64, 32), dtype=np.uint8); y = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    D, R = pair(0.6, True)
    assert np.array_equal(D.descriptor_distance(x, y), R.descriptor_distance(x, y))


def test_dropin_cpp_frame_pack_cache(cuda_lib):
    """The drop-in keeps the device pack of a Frame / KeyFrame in an LRU side table keyed by mId: a frame searched three
    times is uploaded once, different frames never share a pack, and every result still equals the reference class."""
    import ctypes
    import _oracle
    if not (_oracle.dropin_match_available() and _oracle.ref_match_available()):
        pytest.skip("oracle/_ref/libdropin_match.so not built (needs the reference headers at build time)")
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    grid = frame_grid(W, H)
    D, R = _oracle.RefMatcher(0.9, True, lib="libdropin_match.so"), _oracle.RefMatcher(0.9, True)
    hits, misses = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    D.L.ccm_dropin_pack_cache_clear(ctypes.byref(hits), ctypes.byref(misses))
    frames = []
    for seed in (31, 32):
        rng = np.random.default_rng(seed)
        a, b, src = _frames(seed)
        valid, u, v, level, obs = _projection_inputs(a, b, src, rng)
        valid &= ((u >= 0) & (u <= W) & (v >= 0) & (v <= H)).astype(np.uint8)
        frames.append((a, D.frame(b, grid), R.frame(b, grid), valid, u, v, level, obs, len(b["x"])))
    for rep in range(3):                                   # the same two frames again: hits from the second round on
        for a, fd, fr, valid, u, v, level, obs, nb in frames:
            for th in (7.0, 14.0):
                hd, hr = np.full(nb, -1, np.int32), np.full(nb, -1, np.int32)
                nd = D.projection_lastframe(fd, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hd)
                nr = R.projection_lastframe(fr, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hr)
                assert nd == nr > 200 and np.array_equal(hd, hr), (rep, th)
            viewcos = np.full(len(u), 0.9, np.float32)
            hd, hr = np.full(nb, -1, np.int32), np.full(nb, -1, np.int32)
            D.nnratio = R.nnratio = float(np.float32(0.8))
            nd = D.projection_mappoints(fd, valid, u, v, level, viewcos, a["desc"], obs, 1.0, SF, hd)
            nr = R.projection_mappoints(fr, valid, u, v, level, viewcos, a["desc"], obs, 1.0, SF, hr)
            assert nd == nr and np.array_equal(hd, hr)
    D.L.ccm_dropin_pack_cache_clear(ctypes.byref(hits), ctypes.byref(misses))
    assert misses.value == 2 and hits.value == 3 * 2 * 3 - 2, (hits.value, misses.value)


def test_dropin_cpp_fuse_and_sim3_match_reference_class(cuda_lib):
    """Drop-in Fuse (both overloads) and SearchBySim3 — candidate search on the device (CCM_WIN_BEST), map mutations
    replayed on the host — against the reference's own bodies (ORBmatcher.cpp:854-1348) through the same glue: the same
    Replace / AddObservation / AddMapPoint / vpReplacePoint / vpMatches12 state, including map points listed twice,
    occupied slots, bad and mbDoNotReplace points."""
    import _oracle
    import _fuse_cases as fc
    if not (_oracle.dropin_match_available() and _oracle.ref_match_available()):
        pytest.skip("oracle/_ref/libdropin_match.so not built (needs the reference headers at build time)")
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    grid = frame_grid(W, H)
    for seed, th in [(301, 3.0), (302, 8.0), (303, 2.5)]:
        c = fc.fuse_case(seed)
        D, R = _oracle.RefMatcher(0.6, True, lib="libdropin_match.so"), _oracle.RefMatcher(0.6, True)
        d, r = D.fuse(D.frame(c["kf"], grid), c, th, fc.SF, fc.INV_SIGMA2), R.fuse(R.frame(c["kf"], grid), c, th, fc.SF, fc.INV_SIGMA2)
        assert d["n"] == r["n"] > 150
        for k in ("replaced_by", "added_idx", "q_bad", "slot_id", "slot_bad", "occ_replaced_by"):
            assert np.array_equal(d[k], r[k]), k
    for seed, th in [(311, 4.0), (312, 10.0)]:
        c = fc.fuse_case(seed)
        D, R = _oracle.RefMatcher(0.8, True, lib="libdropin_match.so"), _oracle.RefMatcher(0.8, True)
        d, r = D.fuse_sim3(D.frame(c["kf"], grid), c, th, fc.SF), R.fuse_sim3(R.frame(c["kf"], grid), c, th, fc.SF)
        assert d["n"] == r["n"] > 150
        for k in ("replace_point", "added_idx", "slot_id", "slot_bad"):
            assert np.array_equal(d[k], r[k]), k
    for seed, th in [(321, 7.5), (322, 15.0)]:
        c = fc.sim3_case(seed)
        D, R = _oracle.RefMatcher(0.75, True, lib="libdropin_match.so"), _oracle.RefMatcher(0.75, True)
        nd, md = D.search_by_sim3(D.frame(c["kf1"], grid), D.frame(c["kf2"], grid), c, th, fc.SF)
        nr, mr = R.search_by_sim3(R.frame(c["kf1"], grid), R.frame(c["kf2"], grid), c, th, fc.SF)
        assert nd == nr > 150 and np.array_equal(md, mr)


@pytest.mark.parametrize("seed,ori,nodes", [(331, False, 120), (332, True, 120), (333, False, 12), (334, False, 1000), (335, True, 1)])
def test_search_for_triangulation(gm, seed, ori, nodes):
    """ccm_search_triangulation vs the oracle (itself identical to the reference's SearchForTriangulation on these
    seeds, tests/test_match_oracle_vs_ref.py): epipole gate, epipolar-line gate with the double-precision threshold
    folded to a float bound, last-wins ties, features without node, rotation histogram."""
    import _fuse_cases as fc
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    c = fc.tri_case(seed, n_nodes=nodes)
    rng = np.random.default_rng(seed)
    c["node1"][rng.random(len(c["node1"])) < 0.05] = -1
    c["node2"][rng.random(len(c["node2"])) < 0.05] = -1
    a, b = c["kf1"], c["kf2"]
    nm_g, m_g = ORBmatcher(gm, 0.6, ori).SearchForTriangulation(
        a["desc"], c["node1"], c["free1"], a["x"], a["y"], a["angle"], b["desc"], c["node2"], c["free2"], b["x"], b["y"],
        b["octave"], b["angle"], c["f12"], c["ex"], c["ey"], fc.SF, fc.LEVEL_SIGMA2)
    nm_c, m_c = OracleMatcher(0.6, ori).search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, ori)
    assert nm_g == nm_c and np.array_equal(m_g, m_c)
    assert nm_c > (20 if nodes == 1000 else 100)


def test_search_for_triangulation_edges(gm):
    import _fuse_cases as fc
    from motioncheck_ccm_slam_b200.matcher import ORBmatcher
    c = fc.tri_case(340, n=64, n_nodes=4)
    a, b = c["kf1"], c["kf2"]
    g = ORBmatcher(gm, 0.6, False)

    def run(cc):
        aa, bb = cc["kf1"], cc["kf2"]
        got = g.SearchForTriangulation(aa["desc"], cc["node1"], cc["free1"], aa["x"], aa["y"], aa["angle"], bb["desc"], cc["node2"],
                                       cc["free2"], bb["x"], bb["y"], bb["octave"], bb["angle"], cc["f12"], cc["ex"], cc["ey"],
                                       fc.SF, fc.LEVEL_SIGMA2)
        want = OracleMatcher(0.6, False).search_for_triangulation(cc, fc.SF, fc.LEVEL_SIGMA2, False)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
        return got[0]

    # degenerate F12 (den == 0 for every feature, :170-171): nothing matches
    assert run(dict(c, f12=np.zeros(9, np.float32))) == 0
    # no free feature on one side
    assert run(dict(c, free2=np.zeros_like(c["free2"]))) == 0
    # disjoint vocabularies
    assert run(dict(c, node2=c["node2"] + 100)) == 0
    # empty keyframe
    empty = {k: v[:0] for k, v in b.items()}
    assert run(dict(c, kf2=empty, node2=c["node2"][:0], free2=c["free2"][:0])) == 0
    assert run(c) > 5


def test_dropin_cpp_triangulation_matches_reference_class(cuda_lib):
    import _oracle
    import _fuse_cases as fc
    if not (_oracle.dropin_match_available() and _oracle.ref_match_available()):
        pytest.skip("oracle/_ref/libdropin_match.so not built (needs the reference headers at build time)")
    for seed, ori, nodes in [(331, False, 120), (332, True, 120), (333, False, 12)]:
        c = fc.tri_case(seed, n_nodes=nodes)
        nd, md = _oracle.RefMatcher(0.6, ori, lib="libdropin_match.so").search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, ori)
        nr, mr = _oracle.RefMatcher(0.6, ori).search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, ori)
        assert nd == nr > 100 and np.array_equal(md, mr)


@pytest.mark.parametrize("seed,k,L,ragged,levelsup", [(1, 10, 4, 0.0, 2), (2, 10, 3, 0.15, 4), (3, 6, 5, 0.1, 4), (4, 10, 4, 0.0, 0),
                                                      (5, 3, 6, 0.2, 4), (6, 10, 2, 0.0, 1), (7, 20, 3, 0.0, 1), (8, 40, 2, 0.3, 1)])
def test_vocabulary_transform(gm, seed, k, L, ragged, levelsup):
    """ccm_vocabulary_transform vs the oracle (identical to the vendored DBoW2 on seeds 1-6, tests/test_vocab_oracle_vs_ref.py):
    word and node ids per feature, then the BowVector / FeatureVector Frame::ComputeBoW fills, bit for bit."""
    import _oracle
    from motioncheck_ccm_slam_b200.matcher import Vocabulary
    from motioncheck_ccm_slam_b200.synth import make_vocabulary, perturb_descriptors
    voc = make_vocabulary(seed, k, L, ragged)
    rng = np.random.default_rng(seed + 100)
    leaves = np.nonzero(voc["is_leaf"])[0]
    desc = perturb_descriptors(rng, voc["desc"][rng.choice(leaves, 1500)], 0.04)
    desc[:150] = rng.integers(0, 256, (150, 32), dtype=np.uint8)
    V = Vocabulary(gm, voc)
    w_g, n_g = V.transform_features(desc, levelsup)
    w_o, wt_o, n_o = _oracle.oracle_vocab_transform(voc, desc, levelsup)
    assert np.array_equal(w_g, w_o) and np.array_equal(n_g, n_o)
    bw_g, bv_g, fn_g, fo_g = V.ComputeBoW(desc, levelsup)
    bw_o, bv_o, fn_o, fo_o = _oracle.assemble_bow(w_o, wt_o, n_o)
    assert np.array_equal(bw_g, bw_o) and np.array_equal(bv_g.view(np.uint64), bv_o.view(np.uint64))
    assert np.array_equal(fn_g, fn_o) and np.array_equal(fo_g, fo_o)
    assert len(V.transform_features(desc[:0], levelsup)[0]) == 0
    V.close()


def test_vocabulary_transform_vs_reference_dbow2(gm, tmp_path):
    """Straight against the reference's own DBoW2 (compiled unmodified, oracle/_ref/libref_vocab.so) where it was built."""
    import _oracle
    if not _oracle.ref_vocab_available():
        pytest.skip("oracle/_ref/libref_vocab.so not built")
    from motioncheck_ccm_slam_b200.matcher import Vocabulary
    from motioncheck_ccm_slam_b200.synth import make_vocabulary, perturb_descriptors, write_vocabulary_text
    voc = make_vocabulary(21, 10, 4)
    path = str(tmp_path / "voc.txt")
    write_vocabulary_text(voc, path)
    R = _oracle.RefVocabulary(path)
    rng = np.random.default_rng(5)
    desc = perturb_descriptors(rng, voc["desc"][rng.choice(np.nonzero(voc["is_leaf"])[0], 2000)], 0.05)
    V = Vocabulary(gm, voc)
    bw_g, bv_g, fn_g, fo_g = V.ComputeBoW(desc, 2)
    bw_r, bv_r, fn_r, fo_r = R.transform(desc, 2)
    assert np.array_equal(bw_g, bw_r) and np.array_equal(bv_g.view(np.uint64), bv_r.view(np.uint64))
    assert np.array_equal(fn_g, fn_r) and np.array_equal(fo_g, fo_r) and len(bw_r) > 500
    V.close()


@pytest.mark.parametrize("seed,n_points,max_obs", [(5, 400, 40), (6, 50, 300), (7, 3000, 12)])
def test_distinctive_descriptors(gm, seed, n_points, max_obs):
    from _oracle import distinctive_case, oracle_distinctive_descriptors
    from motioncheck_ccm_slam_b200.matcher import distinctive_descriptors
    start, desc = distinctive_case(seed, n_points, max_obs)
    bg, mg = distinctive_descriptors(gm, start, desc)
    bo, mo = oracle_distinctive_descriptors(start, desc)
    assert np.array_equal(bg, bo) and np.array_equal(mg, mo)
    assert (bo == -1).any() and (bo > 0).sum() > n_points // 4
    b0, _ = distinctive_descriptors(gm, np.zeros(1, np.int32), np.zeros((0, 32), np.uint8))
    assert len(b0) == 0


# ---- Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cpp:277-343) ----------------------------------------------------
_UND_CAMERAS = ("euroc", "k3", "barrel", "strong")


def _und_golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "undistort.npz"))


def _camera(U, cam):
    from motioncheck_ccm_slam_b200.matcher import make_camera
    K = U[cam + "_K"]
    return make_camera(K[0, 0], K[1, 1], K[0, 2], K[1, 2], list(U[cam + "_D"]))


@pytest.mark.parametrize("cam", _UND_CAMERAS)
def test_undistort_keypoints(gm, cam):
    """cv::undistortPoints on the device: bit-identical to the cv2 4.13.0 golden vectors and, on 200k random points, to the
    oracle restatement (doubles, no FMA contraction, five iterations, the icdist < 0 bail-out of the 'strong' camera)."""
    from _oracle import oracle_image_bounds, oracle_undistort
    U = _und_golden()
    c = _camera(U, cam)
    got = gm.UndistortKeyPoints(c, U[cam + "_pts"])
    assert np.array_equal(got.view(np.uint32), U[cam + "_und"].view(np.uint32))
    w, h = U[cam + "_wh"]
    rng = np.random.default_rng(5)
    pts = np.concatenate([rng.uniform(-80, [w + 80, h + 80], (200000, 2)), rng.integers(0, [w, h], (50000, 2))]).astype(np.float32)
    assert np.array_equal(gm.UndistortKeyPoints(c, pts).view(np.uint32), oracle_undistort(pts, U[cam + "_K"], U[cam + "_D"]).view(np.uint32))
    assert gm.UndistortKeyPoints(c, np.zeros((0, 2), np.float32)).shape == (0, 2)
    # ComputeImageBounds + grid constants
    g, b = gm.ComputeImageBounds(c, w, h)
    ob = oracle_image_bounds(U[cam + "_K"], U[cam + "_D"], w, h)
    cu = U[cam + "_corners_und"]
    assert np.array_equal(b, np.array([min(cu[0, 0], cu[2, 0]), max(cu[1, 0], cu[3, 0]), min(cu[0, 1], cu[1, 1]), max(cu[2, 1], cu[3, 1])], np.float32))
    assert np.array_equal(np.array([g.min_x, b[1], g.min_y, b[3], g.inv_w, g.inv_h], np.float32), ob)


def test_undistort_zero_k1_is_a_copy(gm):
    """Frame.cpp:279-283 / :311: only k1 is tested; with k1 == 0 the keypoints are copied and the bounds are the image."""
    from motioncheck_ccm_slam_b200.matcher import make_camera
    c = make_camera(458.654, 457.296, 367.215, 248.375, [0.0, 0.07, 0.001, 0.002])
    pts = np.random.default_rng(1).uniform(0, 700, (1000, 2)).astype(np.float32)
    assert np.array_equal(gm.UndistortKeyPoints(c, pts), pts)
    g, b = gm.ComputeImageBounds(c, 752, 480)
    assert np.array_equal(b, np.array([0, 752, 0, 480], np.float32))
    assert g.inv_w == np.float32(75) / np.float32(752) and g.inv_h == np.float32(48) / np.float32(480)
    with pytest.raises(Exception):
        gm.UndistortKeyPoints(make_camera(0.0, 1.0, 0.0, 0.0, [0.1, 0, 0, 0]), pts)


def test_extractor_to_packs_with_distorted_camera(gm):
    """ExtractORB -> UndistortKeyPoints -> AssignFeaturesToGrid without leaving HBM: the undistortion fused into the pack
    fill and the stand-alone batched kernel both give mvKeysUn, and the grid built from them, of the oracle."""
    torch = pytest.importorskip("torch")
    from _oracle import oracle_undistort
    from motioncheck_ccm_slam_b200 import ORBextractor
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    from motioncheck_ccm_slam_b200.matcher import FramePack, packs_from_extractor_batch_device
    from motioncheck_ccm_slam_b200.synth import make_stream
    U = _und_golden()
    cam = _camera(U, "euroc")
    B, w, h = 3, 752, 480
    frames = make_stream(777, B, w, h)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B, device=0)
    cap = ex.max_keypoints()
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(frames).to(dev)
    d_kp = torch.zeros((B, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros((B,), dtype=torch.int32, device=dev)
    d_un = torch.full((B, cap, 2), -1.0, dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    ex.extract_batch_device(d_img.data_ptr(), B, h, w, w, w * h, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    ex.sync()
    kp = d_kp.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
    desc, n = d_desc.cpu().numpy(), d_n.cpu().numpy()
    grid, bounds = gm.ComputeImageBounds(cam, w, h)
    assert bounds[0] < 0 and bounds[1] > w                       # barrel distortion: the undistorted image is larger
    gm.undistort_keypoints_batch_device(cam, B, d_kp.data_ptr(), d_n.data_ptr(), cap, d_un.data_ptr())
    fused = [FramePack(gm, cap) for _ in range(B)]
    staged = [FramePack(gm, cap) for _ in range(B)]
    packs_from_extractor_batch_device(fused, grid, d_kp.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap, cam=cam)
    packs_from_extractor_batch_device(staged, grid, d_kp.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap, d_xy_un=d_un.data_ptr())
    gm.sync()
    un = d_un.cpu().numpy()
    for j in range(B):
        k = kp[j, :n[j]]
        want = oracle_undistort(np.stack([k["x"], k["y"]], 1), U["euroc_K"], U["euroc_D"])
        assert np.array_equal(un[j, :n[j]].view(np.uint32), want.view(np.uint32))
        assert (un[j, n[j]:] == -1).all()                                           # rows past the count are not touched
        of = OracleFrame(desc[j, :n[j]], want[:, 0], want[:, 1], k["octave"], k["angle"], grid)
        ocs, oitems = of.csr()
        for p in (fused[j], staged[j]):
            cs, items = p.get_grid()
            assert np.array_equal(cs, ocs) and np.array_equal(items, oitems)
            assert p.count() == n[j]
    ex.close()


# ---- keyframe wire format (KeyFrame.cpp:1462-1472, 1689-1702; Converter.cc:166-193) --------------------------------------
@pytest.mark.parametrize("seed,n", [(1, 1000), (2, 1), (3, 0), (4, 257), (5, 256), (6, 5000)])
def test_wire_format_encode_decode(gm, seed, n):
    """Client-side encode and server-side decode-into-a-pack on the device against the oracle restatement: byte strings
    identical, decoded keypoints (lossy u8 size / response, i8 octave) identical, the pack's grid that of the decoded frame."""
    from _oracle import oracle_wire_decode, oracle_wire_encode
    from test_match_oracle import _wire_keypoints
    from motioncheck_ccm_slam_b200.matcher import FramePack, frame_grid
    kp, desc = _wire_keypoints(seed, n)
    kb, db = gm.wire_encode_keyframe(kp, desc)
    okb, odb = oracle_wire_encode(kp, desc)
    assert kb == okb and db == odb
    grid = frame_grid(W, H, min_x=-20.0, min_y=-20.0)
    p = FramePack(gm, max(n, 1) + 3)
    got = p.from_wire(grid, kb, db)
    back, d2 = oracle_wire_decode(okb, odb)
    for k in ("x", "y", "size", "angle", "response", "octave"):
        assert np.array_equal(got[k], back[k]), k
    assert p.count() == n
    of = OracleFrame(d2, back["x"], back["y"], back["octave"], back["angle"], grid)
    cs, items = p.get_grid()
    ocs, oitems = of.csr()
    assert np.array_equal(cs, ocs) and np.array_equal(items, oitems)
    if n >= 256:    # the descriptors landed in the pack: a query equal to row 7 finds row 7 at distance 0
        from motioncheck_ccm_slam_b200.matcher import ORBmatcher
        _, res = ORBmatcher(gm).SearchBestInWindow(p, np.ones(1, np.uint8), back["x"][7:8], back["y"][7:8], np.full(1, 3.0, np.float32),
                                                back["octave"][7:8].astype(np.int32), d2[7:8], 256)
        assert res["best_dist"][0] == 0
    with pytest.raises(Exception):
        p.from_wire(grid, kb, db + b"\0")                                  # length does not match the count
    p.close()


def test_wire_format_mvkeysun_override(gm):
    """ConvertToMessage sends mvKeysUn: the encoder takes the undistorted positions from the d_xy_un array."""
    from _oracle import oracle_wire_encode
    from test_match_oracle import _wire_keypoints
    kp, desc = _wire_keypoints(9, 777)
    xy = np.random.default_rng(9).uniform(0, 700, (777, 2)).astype(np.float32)
    kb, db = gm.wire_encode_keyframe(kp, desc, xy_un=xy)
    kp2 = kp.copy(); kp2["x"] = xy[:, 0]; kp2["y"] = xy[:, 1]
    assert (kb, db) == oracle_wire_encode(kp2, desc)


# ---- fused fold + exchange + merge over peer memory (ccm_exchange_*) --------------------------------------------------
def test_allpairs_exchange_world1_equals_merge(gm):
    """One rank: the fused kernel publishes into its own buffer and merges: identical to search + merge, over several calls
    (epoch parity), with fewer queries than the context's capacity and an empty shard."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE, Exchange
    rng = np.random.default_rng(21)
    x = Exchange(gm, 0, 1, 1000)
    for call, (nq, nt) in enumerate([(1000, 50000), (1000, 50000), (300, 7000), (1, 1), (513, 0)]):
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        t = rng.integers(0, 256, (max(nt, 1), 32), dtype=np.uint8)
        if nt:
            t[rng.integers(0, nt, min(nq, nt) // 2)] = q[: min(nq, nt) // 2]          # exact duplicates: distance 0, index ties
        d_q, d_t = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
        d_out = torch.zeros((nq, 8), dtype=torch.uint8, device="cuda")
        d_acc = torch.zeros(nq, dtype=torch.uint8, device="cuda")
        d_ref = torch.zeros((nq, 8), dtype=torch.uint8, device="cuda")
        d_ref2 = torch.zeros((nq, 8), dtype=torch.uint8, device="cuda")
        d_acc2 = torch.zeros(nq, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        x.allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), nt, 1000 * call, 0.6, d_out.data_ptr(), d_acc.data_ptr())
        gm.match_allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), nt, 1000 * call, d_ref.data_ptr())
        gm.match_allpairs_merge_device(d_ref.data_ptr(), 1, nq, 0.6, d_ref2.data_ptr(), d_acc2.data_ptr())
        gm.sync()
        assert np.array_equal(d_out.cpu().numpy().view(BEST2_DTYPE), d_ref2.cpu().numpy().view(BEST2_DTYPE)), call
        assert np.array_equal(d_acc.cpu().numpy(), d_acc2.cpu().numpy())
    x.close()


def test_allpairs_exchange_missing_peer_times_out(cuda_lib, monkeypatch):
    """A rank whose peer never launches must not spin forever (the kernel cannot be cancelled from the host): the wait is
    bounded, the call's results are "no match", and ccm_exchange_check reports CCM_E_CUDA."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200._lib import CCM_E_CUDA, CcmError
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE, Exchange, Matcher
    monkeypatch.setenv("CCM_EXCHANGE_TIMEOUT_MS", "50")
    ms = [Matcher(0), Matcher(0)]                       # two ranks' contexts on one device: rank 1 never calls
    xs = [Exchange(ms[r], r, 2, 512) for r in range(2)]
    bufs = [x.buffer() for x in xs]
    for x in xs:
        x.connect_ptrs(bufs)
    rng = np.random.default_rng(1)
    nq, nt = 300, 5000
    d_q = torch.from_numpy(rng.integers(0, 256, (nq, 32), dtype=np.uint8)).cuda()
    d_t = torch.from_numpy(rng.integers(0, 256, (nt, 32), dtype=np.uint8)).cuda()
    d_out = torch.zeros((nq, 8), dtype=torch.uint8, device="cuda")
    d_acc = torch.ones(nq, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    xs[0].check()                                       # nothing happened yet
    xs[0].allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), nt, 0, 0.6, d_out.data_ptr(), d_acc.data_ptr())
    with pytest.raises(CcmError) as e:
        xs[0].check()
    assert e.value.code == CCM_E_CUDA and "peer rank did not publish" in str(e.value)
    out = d_out.cpu().numpy().view(BEST2_DTYPE).reshape(-1)
    assert np.all(out["idx"] == 0xFFFFFFFF) and np.all(out["d1"] == 256) and not d_acc.cpu().numpy().any()
    for x in xs:
        x.close()
    for m in ms:
        m.close()


def test_allpairs_exchange_two_gpus(cuda_lib):
    """Two ranks on two devices of one process (peer access): every rank ends with the result of the unsharded search."""
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE, Exchange, Matcher
    rng = np.random.default_rng(22)
    nq, nt = 1000, 300000
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    t[rng.integers(0, nt, 400)] = q[:400]
    ms = [Matcher(0), Matcher(1)]
    xs = [Exchange(ms[r], r, 2, nq) for r in range(2)]
    bufs = [x.buffer() for x in xs]
    for x in xs:
        x.connect_ptrs(bufs)
    want = ms[0].match_allpairs_best2(q, t)
    half = nt // 2 + 37
    outs = []
    ins = []
    for r in range(2):
        dev = torch.device("cuda", r)
        lo, hi = (0, half) if r == 0 else (half, nt)
        ins.append((torch.from_numpy(q).to(dev), torch.from_numpy(t[lo:hi]).to(dev), torch.zeros((nq, 8), dtype=torch.uint8, device=dev), lo, hi))
        torch.cuda.synchronize(dev)
    for rep in range(3):
        outs = []
        for r in range(2):                                   # both launches are asynchronous: rank 0 waits on the device for rank 1
            d_q, d_t, d_out, lo, hi = ins[r]
            xs[r].allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), hi - lo, lo, 0.6, d_out.data_ptr())
            outs.append((d_q, d_t, d_out))
        for r in range(2):
            ms[r].sync()
        for r in range(2):
            got = outs[r][2].cpu().numpy().view(BEST2_DTYPE).reshape(-1)
            for f in ("idx", "d1", "d2"):
                assert np.array_equal(got[f], want[f]), (rep, r, f)
    for x in xs:
        x.close()
    for m in ms:
        m.close()


def test_children_outlive_their_matcher(cuda_lib):
    """Frame packs and vocabularies may be destroyed after the matcher they were created on (a Python interpreter tears objects
    down in no particular order; a C++ host may hold packs in Frame objects that outlive the matcher singleton)."""
    from motioncheck_ccm_slam_b200.matcher import FramePack, Matcher, Vocabulary
    from motioncheck_ccm_slam_b200.synth import make_vocabulary_complete
    m = Matcher(0)
    p = FramePack(m, 100)
    v = Vocabulary(m, make_vocabulary_complete(3, 4, 3))
    m.close()
    p.close()
    v.close()
