"""CPU: the matcher oracle (oracle/match_oracle.cpp) against the reference's OWN cslam/src/ORBmatcher.cpp, compiled
unmodified over mock Frame / KeyFrame / MapPoint headers (oracle/_ref/libref_match.so, built by oracle/Makefile where
/root/reference exists; the prebuilt library travels with the repo). Same seeded inputs as tests/test_gpu_match.py, so
GPU == oracle == reference by transitivity. What the reference leaves behind is object state (F.mvpMapPoints,
vpMatched, vnMatches12, ...): compared as holder / match arrays and return counts, bit-exact."""
import numpy as np
import pytest

import _oracle
from _oracle import OracleFrame, OracleMatcher
from motioncheck_ccm_slam_b200.matcher import frame_grid
from motioncheck_ccm_slam_b200.synth import bow_node_ids, make_feature_frame, make_next_frame, perturb_descriptors

pytestmark = pytest.mark.skipif(not _oracle.ref_match_available(), reason="oracle/_ref/libref_match.so not built")

W, H = 752, 480
SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def _frames(seed, n=1000, **kw):
    a = make_feature_frame(seed, n, W, H)
    b, src = make_next_frame(seed + 1, a, W, H, **kw)
    return a, b, src


def _inputs(a, b, src, rng):
    n = len(a["x"])
    inv = np.full(n, -1)
    inv[src[src >= 0]] = np.nonzero(src >= 0)[0]
    u = np.where(inv >= 0, b["x"][np.maximum(inv, 0)], a["x"] + 2).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
    v = np.where(inv >= 0, b["y"][np.maximum(inv, 0)], a["y"] + 1).astype(np.float32) + rng.normal(0, 1.5, n).astype(np.float32)
    level = np.where(inv >= 0, b["octave"][np.maximum(inv, 0)], a["octave"]).astype(np.int32)
    level = np.clip(level + rng.choice([0, 0, 0, 1], n), 0, 7).astype(np.int32)
    valid = (rng.random(n) < 0.9).astype(np.uint8)
    # projections outside the image bounds are host-side `continue`s of the reference (ORBmatcher.cpp:1393-1396): not valid
    valid &= ((u >= 0) & (u <= W) & (v >= 0) & (v <= H)).astype(np.uint8)
    obs = (rng.random(n) < 0.8).astype(np.uint8)
    return valid, u, v, level, obs


def _both(b):
    grid = frame_grid(W, H)
    return OracleFrame(b["desc"], b["x"], b["y"], b["octave"], b["angle"], grid), grid


def test_descriptor_distance():
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8); b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    assert np.array_equal(_oracle.RefMatcher().descriptor_distance(a, b), OracleMatcher().descriptor_distance(a, b))


@pytest.mark.parametrize("seed,th,nnratio", [(21, 1.0, 0.8), (22, 5.0, 0.8), (23, 3.0, 0.6), (24, 12.0, 0.9)])
def test_projection_mappoints(seed, th, nnratio):
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed)
    valid, u, v, level, obs = _inputs(a, b, src, rng)
    viewcos = rng.choice([0.9, 0.9985, 0.998, 0.99801], len(u)).astype(np.float32)
    holder0 = rng.choice([-1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
    of, grid = _both(b)
    R = _oracle.RefMatcher(nnratio, True)
    rf = R.frame(b, grid)
    hc, hr = holder0.copy(), holder0.copy()
    nm_c, _ = OracleMatcher(nnratio, True).projection_mappoints(of, valid, u, v, level, viewcos, a["desc"], obs, th, SF, hc)
    nm_r = R.projection_mappoints(rf, valid, u, v, level, viewcos, a["desc"], obs, th, SF, hr)
    assert nm_c == nm_r > 200 and np.array_equal(hc, hr)


@pytest.mark.parametrize("seed,th,ori", [(31, 7.0, True), (32, 14.0, True), (33, 7.0, False), (34, 30.0, True)])
def test_projection_lastframe(seed, th, ori):
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed)
    valid, u, v, _, obs = _inputs(a, b, src, rng)
    holder0 = rng.choice([-1, -1, -1, -1, -2, -3], len(b["x"])).astype(np.int32)
    of, grid = _both(b)
    R = _oracle.RefMatcher(0.9, ori)
    rf = R.frame(b, grid)
    hc, hr = holder0.copy(), holder0.copy()
    nm_c, _ = OracleMatcher(0.9, ori).projection_lastframe(of, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hc)
    nm_r = R.projection_lastframe(rf, valid, u, v, a["octave"], a["angle"], a["desc"], obs, th, SF, hr)
    assert nm_c == nm_r > 200 and np.array_equal(hc, hr)


@pytest.mark.parametrize("seed,th", [(41, 10), (42, 4)])
def test_projection_kf_sim3(seed, th):
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed, n=3000)
    valid, u, v, level, claims = _inputs(a, b, src, rng)
    holder0 = rng.choice([-1, -1, -1, -2], len(b["x"])).astype(np.int32)
    of, grid = _both(b)
    R = _oracle.RefMatcher(0.75, True)
    rf = R.frame(b, grid)
    hc, hr = holder0.copy(), holder0.copy()
    nm_c, _ = OracleMatcher(0.75, True).projection_kf_sim3(of, valid, u, v, level, a["desc"], claims, th, SF, hc)
    nm_r = R.projection_kf_sim3(rf, valid, u, v, level, a["desc"], claims, th, SF, hr)
    assert nm_c == nm_r > 100 and np.array_equal(hc, hr)


def test_greedy_chain_conflicts():
    rng = np.random.default_rng(61)
    nt, nq = 40, 300
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    b = dict(desc=perturb_descriptors(rng, np.repeat(base, nt, 0), 0.05), x=rng.uniform(100, 130, nt).astype(np.float32),
             y=rng.uniform(100, 130, nt).astype(np.float32), octave=np.zeros(nt, np.int32), angle=rng.uniform(0, 360, nt).astype(np.float32))
    qd = perturb_descriptors(rng, np.repeat(base, nq, 0), 0.02)
    u = np.full(nq, 115, np.float32); v = np.full(nq, 115, np.float32)
    valid = np.ones(nq, np.uint8); obs = (rng.random(nq) < 0.7).astype(np.uint8)
    of, grid = _both(b)
    z = np.zeros(nq)
    R = _oracle.RefMatcher(0.9, False); rf = R.frame(b, grid)
    hc, hr = np.full(nt, -1, np.int32), np.full(nt, -1, np.int32)
    nm_c, _ = OracleMatcher(0.9, False).projection_lastframe(of, valid, u, v, z, z, qd, obs, 30.0, SF, hc)
    assert nm_c == R.projection_lastframe(rf, valid, u, v, z, z, qd, obs, 30.0, SF, hr) >= 20 and np.array_equal(hc, hr)
    R = _oracle.RefMatcher(0.95, False); rf = R.frame(b, grid)
    hc, hr = np.full(nt, -1, np.int32), np.full(nt, -1, np.int32)
    nm_c, _ = OracleMatcher(0.95, False).projection_mappoints(of, valid, u, v, z, np.ones(nq), qd, obs, 8.0, SF, hc)
    assert nm_c == R.projection_mappoints(rf, valid, u, v, z, np.ones(nq), qd, obs, 8.0, SF, hr) >= 20 and np.array_equal(hc, hr)
    R = _oracle.RefMatcher(); rf = R.frame(b, grid)
    hc, hr = np.full(nt, -1, np.int32), np.full(nt, -1, np.int32)
    nm_c, _ = OracleMatcher().projection_kf_sim3(of, valid, u, v, z, qd, obs, 30, SF, hc)
    assert nm_c == R.projection_kf_sim3(rf, valid, u, v, z, qd, obs, 30, SF, hr) >= 20 and np.array_equal(hc, hr)


@pytest.mark.parametrize("seed,window,ori", [(71, 100, True), (72, 30, True), (73, 100, False)])
def test_search_for_initialization(seed, window, ori):
    a, b, src = _frames(seed, n=2000, motion=(6.0, -3.0))
    of, grid = _both(b)
    R = _oracle.RefMatcher(0.9, ori); rf = R.frame(b, grid)
    pc = np.stack([a["x"], a["y"]], 1).astype(np.float32).copy(); pr = pc.copy()
    nm_c, m_c = OracleMatcher(0.9, ori).search_for_initialization(a["desc"], a["octave"], a["angle"], pc, of, window)
    nm_r, m_r = R.search_for_initialization(a["desc"], a["octave"], a["angle"], pr, rf, window)
    assert nm_c == nm_r > 50 and np.array_equal(m_c, m_r) and np.array_equal(pc, pr)


@pytest.mark.parametrize("seed,nnratio,ori,n_nodes", [(81, 0.7, True, 1000), (82, 0.75, False, 100), (83, 0.9, True, 7)])
def test_search_by_bow(seed, nnratio, ori, n_nodes):
    rng = np.random.default_rng(seed)
    a, b, src = _frames(seed, flip_prob=0.03)
    na, nb = bow_node_ids(a["desc"], n_nodes), bow_node_ids(b["desc"], n_nodes)
    na[rng.random(len(na)) < 0.05] = -1; nb[rng.random(len(nb)) < 0.05] = -1
    va = (rng.random(len(na)) < 0.8).astype(np.uint8); vb = (rng.random(len(nb)) < 0.8).astype(np.uint8)
    c, R = OracleMatcher(nnratio, ori), _oracle.RefMatcher(nnratio, ori)
    nm_c, mf_c, _ = c.bow_kf_frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
    nm_r, mf_r = R.bow_kf_frame(a["desc"], na, va, a["angle"], b["desc"], nb, b["angle"])
    assert nm_c == nm_r > 20 and np.array_equal(mf_c, mf_r)
    nm_c, m_c, _ = c.bow_kf_kf(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
    nm_r, m_r = R.bow_kf_kf(a["desc"], na, va, a["angle"], b["desc"], nb, vb, b["angle"])
    assert nm_c == nm_r > 10 and np.array_equal(m_c, m_r)


# ---- Fuse / SearchBySim3: the reference methods against the oracle's best-in-window search + the host replay ----
import _fuse_cases as fc


def _best(of, nnratio, valid, c_u, c_v, c_level, c_desc, th, th_dist, inv_sigma2=None):
    radius = (np.float32(th) * fc.SF[c_level]).astype(np.float32)
    _, res = OracleMatcher(nnratio, True).best_in_window(of, valid.astype(np.uint8), c_u, c_v, radius, c_level, c_desc, th_dist, inv_sigma2)
    return res["accepted"], res["best_idx"]


@pytest.mark.parametrize("seed,th", [(301, 3.0), (302, 8.0), (303, 2.5)])
def test_fuse(seed, th):
    c = fc.fuse_case(seed)
    of, grid = _both(c["kf"])
    R = _oracle.RefMatcher(0.6, True)
    ref = R.fuse(R.frame(c["kf"], grid), c, th, fc.SF, fc.INV_SIGMA2)
    rows, searched = fc.effective_rows(c), fc.searched_fuse(c, first=True)
    acc, best = _best(of, 0.6, searched, c["q_u"][rows], c["q_v"][rows], c["q_level"][rows], c["q_desc"][rows], th, 50, fc.INV_SIGMA2)
    mine = fc.replay_fuse(c, searched, acc, best)
    assert mine["n"] == ref["n"] > 150
    for k in ("replaced_by", "added_idx", "q_bad", "slot_id", "slot_bad", "occ_replaced_by"):
        assert np.array_equal(mine[k], ref[k]), k
    assert (ref["replaced_by"] >= 0).sum() > 10 and (ref["occ_replaced_by"] >= 0).sum() > 10 and (ref["added_idx"] >= 0).sum() > 50


@pytest.mark.parametrize("seed,th", [(311, 4.0), (312, 10.0)])
def test_fuse_sim3(seed, th):
    c = fc.fuse_case(seed)
    of, grid = _both(c["kf"])
    R = _oracle.RefMatcher(0.8, True)
    ref = R.fuse_sim3(R.frame(c["kf"], grid), c, th, fc.SF)
    rows, searched = fc.effective_rows(c), fc.searched_fuse(c, first=False)
    acc, best = _best(of, 0.8, searched, c["q_u"][rows], c["q_v"][rows], c["q_level"][rows], c["q_desc"][rows], th, 50)
    mine = fc.replay_fuse_sim3(c, searched, acc, best)
    assert mine["n"] == ref["n"] > 150
    for k in ("replace_point", "added_idx", "slot_id", "slot_bad"):
        assert np.array_equal(mine[k], ref[k]), k
    assert (ref["replace_point"] >= 0).sum() > 30 and (ref["added_idx"] >= 0).sum() > 50


@pytest.mark.parametrize("seed,th", [(321, 7.5), (322, 15.0)])
def test_search_by_sim3(seed, th):
    c = fc.sim3_case(seed)
    grid = frame_grid(W, H)
    o1 = OracleFrame(*(c["kf1"][k] for k in ("desc", "x", "y", "octave", "angle")), grid)
    o2 = OracleFrame(*(c["kf2"][k] for k in ("desc", "x", "y", "octave", "angle")), grid)
    R = _oracle.RefMatcher(0.75, True)
    n_ref, m_ref = R.search_by_sim3(R.frame(c["kf1"], grid), R.frame(c["kf2"], grid), c, th, fc.SF)
    m1, m2 = fc.sim3_masks(c)
    a1, b1 = _best(o2, 0.75, m1, c["u1"], c["v1"], c["l1"], c["d1"], th, 100)
    a2, b2 = _best(o1, 0.75, m2, c["u2"], c["v2"], c["l2"], c["d2"], th, 100)
    n_mine, m_mine = fc.replay_sim3(c, a1, b1, a2, b2, m1, m2)
    assert n_mine == n_ref > 150 and np.array_equal(m_mine, m_ref)


@pytest.mark.parametrize("seed,ori,nodes", [(331, False, 120), (332, True, 120), (333, False, 12), (334, False, 1000)])
def test_search_for_triangulation(seed, ori, nodes):
    c = fc.tri_case(seed, n_nodes=nodes)
    n_ref, m_ref = _oracle.RefMatcher(0.6, ori).search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, ori)
    n_mine, m_mine = OracleMatcher(0.6, ori).search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2, ori)
    assert n_mine == n_ref and np.array_equal(m_mine, m_ref)
    assert n_ref > (20 if nodes == 1000 else 100), n_ref
    # every gate decides something: opening the epipolar gate, or moving the epipole away, changes the result
    n_line, m_line = OracleMatcher(0.6, ori).search_for_triangulation(c, fc.SF, fc.LEVEL_SIGMA2 * 1e6, ori)
    n_epi, m_epi = OracleMatcher(0.6, ori).search_for_triangulation(dict(c, ex=-1e6, ey=-1e6), fc.SF, fc.LEVEL_SIGMA2, ori)
    assert n_line > n_ref and not np.array_equal(m_epi, m_ref)
