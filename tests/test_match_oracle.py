"""CPU: the matcher oracle (oracle/match_oracle.cpp) against independent numpy restatements of the arithmetic it
rests on — SWAR popcount == bit count, GetFeaturesInArea == brute-force window filter in grid order, all-pairs
best-2 == argmin/partition — and hand-built known-answer cases of the greedy rules of cslam/src/ORBmatcher.cpp.
The oracle is additionally pinned to the reference's own ORBmatcher.cpp compiled unmodified: tests/test_match_oracle_vs_ref.py."""
import numpy as np

from _oracle import OracleFrame, OracleMatcher
from motioncheck_ccm_slam_b200.matcher import frame_grid
from motioncheck_ccm_slam_b200.synth import make_feature_frame, perturb_descriptors

W, H = 752, 480


def test_descriptor_distance_is_bit_count():
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b[0] = a[0]; b[1] = ~a[1]
    d = OracleMatcher().descriptor_distance(a, b)
    assert np.array_equal(d, np.unpackbits(a ^ b, axis=1).sum(1))
    assert d[0] == 0 and d[1] == 256


def test_grid_and_features_in_area_vs_bruteforce():
    fr = make_feature_frame(5, 1500, W, H)
    g = frame_grid(W, H)
    F = OracleFrame(fr["desc"], fr["x"], fr["y"], fr["octave"], fr["angle"], g)
    cs, items = F.csr()
    px = np.round((fr["x"] - np.float32(g.min_x)) * np.float32(g.inv_w)).astype(int)   # Frame.cpp:255-265 (round)
    py = np.round((fr["y"] - np.float32(g.min_y)) * np.float32(g.inv_h)).astype(int)
    ok = (px >= 0) & (px < g.cols) & (py >= 0) & (py < g.rows)
    cell = px * g.rows + py
    order = np.lexsort((np.arange(len(px))[ok], cell[ok]))
    assert np.array_equal(items, np.nonzero(ok)[0][order])
    assert np.array_equal(np.diff(cs), np.bincount(cell[ok], minlength=g.cols * g.rows))
    rng = np.random.default_rng(6)
    for _ in range(200):
        x, y, r = np.float32(rng.uniform(-20, W + 20)), np.float32(rng.uniform(-20, H + 20)), np.float32(rng.uniform(1, 60))
        lo, hi = int(rng.integers(-1, 4)), int(rng.integers(-1, 8))
        got = F.features_in_area(x, y, r, lo, hi)
        inside = (np.abs(fr["x"] - x) < r) & (np.abs(fr["y"] - y) < r) & ok
        if lo > 0 or hi >= 0:
            inside &= fr["octave"] >= lo
            if hi >= 0:
                inside &= fr["octave"] <= hi
        # window cells (Frame.cpp:205-221): a feature inside the window whose cell falls outside the cell range is dropped
        c0x = max(0, int(np.floor((x - np.float32(g.min_x) - r) * np.float32(g.inv_w))))
        c1x = min(g.cols - 1, int(np.ceil((x - np.float32(g.min_x) + r) * np.float32(g.inv_w))))
        c0y = max(0, int(np.floor((y - np.float32(g.min_y) - r) * np.float32(g.inv_h))))
        c1y = min(g.rows - 1, int(np.ceil((y - np.float32(g.min_y) + r) * np.float32(g.inv_h))))
        inside &= (px >= c0x) & (px <= c1x) & (py >= c0y) & (py <= c1y)
        idx = np.nonzero(inside)[0]
        want = idx[np.lexsort((idx, cell[idx]))]
        assert np.array_equal(got, want)


def test_allpairs_best2_vs_numpy():
    rng = np.random.default_rng(2)
    t = rng.integers(0, 256, (3000, 32), dtype=np.uint8)
    q = perturb_descriptors(rng, t[rng.integers(0, 3000, 200)], 0.08)
    t[2000] = t[10]; q[0] = t[10]
    res = OracleMatcher(0.75).allpairs_best2(q, t)
    D = np.unpackbits(q[:, None, :] ^ t[None, :, :], axis=2).sum(2)
    assert np.array_equal(res["best_idx"], D.argmin(1))            # argmin = first minimum = strict-< first-wins
    part = np.sort(D, axis=1)
    assert np.array_equal(res["best_dist"], part[:, 0]) and np.array_equal(res["best_dist2"], part[:, 1])
    acc = (part[:, 0] < 50) & (part[:, 0].astype(np.float32) < np.float32(0.75) * part[:, 1].astype(np.float32))
    assert np.array_equal(res["accepted"], acc.astype(np.int32))
    assert res["best_idx"][0] == 10 and res["best_dist2"][0] == 0 and res["accepted"][0] == 0


def _tiny_frame(descs, xs, ys, octs):
    g = frame_grid(W, H)
    n = len(xs)
    return OracleFrame(np.array(descs, np.uint8), np.array(xs, np.float32), np.array(ys, np.float32), np.array(octs, np.int32),
                       np.zeros(n, np.float32), g)


def _d(nbits):
    d = np.zeros(32, np.uint8)
    d[: nbits // 8] = 0xFF
    if nbits % 8:
        d[nbits // 8] = (1 << (nbits % 8)) - 1
    return d


def test_known_answer_greedy_rules():
    """Hand-checked cases of the rules in ORBmatcher.cpp:113-146 (SearchByProjection, Frame x MapPoints)."""
    sf = np.ones(8, np.float32)
    # targets: t0 dist 10 from zero descriptor, t1 dist 12 (same level), t2 dist 40 (other level)
    F = _tiny_frame([_d(10), _d(12), _d(40)], [100, 101, 102], [100, 100, 100], [1, 1, 2])
    zero = np.zeros((3, 32), np.uint8)
    one = np.ones(3)
    # ratio 0.8: 10 > 0.8*12 -> rejected when both on the same level (:139)
    h = np.full(3, -1, np.int32)
    nm, res = OracleMatcher(0.8).projection_mappoints(F, one[:1], [100], [100], [2], one[:1], zero[:1], one[:1], 1.0, sf, h)
    assert nm == 0 and res[0]["best_idx"] == 0 and res[0]["best_dist"] == 10 and res[0]["best_dist2"] == 12 and h[0] == -1
    # ratio 0.9: accepted, claims t0
    h = np.full(3, -1, np.int32)
    nm, res = OracleMatcher(0.9).projection_mappoints(F, one[:1], [100], [100], [2], one[:1], zero[:1], one[:1], 1.0, sf, h)
    assert nm == 1 and h[0] == 0
    # three identical queries, all observed: q0 takes t0, q1 then sees (t1=12, t2=40) -> different levels -> takes t1,
    # q2 sees only t2 -> 40 <= TH_HIGH -> takes t2
    h = np.full(3, -1, np.int32)
    nm, res = OracleMatcher(0.9).projection_mappoints(F, one, [100] * 3, [100] * 3, [2] * 3, one, zero, one, 1.0, sf, h)
    assert nm == 3 and list(h) == [0, 1, 2]
    # same, but query 0's map point has no observations: it does not block -> q1 re-claims t0 (overwrites), q2 gets t1
    h = np.full(3, -1, np.int32)
    nm, res = OracleMatcher(0.9).projection_mappoints(F, one, [100] * 3, [100] * 3, [2] * 3, one, zero, [0, 1, 1], 1.0, sf, h)
    assert nm == 3 and list(h) == [1, 2, -1]
    # holder -2 blocks, -3 does not
    h = np.array([-2, -3, -1], np.int32)
    nm, res = OracleMatcher(0.9).projection_mappoints(F, one[:1], [100], [100], [2], one[:1], zero[:1], one[:1], 1.0, sf, h)
    assert nm == 1 and list(h) == [-2, 0, -1] and res[0]["best_dist"] == 12


def test_rotation_histogram_quirk():
    """rot in [0,360) * (1/30) rounds into bins 0..12 only (upstream quirk): kept as is."""
    from _oracle import match_lib  # noqa: F401
    sf = np.ones(8, np.float32)
    n = 40
    descs = [np.random.default_rng(i).integers(0, 256, 32, dtype=np.uint8) for i in range(n)]
    xs = [20 + 18 * (i % 20) for i in range(n)]
    ys = [50 + 100 * (i // 20) for i in range(n)]
    F = _tiny_frame(descs, xs, ys, [0] * n)
    ang = np.zeros(n, np.float32)
    ang[:30] = 10.0       # rot 10 -> bin 0 (kept)
    ang[30:34] = 100.0    # bin 3 (kept: 4 >= 0.1*30)
    ang[34:36] = 200.0    # bin 7 (2 < 3 -> dropped)
    ang[36:] = 359.0      # bin 12 (kept as third maximum: 4 entries)
    h = np.full(n, -1, np.int32)
    nm, res = OracleMatcher(0.9, True).projection_lastframe(F, np.ones(n), xs, ys, [0] * n, ang, np.array(descs), np.ones(n), 3.0, sf, h)
    assert res["accepted"].sum() == n
    assert nm == n - 2 and list(np.nonzero(h < 0)[0]) == [34, 35]


def test_distinctive_descriptors_vs_numpy():
    """mo_distinctive_descriptors (MapPoint.cpp:929-994 restated; pinned to the reference's own lines in
    tests/test_pin_oracle_vs_ref.py) against an independent numpy statement: argmin over rows of sorted(row)[int(0.5*(N-1))]."""
    from _oracle import distinctive_case, oracle_distinctive_descriptors
    start, desc = distinctive_case(5)
    best, med = oracle_distinctive_descriptors(start, desc)
    bits = np.unpackbits(desc, axis=1).astype(np.int32)
    for p in range(len(start) - 1):
        b = bits[start[p]:start[p + 1]]
        if len(b) == 0:
            assert best[p] == -1
            continue
        D = (b[:, None, :] != b[None, :, :]).sum(2)
        m = np.sort(D, axis=1)[:, int(0.5 * (len(b) - 1))]
        assert best[p] == int(np.argmin(m)) and med[p] == int(m.min())


def _wire_keypoints(seed, n):
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    rng = np.random.default_rng(seed)
    kp = np.zeros(n, KP_DTYPE)
    kp["octave"] = rng.integers(0, 8, n)
    kp["x"] = rng.uniform(-20, 780, n).astype(np.float32)
    kp["y"] = rng.uniform(-20, 500, n).astype(np.float32)
    kp["size"] = (31 * np.float32(1.2) ** kp["octave"]).astype(np.int32)          # (int)(31 * scale), up to 111
    kp["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    kp["response"] = rng.integers(7, 256, n)
    if n > 4:
        kp["response"][:2] = [37.75, 255.0]                                         # fractional values truncate
        kp["size"][2] = 300.5                                                       # out of u8 range: low byte of the integer part
        kp["octave"][3] = -1
    return kp, rng.integers(0, 256, (n, 32), dtype=np.uint8)


def test_wire_format_vs_numpy_packing():
    """Keyframe wire arrays (KeyFrame.cpp:1462-1472, Converter.cc:166-193; pinned to the reference's statements in
    tests/test_pin_oracle_vs_ref.py):
    the restatement against an independent numpy structured-dtype packing of the .msg definitions, and the round trip."""
    from _oracle import oracle_wire_decode, oracle_wire_encode
    msg = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "u1"), ("angle", "<f4"), ("response", "u1"), ("octave", "i1")])
    assert msg.itemsize == 15
    for seed, n in [(1, 1000), (2, 1), (3, 0), (4, 257)]:
        kp, desc = _wire_keypoints(seed, n)
        kb, db = oracle_wire_encode(kp, desc)
        rec = np.zeros(n, msg)
        for k in ("x", "y", "angle"):
            rec[k] = kp[k]
        rec["size"] = kp["size"].astype(np.int64) & 255
        rec["response"] = kp["response"].astype(np.int64) & 255
        rec["octave"] = kp["octave"].astype(np.int8)
        assert kb == np.uint32(n).tobytes() + rec.tobytes()
        assert db == np.uint32(n).tobytes() + desc.tobytes()
        back, d2 = oracle_wire_decode(kb, db)
        assert np.array_equal(d2, desc)
        for k in ("x", "y", "angle"):
            assert np.array_equal(back[k], kp[k])
        assert np.array_equal(back["size"], rec["size"].astype(np.float32)) and np.array_equal(back["response"], rec["response"].astype(np.float32))
        assert np.array_equal(back["octave"], rec["octave"].astype(np.int32))
    kb, db = oracle_wire_encode(*_wire_keypoints(5, 10))
    assert oracle_wire_decode(kb, db[:4 + 32 * 9].replace(db[:4], np.uint32(9).tobytes(), 1)) is None     # counts disagree
