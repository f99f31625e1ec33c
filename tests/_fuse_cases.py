"""Seeded inputs for the Fuse / SearchBySim3 parity cases, and the host-side replay of those three methods over a
candidate-search result (the oracle's or the device's CCM_WIN_BEST search). The replay restates what the reference does
with each accepted candidate (ORBmatcher.cpp:967-989 Fuse, :1102-1117 Fuse with Scw, :1329-1345 SearchBySim3) on the
state model of the mock KeyFrame / MapPoint in oracle/matchshim (Replace moves the observation and the keyframe slot to
the survivor, AddObservation / AddMapPoint fill the slot)."""
import numpy as np

from motioncheck_ccm_slam_b200.synth import make_feature_frame, make_next_frame, perturb_descriptors

W, H = 752, 480
SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
INV_SIGMA2 = (np.float32(1.0) / (SF * SF)).astype(np.float32)
OCC_ID = 100000


def projections(a, b, src, rng, noise=1.5):
    """Where a's features land in b (b's matching feature position + noise), and b's level there."""
    n = len(a["x"])
    inv = np.full(n, -1)
    inv[src[src >= 0]] = np.nonzero(src >= 0)[0]
    u = np.where(inv >= 0, b["x"][np.maximum(inv, 0)], a["x"] + 2).astype(np.float32) + rng.normal(0, noise, n).astype(np.float32)
    v = np.where(inv >= 0, b["y"][np.maximum(inv, 0)], a["y"] + 1).astype(np.float32) + rng.normal(0, noise, n).astype(np.float32)
    level = np.where(inv >= 0, b["octave"][np.maximum(inv, 0)], a["octave"]).astype(np.int32)
    level = np.clip(level + rng.choice([0, 0, 0, 1], n), 0, 7).astype(np.int32)
    inside = ((u >= 0) & (u <= W) & (v >= 0) & (v <= H))
    return u, v, level, inside


def fuse_case(seed, n=1000):
    rng = np.random.default_rng(seed)
    a = make_feature_frame(seed, n, W, H)
    b, src = make_next_frame(seed + 1, a, W, H)
    u, v, level, inside = projections(a, b, src, rng)
    nb = len(b["x"])
    c = {
        "kf": b, "q_u": u, "q_v": v, "q_level": level, "q_desc": np.ascontiguousarray(a["desc"]),
        "q_valid": ((rng.random(n) < 0.9) & inside).astype(np.uint8),
        "q_obs": rng.integers(0, 6, n).astype(np.int32),
        "q_flags": np.where(rng.random(n) < 0.05, 2, 0).astype(np.uint8),
        "q_alias": np.full(n, -1, np.int32),
        "occ_obs": np.where(rng.random(nb) < 0.4, rng.integers(0, 6, nb), -1).astype(np.int32),
        "occ_flags": (np.where(rng.random(nb) < 0.1, 1, 0) | np.where(rng.random(nb) < 0.1, 2, 0)).astype(np.uint8),
    }
    for i in range(1, n):
        r = rng.random()
        if r < 0.05:
            c["q_alias"][i] = rng.integers(0, i)                 # the same map point listed twice
        elif r < 0.08:
            c["q_alias"][i] = -(int(rng.integers(0, nb)) + 2)    # a point the keyframe already holds
    return c


class _Pt:
    __slots__ = ("id", "bad", "nobs", "dnr", "slot", "replaced_by", "added_idx")

    def __init__(self, id, nobs):
        self.id, self.bad, self.nobs, self.dnr, self.slot, self.replaced_by, self.added_idx = id, False, nobs, False, None, -1, -1


def _build_state(c):
    nb = len(c["kf"]["x"])
    slots = [None] * nb
    for k in range(nb):
        if c["occ_obs"][k] >= 0:
            o = _Pt(OCC_ID + k, int(c["occ_obs"][k]))
            o.bad, o.dnr, o.slot = bool(c["occ_flags"][k] & 1), bool(c["occ_flags"][k] & 2), k
            slots[k] = o
    occ = list(slots)
    pts = []
    for i in range(len(c["q_u"])):
        al = int(c["q_alias"][i])
        if 0 <= al < i:
            pts.append(pts[al]); continue
        p = slots[-al - 2] if al <= -2 and slots[-al - 2] is not None else _Pt(i, int(c["q_obs"][i]))
        if not c["q_valid"][i]:
            p.bad = True
        p.dnr = p.dnr or bool(c["q_flags"][i] & 2)
        pts.append(p)
    return slots, occ, pts


def _replace(slots, p, by):
    """mock MapPoint::Replace (MapPoint.cpp:583-612, common branch)"""
    if by.id == p.id:
        return
    p.replaced_by, p.bad = by.id, True
    k, p.slot = p.slot, None
    if k is not None and by.slot is None:
        slots[k] = by; by.slot = k; by.nobs += 1


def _add(slots, p, k):
    p.added_idx, p.slot = k, k
    p.nobs += 1
    slots[k] = p


def searched_fuse(c, first):
    """Queries that reach the candidate loop: the static checks on entry (ORBmatcher.cpp:875-881 for the first Fuse,
    :1022 for the Scw one)."""
    slots, occ, pts = _build_state(c)
    m = np.zeros(len(pts), bool)
    for i, p in enumerate(pts):
        m[i] = not p.bad and p.slot is None and not (first and p.dnr)
    return m


def effective_rows(c):
    """Row whose (u, v, level, desc) a query uses: map points listed twice keep the LAST row written to them by the
    glue's make_points (an alias row j<i is skipped there, so the point keeps its own row)."""
    rows = np.arange(len(c["q_u"]))
    for i in range(len(rows)):
        al = int(c["q_alias"][i])
        if 0 <= al < i:
            rows[i] = rows[al]
    return rows


def replay_fuse(c, searched, accepted, best_idx):
    slots, occ, pts = _build_state(c)
    n_fused = 0
    for i, p in enumerate(pts):
        if not searched[i] or not accepted[i] or p.bad or p.slot is not None:
            continue
        k = int(best_idx[i])
        o = slots[k]
        if o is not None:
            if not o.bad and not o.dnr:
                if o.nobs > p.nobs:
                    _replace(slots, p, o)
                else:
                    _replace(slots, o, p)
        else:
            _add(slots, p, k)
        n_fused += 1
    return {
        "n": n_fused,
        "replaced_by": np.array([p.replaced_by for p in pts], np.int32),
        "added_idx": np.array([p.added_idx for p in pts], np.int32),
        "q_bad": np.array([p.bad for p in pts], np.int32),
        "slot_id": np.array([s.id if s is not None else -1 for s in slots], np.int32),
        "slot_bad": np.array([int(s.bad) if s is not None else 0 for s in slots], np.int32),
        "occ_replaced_by": np.array([o.replaced_by if o is not None else -1 for o in occ], np.int32),
    }


def replay_fuse_sim3(c, searched, accepted, best_idx):
    slots, occ, pts = _build_state(c)
    n_fused = 0
    rep = np.full(len(pts), -1, np.int32)
    for i, p in enumerate(pts):
        if not searched[i] or not accepted[i]:
            continue
        k = int(best_idx[i])
        o = slots[k]
        if o is not None:
            if not o.bad:
                rep[i] = o.id
        else:
            _add(slots, p, k)
        n_fused += 1
    return {
        "n": n_fused, "replace_point": rep,
        "added_idx": np.array([p.added_idx for p in pts], np.int32),
        "slot_id": np.array([s.id if s is not None else -1 for s in slots], np.int32),
        "slot_bad": np.array([int(s.bad) if s is not None else 0 for s in slots], np.int32),
    }


def sim3_case(seed, n=1000):
    rng = np.random.default_rng(seed)
    a = make_feature_frame(seed, n, W, H)
    b, src = make_next_frame(seed + 1, a, W, H)
    n1, n2 = len(a["x"]), len(b["x"])
    u1, v1, l1, in1 = projections(a, b, src, rng)
    # KF2's points projected back into KF1: the source feature's position
    s = np.maximum(src, 0)
    u2 = np.where(src >= 0, a["x"][s], b["x"] - 2).astype(np.float32) + rng.normal(0, 1.5, n2).astype(np.float32)
    v2 = np.where(src >= 0, a["y"][s], b["y"] - 1).astype(np.float32) + rng.normal(0, 1.5, n2).astype(np.float32)
    l2 = np.clip(np.where(src >= 0, a["octave"][s], b["octave"]) + rng.choice([0, 0, 0, 1], n2), 0, 7).astype(np.int32)
    in2 = (u2 >= 0) & (u2 <= W) & (v2 >= 0) & (v2 <= H)
    st1 = np.where(in1, rng.choice([0, 1, 1, 1, 1, 1, 1, 2], n1), 0).astype(np.uint8)
    st2 = np.where(in2, rng.choice([0, 1, 1, 1, 1, 1, 1, 2], n2), 0).astype(np.uint8)
    prior = np.full(n1, -1, np.int32)
    pick = rng.random(n1) < 0.1
    prior[pick] = rng.integers(0, n2, int(pick.sum()))
    prior[rng.random(n1) < 0.02] = -2
    return {"kf1": a, "kf2": b, "st1": st1, "u1": u1, "v1": v1, "l1": l1, "d1": perturb_descriptors(np.random.default_rng(seed + 7), a["desc"], 0.03),
            "st2": st2, "u2": u2, "v2": v2, "l2": l2, "d2": perturb_descriptors(np.random.default_rng(seed + 8), b["desc"], 0.03), "prior": prior}


def sim3_masks(c):
    """Which slots of KF1 / KF2 reach the candidate loop (ORBmatcher.cpp:1156-1170, :1178-1185, :1255-1262)."""
    n1, n2 = len(c["st1"]), len(c["st2"])
    done1, done2 = np.zeros(n1, bool), np.zeros(n2, bool)
    for i in range(n1):
        p = int(c["prior"][i])
        if p >= 0 and c["st2"][p]:
            done1[i] = True; done2[p] = True
        elif p == -2:
            done1[i] = True
    return (c["st1"] == 1) & ~done1, (c["st2"] == 1) & ~done2


def replay_sim3(c, acc1, best1, acc2, best2, m1, m2):
    n1 = len(c["st1"])
    out = np.full(n1, -1, np.int32)
    for i in range(n1):
        p = int(c["prior"][i])
        if p >= 0 and c["st2"][p]:
            out[i] = p
        elif p == -2:
            out[i] = -2
    vn1 = np.where(m1 & acc1.astype(bool), best1, -1)
    vn2 = np.where(m2 & acc2.astype(bool), best2, -1)
    found = 0
    for i in range(n1):
        j = int(vn1[i])
        if j >= 0 and vn2[j] == i:
            out[i] = j; found += 1
    return found, out


LEVEL_SIGMA2 = (SF * SF).astype(np.float32)


def tri_case(seed, n=1000, n_nodes=120, motion=(9.0, 4.0)):
    """Two keyframes related by an image translation: F12 = [t]x makes the epipolar line of a KF1 feature the line
    through it along the motion. To make every predicate of the loop decide something: a third of KF2's features are
    pushed off their epipolar line (3.84 sigma2 gate, :159-176), 150 look-alikes (same vocabulary node, 0-5 flipped bits,
    0 = an exact tie, which the reference resolves to the LATER feature, :768) are placed along the lines, and the
    epipole sits on a feature cluster of KF2 (:777 gate)."""
    from motioncheck_ccm_slam_b200.synth import bow_node_ids
    rng = np.random.default_rng(seed)
    a = make_feature_frame(seed, n, W, H)
    b, src = make_next_frame(seed + 1, a, W, H, motion=motion)
    tx, ty = np.float32(motion[0]), np.float32(motion[1])
    tn = float(np.hypot(tx, ty))
    off = rng.random(len(b["x"])) < 0.33
    perp = rng.normal(0, 3.0, len(b["x"])) * off
    b["x"] = (b["x"] - perp * ty / tn).astype(np.float32)
    b["y"] = (b["y"] + perp * tx / tn).astype(np.float32)
    cand = np.nonzero(src >= 0)[0]
    twins = rng.choice(cand, min(150, len(cand) // 3), replace=False)
    sl = rng.uniform(-40, 40, len(twins))
    td = b["desc"][twins].copy()
    for k, row in enumerate(td):
        for bit in rng.integers(16, 256, int(rng.choice([0, 0, 1, 2, 5]))):
            row[bit >> 3] ^= np.uint8(1 << (bit & 7))
    b = {"desc": np.concatenate([b["desc"], td]), "x": np.concatenate([b["x"], (b["x"][twins] + sl * tx / tn).astype(np.float32)]),
         "y": np.concatenate([b["y"], (b["y"][twins] + sl * ty / tn).astype(np.float32)]),
         "octave": np.concatenate([b["octave"], b["octave"][twins]]), "angle": np.concatenate([b["angle"], b["angle"][twins]])}
    f12 = np.array([[0, 0, ty], [0, 0, -tx], [-ty, tx, 0]], np.float32) * np.float32(0.37)
    free1 = (rng.random(len(a["x"])) < 0.7).astype(np.uint8)
    free2 = (rng.random(len(b["x"])) < 0.7).astype(np.uint8)
    return {"kf1": a, "kf2": b, "node1": bow_node_ids(a["desc"], n_nodes), "node2": bow_node_ids(b["desc"], n_nodes),
            "free1": free1, "free2": free2, "f12": np.ascontiguousarray(f12.reshape(-1)),
            "ex": float(b["x"][5]), "ey": float(b["y"][5])}
