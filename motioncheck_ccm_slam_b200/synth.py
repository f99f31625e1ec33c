"""Seeded synthetic grayscale frame streams for parity tests and benchmarks.

A stream is one static scene (multi-octave value noise + filled rectangles/discs for strong FAST
corners) viewed through a window that moves 2 px right / 1 px down per frame, plus fresh ±4 pixel
noise per frame, so consecutive frames share most of their corners. numpy only.
"""
import numpy as np


def _value_noise(rng, h, w, cell, amp):
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.uniform(-amp, amp, size=(gh, gw)).astype(np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = ys.astype(np.int32)
    x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def make_scene(seed, h, w, n_shapes=None):
    """Static scene canvas, float32 in [0,255]."""
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128.0, dtype=np.float32)
    for cell, amp in ((64, 60.0), (16, 40.0), (4, 20.0)):
        img += _value_noise(rng, h, w, cell, amp)
    if n_shapes is None:
        n_shapes = max(200, (h * w) // 1800)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n_shapes):
        cx, cy = int(rng.integers(0, w)), int(rng.integers(0, h))
        val = float(rng.integers(0, 256))
        if rng.random() < 0.5:
            rw, rh = int(rng.integers(4, 40)), int(rng.integers(4, 40))
            img[max(0, cy - rh):cy + rh, max(0, cx - rw):cx + rw] = val
        else:
            r = int(rng.integers(3, 24))
            y0, y1, x0, x1 = max(0, cy - r), min(h, cy + r + 1), max(0, cx - r), min(w, cx + r + 1)
            m = (yy[y0:y1, x0:x1] - cy) ** 2 + (xx[y0:y1, x0:x1] - cx) ** 2 <= r * r
            img[y0:y1, x0:x1][m] = val
    return np.clip(img, 0, 255)


def make_stream(seed, n_frames, w=752, h=480):
    """uint8 [n_frames, h, w]; frame t is the scene window at offset (2t, t) plus fresh ±4 noise."""
    scene = make_scene(seed, h + n_frames, w + 2 * n_frames)
    out = np.empty((n_frames, h, w), dtype=np.uint8)
    for t in range(n_frames):
        rng = np.random.default_rng(seed * 7919 + t + 1)
        f = scene[t:t + h, 2 * t:2 * t + w] + rng.integers(-4, 5, size=(h, w)).astype(np.float32)
        out[t] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return out


def make_frame(seed, w=752, h=480):
    return make_stream(seed, 1, w, h)[0]


# ---- synthetic matching workloads -------------------------------------------------------
def random_descriptors(rng, n):
    """i.i.d. uniform 256-bit descriptors, uint8 [n, 32]."""
    return rng.integers(0, 256, size=(n, 32), dtype=np.uint8)


def perturb_descriptors(rng, desc, flip_prob):
    """Flip each bit with probability flip_prob (the same feature seen again)."""
    bits = np.unpackbits(desc, axis=1)
    flips = (rng.random(bits.shape) < flip_prob).astype(np.uint8)
    return np.packbits(bits ^ flips, axis=1)


def make_feature_frame(seed, n=1000, w=752, h=480, nlevels=8):
    """A frame as the matcher sees it: dict(desc, x, y, octave, angle). Positions inside the image, clustered so that
    grid cells hold several features and windows hold tens of candidates; octaves geometric like the extractor's quotas."""
    rng = np.random.default_rng(seed)
    centers = rng.uniform([20, 20], [w - 20, h - 20], size=(max(n // 12, 1), 2))
    pts = centers[rng.integers(0, len(centers), n)] + rng.normal(0, 9.0, size=(n, 2))
    x = np.clip(pts[:, 0], 0.5, w - 1.5).astype(np.float32)
    y = np.clip(pts[:, 1], 0.5, h - 1.5).astype(np.float32)
    p = 0.8333 ** np.arange(nlevels)
    octave = rng.choice(nlevels, size=n, p=p / p.sum()).astype(np.int32)
    angle = rng.uniform(0, 360, n).astype(np.float32)
    return dict(desc=random_descriptors(rng, n), x=x, y=y, octave=octave, angle=angle)


def make_next_frame(seed, fr, w=752, h=480, motion=(2.0, 1.0), flip_prob=0.06, keep=0.85, n_new=None, d_angle=3.0):
    """The scene of `fr` one time step later: most features survive (permuted, moved by `motion` + jitter, a few descriptor
    bits flipped, angle turned by ~d_angle), the rest are new. Returns (frame dict, src) with src[i] = index in `fr` or -1."""
    rng = np.random.default_rng(seed)
    n = len(fr["x"])
    keep_idx = np.nonzero(rng.random(n) < keep)[0]
    n_new = n - len(keep_idx) if n_new is None else n_new
    new = make_feature_frame(seed + 77, max(n_new, 1), w, h, int(fr["octave"].max()) + 1)
    desc = np.concatenate([perturb_descriptors(rng, fr["desc"][keep_idx], flip_prob), new["desc"][:n_new]])
    x = np.concatenate([fr["x"][keep_idx] + motion[0] + rng.normal(0, 0.7, len(keep_idx)), new["x"][:n_new]]).astype(np.float32)
    y = np.concatenate([fr["y"][keep_idx] + motion[1] + rng.normal(0, 0.7, len(keep_idx)), new["y"][:n_new]]).astype(np.float32)
    octave = np.concatenate([np.clip(fr["octave"][keep_idx] + rng.choice([-1, 0, 0, 0, 1], len(keep_idx)), 0, fr["octave"].max()),
                             new["octave"][:n_new]]).astype(np.int32)
    angle = np.concatenate([np.mod(fr["angle"][keep_idx] + rng.normal(d_angle, 2.0, len(keep_idx)), 360.0),
                            new["angle"][:n_new]]).astype(np.float32)
    src = np.concatenate([keep_idx, np.full(n_new, -1)]).astype(np.int64)
    perm = rng.permutation(len(src))
    x = np.clip(x, 0.5, w - 1.5); y = np.clip(y, 0.5, h - 1.5)
    return dict(desc=desc[perm], x=x[perm], y=y[perm], octave=octave[perm], angle=angle[perm]), src[perm]


def bow_node_ids(desc, n_nodes=1000):
    """Stand-in for DBoW2's FeatureVector node ids (the ORB vocabulary blob is absent from the reference checkout,
    SURVEY.md §8c): node = first 13 descriptor bits mod n_nodes, the same rule on both sides."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    return (((d[:, 0].astype(np.int32) << 5) | (d[:, 1].astype(np.int32) >> 3)) % n_nodes).astype(np.int32)


def make_vocabulary(seed, k=10, L=4, ragged=0.0, stop=0.02):
    """A seeded stand-in for the ORB vocabulary (cslam/conf/ORBvoc.txt.zip is a missing blob): a k-ary tree
    of depth L in the node numbering loadFromTextFile produces (TemplatedVocabulary.h:1338-1420: root = 0, nodes 1.. in file
    order, word ids in file order of the leaves). A child descriptor is its parent's with a level-dependent number of bits
    flipped, so descents are decided by small distance differences and ties occur. `ragged` = probability that an inner
    node is cut short into a leaf or gets fewer than k children; `stop` = fraction of words with weight 0 (stopped words).
    Returns dict(k, L, parent [n], is_leaf [n], desc [n][32], weight [n], word [n], child_start [n+1], child_ids) with n
    counting the root (index 0; its descriptor and weight are unused)."""
    rng = np.random.default_rng(seed)
    parent, is_leaf, desc, level = [0], [0], [np.zeros(32, np.uint8)], [0]
    frontier = [0]
    root_desc = rng.integers(0, 256, 32, dtype=np.uint8)
    for lvl in range(1, L + 1):
        nxt = []
        flips = max(256 >> (lvl + 1), 3)
        for p in frontier:
            base = root_desc if p == 0 else desc[p]
            nk = k if (p == 0 or rng.random() >= ragged) else int(rng.integers(1, k + 1))
            for _ in range(nk):
                d = np.unpackbits(base).copy()
                d[rng.choice(256, flips, replace=False)] ^= 1
                leaf = lvl == L or (lvl > 1 and rng.random() < ragged * 0.5)
                parent.append(p); is_leaf.append(int(leaf)); desc.append(np.packbits(d)); level.append(lvl)
                if not leaf:
                    nxt.append(len(parent) - 1)
        frontier = nxt
    n = len(parent)
    parent = np.array(parent, np.int32); is_leaf = np.array(is_leaf, np.uint8); desc = np.stack(desc).astype(np.uint8)
    weight = np.where(is_leaf > 0, rng.uniform(0.5, 9.0, n), 0.0)
    weight[(is_leaf > 0) & (rng.random(n) < stop)] = 0.0
    word = np.full(n, -1, np.int32)
    word[is_leaf > 0] = np.arange(int(is_leaf.sum()), dtype=np.int32)
    order = np.argsort(parent[1:], kind="stable") + 1          # children of a node in file order
    counts = np.bincount(parent[1:], minlength=n)
    child_start = np.zeros(n + 1, np.int32); child_start[1:] = np.cumsum(counts)
    return dict(k=k, L=L, parent=parent, is_leaf=is_leaf, desc=desc, weight=weight, word=word,
                child_start=child_start, child_ids=order.astype(np.int32), level=np.array(level, np.int32))


def write_vocabulary_text(voc, path):
    """ORB-SLAM's vocabulary text format (TemplatedVocabulary.h:1338-1420, 1423-1447): 'k L scoring weighting', then one line
    per node 'parent isLeaf b0 .. b31 weight'. No newline after the last node: the loader's while(!eof) loop would otherwise
    parse one more, empty, node. Scoring 0 = L1_NORM, weighting 0 = TF_IDF (the ORB vocabulary's header)."""
    with open(path, "w") as f:
        f.write("%d %d 0 0" % (voc["k"], voc["L"]))
        for i in range(1, len(voc["parent"])):
            f.write("\n%d %d %s %r" % (voc["parent"][i], voc["is_leaf"][i], " ".join(str(int(b)) for b in voc["desc"][i]), float(voc["weight"][i])))


def make_vocabulary_complete(seed, k=10, L=6):
    """make_vocabulary's complete-tree case, vectorised per level (a k = 10, L = 6 tree has 1.1 M nodes): the same node
    numbering (level by level, children of a node adjacent), word ids in node order of the leaves, weights in [0.5, 9)."""
    rng = np.random.default_rng(seed)
    descs = [np.zeros((1, 32), np.uint8)]
    parents = [np.zeros(1, np.int32)]
    cur = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    first = 0                                     # node id of cur[0]
    n_total = 1
    for lvl in range(1, L + 1):
        flips = max(256 >> (lvl + 1), 3)
        n = len(cur) * k
        child = np.repeat(cur, k, axis=0)
        for c0 in range(0, n, 1 << 16):
            c1 = min(n, c0 + (1 << 16))
            bits = np.zeros((c1 - c0, 256), np.uint8)
            np.put_along_axis(bits, rng.integers(0, 256, (c1 - c0, flips)), 1, axis=1)
            child[c0:c1] ^= np.packbits(bits, axis=1)
        parents.append(np.repeat(np.arange(first, first + len(cur), dtype=np.int32), k))
        descs.append(child)
        first, cur = n_total, child
        n_total += n
    desc = np.concatenate(descs); parent = np.concatenate(parents)
    n = len(parent)
    is_leaf = np.zeros(n, np.uint8); is_leaf[first:] = 1
    weight = np.zeros(n); weight[first:] = rng.uniform(0.5, 9.0, n - first)
    word = np.full(n, -1, np.int32); word[first:] = np.arange(n - first, dtype=np.int32)
    child_start = np.zeros(n + 1, np.int32)
    child_start[1:first + 1] = (np.arange(first, dtype=np.int64) + 1) * k
    child_start[first + 1:] = n - 1
    return dict(k=k, L=L, parent=parent, is_leaf=is_leaf, desc=desc, weight=weight, word=word, child_start=child_start,
                child_ids=np.arange(1, n, dtype=np.int32))
