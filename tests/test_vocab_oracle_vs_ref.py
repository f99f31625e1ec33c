"""CPU: the vocabulary-descent oracle (oracle/match_oracle.cpp mo_vocab_transform) and the BowVector / FeatureVector
assembly restated in tests/_oracle.py against the reference's OWN vocabulary code — the vendored DBoW2 compiled unmodified
(oracle/_ref/libref_vocab.so) — on seeded vocabularies written in ORB-SLAM's text format and loaded with the reference's
loadFromTextFile. What Frame::ComputeBoW (cslam/src/Frame.cpp:268-275) leaves behind is mBowVec (word -> L1-normalised
tf-idf weight, doubles) and mFeatVec (node 4 levels above the leaves -> feature indices): compared bit for bit."""
import os

import numpy as np
import pytest

import _oracle
from motioncheck_ccm_slam_b200.synth import make_vocabulary, perturb_descriptors, write_vocabulary_text

pytestmark = pytest.mark.skipif(not _oracle.ref_vocab_available(), reason="oracle/_ref/libref_vocab.so not built")


def _features(voc, seed, n):
    """Descriptors near vocabulary leaves (a few flipped bits) plus uniform ones, so descents are contested."""
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(voc["is_leaf"])[0]
    d = perturb_descriptors(rng, voc["desc"][rng.choice(leaves, n)], 0.04)
    d[: n // 10] = rng.integers(0, 256, (n // 10, 32), dtype=np.uint8)
    return np.ascontiguousarray(d)


@pytest.mark.parametrize("seed,k,L,ragged,levelsup", [(1, 10, 4, 0.0, 2), (2, 10, 3, 0.15, 4), (3, 6, 5, 0.1, 4), (4, 10, 4, 0.0, 0),
                                                      (5, 3, 6, 0.2, 4), (6, 10, 2, 0.0, 1)])
def test_transform_matches_dbow2(tmp_path, seed, k, L, ragged, levelsup):
    # ragged trees keep L - levelsup <= 2 (their shallowest leaves): when a descent ends ABOVE that level the batch transform
    # reads `nid` uninitialised (TemplatedVocabulary.h:1150-1154 declares it in the loop, :1258 never assigns it) — undefined
    # upstream, impossible with the complete k=10, L=6 ORB vocabulary; the oracle and the CUDA path return 0 there
    voc = make_vocabulary(seed, k, L, ragged)
    path = os.path.join(str(tmp_path), "voc.txt")
    write_vocabulary_text(voc, path)
    R = _oracle.RefVocabulary(path)
    assert R.words() == int(voc["is_leaf"].sum())
    desc = _features(voc, seed + 100, 1000)
    # per feature: word id, weight, node `levelsup` levels up
    w_r, wt_r, nd_r = R.transform_features(desc, levelsup)
    w_o, wt_o, nd_o = _oracle.oracle_vocab_transform(voc, desc, levelsup)
    assert np.array_equal(w_o, w_r) and np.array_equal(wt_o, wt_r) and np.array_equal(nd_o, nd_r)
    assert len(np.unique(w_r)) > 50 and (wt_r == 0).any()
    # the containers ComputeBoW fills
    bw_r, bv_r, fn_r, fo_r = R.transform(desc, levelsup)
    bw_o, bv_o, fn_o, fo_o = _oracle.assemble_bow(w_o, wt_o, nd_o)
    assert np.array_equal(bw_o, bw_r) and np.array_equal(bv_o.view(np.uint64), bv_r.view(np.uint64))
    assert np.array_equal(fn_o, fn_r) and np.array_equal(fo_o, fo_r)
    assert abs(bv_r.sum() - 1.0) < 1e-9


def test_distance_ties_take_the_first_child(tmp_path):
    """Two children at the same distance: DBoW2 keeps the first (`d < best_d`, TemplatedVocabulary.h:1244)."""
    voc = make_vocabulary(9, 4, 2)
    first = voc["child_ids"][voc["child_start"][0]:voc["child_start"][1]]
    voc["desc"][first[2]] = voc["desc"][first[0]]                     # root children 0 and 2 identical
    path = os.path.join(str(tmp_path), "voc.txt")
    write_vocabulary_text(voc, path)
    R = _oracle.RefVocabulary(path)
    desc = np.repeat(voc["desc"][first[0]][None], 4, 0)
    w_r, _, nd_r = R.transform_features(desc, 1)
    w_o, _, nd_o = _oracle.oracle_vocab_transform(voc, desc, 1)
    assert np.array_equal(w_o, w_r) and np.array_equal(nd_o, nd_r) and (nd_r == first[0]).all()
