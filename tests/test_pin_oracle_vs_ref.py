"""CPU: the two "next" rows whose reference files do not build outside the ROS workspace, pinned to the reference's own
statements. oracle/Makefile cuts MapPoint.cpp:929-994 (ComputeDistinctiveDescriptors), ORBmatcher.cpp:1653-1669
(DescriptorDistance), Converter.cc:166-193 (to/fromCvKeyPointMsg) and KeyFrame.cpp:1462-1472 / 1688-1704 (the keypoint and
descriptor loops of ConvertToMessage / WriteMembersFromMessage) out of /root/reference at build time and compiles them
unmodified over oracle/pinshim/pin_prelude.h (oracle/_ref/libref_pin.so; the prebuilt library travels with the repo).
oracle/match_oracle.cpp's mo_distinctive_descriptors and mo_wire_* must agree with them; the GPU tests compare the kernels
with those restatements on the same seeds, so GPU == oracle == reference by transitivity."""
import numpy as np
import pytest

import _oracle
from _oracle import (distinctive_case, oracle_distinctive_descriptors, oracle_wire_decode, oracle_wire_encode, ref_distinctive_descriptors,
                     ref_wire_decode, ref_wire_encode)

pytestmark = pytest.mark.skipif(not _oracle.ref_pin_available(), reason="oracle/_ref/libref_pin.so not built")


@pytest.mark.parametrize("seed,n_points,max_obs", [(5, 400, 40), (6, 60, 130), (7, 300, 8)])
def test_distinctive_descriptors_vs_reference(seed, n_points, max_obs):
    """Same cases as tests/test_gpu_match.py::test_distinctive_descriptors. The reference leaves the chosen descriptor in
    mDescriptor (no index): compared byte for byte with the row the oracle picks; points without observations keep an
    empty mDescriptor (the early return at :943) and the oracle reports -1."""
    start, desc = distinctive_case(seed, n_points, max_obs)
    best, med = oracle_distinctive_descriptors(start, desc)
    chosen, has = ref_distinctive_descriptors(start, desc)
    assert len(best) == n_points
    for p in range(n_points):
        n = start[p + 1] - start[p]
        assert has[p] == (1 if n > 0 else 0) and (best[p] >= 0) == (n > 0), p
        if n > 0:
            assert np.array_equal(chosen[p], desc[start[p] + best[p]]), (p, n, best[p])


def test_distinctive_descriptors_ties_and_bad_keyframes():
    """Equal medians: the first row in the reference's iteration order wins (:979 strict <). Bad keyframes are left out
    (:952) — the caller of the batched entry point passes only the rows of good keyframes."""
    rng = np.random.default_rng(11)
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    rows = np.repeat(base, 6, 0)
    rows[3, 0] ^= 1; rows[4, 0] ^= 3
    start = np.array([0, 6], np.int32)
    best, med = oracle_distinctive_descriptors(start, rows)
    chosen, has = ref_distinctive_descriptors(start, rows)
    assert has[0] == 1 and best[0] == 0 and med[0] == 0 and np.array_equal(chosen[0], rows[0])
    # drop rows 0..2 through isBad(): the reference then picks among 3, 4, 5 = the oracle on those rows alone
    bad = np.array([1, 1, 1, 0, 0, 0], np.uint8)
    chosen, has = ref_distinctive_descriptors(start, rows, bad)
    b2, _ = oracle_distinctive_descriptors(np.array([0, 3], np.int32), rows[3:])
    assert has[0] == 1 and np.array_equal(chosen[0], rows[3 + b2[0]])
    chosen, has = ref_distinctive_descriptors(start, rows, np.ones(6, np.uint8))      # every keyframe bad: :956 return
    assert has[0] == 0


def _wire_keypoints(seed, n):
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    rng = np.random.default_rng(seed)
    kp = np.zeros(n, KP_DTYPE)
    kp["octave"] = rng.integers(0, 8, n)
    kp["x"] = rng.uniform(-20, 780, n).astype(np.float32)
    kp["y"] = rng.uniform(-20, 500, n).astype(np.float32)
    kp["size"] = (31 * np.float32(1.2) ** kp["octave"]).astype(np.int32)
    kp["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    kp["response"] = rng.integers(7, 256, n)
    if n > 4:
        kp["response"][:2] = [37.75, 255.0]          # fractional values truncate in (u_int8_t)kp.response
        kp["octave"][3] = -1
    return kp, rng.integers(0, 256, (n, 32), dtype=np.uint8)


@pytest.mark.parametrize("seed,n", [(1, 1000), (2, 1), (4, 257), (9, 2000)])
def test_wire_format_vs_reference(seed, n):
    """Same cases as the GPU wire tests (in-range sizes and responses; a float outside [0, 256) through (u_int8_t) is
    undefined behaviour upstream and is not compared)."""
    kp, desc = _wire_keypoints(seed, n)
    kb_o, db_o = oracle_wire_encode(kp, desc)
    kb_r, db_r = ref_wire_encode(kp, desc)
    assert kb_o == kb_r and db_o == db_r
    back_o, d_o = oracle_wire_decode(kb_o, db_o)
    back_r, d_r = ref_wire_decode(kb_r, db_r)
    assert np.array_equal(d_o, d_r) and np.array_equal(d_r, desc)
    for k in ("x", "y", "size", "angle", "response", "octave"):
        assert np.array_equal(back_o[k], back_r[k]), k
