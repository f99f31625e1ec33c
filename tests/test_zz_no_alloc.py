"""The allocation-free database search (ccm_match_allpairs_best2_noalloc_device, include/ccm_match.h; no counterpart in
the reference) against the existing entry point
ccm_match_allpairs_best2_device, through the C ABI. Integer / index work: bit-exact. Collected last on
purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(rng, nq, nt):
    from motioncheck_ccm_slam_b200.synth import perturb_descriptors
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    if nt > 10:
        k = min(nq, nt) // 2
        q[:k] = perturb_descriptors(rng, t[rng.integers(0, nt, k)], 0.05)
        t[nt // 2] = t[3]; t[nt - 1] = t[3]        # duplicates: first index wins, second distance == best
        q[0] = t[3]
    return q, t


def test_noalloc_equals_existing_entry_point_twice(cuda_lib):
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200._lib import CCM_E_CAPACITY, CCM_OK
    from motioncheck_ccm_slam_b200.matcher import BEST2_DTYPE, Matcher
    torch.cuda.set_device(0)
    lib = cuda_lib
    m = Matcher(0)
    rng = np.random.default_rng(2026)
    # second call is smaller than the first (the shared scratch still holds the first call's records), third is the
    # largest query count the entry point takes, fourth an empty shard
    cases = [(1000, 60000, 0), (300, 7001, 123456), (4096, 2049, 7), (5, 0, 0)]
    bufs = []
    for nq, nt, base in cases:
        q, t = _case(rng, nq, nt)
        d_q = torch.from_numpy(q).cuda()
        d_t = torch.from_numpy(t).cuda() if nt else torch.zeros((1, 32), dtype=torch.uint8, device="cuda")
        d_new = torch.full((nq, 8), 0xAB, dtype=torch.uint8, device="cuda")
        d_old = torch.full((nq, 8), 0xCD, dtype=torch.uint8, device="cuda")
        bufs.append((nq, nt, base, d_q, d_t, d_new, d_old))
    torch.cuda.synchronize()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    free_before = None
    try:
        for round_ in range(2):                                   # two calls in a row per case, back to back
            for nq, nt, base, d_q, d_t, d_new, d_old in bufs:
                d_new.fill_(0xAB)
                torch.cuda.synchronize()
                rc = lib.ccm_match_allpairs_best2_noalloc_device(stream, d_q.data_ptr(), nq, d_t.data_ptr(), nt, base, d_new.data_ptr())
                assert rc == CCM_OK, lib.ccm_matcher_last_error(None)
                rc = lib.ccm_match_allpairs_best2_noalloc_device(stream, d_q.data_ptr(), nq, d_t.data_ptr(), nt, base, d_new.data_ptr())
                assert rc == CCM_OK, lib.ccm_matcher_last_error(None)
                torch.cuda.synchronize()
                m.match_allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), nt, base, d_old.data_ptr())
                m.sync()
                new = d_new.cpu().numpy().view(BEST2_DTYPE).reshape(-1)
                old = d_old.cpu().numpy().view(BEST2_DTYPE).reshape(-1)
                assert np.array_equal(new, old), (nq, nt, np.nonzero(new != old)[0][:10])
                if nt > 10:
                    assert new["idx"][0] == base + 3 and new["d1"][0] == 0 and new["d2"][0] == 0
                else:
                    assert np.all(new["idx"] == 0xFFFFFFFF) and np.all(new["d1"] == 256)
            if round_ == 0:
                free_before = torch.cuda.mem_get_info()[0]
        # the second round allocated no device memory (torch's caching allocator serves the fills and copies above)
        assert torch.cuda.mem_get_info()[0] == free_before
        # beyond the stated maximum: rejected with a message, nothing launched
        nq, nt, base, d_q, d_t, d_new, d_old = bufs[2]
        rc = lib.ccm_match_allpairs_best2_noalloc_device(stream, d_q.data_ptr(), 4097, d_t.data_ptr(), nt, base, d_new.data_ptr())
        assert rc == CCM_E_CAPACITY and b"CCM_NOALLOC_MAX_QUERIES" in lib.ccm_matcher_last_error(None)
        rc = lib.ccm_match_allpairs_best2_noalloc_device(stream, d_q.data_ptr(), 4096, d_t.data_ptr(), (128 << 23) + 1, base, d_new.data_ptr())
        assert rc == CCM_E_CAPACITY and b"nt too large" in lib.ccm_matcher_last_error(None)
        torch.cuda.synchronize()
    finally:
        m.close()
