"""Parity of the CUDA extractor (through the C ABI) with the CPU oracle / reference fixtures. Bars:
keypoint set, order, coordinates, FAST scores, quadtree selection, octave, size and ANGLE bit-exact;
descriptor bits >= 99.9 % identical (north_star tolerance for float cos/sin; measured 100 % so far)."""
import ctypes
import os

import numpy as np
import pytest

from _oracle import OracleExtractor
from motioncheck_ccm_slam_b200.synth import make_frame, make_stream

pytestmark = pytest.mark.gpu
DESC_BITS_TOL = 0.999
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_extract.npz"))


def make(cuda_lib, *args, **kw):
    from motioncheck_ccm_slam_b200 import ORBextractor
    return ORBextractor(*args, **kw)


def assert_same(kp, desc, ko, do, tag=""):
    assert len(kp) == len(ko), (tag, len(kp), len(ko))
    for f in ("octave", "x", "y", "response", "size", "angle"):
        assert np.array_equal(kp[f], ko[f]), (tag, f)
    if len(kp):
        same = 1.0 - np.unpackbits(desc ^ do, axis=1).mean()
        assert same >= DESC_BITS_TOL, (tag, same)


def test_c1_stream_batch_bit_exact(cuda_lib):
    frames = make_stream(1000, 8)
    ex = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=8)
    orc = OracleExtractor(1000, 1.2, 8, 20, 7)
    kps, descs = ex.extract_batch(list(frames))
    for f in range(8):
        ko, do = orc(frames[f])
        assert_same(kps[f], descs[f], ko, do, "frame %d" % f)
        if f in (0, 7):     # stage taps
            for l in range(8):
                assert np.array_equal(ex.pyramid_level(l, f), orc.level_image(l)), ("pyr", f, l)
                assert np.array_equal(ex.pyramid_level(l, f, blurred=True), orc.level_image(l, True)), ("blur", f, l)
                assert np.array_equal(ex.level_corners(l, f), orc.level_corners(l)), ("corners", f, l)
    assert ex.launch_count() > 0


@pytest.mark.parametrize("name", ["c1", "small", "kitti_half"])
def test_reference_fixtures(cuda_lib, name):
    """Against the reference's own ORBextractor.cpp output (committed fixture, tests/golden/make_golden.py)."""
    w, h, nf, nl, seed = [int(v) for v in G[name + "_cfg"]]
    ex = make(cuda_lib, nf, 1.2, nl, 20, 7, max_width=w, max_height=h)
    kp, desc = ex(make_frame(seed, w, h))
    assert_same(kp, desc, G[name + "_kp"], G[name + "_desc"], name)


@pytest.mark.parametrize("w,h,nf,nl,seed", [(1241, 376, 2000, 8, 3000), (752, 480, 2000, 8, 11), (641, 479, 700, 6, 5),
                                             (320, 240, 300, 4, 7), (400, 300, 20, 5, 2), (752, 480, 5000, 8, 13),
                                             (512, 512, 1500, 7, 17), (380, 560, 600, 5, 19),
                                             (1920, 1080, 3000, 8, 23),      # full HD: 2.4 M pyramid pixels, ~650 nodes on level 0
                                             (4095, 131, 1500, 3, 29)])      # the widest image the handle accepts, three cell rows
def test_shapes_and_quotas(cuda_lib, w, h, nf, nl, seed):
    ex = make(cuda_lib, nf, 1.2, nl, 20, 7, max_width=w, max_height=h, max_batch=2)
    orc = OracleExtractor(nf, 1.2, nl, 20, 7)
    frames = make_stream(seed, 2, w, h)
    kps, descs = ex.extract_batch(list(frames))
    for f in range(2):
        ko, do = orc(frames[f])
        assert_same(kps[f], descs[f], ko, do, (w, h, nf, f))
    k1, d1 = ex(frames[1])                       # the single-frame entry point (one CUDA graph, a branch per level)
    assert np.array_equal(k1, kps[1]) and np.array_equal(d1, descs[1])


def test_other_scale_and_thresholds(cuda_lib):
    img = make_frame(23, 640, 480)
    for args in [(800, 1.5, 4, 20, 7), (800, 1.1, 10, 30, 10), (800, 1.2, 8, 12, 12), (300, 2.0, 3, 40, 5),
                 (300, 2.6, 3, 20, 7), (400, 3.0, 2, 20, 7)]:      # factors > 2 take the generic resize kernel
        kp, desc = make(cuda_lib, *args, max_width=640, max_height=480)(img)
        ko, do = OracleExtractor(*args)(img)
        assert_same(kp, desc, ko, do, args)


def test_sparse_and_flat_images(cuda_lib):
    ex = make(cuda_lib, 500, 1.2, 6, 20, 7, max_width=400, max_height=300)
    orc = OracleExtractor(500, 1.2, 6, 20, 7)
    flat = np.full((300, 400), 128, np.uint8)
    kp, desc = ex(flat)
    assert len(kp) == 0 and desc.shape == (0, 32)
    # weak texture: most cells only fire at minThFAST (fallback path, ORBextractor.cpp:981-985)
    rng = np.random.default_rng(3)
    weak = (128 + rng.integers(-9, 10, (300, 400))).astype(np.uint8)
    weak[100:140, 50:90] = 200
    kp, desc = ex(weak)
    ko, do = orc(weak)
    assert len(ko) > 0
    assert_same(kp, desc, ko, do, "weak")
    # saturated checkerboard: plateaus of equal scores annihilate under strict NMS
    cb = (np.indices((300, 400)).sum(0) // 16 % 2 * 255).astype(np.uint8)
    kp, desc = ex(cb)
    ko, do = orc(cb)
    assert_same(kp, desc, ko, do, "checker")


def test_dense_candidates_and_high_thresholds(cuda_lib):
    """k_fast_chunks corner cases: (1) white noise and a 1-px checkerboard make almost every pixel a prefilter candidate, so a
    warp's candidate list overflows, is drained, and the NMS walks the score map instead of the list; (2) FAST thresholds
    >= 128 take the other form of the byte compare (fc_gt<true>), also mixed with a fallback threshold below 128."""
    rng = np.random.default_rng(9)
    noise = rng.integers(0, 256, (300, 400), dtype=np.uint8)
    fine = (np.indices((300, 400)).sum(0) % 2 * 255).astype(np.uint8)
    salt = np.full((300, 400), 30, np.uint8)
    salt[rng.random((300, 400)) < 0.02] = 250                      # isolated bright pixels on a dark field: |d| = 220
    blocks = (rng.integers(0, 2, (19, 25)) * 255).astype(np.uint8).repeat(16, 0).repeat(16, 1)[:300, :400]
    for name, img, args in [("noise", noise, (500, 1.2, 4, 20, 7)), ("fine checkerboard", fine, (500, 1.2, 4, 20, 7)),
                            ("noise, th 60/30", noise, (800, 1.2, 3, 60, 30)),
                            ("salt, th 200/150", salt, (500, 1.2, 3, 200, 150)), ("salt, th 140/100", salt, (500, 1.2, 3, 140, 100)),
                            ("blocks, th 250/128", blocks, (500, 1.2, 3, 250, 128)), ("noise, th 128/127", noise, (500, 1.2, 2, 128, 127))]:
        ex = make(cuda_lib, *args, max_width=400, max_height=300)
        orc = OracleExtractor(*args)
        kp, desc = ex(img)
        ko, do = orc(img)
        assert_same(kp, desc, ko, do, name)
        for l in range(args[2]):
            assert np.array_equal(ex.level_corners(l, 0), orc.level_corners(l)), (name, "corners", l)
        ex.close()


def test_strided_input_and_empty(cuda_lib):
    ex = make(cuda_lib, 300, 1.2, 4, 20, 7, max_width=320, max_height=240)
    big = np.zeros((260, 400), np.uint8)
    img = make_frame(7, 320, 240)
    big[10:250, 40:360] = img
    view = big[10:250, 40:360]                      # step 400 != cols 320 (a cv::Mat ROI)
    kp, desc = ex(view)
    kc, dc = ex(img)
    assert np.array_equal(kp, kc) and np.array_equal(desc, dc)
    kp, desc = ex(np.zeros((0, 0), np.uint8))       # ORBextractor.cpp:1219: empty image returns silently
    assert len(kp) == 0


def test_batch64_equals_single_and_is_deterministic(cuda_lib):
    frames = make_stream(2000, 64)
    ex = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=64)
    k1, d1 = ex.extract_batch(list(frames))
    k2, d2 = ex.extract_batch(list(frames))
    one = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=1)
    orc = OracleExtractor(1000, 1.2, 8, 20, 7)
    for f in range(64):
        assert np.array_equal(k1[f], k2[f]) and np.array_equal(d1[f], d2[f])
        assert abs(len(k1[f]) - 1000) <= 3 * 8 + 8, len(k1[f])       # every level met its quota
        if f % 8 == 0:
            ks, ds = one(frames[f])
            assert np.array_equal(k1[f], ks) and np.array_equal(d1[f], ds)
        if f in (31, 63):
            ko, do = orc(frames[f])
            assert_same(k1[f], d1[f], ko, do, f)


def test_capacity_and_errors(cuda_lib):
    from motioncheck_ccm_slam_b200._lib import CCM_E_CAPACITY, CcmError
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    ex = make(cuda_lib, 300, 1.2, 4, 20, 7, max_width=320, max_height=240)
    img = make_frame(7, 320, 240)
    kfull, dfull = ex(img)
    kp = np.zeros(50, KP_DTYPE)
    desc = np.zeros((50, 32), np.uint8)
    n = ctypes.c_int()
    rc = cuda_lib.ccm_orb_extract(ex._h, img.ctypes.data, 240, 320, 320, kp.ctypes.data, desc.ctypes.data, 50, ctypes.byref(n))
    assert rc == 0 and n.value == len(kfull)
    assert np.array_equal(kp, kfull[:50]) and np.array_equal(desc, dfull[:50])
    with pytest.raises(CcmError) as e:
        ex(make_frame(1, 400, 300))
    assert e.value.code == CCM_E_CAPACITY
    with pytest.raises(CcmError):
        make(cuda_lib, 300, 1.2, 8, 20, 7, max_width=100, max_height=100)   # level 7 has no 30-px cell


def test_pyramid_member_with_border(cuda_lib):
    """mvImagePyramid (ORBextractor.h:138) incl. the 19-px BORDER_REFLECT_101 apron of ComputePyramid."""
    img = make_frame(9, 320, 240)
    ex = make(cuda_lib, 300, 1.2, 4, 20, 7, max_width=320, max_height=240)
    ex(img)
    orc = OracleExtractor(300, 1.2, 4, 20, 7)
    orc(img)
    for l in range(4):
        lv = orc.level_image(l)
        want = np.pad(lv, 19, mode="reflect")
        assert np.array_equal(ex.pyramid_level(l, 0, with_border=True), want)
        assert np.array_equal(ex.mvImagePyramid[l], lv)


def test_device_resident_path_matches_host_path(cuda_lib):
    torch = pytest.importorskip("torch")
    frames = make_stream(3000, 4)
    ex = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=4)
    kh, dh = ex.extract_batch(list(frames))
    cap = ex.max_keypoints()
    d_img = torch.from_numpy(frames).cuda()                          # [4, 480, 752] u8, pitch 752
    d_kp = torch.zeros((4, cap, 6), dtype=torch.float32, device="cuda")
    d_desc = torch.zeros((4, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(4, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ex.extract_batch_device(d_img.data_ptr(), 4, 480, 752, 752, 752 * 480, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
    ex.sync()
    n = d_n.cpu().numpy()
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    for f in range(4):
        assert n[f] == len(kh[f])
        kp = d_kp[f, :n[f]].cpu().numpy().view(np.uint8).reshape(-1).view(KP_DTYPE)
        assert np.array_equal(kp, kh[f])
        assert np.array_equal(d_desc[f, :n[f]].cpu().numpy(), dh[f])


def test_device_buffers_tma_cannot_address_take_the_word_load_kernel(cuda_lib):
    """Level 0 is read in place from the caller's buffer. A pitch that is not a 16-byte multiple (756) or a base 4 bytes off a
    16-byte boundary cannot be described to TMA: those calls run k_fast_chunks<false> (word loads) and must give the same frames."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    frames = make_stream(3100, 3)
    ex = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=3)
    kh, dh = ex.extract_batch(list(frames))
    cap = ex.max_keypoints()
    for pitch, lead in ((756, 0), (768, 4), (752, 0)):
        buf = torch.zeros(lead + 3 * 480 * pitch + 64, dtype=torch.uint8, device="cuda")
        view = buf[lead:lead + 3 * 480 * pitch].view(3, 480, pitch)
        view[:, :, :752] = torch.from_numpy(frames).cuda()
        d_kp = torch.zeros((3, cap, 6), dtype=torch.float32, device="cuda")
        d_desc = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(3, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        ex.extract_batch_device(buf.data_ptr() + lead, 3, 480, 752, pitch, pitch * 480, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        ex.sync()
        n = d_n.cpu().numpy()
        for f in range(3):
            assert n[f] == len(kh[f]), (pitch, lead, f)
            kp = d_kp[f, :n[f]].cpu().numpy().view(np.uint8).reshape(-1).view(KP_DTYPE)
            assert np.array_equal(kp, kh[f]) and np.array_equal(d_desc[f, :n[f]].cpu().numpy(), dh[f]), (pitch, lead, f)


def test_dropin_cpp_class_matches_reference_class(cuda_lib):
    """The drop-in C++ class (motioncheck_ccm_slam_b200/host/ORBextractor.cpp, compiled against the reference's own
    <cslam/ORBextractor.h>) and the reference's ORBextractor.cpp, both driven through the same glue
    (cslam::ORBextractor::operator() on a cv::Mat): identical keypoints, descriptors within tolerance, identical
    mvImagePyramid levels. Built only where /root/reference exists; the prebuilt .so travels to the GPU box."""
    import _oracle
    if not (_oracle.dropin_available() and _oracle.ref_available(mono=True)):
        pytest.skip("oracle/_ref/libdropin_orb.so not built (needs the reference header at build time)")
    os.environ["CCM_ORB_FILL_PYRAMID"] = "1"
    from motioncheck_ccm_slam_b200.synth import make_frame
    args = (1000, 1.2, 8, 20, 7)
    drop = _oracle.RefExtractor(*args, lib="libdropin_orb.so")
    ref = _oracle.RefExtractor(*args, mono=True)
    for seed, (w, h) in [(101, (752, 480)), (102, (640, 400)), (103, (752, 480))]:
        img = make_frame(seed, w, h)
        kd, dd = drop(img)
        kr, dr = ref(img)
        assert len(kd) == len(kr) > 900
        for name in ("x", "y", "size", "angle", "response", "octave"):
            assert np.array_equal(kd[name], kr[name]), name
        assert 1.0 - np.unpackbits(dd ^ dr, axis=1).mean() >= 0.999
        for lvl in (0, 3, 7):
            assert np.array_equal(drop.pyramid_level(lvl), ref.pyramid_level(lvl))
    # strided ROI input (cv::Mat with step > cols), as Frame.cpp may hand over
    big = make_frame(104, 800, 500)
    roi = big[10:490, 20:772]
    kd, dd = drop(np.ascontiguousarray(roi))
    kr, dr = ref(np.ascontiguousarray(roi))
    assert len(kd) == len(kr) and np.array_equal(kd["x"], kr["x"]) and np.array_equal(kd["angle"], kr["angle"])


def test_async_submit_wait_pipeline(cuda_lib):
    """ccm_orb_extract_batch_submit / _wait: two batches in flight on alternating frame slots of the arena, results equal to
    the synchronous call for every batch of a stream (pinned outputs in the arena layout take the direct copy-out path,
    pageable ones of another capacity the staged one), slot overflow and a third ticket are refused."""
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200._lib import CCM_E_CAPACITY, CcmError
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    w, h, B, NB = 640, 400, 6, 5
    frames = make_stream(4321, B * NB, w, h).reshape(NB, B, h, w)
    ex = make(cuda_lib, 600, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2 * B + 1)
    cap = ex.max_keypoints()
    want = [ex.extract_batch(list(frames[b])) for b in range(NB)]
    for pinned, capacity in ((True, cap), (False, cap - 5), (True, 100)):
        pin = torch.from_numpy(frames.copy()).pin_memory() if pinned else torch.from_numpy(frames.copy())
        base = pin.numpy()
        mk = (lambda *s, dt: torch.zeros(s, dtype=dt).pin_memory()) if pinned else (lambda *s, dt: torch.zeros(s, dtype=dt))
        out_kp = [mk(B, capacity, 6, dt=torch.float32) for _ in range(2)]
        out_desc = [mk(B, capacity, 32, dt=torch.uint8) for _ in range(2)]
        out_n = [(ctypes.c_int * B)() for _ in range(2)]
        ptrs = [(ctypes.c_void_p * B)(*[base[b, f].ctypes.data for f in range(B)]) for b in range(NB)]

        def check(b):
            s = b & 1
            kp = out_kp[s].numpy().view(KP_DTYPE).reshape(B, capacity)
            for f in range(B):
                n = out_n[s][f]
                assert n == len(want[b][0][f])
                k = min(n, capacity)
                assert np.array_equal(kp[f, :k], want[b][0][f][:k]) and np.array_equal(out_desc[s].numpy()[f, :k], want[b][1][f][:k])

        tickets = {}
        for b in range(NB):
            s = b & 1
            tickets[b] = ex.extract_batch_submit(ptrs[b], B, h, w, w, out_kp[s].data_ptr(), out_desc[s].data_ptr(), capacity, out_n[s])
            if b >= 1:
                ex.extract_batch_wait(tickets[b - 1])
                check(b - 1)
        # both slots busy? no: one ticket open. A second fits, a third does not; nor does a batch larger than the free slots.
        t2 = ex.extract_batch_submit(ptrs[0], B, h, w, w, out_kp[NB & 1].data_ptr(), out_desc[NB & 1].data_ptr(), capacity, out_n[NB & 1])
        with pytest.raises(CcmError) as e:
            ex.extract_batch_submit(ptrs[1], 1, h, w, w, out_kp[0].data_ptr(), out_desc[0].data_ptr(), capacity, out_n[0])
        assert e.value.code == CCM_E_CAPACITY
        ex.extract_batch_wait(tickets[NB - 1])
        check(NB - 1)
        ex.extract_batch_wait(t2)
        with pytest.raises(CcmError):
            ex.extract_batch_wait(t2)                                       # already collected
    # frames in flight must fit max_batch
    small = make(cuda_lib, 600, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B + 2)
    o_kp, o_d, o_n = np.zeros((2, B, cap), KP_DTYPE), np.zeros((2, B, cap, 32), np.uint8), [(ctypes.c_int * B)() for _ in range(2)]
    fr = np.ascontiguousarray(frames[0])
    p0 = (ctypes.c_void_p * B)(*[fr[f].ctypes.data for f in range(B)])
    t = small.extract_batch_submit(p0, B, h, w, w, o_kp[0].ctypes.data, o_d[0].ctypes.data, cap, o_n[0])
    with pytest.raises(CcmError) as e:
        small.extract_batch_submit(p0, B, h, w, w, o_kp[1].ctypes.data, o_d[1].ctypes.data, cap, o_n[1])
    assert e.value.code == CCM_E_CAPACITY
    small.extract_batch_wait(t)
    assert [o_n[0][f] for f in range(B)] == [len(k) for k in want[0][0]]
    # a synchronous call afterwards still works and the stage taps still see frame 0 in slot 0
    k, d = small.extract_batch(list(frames[1]))
    assert all(np.array_equal(k[f], want[1][0][f]) for f in range(B))


def test_four_agents_share_one_gpu(cuda_lib):
    """BASELINE configs[3] on one GPU: four agent streams (seeds 1000..4000), one extractor handle and one host thread per agent
    (ClientSystem = one process per agent upstream), all running at once through the asynchronous pair. Every agent's output
    must be what its handle gives when it runs alone, and agent 0's first frame what the oracle gives."""
    import threading
    torch = pytest.importorskip("torch")
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    w, h, B, NB, A = 752, 480, 4, 3, 4
    streams = [make_stream(1000 * (a + 1), B * NB, w, h).reshape(NB, B, h, w) for a in range(A)]
    exs = [make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=2 * B) for _ in range(A)]
    cap = exs[0].max_keypoints()
    alone = [[exs[a].extract_batch(list(streams[a][b])) for b in range(NB)] for a in range(A)]
    got = [[None] * NB for _ in range(A)]
    errs = []

    def agent(a):
        try:
            pin = torch.from_numpy(streams[a].copy()).pin_memory()
            base = pin.numpy()
            okp = [torch.zeros((B, cap, 6), dtype=torch.float32).pin_memory() for _ in range(2)]
            ode = [torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory() for _ in range(2)]
            on = [(ctypes.c_int * B)() for _ in range(2)]
            ptrs = [(ctypes.c_void_p * B)(*[base[b, f].ctypes.data for f in range(B)]) for b in range(NB)]

            def collect(b):
                s = b & 1
                kp = okp[s].numpy().view(KP_DTYPE).reshape(B, cap)
                got[a][b] = ([kp[f, :on[s][f]].copy() for f in range(B)], [ode[s].numpy()[f, :on[s][f]].copy() for f in range(B)])

            for rep in range(3):                                   # keep the GPU shared for a while
                t_prev = None
                for b in range(NB):
                    s = b & 1
                    t = exs[a].extract_batch_submit(ptrs[b], B, h, w, w, okp[s].data_ptr(), ode[s].data_ptr(), cap, on[s])
                    if t_prev is not None:
                        exs[a].extract_batch_wait(t_prev)
                        collect(b - 1)
                    t_prev = t
                exs[a].extract_batch_wait(t_prev)
                collect(NB - 1)
        except Exception as e:                                     # noqa: BLE001
            errs.append((a, repr(e)))

    ths = [threading.Thread(target=agent, args=(a,)) for a in range(A)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    for a in range(A):
        for b in range(NB):
            for f in range(B):
                assert np.array_equal(got[a][b][0][f], alone[a][b][0][f]) and np.array_equal(got[a][b][1][f], alone[a][b][1][f]), (a, b, f)
    ko, do = OracleExtractor(1000, 1.2, 8, 20, 7)(streams[0][0, 0])
    assert_same(got[0][0][0][0], got[0][0][1][0], ko, do, "agent 0 frame 0")
    for e in exs:
        e.close()


def test_opencv2_blur_switch(cuda_lib, monkeypatch):
    """CCM_ORB_BLUR=opencv2 selects OpenCV 2.4.13's 8-bit Gaussian taps {18,34,49,55,..} (sum 257, saturated) in k_blur_levels:
    blurred levels and descriptors then equal the oracle run with the same variant; keypoints do not depend on the blur."""
    from _oracle import oracle_lib
    monkeypatch.setenv("CCM_ORB_BLUR", "opencv2")
    frames = make_stream(1000, 2)
    ex = make(cuda_lib, 1000, 1.2, 8, 20, 7, max_batch=2)
    orc = OracleExtractor(1000, 1.2, 8, 20, 7)
    L = oracle_lib()
    kps, descs = ex.extract_batch(list(frames))
    try:
        L.ocv_set_blur_variant(1)
        for f in range(2):
            ko, do = orc(frames[f])
            assert_same(kps[f], descs[f], ko, do, "opencv2 blur, frame %d" % f)
            for l in (0, 3, 7):
                assert np.array_equal(ex.pyramid_level(l, f, blurred=True), orc.level_image(l, True)), ("blur", f, l)
        L.ocv_set_blur_variant(0)
        ko, do = orc(frames[0])
        assert not np.array_equal(descs[0], do)            # and it is a different result from the default taps
    finally:
        L.ocv_set_blur_variant(0)
    ex.close()


@pytest.mark.gpu
def test_one_handle_changing_shapes_and_entry_points(cuda_lib):
    """One handle fed images of different sizes and through different entry points in turn: the single-frame CUDA graph is
    rebuilt when the shape (arena) changes, and a batch call in between leaves it valid."""
    ex = make(cuda_lib, 500, 1.2, 6, 20, 7, max_width=752, max_height=480, max_batch=4)
    orc = OracleExtractor(500, 1.2, 6, 20, 7)
    shapes = [(752, 480), (640, 400), (752, 480), (401, 303), (640, 400)]
    for k, (w, h) in enumerate(shapes):
        img = make_frame(50 + k, w, h)
        ko, do = orc(img)
        kp, desc = ex(img)                                   # graph path
        assert_same(kp, desc, ko, do, ("single", w, h, k))
        kps, descs = ex.extract_batch([img, img])            # batch path on the same handle and shape
        for f in range(2):
            assert_same(kps[f], descs[f], ko, do, ("batch", w, h, k, f))
        kp2, desc2 = ex(img)                                 # and the graph again
        assert np.array_equal(kp2, kp) and np.array_equal(desc2, desc)
    ex.close()
