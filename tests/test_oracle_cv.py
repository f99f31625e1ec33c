"""Pins oracle/cv_primitives.cpp (CPU restatement of the OpenCV arithmetic on the hot path) bit-for-bit:
(a) against the committed cv2 4.13.0 golden vectors (runs anywhere), (b) live against cv2 when it is importable."""
import ctypes
import os

import numpy as np
import pytest

from _oracle import oracle_image_bounds, oracle_lib, oracle_undistort

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv_primitives.npz"))
u8p = ctypes.c_void_p


class Corner(ctypes.Structure):
    _fields_ = [("x", ctypes.c_int), ("y", ctypes.c_int), ("s", ctypes.c_int)]


@pytest.fixture(scope="module")
def L():
    lib = oracle_lib()
    lib.ocv_fast_atan2.restype = ctypes.c_float
    lib.ocv_fast_atan2.argtypes = [ctypes.c_float, ctypes.c_float]
    lib.ocv_resize_linear_u8.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t]
    lib.ocv_gaussian7_sigma2_u8.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_size_t]
    lib.ocv_border_reflect101_u8.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p, ctypes.c_size_t] + [ctypes.c_int] * 4
    lib.ocv_fast9_16.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    return lib


def resize(L, src, dw, dh):
    src = np.ascontiguousarray(src)
    out = np.zeros((dh, dw), np.uint8)
    L.ocv_resize_linear_u8(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], out.ctypes.data, dw, dh, dw)
    return out


def blur(L, src):
    src = np.ascontiguousarray(src)
    out = np.zeros_like(src)
    L.ocv_gaussian7_sigma2_u8(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], out.ctypes.data, out.strides[0])
    return out


def fast(L, src, th):
    buf = (Corner * (src.size + 1))()
    n = L.ocv_fast9_16(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], th, 1, buf, src.size)
    return np.array([[buf[i].x, buf[i].y, buf[i].s] for i in range(n)], np.int32).reshape(-1, 3)


def test_resize_golden(L):
    for i in range(3):
        dst = G["resize_dst_%d" % i]
        assert np.array_equal(resize(L, G["resize_src_%d" % i], dst.shape[1], dst.shape[0]), dst)


def test_blur_golden(L):
    for i in range(4):
        assert np.array_equal(blur(L, G["blur_src_%d" % i]), G["blur_dst_%d" % i])


def test_border_golden(L):
    src = np.ascontiguousarray(G["border_src"])
    out = np.zeros((src.shape[0] + 38, src.shape[1] + 38), np.uint8)
    L.ocv_border_reflect101_u8(src.ctypes.data, src.shape[1], src.shape[0], src.strides[0], out.ctypes.data, out.strides[0], 19, 19, 19, 19)
    assert np.array_equal(out, G["border_dst"])


def test_fast_golden_ordered(L):
    for i in range(5):
        src = np.ascontiguousarray(G["fast_src_%d" % i])
        for th in (20, 7):
            assert np.array_equal(fast(L, src, th), G["fast_kp_%d_%d" % (i, th)]), (i, th)


def test_fast_on_strided_roi(L):
    """cv::FAST is called on cell ROIs of a larger image (ORBextractor.cpp:978): step != width."""
    src = np.ascontiguousarray(G["fast_src_0"])
    big = np.zeros((src.shape[0] + 10, src.shape[1] + 20), np.uint8)
    big[5:5 + src.shape[0], 7:7 + src.shape[1]] = src
    roi = big[5:5 + src.shape[0], 7:7 + src.shape[1]]
    buf = (Corner * src.size)()
    n = L.ocv_fast9_16(roi.ctypes.data, roi.shape[1], roi.shape[0], roi.strides[0], 20, 1, buf, src.size)
    got = np.array([[buf[i].x, buf[i].y, buf[i].s] for i in range(n)], np.int32).reshape(-1, 3)
    assert np.array_equal(got, G["fast_kp_0_20"])


def test_atan_golden(L):
    got = np.array([L.ocv_fast_atan2(float(y), float(x)) for y, x in G["atan_yx"]], np.float32)
    assert np.array_equal(got, G["atan_deg"])
    true = np.degrees(np.arctan2(G["atan_yx"][:, 0].astype(np.float64), G["atan_yx"][:, 1].astype(np.float64))) % 360
    err = np.abs((got - true + 180) % 360 - 180)
    assert err.max() < 0.02   # fastAtan2 is an approximation (up to ~1.7e-4 rad off true atan2): replicate it, don't "fix" it


U = np.load(os.path.join(os.path.dirname(__file__), "golden", "undistort.npz"))
UNDISTORT_CAMERAS = ("euroc", "k3", "barrel", "strong")


@pytest.mark.parametrize("cam", UNDISTORT_CAMERAS)
def test_undistort_golden(cam):
    """cv::undistortPoints as Frame::UndistortKeyPoints / ComputeImageBounds call it (Frame.cpp:295, :321), bit for bit."""
    got = oracle_undistort(U[cam + "_pts"], U[cam + "_K"], U[cam + "_D"])
    assert np.array_equal(got.view(np.uint32), U[cam + "_und"].view(np.uint32))
    assert np.abs(got - U[cam + "_pts"]).max() > 1.0          # the model does move points
    w, h = U[cam + "_wh"]
    c = U[cam + "_corners_und"]
    b = oracle_image_bounds(U[cam + "_K"], U[cam + "_D"], w, h)
    assert np.array_equal(b[:4], np.array([min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])], np.float32))


def test_undistort_live_against_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(11)
    for cam in UNDISTORT_CAMERAS:
        K, D = U[cam + "_K"], U[cam + "_D"]
        w, h = U[cam + "_wh"]
        pts = rng.uniform(-40, [w + 40, h + 40], (50000, 2)).astype(np.float32)
        ref = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D.reshape(-1, 1), None, K).reshape(-1, 2)
        assert np.array_equal(oracle_undistort(pts, K, D).view(np.uint32), ref.view(np.uint32))


def test_live_against_cv2(L):
    cv2 = pytest.importorskip("cv2")
    cv2.setNumThreads(1)
    rng = np.random.default_rng(5)
    for (sw, sh, dw, dh) in [(752, 480, 627, 400), (627, 400, 522, 333), (1241, 376, 1034, 313), (302, 193, 252, 161)]:
        img = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        assert np.array_equal(resize(L, img, dw, dh), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR))
    for (w, h) in [(752, 480), (210, 134), (41, 9)]:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        assert np.array_equal(blur(L, img), cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101))
    for t in range(120):
        h, w = int(rng.integers(7, 46)), int(rng.integers(7, 46))
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        if t % 2:
            img = cv2.GaussianBlur(img, (5, 5), 1.2)
        if t % 5 == 0:
            img = (img // 64 * 64).astype(np.uint8)
        for th in (20, 7):
            det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True)
            ref = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in det.detect(img)], np.int32).reshape(-1, 3)
            assert np.array_equal(fast(L, img, th), ref)
    ys = rng.integers(-3000000, 3000000, 3000)
    xs = rng.integers(-3000000, 3000000, 3000)
    for y, x in zip(ys, xs):
        assert np.float32(L.ocv_fast_atan2(float(y), float(x))) == np.float32(cv2.fastAtan2(float(y), float(x)))


def test_opencv_2_4_13_blur_variant_distance():
    """The reference's CMakeLists pins OpenCV 2.4.13 (source zip missing); this repo pins the arithmetic to the cv2 4.13 wheel.
    The one primitive where the two differ is the 8-bit GaussianBlur: 2.4.13 rounds each tap times 256 on its own
    ({18,34,49,55,...}, sum 257) where 4.x normalises the 8.8 kernel to 256 ({18,34,48,56,...}). This test states the
    2.4.13 variant (from knowledge of its sources; no build here can confirm it) and MEASURES the distance on the test
    stream: blurred pixels and descriptor bits of the same keypoints. The figures are quoted in DESIGN.md section 6."""
    from _oracle import OracleExtractor, oracle_lib
    from motioncheck_ccm_slam_b200.synth import make_stream
    L = oracle_lib()
    frames = make_stream(1000, 4)
    ex = OracleExtractor(1000, 1.2, 8, 20, 7)
    px_diff = px_tot = bit_diff = bit_tot = kp_same = 0
    maxabs = 0
    try:
        for f in frames:
            L.ocv_set_blur_variant(0)
            k0, d0 = ex(f)
            b0 = [ex.level_image(l, True).astype(np.int32) for l in range(8)]
            L.ocv_set_blur_variant(1)
            k1, d1 = ex(f)
            b1 = [ex.level_image(l, True).astype(np.int32) for l in range(8)]
            assert len(k0) == len(k1) and all(np.array_equal(k0[n], k1[n]) for n in k0.dtype.names)   # the blur only feeds the descriptors
            kp_same += len(k0)
            for a, b in zip(b0, b1):
                px_diff += int((a != b).sum()); px_tot += a.size
                maxabs = max(maxabs, int(np.abs(a - b).max()))
            bit_diff += int(np.unpackbits(d0 ^ d1, axis=1).sum()); bit_tot += d0.size * 8
    finally:
        L.ocv_set_blur_variant(0)
    px_rate, bit_rate = px_diff / px_tot, bit_diff / bit_tot
    print("2.4.13 vs 4.13 blur: %.1f %% of blurred pixels differ (max |d| = %d), %.3f %% of descriptor bits differ on %d keypoints"
          % (100 * px_rate, maxabs, 100 * bit_rate, kp_same))
    # a gain of 257/256 per pass: most pixels move by 1-4 grey levels, but a monotone change rarely flips a BRIEF comparison
    assert 0.3 < px_rate < 1.0 and maxabs <= 5
    assert bit_rate < 0.02
