"""Generates the committed golden fixtures. Run in the build container (needs cv2 4.13.0 and /root/reference):
    python tests/golden/make_golden.py
* cv_primitives.npz  — inputs and cv2 4.13.0 outputs for resize / GaussianBlur / copyMakeBorder / FAST / fastAtan2
* undistort.npz      — cv2 4.13.0 cv::undistortPoints(pts, K, D, None, K) on keypoint-like points for four camera models
                       (EuRoC vi_euroc.yaml, a 5-coefficient one, barrel, strong pincushion) + their image-corner bounds
* ref_extract.npz    — keypoints + descriptors from the reference's own ORBextractor.cpp compiled unmodified
                       (oracle/_ref/libref_orb_mono.so, monotone allocator; see DESIGN.md quadtree tie rule)
Seeds are fixed; re-running reproduces the files bit for bit."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def cv_fixtures():
    import cv2
    cv2.setNumThreads(1)
    assert cv2.__version__ == "4.13.0", cv2.__version__
    rng = np.random.default_rng(42)
    out = {}
    # resize: the last C1 level pair and a small odd pair, on noise and on a smooth gradient
    for i, (sw, sh, dw, dh) in enumerate([(252, 161, 210, 134), (97, 61, 81, 51), (160, 120, 133, 100)]):
        src = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        if i == 2:
            yy, xx = np.mgrid[0:sh, 0:sw]
            src = ((xx * 3 + yy * 5) % 256).astype(np.uint8)
        out["resize_src_%d" % i] = src
        out["resize_dst_%d" % i] = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
    # blur: noise, impulse, step, tiny
    imgs = [rng.integers(0, 256, (70, 90), dtype=np.uint8), np.zeros((21, 21), np.uint8), np.zeros((16, 40), np.uint8),
            rng.integers(0, 256, (8, 9), dtype=np.uint8)]
    imgs[1][10, 10] = 255
    imgs[2][:, 20:] = 200
    for i, im in enumerate(imgs):
        out["blur_src_%d" % i] = im
        out["blur_dst_%d" % i] = cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    b = rng.integers(0, 256, (30, 41), dtype=np.uint8)
    out["border_src"] = b
    out["border_dst"] = cv2.copyMakeBorder(b, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    # FAST on cell-sized ROIs: noise, blurred noise, flat, plateau ties, posterised
    rois = [rng.integers(0, 256, (38, 36), dtype=np.uint8),
            cv2.GaussianBlur(rng.integers(0, 256, (38, 36), dtype=np.uint8), (5, 5), 1.2),
            np.full((36, 36), 77, np.uint8),
            (rng.integers(0, 4, (40, 37)) * 64).astype(np.uint8),
            np.kron(rng.integers(0, 2, (10, 10)) * 180 + 30, np.ones((4, 4))).astype(np.uint8)]
    for i, r in enumerate(rois):
        out["fast_src_%d" % i] = r
        for th in (20, 7):
            det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True, type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            k = det.detect(r)
            out["fast_kp_%d_%d" % (i, th)] = np.array([[int(p.pt[0]), int(p.pt[1]), int(p.response)] for p in k], np.int32).reshape(-1, 3)
    # fastAtan2: dense small lattice + random moments in the IC_Angle range
    ys, xs = np.mgrid[-24:25, -24:25]
    yx = np.concatenate([np.stack([ys.ravel(), xs.ravel()], 1), rng.integers(-2900000, 2900000, (4000, 2))]).astype(np.float32)
    out["atan_yx"] = yx
    out["atan_deg"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in yx], np.float32)
    np.savez_compressed(os.path.join(HERE, "cv_primitives.npz"), **out)


UNDISTORT_CAMERAS = [
    # fx, fy, cx, cy, dist, image (w, h)
    ("euroc", 458.654, 457.296, 367.215, 248.375, [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], (752, 480)),   # conf/vi_euroc.yaml
    ("k3", 458.654, 457.296, 367.215, 248.375, [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.013], (752, 480)),
    ("barrel", 718.856, 718.856, 607.1928, 185.2157, [0.11, -0.23, 0.0012, -0.0021], (1241, 376)),
    ("strong", 300.0, 310.0, 320.5, 239.5, [-0.9, 0.5, 0.01, 0.01, 0.3], (640, 480)),                                    # drives icdist < 0 at the rim
]


def undistort_fixtures():
    import cv2
    assert cv2.__version__ == "4.13.0", cv2.__version__
    rng = np.random.default_rng(77)
    out = {}
    for name, fx, fy, cx, cy, dist, (w, h) in UNDISTORT_CAMERAS:
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        D = np.array(dist, np.float32).reshape(-1, 1)
        # keypoint-like positions: integer level-0 pixels, higher-level pixels times a scale factor, and a few out-of-image points
        pts = np.concatenate([rng.integers(0, [w, h], (1500, 2)).astype(np.float32),
                              (rng.integers(0, [w // 2, h // 2], (1500, 2)) * np.float32(1.2) ** rng.integers(1, 8, (1500, 1))).astype(np.float32),
                              rng.uniform(-60, [w + 60, h + 60], (200, 2)).astype(np.float32)])
        corners = np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32)          # Frame::ComputeImageBounds, Frame.cpp:313-317
        out[name + "_K"] = K
        out[name + "_D"] = D.ravel()
        out[name + "_wh"] = np.array([w, h], np.int32)
        out[name + "_pts"] = pts
        out[name + "_und"] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
        out[name + "_corners_und"] = cv2.undistortPoints(corners.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    np.savez_compressed(os.path.join(HERE, "undistort.npz"), **out)


def ref_fixtures():
    from _oracle import RefExtractor
    from motioncheck_ccm_slam_b200.synth import make_frame
    out = {}
    cases = [("c1", 752, 480, 1000, 8, 1000), ("small", 320, 240, 300, 4, 7), ("kitti_half", 620, 188, 500, 5, 3000)]
    for name, w, h, nf, nl, seed in cases:
        img = make_frame(seed, w, h)
        kp, desc = RefExtractor(nf, 1.2, nl, 20, 7, mono=True)(img)
        out[name + "_cfg"] = np.array([w, h, nf, nl, seed], np.int32)
        out[name + "_kp"] = kp
        out[name + "_desc"] = desc
    np.savez_compressed(os.path.join(HERE, "ref_extract.npz"), **out)


if __name__ == "__main__":
    cv_fixtures()
    undistort_fixtures()
    ref_fixtures()
    for f in ("cv_primitives.npz", "undistort.npz", "ref_extract.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
