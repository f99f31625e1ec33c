"""Single-frame latency of the host entry point (the call Frame::ExtractORB makes once per image): ms per ccm_orb_extract."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from motioncheck_ccm_slam_b200 import ORBextractor
from motioncheck_ccm_slam_b200.synth import make_stream

w, h = 752, 480
frames = make_stream(1000, 16, w, h)
pin = torch.from_numpy(frames).pin_memory()
pn = pin.numpy()
for mb in (1, 8):
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=mb, device=0)
    cap = ex.max_keypoints()
    okp = torch.zeros((1, cap, 6), dtype=torch.float32).pin_memory()
    ode = torch.zeros((1, cap, 32), dtype=torch.uint8).pin_memory()
    on = (ctypes.c_int * 1)()
    ptrs = [(ctypes.c_void_p * 1)(pn[f].ctypes.data) for f in range(16)]
    N = 300
    for name, call in (("ccm_orb_extract (graph)", lambda i: ex.extract_raw(pn[i % 16].ctypes.data, h, w, w, okp.data_ptr(), ode.data_ptr(), cap, on)),
                       ("ccm_orb_extract_batch n=1", lambda i: ex.extract_batch_raw(ptrs[i % 16], 1, h, w, w, okp.data_ptr(), ode.data_ptr(), cap, on))):
        for i in range(20):
            call(i)
        per = np.empty(N)
        t0 = time.perf_counter()
        for i in range(N):
            ta = time.perf_counter()
            call(i)
            per[i] = time.perf_counter() - ta
        dt = time.perf_counter() - t0
        q = np.percentile(per, [10, 50, 90]) * 1e6
        print("max_batch %d, %s: %.1f us per frame (p10 %.1f, p50 %.1f, p90 %.1f), %.0f frames/s, %d keypoints"
              % (mb, name, 1e6 * dt / N, q[0], q[1], q[2], N / dt, on[0]))
    ex.close()
