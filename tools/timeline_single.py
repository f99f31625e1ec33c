"""Device-side timeline of the single-frame graph: kernel start stamps (globaltimer) written by a DEBUG build of the library
(CCM_NVCC_EXTRA=-DCCM_DEBUG_TS python -c 'from motioncheck_ccm_slam_b200 import build as b; b.build(force=True)'), median over
40 calls, us relative to the start of FAST on level 0. Rebuild without the flag afterwards."""
import ctypes, sys
import numpy as np, torch
sys.path.insert(0, ".")
from motioncheck_ccm_slam_b200 import ORBextractor, _lib
from motioncheck_ccm_slam_b200.synth import make_stream
w, h = 752, 480
frames = make_stream(1000, 4, w, h)
pin = torch.from_numpy(frames).pin_memory(); pn = pin.numpy()
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1, device=0)
cap = ex.max_keypoints()
okp = torch.zeros((1, cap, 6), dtype=torch.float32).pin_memory(); ode = torch.zeros((1, cap, 32), dtype=torch.uint8).pin_memory()
on = (ctypes.c_int * 1)()
L = ctypes.CDLL(_lib.LIB_PATH)
ts = (ctypes.c_ulonglong * 128)()
acc = []
for i in range(60):
    ex.extract_raw(pn[i % 4].ctypes.data, h, w, w, okp.data_ptr(), ode.data_ptr(), cap, on)
    L.ccm_debug_ts(ts)
    t = np.array(ts[:], dtype=np.float64)
    if i >= 20: acc.append(t)
t = np.median(np.array(acc) - np.array(acc)[:, 16:17], axis=0) / 1e3     # us relative to FAST level 0 start
print("resize start  ", [round(x, 1) for x in t[1:8]])
print("FAST start    ", [round(x, 1) for x in t[16:24]])
print("quadtree start", [round(x, 1) for x in t[32:40]])
print("quadtree end  ", [round(x, 1) for x in t[48:56]])
print("blur start", round(t[65], 1), "describe start", round(t[64], 1))
