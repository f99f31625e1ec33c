"""ms per ccm_match_allpairs_best2_device call for a lone shard of NT descriptors (1000 queries): python tools/time_allpairs.py NT..."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from motioncheck_ccm_slam_b200.matcher import Matcher

gm = Matcher(0)
dev = torch.device("cuda", 0)
rng = np.random.default_rng(5)
q = torch.from_numpy(rng.integers(0, 256, (1000, 32), dtype=np.uint8)).to(dev)
out = torch.zeros((1000, 2), dtype=torch.int32, device=dev)
for nt in [int(a) for a in sys.argv[1:]] or [500000, 1000000, 2000000, 4000000]:
    t = torch.from_numpy(rng.integers(0, 256, (nt, 32), dtype=np.uint8)).to(dev)
    for _ in range(3):
        gm.match_allpairs_best2_device(q.data_ptr(), 1000, t.data_ptr(), nt, 0, out.data_ptr())
    gm.sync()
    import time
    t0 = time.perf_counter()
    for _ in range(20):
        gm.match_allpairs_best2_device(q.data_ptr(), 1000, t.data_ptr(), nt, 0, out.data_ptr())
    gm.sync()
    ms = 1e3 * (time.perf_counter() - t0) / 20
    print("nt %d: %.4f ms, %.1f Gpair/s" % (nt, ms, 1000 * nt / ms / 1e6))
