"""BASELINE.json configs[3] (C4): four agent streams sharing one GPU, each driven frame by frame through the host entry
points a drop-in Frame / Tracking would call — ccm_orb_extract, frame pack + grid, vocabulary transform,
SearchByBoW(KF,F) (nnratio 0.7, Tracking.cpp:521-524), SearchByProjection(F,LastFrame,7) (:582) and
SearchByProjection(F,MapPoints,1) (:797-805) — from four host threads at once (ClientSystem.cpp:29-69: one process per
agent upstream). Every search output of the sampled frames must equal the CPU oracle's: handles of different agents must
not disturb each other."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_four_agents_tracking_step_on_one_gpu(cuda_lib):
    torch = pytest.importorskip("torch")
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    import threading
    import bench_legs
    from motioncheck_ccm_slam_b200.synth import make_stream, make_vocabulary
    voc = make_vocabulary(77, 10, 4)
    T = 6
    agents = [bench_legs._Agent(a, 0, make_stream(1000 * (a + 1), T, 752, 480), voc, 1000, 8) for a in range(4)]
    for ag in agents:
        for t in range(T):
            ag.frame(t, record=True)                       # builds the queries of frames t+1, t+2 from frame t's keypoints
    sample = (2, 3, T - 1)
    ths = [threading.Thread(target=ag.run, kwargs={"sample": sample}) for ag in agents]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    for ag in agents:
        assert ag.done == 2 * T
        assert ag.check_against_oracle(voc, sample) == len(sample)
        nm = ag.nm[2:]
        assert nm[:, 0].mean() > 100 and nm[:, 1].mean() > 300 and nm[:, 2].mean() > 200, nm.mean(0)
    # the four streams are different scenes: their keypoints differ
    assert not np.array_equal(agents[0].kp[0]["x"][:200], agents[1].kp[0]["x"][:200])
    for ag in agents:
        ag.close()
