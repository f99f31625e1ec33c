#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 ORB front-end (BASELINE.json metric / config).

One "step" = one pass of the hot path over one batch of synthetic frames of one agent stream:
ORBextractor::operator() on BATCH frames (752x480, 1000 features, 8 levels, 1.2 — EuRoC settings, configs[1]),
followed (when the matcher library is present) by the tracking-side Hamming searches on the extracted frames.
  value        frames/s with the input frames already resident in HBM (device-resident C-ABI entry point)
  e2e          frames/s through the host-buffer C-ABI call a user of the reference interface makes
               (pinned host images in, keypoints + descriptors out, copies inside the timed region)
  roofline     dominant kernel: algorithmic bytes per launch / its CUDA-event time, against MEASURED_PEAKS.json
  cpu_baseline the reference's own ORBextractor.cpp (oracle/_ref, compiled unmodified) on the host cores
Launch: python bench.py [--gpus N --steps K --warmup W]; for N>1 via torch.distributed.run (one rank per GPU,
one agent stream per rank, no data-path collective: "scaling": "weak").  --impl reference times the CPU reference.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (width, height, nfeatures, nlevels)
    "c1_euroc_752x480_1000kp": (752, 480, 1000, 8),
    "c3_kitti_1241x376_2000kp": (1241, 376, 2000, 8),
}
SCALE, INI_TH, MIN_TH = 1.2, 20, 7
METRIC = "ORB frames/s @752x480,1k kp"
STAGES = ["pyramid_resize", "fast_cells", "gaussian_blur", "quadtree", "orient_describe"]


def level_sizes(w, h, nlevels, scale=SCALE):
    """Level geometry exactly as ORBextractor.cpp:579-592,1284-1286 computes it (float/double mix kept)."""
    sf = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(scale))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(w, h, nfeatures, nlevels):
    """Bytes per frame each stage must move at minimum, and the pipeline total."""
    px = [a * b for a, b in level_sizes(w, h, nlevels)]
    tot = sum(px)
    per = {
        "pyramid_resize": sum(px[:-1]) + sum(px[1:]),        # read level l-1, write level l
        "fast_cells": tot,                                   # read every level once
        "gaussian_blur": 2 * tot,                            # read level, write blurred level
        "quadtree": 0,                                       # corner lists: latency-bound, no roofline claimed
        "orient_describe": nfeatures * (749 + 512 + 56),     # patch + taps + outputs per keypoint
    }
    pipeline = w * h + 4 * tot + nfeatures * 56              # SURVEY §8d un-fused figure (4.89 MB at C1)
    return per, pipeline


def issue_roofline(workload, B, stage_ms, sms, w, h, nl):
    """The roofline the kernels actually live under: warp instructions per launch (from the committed ncu --set full capture,
    profiles/issue.json) / (SMs x 4 schedulers x SM clock x the stage's live event time)."""
    p = os.path.join(ROOT, "profiles", "issue.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        ij = json.load(f)
    if ij.get("workload") != workload or ij.get("frames_per_launch") != B:
        return None
    px = sum(a * b for a, b in level_sizes(w, h, nl)) * B
    mhz = float(ij.get("sm_mhz", 1965.0))
    out = {"source": ij.get("source"), "unit": "fraction of issue slots (SMs x 4 x %.0f MHz)" % mhz, "stages": {}}
    tot_i = tot_slots = 0.0
    for st, wi in ij["warp_instructions"].items():
        ms = stage_ms.get(st, 0.0)
        if ms <= 0:
            continue
        slots = sms * 4 * mhz * 1e6 * ms * 1e-3
        out["stages"][st] = {"warp_instructions": wi, "frac": wi / slots, "thread_instructions_per_pixel": 32.0 * wi / px}
        tot_i += wi; tot_slots += slots
    out["warp_instructions_per_step"] = tot_i
    out["frac"] = tot_i / tot_slots if tot_slots else None
    out["budget"] = ij.get("budget")
    return out


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# CPU leg: the reference's own ORBextractor.cpp (oracle/_ref/libref_orb.so, compiled unmodified over the
# OpenCV-4.13-exact shim), one extractor per host thread, one agent stream per thread. This is the ONLY place
# bench.py touches oracle/: as the reported baseline, never as the measured product.
# ----------------------------------------------------------------------------------------------------------
def cpu_reference_fps(frames, nfeatures, nlevels, threads, frames_per_thread):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    if _oracle.ref_available(mono=False):
        kind, mk = "reference", lambda: _oracle.RefExtractor(nfeatures, SCALE, nlevels, INI_TH, MIN_TH, mono=False)
    else:
        kind, mk = "port", lambda: _oracle.OracleExtractor(nfeatures, SCALE, nlevels, INI_TH, MIN_TH)
    exs = [mk() for _ in range(threads)]
    exs[0](frames[0])                                   # warm-up (page-in, allocator)
    done = [0] * threads

    def work(t):
        for i in range(frames_per_thread):
            exs[t](frames[(t * frames_per_thread + i) % len(frames)])   # ctypes releases the GIL during the call
            done[t] += 1

    ths = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    return sum(done) / dt, dt, kind


def run_reference(args, wl, rank):
    """--impl reference: the reference's CPU implementation of the path on all host threads; rank 0 only."""
    if rank != 0:
        return
    from motioncheck_ccm_slam_b200.synth import make_stream
    w, h, nf, nl = WORKLOADS[wl]
    cores = os.cpu_count() or 1
    frames = make_stream(1000, 16, w, h)
    per_step = 8                                         # frames per thread per step: a bounded sample of the 64-frame batch
    for _ in range(max(args.warmup, 0)):
        cpu_reference_fps(frames, nf, nl, cores, 1)
    tot_frames, tot_t, kind = 0, 0.0, "reference"
    for _ in range(args.steps):
        fps, dt, kind = cpu_reference_fps(frames, nf, nl, cores, per_step)
        tot_frames += per_step * cores
        tot_t += dt
    v = tot_frames / tot_t
    sample = "%d steps x %d threads x %d frames of the %s stream (seed 1000)" % (args.steps, cores, per_step, wl)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl, "frames_per_step": per_step * cores, "host": "cpu"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------------------------------
# Hamming legs of the metric ("Hamming Gmatch/s"): the server keyframe-database search (configs[4]: 1000 query
# descriptors against 8 x 500 KF x 1000 = 4.0 M target descriptors, sharded by rank + NCCL allgather of the per-shard
# best-2 records + merge) and the tracking-side windowed searches on the extractor's own output (configs[1]/[3]).
# ----------------------------------------------------------------------------------------------------------
def int_pipe_peak_gpairs(sms, sm_mhz):
    """POPC-pipe roofline for the textbook loop (8 XOR + 8 POPC per pair): POPC issues 16 lanes/clk/SM on B200
    (tools/ubench_int.cu -> profiles/r1_ubench_int.json: 15.9 measured)."""
    return sms * 16 * sm_mhz * 1e6 / 8 / 1e9


def hamming_legs(args, torch, dist, dev, rank, world, ex, ext, d_frames, d_kp, d_desc, d_n, cap, w, h, pitch, barrier):
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    from motioncheck_ccm_slam_b200.matcher import (Exchange, FramePack, Matcher, QUERY_DTYPE, Q_OBS, Q_VALID, WIN_PROJ_LASTFRAME,
                                                   WIN_PROJ_MAPPOINTS, frame_grid, packs_from_extractor_batch_device,
                                                   search_window_batch_device)
    out = {}
    m = Matcher(dev.index)
    ms = torch.cuda.ExternalStream(m.stream(), device=dev)
    B = d_kp.shape[0]

    # ---- (1) all-pairs best-2 over the keyframe database, target shard per rank ----
    nq, nt_total = 1000, args.db
    from motioncheck_ccm_slam_b200.sharding import shard_range
    lo, hi = shard_range(nt_total, rank, world)                    # database sharded across ranks
    g = torch.Generator(device=dev); g.manual_seed(7)
    d_q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)     # same queries on every rank
    g.manual_seed(100 + rank)
    d_t = torch.randint(0, 256, (hi - lo, 32), dtype=torch.uint8, device=dev, generator=g)
    d_part = torch.zeros((nq, 8), dtype=torch.uint8, device=dev)
    d_all = torch.zeros((world, nq, 8), dtype=torch.uint8, device=dev)
    d_best = torch.zeros((nq, 8), dtype=torch.uint8, device=dev)
    d_acc = torch.zeros((nq,), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()

    def ap_step_nccl():
        m.match_allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), hi - lo, lo, d_part.data_ptr())
        if world > 1:
            with torch.cuda.stream(ms):
                dist.all_gather_into_tensor(d_all, d_part)       # 8 KB per rank
            m.match_allpairs_merge_device(d_all.data_ptr(), world, nq, 0.75, d_best.data_ptr(), d_acc.data_ptr())
        else:
            m.match_allpairs_merge_device(d_part.data_ptr(), 1, nq, 0.75, d_best.data_ptr(), d_acc.data_ptr())

    # world > 1: the search, the exchange of the per-shard best-two records (stores into every peer's buffer over
    # NVLink, release/acquire flags) and the merge are ONE kernel; the NCCL all_gather + merge pair is timed beside it
    xch = None
    if world > 1:
        xch = Exchange(m, rank, world, nq)
        hs = [None] * world
        dist.all_gather_object(hs, xch.handle.tobytes())
        xch.connect_ipc(np.frombuffer(b"".join(hs), np.uint8).reshape(world, 64))
        d_best_x = torch.zeros_like(d_best); d_acc_x = torch.zeros_like(d_acc)
        barrier()

    def ap_step_fused():
        xch.allpairs_best2_device(d_q.data_ptr(), nq, d_t.data_ptr(), hi - lo, lo, 0.75, d_best_x.data_ptr(), d_acc_x.data_ptr())

    def time_ap(step):
        for _ in range(3):
            step()
        m.sync(); barrier()
        l0 = m.launch_count()
        with torch.cuda.stream(ms):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(ksteps):
                step()
            e1.record()
        m.sync(); barrier()
        return e0.elapsed_time(e1) / ksteps, m.launch_count() - l0

    ksteps = max(3, min(args.steps, 10))
    ap_ms, ap_launches = time_ap(ap_step_nccl)
    collective = "none (1 shard)"
    if world > 1:
        nccl_ms = ap_ms
        ap_ms, ap_launches = time_ap(ap_step_fused)
        same = bool(torch.equal(d_best, d_best_x) and torch.equal(d_acc, d_acc_x))
        if not same:
            raise SystemExit("fused exchange search differs from all_gather + merge on rank %d" % rank)
        xch.check()                                       # raises if a wait for a peer ever timed out
        collective = ("chunk fold + peer-memory exchange + merge in ONE kernel after the search launches: %d B per rank stored into every "
                      "peer's buffer over NVLink + flags; nccl all_gather + merge kernel beside it: %.4f ms on rank 0 (identical result)" % (nq * 8, nccl_ms))
        out["_ap_nccl_ms"] = nccl_ms
    out["_ap_ms"] = ap_ms
    out["allpairs"] = {"nq": nq, "nt_total": nt_total, "nt_per_gpu": hi - lo, "steps": ksteps,
                       "collective": collective,
                       "l2": "target shard %.0f MB streamed once per step" % ((hi - lo) * 32 / 1e6)}
    if world > 1:
        out["allpairs"]["nccl_allgather_merge_ms_rank0"] = nccl_ms

    # ---- (2) tracking-side searches on the extractor's output: pack build + LastFrame(th 7) + MapPoints(th 1) ----
    R = min(d_frames.shape[0], 2)
    grid = frame_grid(w, h)
    packs = [FramePack(m, cap) for _ in range(B)]
    sf = ex.GetScaleFactors()
    q_sets = []
    for r in range(R):
        ex.extract_batch_device(d_frames[r].data_ptr(), B, h, w, pitch, pitch * h, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        ex.sync()
        kp = d_kp.cpu().numpy().view(KP_DTYPE).reshape(B, cap)
        n = d_n.cpu().numpy()
        q1 = np.zeros((B, cap), QUERY_DTYPE); q2 = np.zeros((B, cap), QUERY_DTYPE)
        for j in range(B):
            k = kp[j, :n[j]]
            # host geometry of the reference loops: last frame's points moved by the stream's motion (-2,-1) per frame
            q1["u"][j, :n[j]] = k["x"] - 2; q1["v"][j, :n[j]] = k["y"] - 1; q1["angle"][j, :n[j]] = k["angle"]
            q1["r"][j, :n[j]] = np.float32(7.0) * sf[k["octave"]]
            q1["min_level"][j, :n[j]] = k["octave"] - 1; q1["max_level"][j, :n[j]] = k["octave"] + 1
            q1["flags"][j, :n[j]] = Q_VALID | Q_OBS
            q2["u"][j, :n[j]] = k["x"] - 4; q2["v"][j, :n[j]] = k["y"] - 2
            q2["r"][j, :n[j]] = np.float32(2.5) * sf[k["octave"]]            # RadiusByViewingCos(1.0) * th(1) * scale
            q2["min_level"][j, :n[j]] = k["octave"] - 1; q2["max_level"][j, :n[j]] = k["octave"]
            q2["flags"][j, :n[j]] = Q_VALID | Q_OBS
        q_sets.append((torch.from_numpy(q1.view(np.uint8).reshape(B, cap, 24)).to(dev),
                       torch.from_numpy(q2.view(np.uint8).reshape(B, cap, 24)).to(dev)))
    d_holder = torch.zeros((B, cap), dtype=torch.int32, device=dev)
    d_nm = torch.zeros((2, B), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def track_step(i):
        r = i % R
        ex.extract_batch_device(d_frames[r].data_ptr(), B, h, w, pitch, pitch * h, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())
        ms.wait_stream(ext)
        with torch.cuda.stream(ms):
            d_holder.fill_(-1)
        packs_from_extractor_batch_device(packs, grid, d_kp.data_ptr(), d_desc.data_ptr(), d_n.data_ptr(), cap)
        # frame t searched with frame t-1's features (jobs 1..B-1), then with frame t-2's as "map points" (jobs 2..B-1)
        search_window_batch_device(m, packs[1:], WIN_PROJ_LASTFRAME, 0.9, 100, True, d_n.data_ptr(), cap, q_sets[r][0].data_ptr(),
                                   d_desc.data_ptr(), d_holder[1:].data_ptr(), cap, None, d_nm[0, 1:].data_ptr())
        search_window_batch_device(m, packs[2:], WIN_PROJ_MAPPOINTS, 0.8, 100, True, d_n.data_ptr(), cap, q_sets[r][1].data_ptr(),
                                   d_desc.data_ptr(), d_holder[2:].data_ptr(), cap, None, d_nm[1, 2:].data_ptr())
        ext.wait_stream(ms)       # the next batch overwrites d_kp / d_desc

    for i in range(3):
        track_step(i)
    m.sync(); ex.sync(); barrier()
    l0 = m.launch_count()
    with torch.cuda.stream(ext):
        e0 = torch.cuda.Event(enable_timing=True); e0.record()
    for i in range(args.steps):
        track_step(3 + i)
    with torch.cuda.stream(ext):
        e1 = torch.cuda.Event(enable_timing=True); e1.record()
    m.sync(); ex.sync(); barrier()
    tr_ms = e0.elapsed_time(e1) / args.steps
    nm = d_nm.cpu().numpy()
    out["_tr_ms"] = tr_ms
    out["tracking"] = {"what": "extract + device frame packs (75x48 grid) + SearchByProjection(F,LastF,th=7) + SearchByProjection(F,mappoints,th=1), "
                               "%d frames per step, queries = previous frames' keypoints moved by the stream motion" % B,
                       "matches_per_frame_lastframe": float(nm[0, 1:].mean()), "matches_per_frame_mappoints": float(nm[1, 2:].mean()),
                       "matcher_launches_per_step": (m.launch_count() - l0) / args.steps}
    out["_launches"] = ap_launches + (m.launch_count() - l0)

    # ---- (3) Frame::ComputeBoW's vocabulary descent on the last batch's descriptor rows, in place ----
    from motioncheck_ccm_slam_b200.matcher import Vocabulary
    from motioncheck_ccm_slam_b200.synth import make_vocabulary_complete
    voc = make_vocabulary_complete(77, 10, 6)           # the ORB vocabulary's shape (k = 10, L = 6); the blob itself is missing
    V = Vocabulary(m, voc)
    d_word = torch.zeros((B * cap,), dtype=torch.int32, device=dev)
    d_node = torch.zeros((B * cap,), dtype=torch.int32, device=dev)
    n_feat = int(d_n.sum().item())
    for _ in range(3):
        V.transform_device(B * cap, d_desc.data_ptr(), d_word.data_ptr(), d_node.data_ptr(), 4)
    m.sync(); barrier()
    l0 = m.launch_count()
    with torch.cuda.stream(ms):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(ksteps):
            V.transform_device(B * cap, d_desc.data_ptr(), d_word.data_ptr(), d_node.data_ptr(), 4)
        e1.record()
    m.sync(); barrier()
    bow_ms = e0.elapsed_time(e1) / ksteps
    out["_launches"] += m.launch_count() - l0
    out["bow_transform"] = {"what": "DBoW2 transform(levelsup 4) descent of every descriptor row of a %d-frame batch (%d rows, %d keypoints), "
                                    "descriptors read in place from the extractor's output" % (B, B * cap, n_feat),
                            "vocabulary": "synthetic k=10 L=6, %d nodes, %.1f MB of node descriptors (L2-resident)" % (len(voc["word"]), (len(voc["word"]) - 1) * 32 / 1e6),
                            "ms_per_step": bow_ms, "descriptors_per_s": B * cap * world / (bow_ms * 1e-3),
                            "distances_per_s": B * cap * world * 60 / (bow_ms * 1e-3)}
    out["_voc"] = voc
    V.close()
    for p in packs:
        p.close()
    if xch is not None:
        barrier()                                         # no peer may still be storing into this rank's buffer
        xch.close()
    m.close()
    return out


def cpu_matcher_baseline(frames, nf, nl, n_frames=8):
    """The two tracking-side searches of the GPU leg on the reference's own ORBmatcher.cpp (oracle/_ref/libref_match.so), one
    thread, SAME queries as hamming_legs: frame t searched with frame t-1's points moved by (-2,-1) (LastFrame, th 7) and with
    frame t-2's points moved by (-4,-2) (MapPoints, th 1). Falls back to the oracle port when the library is not built."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    from motioncheck_ccm_slam_b200.matcher import frame_grid
    h, w = frames.shape[1:]
    ex = _oracle.OracleExtractor(nf, SCALE, nl, INI_TH, MIN_TH)
    sf = ex.tables()[0][0]
    grid = frame_grid(w, h)
    ext = [ex(frames[i]) for i in range(n_frames)]
    ref = _oracle.ref_match_available()
    M9, M8 = (_oracle.RefMatcher(0.9, True), _oracle.RefMatcher(0.8, True)) if ref else (_oracle.OracleMatcher(0.9, True), _oracle.OracleMatcher(0.8, True))
    mk = (lambda k, d: M9.frame(dict(desc=d, x=k["x"], y=k["y"], octave=k["octave"], angle=k["angle"]), grid)) if ref else \
         (lambda k, d: _oracle.OracleFrame(d, k["x"], k["y"], k["octave"], k["angle"], grid))
    Fs = [mk(*ext[i]) for i in range(n_frames)]
    t0 = time.perf_counter()
    nm1 = nm2 = 0
    for i in range(2, n_frames):
        (k1, d1), (k2, d2) = ext[i - 1], ext[i - 2]
        hold = np.full(len(ext[i][0]), -1, np.int32)
        one = np.ones(len(k1), np.uint8)
        r = M9.projection_lastframe(Fs[i], one, k1["x"] - 2, k1["y"] - 1, k1["octave"], k1["angle"], d1, one, 7.0, sf, hold)
        nm1 += r if ref else r[0]
        hold[:] = -1
        one = np.ones(len(k2), np.uint8)
        r = M8.projection_mappoints(Fs[i], one, k2["x"] - 4, k2["y"] - 2, k2["octave"], one.astype(np.float32), d2, one, 1.0, sf, hold)
        nm2 += r if ref else r[0]
    dt = time.perf_counter() - t0
    n = n_frames - 2
    return n / dt, nm1 / n, nm2 / n, "reference" if ref else "port"


def cpu_allpairs_baseline(nq=64, nt=200000):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    rng = np.random.default_rng(7)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    om = _oracle.OracleMatcher(0.75)
    t0 = time.perf_counter()
    om.allpairs_best2(q, t)
    return nq * nt / (time.perf_counter() - t0) / 1e9


def bind_to_gpu_numa(local_rank):
    """Agent-to-GPU pinning on the host side: run this rank (and first-touch its pinned buffers) on the CPUs of the NUMA node
    its GPU hangs off, so that 8 ranks do not push their copies across the socket link. Best effort; returns the node or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local_rank]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else local_rank
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as f:
            node = int(f.read())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


# ----------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------
def run_ours(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from motioncheck_ccm_slam_b200 import ORBextractor
    from motioncheck_ccm_slam_b200.extractor import KP_DTYPE
    from motioncheck_ccm_slam_b200.synth import make_stream

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU baseline)")
    numa_node = bind_to_gpu_numa(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w, h, nf, nl = WORKLOADS[wl]
    B, R = args.batch, args.rotate
    per_stage_bytes, pipeline_bytes = algorithmic_bytes(w, h, nf, nl)

    # one agent stream per rank: R rotating batches of B frames, so that inputs exceed the 126 MB L2
    frames = np.stack([make_stream(1000 * (rank + 1) + 17 * r, B, w, h) for r in range(R)])        # [R,B,h,w] u8
    ex = ORBextractor(nf, SCALE, nl, INI_TH, MIN_TH, max_width=w, max_height=h, max_batch=2 * B, device=local_rank)   # two host batches in flight
    cap = ex.max_keypoints()
    ext = torch.cuda.ExternalStream(ex.stream(), device=dev)

    pitch = (w + 31) // 32 * 32                                                                      # device rows start on 32-B sectors
    d_frames = torch.zeros((R, B, h, pitch), dtype=torch.uint8, device=dev)
    d_frames[..., :w] = torch.from_numpy(frames).to(dev)
    d_kp = torch.zeros((B, cap, 6), dtype=torch.float32, device=dev)
    d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros((B,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def step_device(i):
        src = d_frames[i % R]
        ex.extract_batch_device(src.data_ptr(), B, h, w, pitch, pitch * h, d_kp.data_ptr(), d_desc.data_ptr(), cap, d_n.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (value) ----
    for i in range(args.warmup):
        step_device(i)
    ex.sync()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = ex.launch_count()
    barrier()
    with torch.cuda.stream(ext):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            step_device(args.warmup + i)
        e1.record()
    ex.sync()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    launches = ex.launch_count() - l0
    # per-stage attribution (roofline of the dominant kernel): the same steps again with the library's CUDA events around each
    # stage; in this mode the stages run back to back on one stream (the timed region above overlaps blur with FAST + quadtree)
    ex.set_stage_timing(True)
    for i in range(args.steps):
        step_device(args.warmup + i)
    ex.sync()
    stage_ms, stage_calls = ex.get_stage_ms(reset=True)
    ex.set_stage_timing(False)
    n_kp = int(d_n.sum().item())

    # ---- end to end through the host-buffer C-ABI call (pinned host images in, host keypoints/descriptors out) ----
    pin = torch.from_numpy(frames).pin_memory()
    pin_np = pin.numpy()
    out_kp = torch.zeros((B, cap, 6), dtype=torch.float32).pin_memory()
    out_desc = torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory()
    out_n = (ctypes.c_int * B)()

    host_ptrs = [(ctypes.c_void_p * B)(*[pin_np[r, f].ctypes.data for f in range(B)]) for r in range(R)]

    # The stream of batches goes through the asynchronous pair of the host entry point (ccm_orb_extract_batch_submit / _wait):
    # batch i+1 is submitted before batch i is collected, so its copy-in runs while batch i is still in the kernels. Every
    # step's H2D and D2H copies are inside the timed region; outputs alternate between two pinned buffer sets.
    out_kp2 = [out_kp, torch.zeros((B, cap, 6), dtype=torch.float32).pin_memory()]
    out_desc2 = [out_desc, torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory()]
    out_n2 = [out_n, (ctypes.c_int * B)()]

    def submit_host(i):
        s = i & 1
        return ex.extract_batch_submit(host_ptrs[i % R], B, h, w, w, out_kp2[s].data_ptr(), out_desc2[s].data_ptr(), cap, out_n2[s])

    def run_host(first, count):
        t = submit_host(first)
        for i in range(first + 1, first + count):
            t_next = submit_host(i)
            ex.extract_batch_wait(t)
            t = t_next
        ex.extract_batch_wait(t)

    run_host(0, max(args.warmup, 1))
    barrier()
    t0 = time.perf_counter()
    run_host(args.warmup, args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    # the same stream through the synchronous call (one batch at a time, nothing overlaps across calls)
    t1 = time.perf_counter()
    for i in range(args.steps):
        ex.extract_batch_raw(host_ptrs[i % R], B, h, w, w, out_kp.data_ptr(), out_desc.data_ptr(), cap, out_n)
    e2e_sync_s = time.perf_counter() - t1
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    n_kp_host = int(sum(out_n[:]))

    import bench_legs
    # the e2e copy pattern alone, every rank at once: what the host link of this box allows for this traffic
    link_ms = bench_legs.link_ceiling(torch, dev, pin, out_kp2, out_desc2, max(10, min(args.steps, 50)), barrier)
    barrier()
    single, single_launches = (bench_legs.single_frame(torch, local_rank, w, h, nf, nl, frames[0]) if rank == 0 else ({}, 0))
    barrier()

    ham = None
    if not args.no_match:
        ham = hamming_legs(args, torch, dist, dev, rank, world, ex, ext, d_frames, d_kp, d_desc, d_n, cap, w, h, pitch, barrier)
        c4, c4_launches = bench_legs.c4_leg(torch, dist, rank, world, local_rank, barrier, ham["_voc"], nf, nl, w, h)
        ham["c4"] = c4; ham["_launches"] += c4_launches
        if rank == 0:
            ham["searches"], n_calls = bench_legs.searches_leg(calls=64)
        barrier()
    t = torch.tensor([dev_ms, e2e_s * 1e3, ham["_ap_ms"] if ham else 0.0, ham["_tr_ms"] if ham else 0.0, link_ms],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)          # max over ranks
    dev_ms, e2e_ms, ap_ms, tr_ms, link_ms = float(t[0]), float(t[1]), float(t[2]), float(t[3]), float(t[4])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    frames_total = B * args.steps * world
    value = frames_total / (dev_ms * 1e-3)
    e2e = frames_total / (e2e_ms * 1e-3)
    peak, peak_kind = peaks()
    per_launch_ms = [m / max(stage_calls, 1) for m in stage_ms]
    dom = int(np.argmax(per_launch_ms))
    dom_bytes = per_stage_bytes[STAGES[dom]] * B
    achieved = dom_bytes / (per_launch_ms[dom] * 1e-3) / 1e9 if per_launch_ms[dom] > 0 else 0.0
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tp):
        with open(tp) as f:
            tj = json.load(f)
        if tj.get("workload") == args.workload and tj.get("frames_per_launch") == B:   # only for the captured configuration
            traffic = tj["bytes"].get(STAGES[dom])
    issue = issue_roofline(args.workload, B, dict(zip(STAGES, per_launch_ms)), torch.cuda.get_device_properties(dev).multi_processor_count,
                           w, h, nl)
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl, "frames_per_step_per_gpu": B, "nfeatures": nf, "nlevels": nl, "scale": SCALE,
                   "agents": world, "parallelism": "agent-stream per GPU, no data-path collective",
                   "l2": "inputs larger than L2: %d rotating %d-frame batches (%.0f MB) + %.0f MB of pyramid scratch rewritten per step"
                         % (R, B, frames.nbytes / 1e6, 2 * sum(a * b for a, b in level_sizes(w, h, nl)) * B / 1e6),
                   "keypoints_per_frame": n_kp / B},
        "roofline": {"bound": "hbm", "kernel": STAGES[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_kind": peak_kind,
                     "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": per_launch_ms[dom]},
        "roofline_pipeline": {"bound": "hbm", "achieved": value / world * pipeline_bytes / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": value / world * pipeline_bytes / 1e9 / peak, "algorithmic_bytes_per_frame": pipeline_bytes},
        "roofline_issue": issue,
        "stage_ms_per_step": dict(zip(STAGES, per_launch_ms)),
        "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": B * w * h,
                "d2h_bytes_per_step": B * cap * 56 + 4 * B, "keypoints_per_frame": n_kp_host / B,
                "api": "ccm_orb_extract_batch_submit / _wait, two batches in flight (copy-in of batch i+1 overlaps the kernels of batch i)",
                "synchronous_call_frames_per_s": B * args.steps / e2e_sync_s, "host_numa_node_rank0": numa_node,
                "h2d_ceiling_frames_per_s": B * world / (link_ms * 1e-3), "frac_of_link_ceiling": e2e / (B * world / (link_ms * 1e-3)),
                "link_ceiling": "the same copies with no kernel in between (pinned host -> device on one stream, device -> pinned host on another), "
                                "all %d ranks at once, max over ranks: %.3f ms per step; the limiter of the host entry point is cudaMemcpyAsync "
                                "from pinned host memory, there is no collective on this path" % (world, link_ms)},
        "gpu_launches": int(launches) + int(single_launches),
        "clocks": clocks,
    }
    line["e2e"].update(single)
    if ham:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        sm_max = (clocks or {}).get("sm_max_mhz") or 1965.0
        peak_pairs = int_pipe_peak_gpairs(sms, sm_max)
        gp = ham["allpairs"]["nq"] * ham["allpairs"]["nt_total"] / (ap_ms * 1e-3) / 1e9
        ham["allpairs"].update({"gpairs_per_s": gp, "ms_per_query_keyframe": ap_ms})
        ham["allpairs"]["roofline"] = {
            "bound": "int-pipe (POPC)", "achieved": gp / world, "peak": peak_pairs, "unit": "Gpair/s per GPU", "frac": gp / world / peak_pairs,
            "peak_kind": "textbook loop, 8 POPC per pair: SMs x 16 POPC/clk (measured 15.9, profiles/r1_ubench_int.json) x %.0f MHz / 8. The kernel's "
                         "carry-save LOP3 tree needs 4 POPC per pair, so frac exceeds 1 against this denominator" % sm_max,
            "frac_of_4popc_bound": gp / world / (2 * peak_pairs),
            "own_bound": "4 POPC + 18 LOP3/XOR + 3 min/max per pair: POPC bound %.0f Gpair/s per GPU; ncu (profiles/r1q): ALU pipe 83 %% busy" % (2 * peak_pairs)}
        loop = measured_loop_ceiling()
        if loop:
            ham["allpairs"]["roofline"].update({
                "measured_loop_ceiling": loop, "frac_of_measured_loop_ceiling": gp / world / loop,
                "measured_loop_ceiling_kind": "the product kernel's inner loop (LOP3 carry-save tree, 4 POPC, packed best/second keys, register "
                                              "prefetch) in tools/ubench_int.cu on an L2-resident random database, Gpair/s per GPU (profiles/r2_ubench_int.json)"})
        ham["tracking"].update({"frames_per_s": B * world / (tr_ms * 1e-3), "ms_per_step": tr_ms})
        line["hamming"] = {k: v for k, v in ham.items() if not k.startswith("_")}
        line["gpu_launches"] += int(ham["_launches"])
    if not args.no_cpu and world == 1:
        cores = os.cpu_count() or 1
        fpt = 48                                             # ~15 s of CPU work on 16 threads
        fps, dt, kind = cpu_reference_fps(frames[0], nf, nl, cores, fpt)
        fps1, dt1, _ = cpu_reference_fps(frames[0], nf, nl, 1, 96)      # how the reference runs: one spinOnce thread per agent (ClientNode.cpp:48-52)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                                "sample": "%d threads x %d frames of the same stream, %.1f s" % (cores, fpt, dt),
                                "single_thread": {"value": fps1, "unit": "frames/s", "cores": 1, "sample": "96 frames, %.1f s" % dt1},
                                "note": "cslam/src/ORBextractor.cpp compiled unmodified over a scalar restatement of the five OpenCV primitives "
                                        "(oracle/cv_primitives.cpp, no SIMD): an upper bound on the speed-up over a production OpenCV build"}
        if ham:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from _oracle import oracle_vocab_transform
            rngb = np.random.default_rng(3)
            db = rngb.integers(0, 256, (20000, 32), dtype=np.uint8)
            t0 = time.perf_counter()
            oracle_vocab_transform(ham["_voc"], db, 4)
            line["cpu_baseline"]["bow_transform"] = {"descriptors_per_s": len(db) / (time.perf_counter() - t0), "cores": 1, "kind": "port",
                                                     "sample": "20000 uniform descriptors through the same vocabulary"}
            pps, nm1, nm2, mkind = cpu_matcher_baseline(frames[0], nf, nl)
            line["cpu_baseline"]["matcher"] = {
                "tracking_searches_frames_per_s": pps, "matches_per_frame_lastframe": nm1, "matches_per_frame_mappoints": nm2,
                "cores": 1, "kind": mkind, "sample": "6 frames of the stream, both SearchByProjection calls per frame, the GPU leg's queries",
                "allpairs_gpairs_per_s": cpu_allpairs_baseline(), "allpairs_kind": "port (the database search has no reference function: it is the inner "
                "loop of SearchByBoW(KF,KF) restated; 64 x 200000 descriptor pairs)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def measured_loop_ceiling():
    """Gpair/s of the all-pairs inner loop as tools/ubench_int.cu measured it on a B200 (committed evidence; None if absent)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ubench_int.json")) as f:
            return float(json.load(f)["hamming_q4_csa_Gpairs_per_s"])
    except (OSError, KeyError, ValueError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c1_euroc_752x480_1000kp", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=64, help="frames in flight per step (configs[1]: 64)")
    ap.add_argument("--rotate", type=int, default=8, help="distinct input batches cycled through (inputs > L2)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-match", action="store_true", help="skip the Hamming-search legs")
    ap.add_argument("--db", type=int, default=4000000, help="keyframe-database descriptors (configs[4]: 8 x 500 x 1000)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, args.workload, rank)
    else:
        run_ours(args, args.workload, rank, world, local_rank)


if __name__ == "__main__":
    main()
