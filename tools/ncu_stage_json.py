"""profiles/issue.json and profiles/traffic.json from a summary CSV written by tools/ncu_summary.py (one 64-frame extraction
step): warp instructions and DRAM bytes per stage (a stage = all launches of its kernel in the step).
usage: python tools/ncu_stage_json.py profiles/rX_ncu_full_orb_summary.csv WORKLOAD FRAMES"""
import csv
import json
import os
import sys

STAGE_OF = {"k_resize": "pyramid_resize", "k_fast": "fast_cells", "k_blur": "gaussian_blur", "k_quadtree": "quadtree", "k_orient": "orient_describe"}
# round-2 instruction budget, warp instructions per 64-frame step at C1
BUDGET = {"pyramid_resize": 30e6, "fast_cells": 100e6, "gaussian_blur": 55e6, "quadtree": 15e6, "orient_describe": 40e6}


def main():
    src, workload, frames = sys.argv[1], sys.argv[2], int(sys.argv[3])
    rows = list(csv.reader(open(src)))
    h = rows[0]
    inst, dram, dur = {}, {}, {}
    for r in rows[2:]:
        d = dict(zip(h, r))
        st = next((v for k, v in STAGE_OF.items() if k in d["Kernel Name"]), None)
        if st is None:
            continue
        inst[st] = inst.get(st, 0) + int(float(d["smsp__inst_executed.sum"]))
        # ncu prints these two columns in MB (scaled unit row); fall back to bytes when large
        rd, wr = float(d["dram__bytes_read.sum"]), float(d["dram__bytes_write.sum"])
        scale = 1e6 if rd < 1e5 else 1.0
        dram[st] = dram.get(st, 0) + int((rd + wr) * scale)
        dur[st] = dur.get(st, 0.0) + float(d["gpu__time_duration.sum"])
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rel = os.path.relpath(os.path.abspath(src), root)
    with open(os.path.join(root, "profiles", "issue.json"), "w") as f:
        json.dump({"workload": workload, "frames_per_launch": frames, "sm_mhz": 1965.0,
                   "source": rel + " (ncu --set full --clock-control none, smsp__inst_executed.sum per launch, summed per stage)",
                   "warp_instructions": inst, "ncu_duration_us": dur, "budget": BUDGET}, f, indent=1)
    with open(os.path.join(root, "profiles", "traffic.json"), "w") as f:
        json.dump({"workload": workload, "frames_per_launch": frames,
                   "source": rel + " (ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum per launch; resize = sum of its 7 launches)",
                   "bytes": dram}, f, indent=1)
    print(json.dumps({"warp_instructions": inst, "dram_bytes": dram, "total_warp_instructions": sum(inst.values())}))


if __name__ == "__main__":
    main()
