"""Per-source-line instruction and stall-sample shares of one kernel from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [min_percent]"""
import csv
import io
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    floor = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                          "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    h = rows[hdr]
    i_inst, i_smp, i_thr = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
    lines = [r for r in rows[hdr + 1:] if r and r[0].isdigit() and r[i_inst].isdigit()]
    tot = sum(int(r[i_inst]) for r in lines) or 1
    for r in lines:
        r[i_smp] = r[i_smp] if r[i_smp].isdigit() else "0"
    tots = sum(int(r[i_smp]) for r in lines) or 1
    totthr = sum(int(r[i_thr]) for r in lines)
    print("warp instructions %d, thread instructions %d (%.1f lanes/inst), samples %d" % (tot, totthr, totthr / tot, tots))
    print("line  inst%  smpl%  lanes  source")
    for r in lines:
        a, b = 100.0 * int(r[i_inst]) / tot, 100.0 * int(r[i_smp]) / tots
        if a >= floor or b >= floor:
            print("%4s %6.2f %6.2f %5.1f  %s" % (r[0], a, b, int(r[i_thr]) / max(int(r[i_inst]), 1), r[1].strip()[:120]))


if __name__ == "__main__":
    main()
