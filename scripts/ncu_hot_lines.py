"""Summarise an `ncu --page source --csv --print-source cuda,sass` export by SOURCE LINE: warp-instructions executed per warp
and stall samples per launch, hottest lines first.  Usage:
    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > src.csv
    python scripts/ncu_hot_lines.py src.csv WARPS_PER_LAUNCH LAUNCHES [min_samples] > profiles/rNN_<kernel>_hot_lines.txt"""
import csv
import sys


def main():
    path, warps, launches = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    min_s = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    cur, hdr, agg, src = None, None, {}, {}
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0] != "" and r[0] != "Function Name" and hdr:
            d = dict(zip(hdr, r))
            try:
                k = (cur, int(r[0]))
                a = agg.setdefault(k, [0] * 8)
                for n, key in enumerate(("Instructions Executed", "# Samples", "stall_long_sb", "stall_wait", "stall_short_sb",
                                         "stall_barrier", "stall_not_selected", "stall_branch_resolving")):
                    a[n] += int(d[key])
                src[k] = r[1].strip()[:96]
            except (ValueError, KeyError):
                pass
    tot_i = sum(v[0] for v in agg.values()) / launches / warps
    tot_s = sum(v[1] for v in agg.values()) / launches
    print("# %s: %d launches, %d warps per launch; %.0f warp-instructions per warp, %.0f stall samples per launch" % (path.split("/")[-1], launches, warps, tot_i, tot_s))
    print("# columns: file line | instr/warp | samples/launch (share) | of which long_scoreboard wait short_scoreboard barrier not_selected branch | source")
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    for (f, l), v in rows:
        if v[1] / launches < min_s:
            break
        print("%-20s %5d | %6.1f | %5.0f (%4.1f%%) | %4.0f %4.0f %4.0f %4.0f %4.0f %4.0f | %s" % (
            f, l, v[0] / launches / warps, v[1] / launches, 100.0 * v[1] / launches / tot_s, v[2] / launches, v[3] / launches, v[4] / launches,
            v[5] / launches, v[6] / launches, v[7] / launches, src[(f, l)]))
    print("# by instruction count (>= 1 %% of the warp's instructions)")
    for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        if v[0] / launches / warps < 0.01 * tot_i:
            break
        print("%-20s %5d | %6.1f (%4.1f%%) | %s" % (f, l, v[0] / launches / warps, 100.0 * v[0] / launches / warps / tot_i, src[(f, l)]))


if __name__ == "__main__":
    main()
