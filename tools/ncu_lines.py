"""Aggregates an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line.

    ncu -i prof.ncu-rep --page source --csv --print-source cuda,sass > src.csv
    python tools/ncu_lines.py src.csv [top_n]
Prints, per source line: SASS instruction count (static), warp instructions executed, thread
instructions executed, mean active threads, stall samples."""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    cur_file = None
    hdr = None
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0, ""])   # static, inst, thread inst, samples, text
    stall_cols = {}
    stalls = defaultdict(lambda: defaultdict(float))
    last_line = None
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "File Path":
            cur_file = row[1].split("/")[-1]; hdr = None; continue
        if row[0] == "Function Name":
            continue
        if row[0] == "Line No":
            hdr = {h: i for i, h in enumerate(row)}
            # two "Source" columns: first = cuda text, second = sass text
            src_idx = [i for i, h in enumerate(row) if h == "Source"]
            stall_cols = {h: i for i, h in enumerate(row) if h.startswith("stall_")}
            continue
        if hdr is None:
            continue
        if row[0] != "":
            last_line = (cur_file, int(row[0]))
            agg[last_line][4] = row[src_idx[0]].strip()[:90]
        addr = row[hdr["Address"]]
        if addr == "" or last_line is None:
            continue
        a = agg[last_line]
        a[0] += 1
        def f(name):
            v = row[hdr[name]]
            try:
                return float(v)
            except ValueError:
                return 0.0
        a[1] += f("Instructions Executed"); a[2] += f("Thread Instructions Executed"); a[3] += f("# Samples")
        for h, i in stall_cols.items():
            try:
                stalls[last_line][h] += float(row[i])
            except ValueError:
                pass
    tot_static = sum(a[0] for a in agg.values()); tot_inst = sum(a[1] for a in agg.values())
    tot_thr = sum(a[2] for a in agg.values()); tot_s = sum(a[3] for a in agg.values())
    print(f"total: static SASS {tot_static}  warp-inst {tot_inst:.3g}  thread-inst {tot_thr:.3g}  avg threads {tot_thr/max(1,tot_inst):.2f}  samples {tot_s:.0f}")
    tot_stall = defaultdict(float)
    for d in stalls.values():
        for h, v in d.items():
            tot_stall[h] += v
    print("stall totals:", {h: int(v) for h, v in sorted(tot_stall.items(), key=lambda kv: -kv[1]) if v > 0})
    print("--- by samples")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][3])[:top]:
        top_st = sorted(stalls[k].items(), key=lambda kv: -kv[1])[:2]
        print(f"{k[0]}:{k[1]:5d} static {a[0]:5d} inst {a[1]:10.3g} thr/inst {a[2]/max(1,a[1]):5.1f} samples {a[3]:8.0f} ({100*a[3]/max(1,tot_s):4.1f}%) {[(h[6:], int(v)) for h, v in top_st]} | {a[4]}")
    print("--- by static size")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{k[0]}:{k[1]:5d} static {a[0]:5d} inst {a[1]:10.3g} thr/inst {a[2]/max(1,a[1]):5.1f} samples {a[3]:8.0f} | {a[4]}")


if __name__ == "__main__":
    main()
