"""Aggregates an `ncu --page source --csv --print-source cuda,sass` export per enclosing CUDA function
(function = nearest preceding line that looks like a definition in the same file).

    python tools/ncu_funcs.py src.csv path/to/file.cuh [path/to/other.cu ...]"""
import csv
import re
import sys
from collections import defaultdict


def func_table(path):
    starts = []
    pat = re.compile(r"^(template\s*<[^>]*>\s*)?(GO1_DEV|GO1_NOINL|__global__|__device__|static|inline)\b.*\b(\w+)\s*\(")
    lines = open(path).read().split("\n")
    for i, l in enumerate(lines, 1):
        m = pat.match(l.strip())
        if m and not l.strip().endswith(";"):
            starts.append((i, m.group(3)))
        # region markers inside long functions: comment lines with dashes
        m2 = re.match(r"\s*// -{10,} (.*)", l)
        if m2:
            starts.append((i, "  > " + m2.group(1)[:50]))
    return starts


def main():
    src = sys.argv[1]
    tabs = {p.split("/")[-1]: func_table(p) for p in sys.argv[2:]}
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    cur_file, hdr, last = None, None, None
    for row in csv.reader(open(src)):
        if not row:
            continue
        if row[0] == "File Path":
            cur_file = row[1].split("/")[-1]; hdr = None; continue
        if row[0] == "Line No":
            hdr = {h: i for i, h in enumerate(row)}; continue
        if hdr is None or row[0] == "Function Name":
            continue
        if row[0] != "":
            last = (cur_file, int(row[0]))
        if row[hdr["Address"]] == "" or last is None:
            continue
        f, ln = last
        name = f
        if f in tabs:
            name = f + ":?"
            for s, n in tabs[f]:
                if s <= ln:
                    name = f"{f}:{n}"
                else:
                    break
        def g(k):
            try:
                return float(row[hdr[k]])
            except ValueError:
                return 0.0
        a = agg[name]
        a[0] += 1; a[1] += g("Instructions Executed"); a[2] += g("Thread Instructions Executed"); a[3] += g("# Samples")
    ti = sum(a[1] for a in agg.values()); ts = sum(a[3] for a in agg.values())
    print(f"{'function':60s} {'static':>7s} {'warp-inst':>10s} {'%inst':>6s} {'thr/inst':>8s} {'%samples':>8s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:60s} {a[0]:7d} {a[1]:10.3g} {100*a[1]/ti:6.1f} {a[2]/max(1,a[1]):8.1f} {100*a[3]/max(1,ts):8.1f}")


if __name__ == "__main__":
    main()
