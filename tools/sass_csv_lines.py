#!/usr/bin/env python
"""Per-source-line view of an ncu SASS page that was exported ON THE GPU BOX
(`ncu -i X.ncu-rep --page source --csv | gzip > X_source.csv.gz`: the report itself is too big to travel).
Joins it to `nvdisasm -g` of the library by instruction index, like tools/sass_lines.py does for a report.

usage: tools/sass_csv_lines.py X_source.csv.gz liblajolla_b200.so kernel_substring [top_n]
       tools/sass_csv_lines.py X_source.csv.gz liblajolla_b200.so kernel_substring --line lj_math.cuh:38   (the SASS of one line)
       tools/sass_csv_lines.py X_source.csv.gz liblajolla_b200.so kernel_substring --calls                 (executed CALL sites)
--calls is how the f64-division slow path was found (profiles/r02_experiments.md): the subroutine has no line info of
its own, so its instructions show up under whatever line precedes it; its CALL sites carry the real lines and counts.
The library must be the build the capture was taken from (the opcode sequence is checked)."""
import collections
import csv
import gzip
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sass_lines  # noqa: E402


def main():
    src, lib, kernel = sys.argv[1:4]
    rest = sys.argv[4:]
    rows = list(csv.reader(gzip.open(src, "rt")))
    hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
    prof = []
    for r in rows[2:]:
        if len(r) < len(hdr) - 5:
            continue
        s = r[ix["Source"]].strip()
        op = s.split()[1] if s.startswith("@") else s.split()[0]
        prof.append((op.rstrip(";"), int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]]), s))
    dis = sass_lines.disasm_lines(lib, kernel)
    n = min(len(prof), len(dis))
    mism = sum(1 for i in range(n) if prof[i][0].split(".")[0] != dis[i][0].split(".")[0])
    print(rows[0][1])
    print(f"ncu instructions {len(prof)}, nvdisasm instructions {len(dis)}, opcode mismatches {mism}")
    tot_i = sum(p[1] for p in prof[:n]) or 1
    tot_s = sum(p[3] for p in prof[:n]) or 1
    if rest and rest[0] == "--line":
        f, l = rest[1].rsplit(":", 1)
        for i in range(n):
            if dis[i][1] == (f, int(l)) and prof[i][1]:
                print(f"{i:6d} {prof[i][4][:72]:72s} {prof[i][1]:10d} {100 * prof[i][1] / tot_i:5.2f}% thr {prof[i][2] / prof[i][1]:4.1f} smp {prof[i][3]}")
        return
    if rest and rest[0] == "--calls":
        for i in range(n):
            if "CALL" in prof[i][4] and prof[i][1]:
                print(f"{i:6d} {prof[i][4][:60]:60s} {prof[i][1]:10d} thr {prof[i][2] / prof[i][1]:4.1f}  {dis[i][1][0]}:{dis[i][1][1]}")
        return
    top = int(rest[0]) if rest else 40
    agg = collections.defaultdict(lambda: [0, 0, 0]); byfile = collections.defaultdict(lambda: [0, 0, 0])
    for i in range(n):
        for a in (agg[dis[i][1]], byfile[dis[i][1][0]]):
            a[0] += prof[i][1]; a[1] += prof[i][2]; a[2] += prof[i][3]
    print("per file:")
    for f, a in sorted(byfile.items(), key=lambda x: -x[1][2]):
        print(f"  {100 * a[0] / tot_i:5.1f}% inst {100 * a[2] / tot_s:5.1f}% samples  lanes {a[1] / max(a[0], 1):5.1f}  {f}")
    print("per line (by stall samples):")
    for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
        print(f"  {100 * a[0] / tot_i:5.1f}% inst {100 * a[2] / tot_s:5.1f}% samples  lanes {a[1] / max(a[0], 1):5.1f}  {f}:{l}")


if __name__ == "__main__":
    main()
