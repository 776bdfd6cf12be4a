#!/usr/bin/env python
"""Attribute an ncu report's per-instruction counters to SOURCE LINES.
ncu's CSV export of the source page carries metrics only per SASS instruction; nvdisasm -g gives the
file:line of every SASS instruction of the same cubin.  Both list the kernel in address order, so they
join by instruction index (the opcode sequence is checked).
usage: tools/sass_lines.py report.ncu-rep object_or_so kernel_name_substring [top_n]
The object must be the exact build the report was captured from."""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def ncu_sass(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        if len(r) < len(hdr) - 5:
            continue
        src = r[ix["Source"]].strip()
        op = src.split()[1] if src.startswith("@") else src.split()[0]
        out.append((op.rstrip(";"), int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]])))
    return rows[0][1], out


def disasm_lines(obj, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    res = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin"):
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        cur_fn = None; cur_line = ("?", 0); take = False
        for line in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", line)
            if m:
                cur_fn = m.group(1); take = kernel in cur_fn; cur_line = ("?", 0)
                if take and res:
                    return res      # first match only
                continue
            if not take:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', line)
            if m:
                cur_line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s+(?:/\*[0-9a-f]+\*/\s+)?(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)\b.*;", line)
            if m and not line.strip().startswith("//") and not line.strip().startswith("."):
                res.append((m.group(1), cur_line))
        if res:
            return res
    return res


def main():
    rep, obj, kernel = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    name, prof = ncu_sass(rep)
    dis = disasm_lines(obj, kernel)
    print(name)
    print(f"ncu instructions {len(prof)}, nvdisasm instructions {len(dis)}")
    n = min(len(prof), len(dis))
    mism = sum(1 for i in range(n) if prof[i][0].split(".")[0] != dis[i][0].split(".")[0])
    print(f"opcode mismatches in the first {n}: {mism}")
    agg = collections.defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    for i in range(n):
        _, ie, te, smp = prof[i]
        a = agg[dis[i][1]]; a[0] += ie; a[1] += te; a[2] += smp
        tot[0] += ie; tot[1] += te; tot[2] += smp
    print(f"total warp-inst {tot[0]}, avg threads {tot[1] / max(tot[0], 1):.1f}")
    byfile = collections.defaultdict(lambda: [0, 0, 0])
    for (f, l), a in agg.items():
        b = byfile[f]; b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
    print("\nper file:")
    for f, a in sorted(byfile.items(), key=lambda x: -x[1][0]):
        print(f"  {100 * a[0] / tot[0]:5.1f}% inst {100 * a[2] / max(tot[2], 1):5.1f}% smp  thr {a[1] / max(a[0], 1):5.1f}  {f}")
    print("\nper line:")
    for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print(f"  {100 * a[0] / tot[0]:5.1f}% inst {100 * a[2] / max(tot[2], 1):5.1f}% smp  thr {a[1] / max(a[0], 1):5.1f}  {f}:{l}")


if __name__ == "__main__":
    main()
