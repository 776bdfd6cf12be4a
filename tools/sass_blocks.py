#!/usr/bin/env python
"""Basic-block view of an ncu report's SASS page: share of warp instructions, stall samples and
average active threads per run of instructions with equal execution count.
usage: tools/sass_blocks.py report.ncu-rep [min_share_pct]"""
import csv, subprocess, sys
from collections import Counter
rep = sys.argv[1]; min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
blocks = []; tot = 0
for r in rows[2:]:
    if len(r) < len(hdr) - 5: continue
    ie = int(r[ix['Instructions Executed']]); te = int(r[ix['Thread Instructions Executed']]); src = r[ix['Source']].strip()
    smp = int(r[ix['# Samples']]); tot += ie
    op = src.split()[1] if src.startswith('@') else src.split()[0]
    if blocks and blocks[-1]['ie'] == ie and abs(blocks[-1]['te'] / max(ie, 1) - te / max(ie, 1)) < 0.01:
        b = blocks[-1]; b['n'] += 1; b['smp'] += smp; b['ops'].append(op)
    else:
        blocks.append({'ie': ie, 'te': te, 'n': 1, 'smp': smp, 'ops': [op]})
tots = sum(b['smp'] for b in blocks)
print(rows[0][1]); print('total warp-inst', tot, ' avg threads', sum(b['te'] * b['n'] for b in blocks) / max(tot, 1))
for b in blocks:
    share = 100 * b['ie'] * b['n'] / tot
    if share < min_share: continue
    c = Counter(o.split('.')[0] for o in b['ops'])
    print(f"{share:5.1f}% inst {100 * b['smp'] / tots:5.1f}% smp  n={b['n']:4d} exec={b['ie']:9d} thr={b['te'] / max(b['ie'], 1):5.1f}  {dict(c.most_common(6))}")
