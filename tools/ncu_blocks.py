#!/usr/bin/env python
"""Basic-block level summary of an `ncu --page source --csv` dump (SASS): instruction share and
stall-sample share per run of instructions with the same execution count."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iI, iSrc = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Source')
data = []
for n, r in enumerate(rows[2:]):
    try: data.append((n, int(r[iS]), int(r[iI]), r[iSrc].strip()))
    except Exception: pass
ts = sum(d[1] for d in data); ti = sum(d[2] for d in data)
print(len(data), "SASS instrs; samples", ts, "warp-inst", ti)
blocks, cur = [], [data[0]]
for d in data[1:]:
    if abs(d[2] - cur[-1][2]) <= 0.02 * max(d[2], cur[-1][2], 1): cur.append(d)
    else: blocks.append(cur); cur = [d]
blocks.append(cur)
big = [(sum(x[2] for x in b), sum(x[1] for x in b), b) for b in blocks]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 24
for tot, smp, b in sorted(big, key=lambda x: -(x[0] / ti + x[1] / ts))[:top]:
    ops = collections.Counter((x[3].split()[1] if x[3].startswith('@') else x[3].split()[0]) for x in b)
    print(f"[{b[0][0]:4d}-{b[-1][0]:4d}] n={len(b):3d} exec={b[0][2]:9d} inst%={100*tot/ti:5.1f} smp%={100*smp/ts:5.1f} {ops.most_common(7)}")
