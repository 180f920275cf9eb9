"""Warp-stall samples of one kernel aggregated by the instruction the warp was waiting to issue, from the source page
of an ncu report:  ncu -i <rep> --page source --csv > x.csv;  python scripts/ncu_stalls_by_op.py x.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
print(rows[0][1] if len(rows[0]) > 1 else rows[0])
h, data = rows[1], rows[2:]
isrc, isamp = h.index('Source'), h.index('# Samples')
stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
tot, byop, n = collections.Counter(), collections.defaultdict(collections.Counter), 0
for r in data:
    if len(r) < len(h) or not r[isamp].strip().isdigit():
        continue
    op = re.sub(r'^@!?U?P\w+\s+', '', r[isrc]).split()[0] if r[isrc] else '?'
    key = 'IMAD.WIDE/HI' if op.startswith(('IMAD.WIDE', 'IMAD.HI')) else 'IMAD (other)' if op.startswith('IMAD') else op.split('.')[0]
    for c in stalls:
        v = int(r[h.index(c)] or 0)
        tot[c] += v
        byop[key][c] += v
    n += int(r[isamp])
all_ = sum(tot.values())
print('%d samples; by reason: %s' % (n, ', '.join('%s %.1f%%' % (k[6:], 100.0 * v / all_) for k, v in tot.most_common(9))))
print('by the instruction waiting to issue:')
for k, c in sorted(byop.items(), key=lambda kv: -sum(kv[1].values()))[:12]:
    t = sum(c.values())
    print('  %-14s %5.1f%%   %s' % (k, 100.0 * t / all_, ', '.join('%s %.1f%%' % (a[6:], 100.0 * b / all_) for a, b in c.most_common(4))))
