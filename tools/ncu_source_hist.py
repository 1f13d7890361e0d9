"""Engineering: opcode histogram (executed warp instructions, stall samples) from `ncu --page source --csv` output.
    ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/ncu_source_hist.py src.csv [top-N hot lines]"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[1], rows[2:]
isrc, isamp, iex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
ish, ishi = hdr.index('L1 Wavefronts Shared'), hdr.index('L1 Wavefronts Shared Ideal')
tot = sum(int(r[isamp]) for r in data)
c, s = Counter(), Counter()
for r in data:
    t = r[isrc].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    c[op] += int(r[iex]); s[op] += int(r[isamp])
te = sum(c.values())
print(f"{len(data)} SASS instructions, {te} warp instructions executed, {tot} stall samples")
for op, n in c.most_common(24):
    print(f"{op:10s} exec {n:10d} {100 * n / te:5.1f}%  samples {100 * s[op] / tot:5.1f}%")
print('shared wavefronts', sum(int(r[ish] or 0) for r in data), 'ideal', sum(int(r[ishi] or 0) for r in data))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if top:
    stall = [h for h in hdr if h.startswith('stall_') and 'Not' not in h]
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:top]:
        why = sorted(((int(r[hdr.index(h)] or 0), h) for h in stall), reverse=True)[:2]
        print(f"{100 * int(r[isamp]) / tot:5.2f}% {r[0][-5:]} {r[isrc][:70]:70s} {why}")
