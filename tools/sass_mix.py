"""Opcode mix / stall share / lane utilisation of one kernel from `ncu --page source --print-source sass --csv`."""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
iA, iS, iI, iT = h.index('Source'), h.index('Warp Stall Sampling (All Samples)'), h.index('Instructions Executed'), h.index('Avg. Threads Executed')
data = [r for r in rows[2:] if len(r) > max(iA, iS, iI, iT) and r[iI].isdigit()]
tot = sum(int(r[iI]) for r in data)
ts = sum(int(r[iS]) for r in data)
print(rows[0][1][:80])
print('warp instructions', tot, 'stall samples', ts, 'sass lines', len(data))
c, s = Counter(), Counter()
for r in data:
    t = r[iA].split()
    op = t[1] if t[0].startswith('@') else t[0]
    op = op.split('.')[0]
    c[op] += int(r[iI])
    s[op] += int(r[iS])
for op, n in c.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print('%-10s %6.2f%% instr  %6.2f%% stall-samples' % (op, 100 * n / tot, 100 * s[op] / max(ts, 1)))
print('avg active threads per instruction %.2f' % (sum(int(r[iI]) * float(r[iT]) for r in data) / tot))
