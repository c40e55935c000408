"""Summarise an `ncu --page source --csv` dump: samples by opcode, top stalled instructions."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
if len(rows) < 3:
    print('kernel not in this capture')
    sys.exit(0)
h = rows[1]
data = []
for r in rows[2:]:          # first launch only: a second launch repeats the two header rows
    if r and r[0] == 'Kernel Name':
        break
    if len(r) == len(h) and r != h:
        data.append(r)
si, ai = h.index('Source'), h.index('Address')
allc, nic, exe = h.index('Warp Stall Sampling (All Samples)'), h.index('Warp Stall Sampling (Not-issued Samples)'), h.index('Instructions Executed')
f = lambda x: float(x or 0)
tot, totni, totexe = sum(f(r[allc]) for r in data), sum(f(r[nic]) for r in data), sum(f(r[exe]) for r in data)
print('instr', len(data), 'samples', tot, 'not-issued', totni, 'executed warp-instr', totexe)
op = collections.Counter(); opn = collections.Counter(); ope = collections.Counter()
for r in data:
    toks = r[si].split()
    o = toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '')
    o = o.split('.')[0]
    op[o] += f(r[allc]); opn[o] += 1; ope[o] += f(r[exe])
for o, v in op.most_common(14):
    print('%-10s samples %7.0f (%4.1f%%)  static %4d  executed %10.0f (%4.1f%%)' % (o, v, 100 * v / tot, opn[o], ope[o], 100 * ope[o] / totexe))
print('top stalled instructions:')
for r in sorted(data, key=lambda r: -f(r[allc]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print('  %6.0f %6.0f  %s' % (f(r[allc]), f(r[nic]), r[si][:100]))
