"""Summarise an `ncu --page source --csv` export: stall reasons and opcode mix."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_')]
tot = {s: 0 for s in stalls}
n = inst = 0
data = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[ix['# Samples']])
    except ValueError:
        continue
    ie = int(r[ix['Instructions Executed']] or 0)
    n += s
    inst += ie
    for st in stalls:
        try:
            tot[st] += int(r[ix[st]] or 0)
        except ValueError:
            pass
    data.append((s, r[ix['Source']], ie, r))
print('samples', n, 'warp-inst', inst, 'SASS lines', len(data))
ts = max(1, sum(tot.values()))
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:9]:
    print('  %-26s %6.1f%%' % (k, 100 * v / ts))
op, ops = collections.Counter(), collections.Counter()
for s, src, ie, _ in data:
    t = src.split()
    o = t[1] if t[0].startswith('@') else t[0]
    o = o.split('.')[0]
    op[o] += ie
    ops[o] += s
print('--- opcode: %inst  %samples')
for k, v in op.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 22):
    print('  %-10s %5.1f%%  %5.1f%%' % (k, 100 * v / inst, 100 * ops[k] / max(n, 1)))
if len(sys.argv) > 3:
    print('--- top SASS lines by samples')
    for s, src, ie, r in sorted(data, key=lambda x: -x[0])[:int(sys.argv[3])]:
        print('  %5.2f%% %s' % (100 * s / n, src[:90]))
