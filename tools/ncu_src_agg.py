#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump: stall reasons, per-opcode samples, top source lines."""
import csv, sys, re, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if r and r[0] == 'Address')
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows if r and r[0].startswith('0x') and len(r) >= len(hdr) - 2]
I = lambda r, k: int(float(r[ix[k]] or 0))
tot = sum(I(r, '# Samples') for r in data)
print('SASS lines', len(data), 'samples', tot, 'warp instr executed', sum(I(r, 'Instructions Executed') for r in data))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(I(r, s) for r in data) for s in stalls}
for s, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
    print('  %-24s %8d %5.1f%%' % (s, v, 100*v/max(tot, 1)))
ops = collections.Counter(); ex = collections.Counter(); lsb = collections.Counter()
for r in data:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)', r[ix['Source']])
    op = m.group(2) if m else r[ix['Source']][:10]
    ops[op] += I(r, '# Samples'); ex[op] += I(r, 'Instructions Executed'); lsb[op] += I(r, 'stall_long_sb')
print('opcode   samples   share   executed   long_sb')
for op, v in ops.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 20):
    print('  %-8s %8d %5.1f%% %12d %8d' % (op, v, 100*v/max(tot, 1), ex[op], lsb[op]))
