"""Summarise an `ncu --page source --csv` dump: stall reasons, opcode mix, hottest SASS lines (by samples).
  python scripts/ncu_src_summary.py file.csv [top]"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr)]
def I(r, k):
    try: return int(float(r[idx[k]] or 0))
    except Exception: return 0
tot = sum(I(r, '# Samples') for r in data); ti = sum(I(r, 'Instructions Executed') for r in data)
print('samples', tot, 'warp inst', ti, 'sass lines', len(data))
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(I(r, h) for r in data) for h in st}
print('stalls:', ', '.join('%s %.1f%%' % (h[6:], 100 * v / max(tot, 1)) for h, v in sorted(agg.items(), key=lambda x: -x[1])[:7]))
op = collections.Counter(); ops = collections.Counter()
for r in data:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[idx['Source']])
    o = m.group(2).split('.')[0] if m else '?'
    op[o] += I(r, 'Instructions Executed'); ops[o] += I(r, '# Samples')
print('opcodes:', ', '.join('%s %.1f%%/%.1f%%' % (o, 100 * v / ti, 100 * ops[o] / max(tot, 1)) for o, v in op.most_common(14)))
print('hottest lines (samples%, inst%, top stall):')
for r in sorted(data, key=lambda r: -I(r, '# Samples'))[:top]:
    s = max(st, key=lambda h: I(r, h))
    print('  %5.1f%% %5.1f%% %-14s %s' % (100 * I(r, '# Samples') / max(tot, 1), 100 * I(r, 'Instructions Executed') / ti, s[6:], r[idx['Source']][:90]))
