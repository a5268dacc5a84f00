"""Summarise an ncu source-page CSV export by SASS region: share of warp instructions, average
active threads and stall samples (helper used while tuning; reads `ncu --page source --csv`)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
step = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) < 10 or r[0] in ('Kernel Name', 'Address'): break
    data.append(r)
I = lambda r, k: int(r[idx[k]])
ti = sum(I(r, 'Instructions Executed') for r in data); tt = sum(I(r, 'Thread Instructions Executed') for r in data); ts = sum(I(r, '# Samples') for r in data)
print(rows[0][1][:80], '| instr', len(data), '| warp-inst', ti, '| avg threads %.1f' % (tt / ti))
for c in range(0, len(data), step):
    ch = data[c:c + step]
    wi = sum(I(r, 'Instructions Executed') for r in ch); th = sum(I(r, 'Thread Instructions Executed') for r in ch); sm = sum(I(r, '# Samples') for r in ch)
    ops = ' '.join(sorted(set(r[idx['Source']].split()[0].split('.')[0] for r in ch if r[idx['Source']].split()))[:12])
    print('%4d-%4d  warp-inst %5.1f%%  avg-thr %5.1f  samples %5.1f%%  %s' % (c, c + len(ch), 100 * wi / ti, th / max(wi, 1), 100 * sm / ts, ops))
