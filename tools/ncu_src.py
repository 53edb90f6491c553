#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: top stalled SASS instructions (first launch only)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
his = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[his[0]]
end = his[1] - 1 if len(his) > 1 else len(rows)
data = [r for r in rows[his[0] + 1:end] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
S, src, ie = ci['# Samples'], ci['Source'], ci['Instructions Executed']
tot = sum(int(r[S] or 0) for r in data)
print('instructions', len(data), 'total samples', tot)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[ci[h]] or 0) for r in data) for h in stalls}
print('stall totals', sorted(agg.items(), key=lambda x: -x[1])[:8])
for idx, r in sorted(enumerate(data), key=lambda x: -int(x[1][S] or 0))[:n]:
    st = sorted(((h, int(r[ci[h]] or 0)) for h in stalls), key=lambda x: -x[1])[:2]
    print(idx, r[S], r[ie], r[src][:100], st)
