#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source sass` export: samples per block of N SASS instructions with the
dominant opcodes and stall reasons.   python scripts/ncu_src_summary.py file.csv [block=40] [lo hi]"""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}; data = rows[2:]
lo = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi = int(sys.argv[4]) if len(sys.argv) > 4 else len(data)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
def op(r):
    t = r[ix['Source']].split()
    return t[1] if t and t[0].startswith('@') else (t[0] if t else '')
tot = sum(int(r[ix['# Samples']] or 0) for r in data)
print('total samples', tot, 'rows', len(data))
for a in range(lo, hi, blk):
    seg = data[a:a + blk]
    n = sum(int(r[ix['# Samples']] or 0) for r in seg)
    st = Counter()
    for r in seg:
        for h in stalls:
            st[h[6:]] += int(r[ix[h]] or 0)
    ex = max(int(r[ix['Instructions Executed']] or 0) for r in seg)
    print(a, n, 'exec', ex, Counter(op(r) for r in seg).most_common(3), [(k, v) for k, v in st.most_common(3) if v])
