#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output: top sampled SASS lines with their dominant stall reasons.
usage: ncu_top_stalls.py file.csv [n_lines] [section_index]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
want = int(sys.argv[3]) if len(sys.argv) > 3 else 0
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = starts[want]
end = starts[want + 1] - 1 if want + 1 < len(starts) else len(rows)
print("sections:", len(starts), "| kernel:", rows[hi - 1][1][:90] if hi > 0 else "?")
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:end] if len(r) == len(hdr) and r[idx['# Samples']].isdigit()]
tot = sum(int(r[idx['# Samples']]) for r in data)
print('total samples', tot, 'lines', len(data))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {}
for r in data:
    for h in stalls:
        agg[h] = agg.get(h, 0) + int(r[idx[h]])
print('stall totals', dict(sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(data, key=lambda r: -int(r[idx['# Samples']]))[:n]:
    s = {h[6:]: int(r[idx[h]]) for h in stalls if int(r[idx[h]]) > 0}
    s = dict(sorted(s.items(), key=lambda kv: -kv[1])[:3])
    print(f"{r[idx['# Samples']]:>7} {r[idx['Instructions Executed']]:>9}  {r[idx['Source']].strip()[:64]:<64} {s}")
