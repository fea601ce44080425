"""Stall-reason totals and the hottest SASS lines of one kernel from `ncu -i X.ncu-rep --page source --csv`:
   python profiles/ncu_stalls.py src.csv [n_top]"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
iS, iSrc = hdr.index('# Samples'), hdr.index('Source')
data = [r for r in rows[2:] if len(r) > iS and r[iS].isdigit()]
seen, uniq = set(), []
for r in data:
    if r[0] not in seen:
        seen.add(r[0]); uniq.append(r)
data = uniq
cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[iS]) for r in data)
print('samples', tot)
print(' '.join(f"{c[6:]}={100 * sum(int(r[hdr.index(c)] or 0) for r in data) / tot:.1f}%" for c in cols
               if sum(int(r[hdr.index(c)] or 0) for r in data) > 0.01 * tot))
ops = collections.Counter()
for r in data:
    t = [o for o in r[iSrc].split() if not o.startswith('@')]
    ops[t[0].split('.')[0] if t else ''] += int(r[iS])
print(ops.most_common(10))
for r in sorted(data, key=lambda r: -int(r[iS]))[:ntop]:
    why = max(cols, key=lambda c: int(r[hdr.index(c)] or 0))
    print(f"{int(r[iS]):6d} {why[6:]:10s} {r[iSrc][:100]}")
