"""Top stall lines of one kernel from an `ncu --page source --csv` export (helper for reading profiles here)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
iS = hdr.index('# Samples'); iSrc = hdr.index('Source'); iEx = hdr.index('Instructions Executed')
stalls = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for r in rows[2:]:
    if len(r) != len(hdr):
        continue
    if r[iS] == '# Samples':
        break
    data.append(r)
tot = sum(int(r[iS]) for r in data)
print('total samples', tot, 'ninstr', len(data))
agg = {hdr[i]: sum(int(r[i]) for r in data) for i in stalls}
print(sorted(agg.items(), key=lambda kv: -kv[1])[:8])
top = sorted(range(len(data)), key=lambda k: -int(data[k][iS]))[:ntop]
for k in sorted(top):
    r = data[k]
    st = sorted([(int(r[i]), hdr[i]) for i in stalls], reverse=True)[:2]
    print(k, r[iS], r[iEx], r[iSrc].strip()[:70], st)
