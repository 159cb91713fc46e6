"""Summarise an ncu --page source --csv dump: top stall lines and shared-memory conflict lines.
usage: ncu -i X.ncu-rep --page source --csv > /tmp/x.csv ; python tools/ncu_src.py /tmp/x.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr)]
def num(r, name):
    try:
        return float(r[col[name]])
    except Exception:
        return 0.0
tot = sum(num(r, '# Samples') for r in body)
print("total samples", tot)
print("== top stall lines")
for r in sorted(body, key=lambda r: -num(r, '# Samples'))[:top]:
    stalls = {h[6:]: num(r, h) for h in hdr if h.startswith('stall_') and 'Not Issued' not in h and num(r, h) > 0}
    s = sorted(stalls.items(), key=lambda kv: -kv[1])[:3]
    print(f"{num(r,'# Samples'):7.0f} {100*num(r,'# Samples')/max(tot,1):5.1f}%  {r[col['Source']].strip()[:70]:70s} {s}")
print("== shared-memory excessive wavefronts")
for r in sorted(body, key=lambda r: -num(r, 'L1 Wavefronts Shared Excessive'))[:10]:
    if num(r, 'L1 Wavefronts Shared Excessive') > 0:
        print(f"{num(r,'L1 Wavefronts Shared Excessive'):10.0f} of {num(r,'L1 Wavefronts Shared'):10.0f} ideal {num(r,'L1 Wavefronts Shared Ideal'):10.0f}  {r[col['Source']].strip()[:80]}")
