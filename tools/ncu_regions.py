"""Aggregate warp-state samples of an ncu source-page CSV by SASS region (split at barriers / mbarrier ops / branches targets).
usage: ncu -i X.ncu-rep --page source --csv > /tmp/x.csv ; python tools/ncu_regions.py /tmp/x.csv [chunk]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 32
hdr = rows[1]; col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr)]
def num(r, n):
    try: return float(r[col[n]])
    except Exception: return 0.0
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(num(r, '# Samples') for r in body)
print("total samples", tot, "instructions", len(body))
KEY = ('BAR', 'SYNCS', 'IMMA', 'LDG', 'REDUX', 'UTMALDG', 'ACQBULK', 'STG', 'LDS', 'STS', 'MUFU', 'CALL', 'EXIT', 'WARPSYNC')
for i in range(0, len(body), chunk):
    ch = body[i:i + chunk]
    sm = sum(num(r, '# Samples') for r in ch)
    if sm < 0.004 * tot: continue
    agg = {h[6:]: sum(num(r, h) for r in ch) for h in st}
    top = [(k, int(v)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:3]]
    ex = max(num(r, 'Instructions Executed') for r in ch)
    ops = {}
    for r in ch:
        toks = r[col['Source']].split()
        o = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
        if o in KEY: ops[o] = ops.get(o, 0) + 1
    hot = max(ch, key=lambda r: num(r, '# Samples'))
    print(f"{i:5d} smp {sm:7.0f} {100*sm/tot:5.1f}% maxexec {ex:7.0f} {top} {ops} | hot: {hot[col['Source']].strip()[:60]} ({num(hot,'# Samples'):.0f})")
