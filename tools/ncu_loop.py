"""Hot-loop view of an ncu report: instruction mix and top stall instructions of the region whose per-warp execution
count lies in [lo, hi].  usage: python tools/ncu_loop.py X.ncu-rep LO HI [N]"""
import collections, csv, subprocess, sys
rep, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
top = int(sys.argv[4]) if len(sys.argv) > 4 else 20
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi_ = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
h = rows[hi_[0]]; col = {n: i for i, n in enumerate(h)}
end = hi_[1] - 1 if len(hi_) > 1 else len(rows)
body = [r for r in rows[hi_[0] + 1:end] if len(r) == len(h)]
S = lambda r: float(r[col['# Samples']] or 0)
tot = sum(S(r) for r in body)
loop = [(i, r) for i, r in enumerate(body) if lo <= int(r[col['Instructions Executed']] or 0) <= hi]
op = lambda r: (r[col['Source']].split()[1] if r[col['Source']].strip().startswith('@') else r[col['Source']].split()[0])
print(f"{len(loop)} instructions in loop (idx {loop[0][0]}..{loop[-1][0]}), {sum(S(r) for _, r in loop):.0f} of {tot:.0f} samples")
print(collections.Counter(op(r) for _, r in loop).most_common(30))
stalls = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
agg = collections.Counter()
for _, r in loop:
    for n in stalls: agg[n[6:]] += float(r[col[n]] or 0)
print("stall totals:", [(a, int(v)) for a, v in agg.most_common(8)])
for i, r in sorted(loop, key=lambda ir: -S(ir[1]))[:top]:
    st = sorted(((n[6:], float(r[col[n]] or 0)) for n in stalls), key=lambda kv: -kv[1])[:2]
    print(i, int(S(r)), r[col['Source']].strip()[:80], [(a, int(v)) for a, v in st if v > 0])
