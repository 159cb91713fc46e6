"""Where do warps wait?  Lists every mbarrier try_wait / bar / long-latency stall site of an ncu report with its
execution count, samples and barrier offset.  usage: python tools/ncu_waits.py X.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi_ = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
h = rows[hi_[0]]; col = {n: i for i, n in enumerate(h)}
end = hi_[1] - 1 if len(hi_) > 1 else len(rows)
body = [r for r in rows[hi_[0] + 1:end] if len(r) == len(h)]
S = lambda r: float(r[col['# Samples']] or 0)
tot = sum(S(r) for r in body)
print("total samples", tot)
for i, r in enumerate(body):
    src = r[col['Source']].strip()
    if 'TRYWAIT' in src or 'BAR.' in src or 'ACQBULK' in src or 'griddep' in src.lower():
        nxt = sum(S(x) for x in body[i:i + 3])
        print(f"{i:5d} exec={r[col['Instructions Executed']]:>8s} samples(+2)={nxt:6.0f} ({100*nxt/tot:4.1f}%)  {src[:90]}")
