"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel time of the last training step."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
cols = rows[hdr]
ki, vi = cols.index('Kernel Name'), cols.index('Metric Value')
data = [(r[ki], float(r[vi].replace(',', ''))) for r in rows[hdr + 2:] if len(r) > vi]
key = sys.argv[2] if len(sys.argv) > 2 else 'adam'
idx = [i for i, (k, v) in enumerate(data) if key in k]
a, b = idx[-2] + 1, idx[-1] + 1
tot = 0
for k, v in data[a:b]:
    tot += v
    print(f'{v/1000:9.2f} us  {k[:110]}')
print(f'{tot/1000:9.2f} us  TOTAL ({b-a} launches; cold-cache, serialised: compare shares)')
