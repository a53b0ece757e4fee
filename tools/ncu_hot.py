"""Top stall-sample SASS lines of a kernel from an .ncu-rep (source page)."""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kern}'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = next(i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r)
hdr = rows[h]
si, ii, src = hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed'), hdr.index('Source')
stall_cols = [i for i, c in enumerate(hdr) if c.startswith('stall_')]
data = []
for k, r in enumerate(rows[h + 1:]):
    if len(r) <= si: continue
    try: s = int(r[si]); ie = int(r[ii])
    except ValueError: continue
    top = sorted(((int(r[i]) if r[i].isdigit() else 0, hdr[i]) for i in stall_cols), reverse=True)[:2]
    data.append((s, ie, k, r[src].strip(), top))
tot = sum(d[0] for d in data)
print(f'total samples {tot}, SASS lines {len(data)}, warp-instr {sum(d[1] for d in data)}')
for s, ie, k, text, top in sorted(data, reverse=True)[:n]:
    print(f'{100*s/max(tot,1):5.1f}%  line {k:4d}  exec {ie:8d}  {text[:70]:70s} {top}')
