"""Per-kernel SASS evidence: one row per kernel of libddls_b200.so with the counts of the mnemonics that prove the
Blackwell-native paths (cuobjdump -sass split at `Function :`).  Runs on the CPU box.
    python tools/sass_per_kernel.py > profiles/sass_per_kernel.txt"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, 'distributeddeeplearningspark_b200', '_lib', 'libddls_b200.so')
COLS = ['UTCHMMA', 'UTMALDG', 'UBLKCP', 'LDTM', 'STTM', 'UTCBAR', 'SYNCS', 'LDGMC', 'REDG', 'LDGSTS', 'HMMA', 'ACQBULK', 'MEMBAR']
out = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(['c++filt', n], capture_output=True, text=True).stdout.strip() or n
rows, cur, counts = [], None, None
for line in out.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        if cur:
            rows.append((cur, counts))
        cur, counts = m.group(1), dict.fromkeys(COLS, 0)
        continue
    if cur:
        m = re.search(r'^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', line)
        if m:
            op = m.group(1)
            for c in COLS:
                if op.startswith(c):
                    counts[c] += 1
if cur:
    rows.append((cur, counts))
print('# SASS mnemonics per kernel of libddls_b200.so (sm_100a).  UTCHMMA = tcgen05.mma, UTMALDG = cp.async.bulk.tensor (TMA),')
print('# UBLKCP = cp.async.bulk, LDTM/STTM = tcgen05.ld/st, UTCBAR = tcgen05.commit, SYNCS = mbarrier, LDGMC = multimem.ld_reduce,')
print('# REDG = red/multimem.red, LDGSTS = cp.async, HMMA = legacy mma.sync (none expected), ACQBULK = griddepcontrol.wait, MEMBAR = fences')
print(f'{"kernel":100s} ' + ' '.join(f'{c:>7s}' for c in COLS))
for name, c in sorted(rows, key=lambda r: demangle(r[0])):
    d = re.sub(r'\(anonymous namespace\)::', '', demangle(name))
    d = re.sub(r'\(.*$', '', d)
    print(f'{d[:100]:100s} ' + ' '.join(f'{c[k]:7d}' for k in COLS))
