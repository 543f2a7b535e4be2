"""profiles/sass_summary.txt: per kernel of libnmpc_b200.so the SASS instruction count, registers and
the counts of the mnemonics that prove what the code uses (UBLKCP = TMA bulk copy, LDGSTS = cp.async,
SYNCS = mbarrier, DFMA / FFMA, REDUX, no UTC*MMA / LDTM tensor-core ops).  No GPU needed.

    python scripts/sass_summary.py > profiles/sass_summary.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
lib = ROOT / 'rotors_mpc_controller_b200' / 'libnmpc_b200.so'
sass = subprocess.run(['cuobjdump', '-sass', str(lib)], capture_output=True, text=True).stdout
res = subprocess.run(['cuobjdump', '-res-usage', str(lib)], capture_output=True, text=True).stdout
regs = {}
cur = None
for line in res.splitlines():
    m = re.search(r'Function (\S+):', line)
    if m:
        cur = m.group(1)
    m = re.search(r'REG:(\d+).*?SHARED:(\d+)', line)
    if m and cur:
        regs[cur] = (int(m.group(1)), int(m.group(2)))
KEYS = ['UBLKCP', 'LDGSTS', 'SYNCS', 'DFMA', 'DMUL', 'DADD', 'FFMA', 'MUFU', 'REDUX', 'LDS', 'STS',
        'LDG', 'STG', 'BAR', 'WARPSYNC', 'UTCHMMA', 'UTCQMMA', 'UTCIMMA', 'LDTM', 'HMMA', 'DMMA']
fn, counts, total = None, None, 0
print(f'# SASS summary of {lib.name} (cuobjdump -sass / -res-usage), arch from the ELF: '
      + ', '.join(sorted(set(re.findall(r'arch = (sm_\w+)', sass)))))


def flush():
    if fn is None:
        return
    name = subprocess.run(['c++filt', fn], capture_output=True, text=True).stdout.strip()
    name = re.sub(r'\(anonymous namespace\)::', '', name)
    name = name.split('(')[0] if not name.startswith('void') else name[5:].split('(')[0]
    r = regs.get(fn, ('?', '?'))
    mn = '  '.join(f'{k} {counts[k]}' for k in KEYS if counts[k])
    print(f'{name}\n    instructions {total}, registers {r[0]}, static smem {r[1]} B\n    {mn}')


for line in sass.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        flush()
        fn, counts, total = m.group(1), collections.Counter(), 0
        continue
    m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
    if m and fn:
        total += 1
        op = m.group(1)
        for k in KEYS:
            if op.startswith(k):
                counts[k] += 1
                break
flush()
