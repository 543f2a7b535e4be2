"""Attribute ncu warp-stall samples of a kernel to CUDA source lines (with inlining chains).

  cuobjdump -xelf all lib.so ; nvdisasm -gi x.cubin > dis.txt
  ncu -i rep.ncu-rep --page source --csv > src.csv
  python scripts/ncu_by_line.py dis.txt src.csv <kernel-substring> [file-substring ...]

For every SASS instruction the inline chain is taken from nvdisasm; samples are then summed per
(file, line) at the innermost level and per line of each requested file (first frame of the
chain that lies in that file).
"""
import collections
import csv
import re
import sys

dis, src, kern = sys.argv[1:4]
files = sys.argv[4:] or ['nmpc_riccati.cuh', 'nmpc_solver.cuh']

chains = []          # one chain (list of (file, line)) per instruction, in order
cur, pending, infn = [], [], False
mark = re.compile(r'//## File "([^"]+)", line (\d+)')
for line in open(dis):
    if '.section' in line and '.text.' in line:
        infn = kern in line
        continue
    if not infn:
        continue
    m = mark.search(line)
    if m:
        pending.append((m.group(1).split('/')[-1], int(m.group(2))))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,6}\*/', line):
        if pending:
            cur, pending = pending, []
        chains.append(cur)

rows = list(csv.reader(open(src)))
hdr = rows[1]
isamp = hdr.index('# Samples')
iexec = hdr.index('Instructions Executed')
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
body = rows[2:]
n = min(len(body), len(chains))
print(f'instructions: sass {len(body)} / disasm {len(chains)}')
tot = sum(int(r[isamp]) for r in body[:n])
inner = collections.Counter()
per_file = {f: collections.Counter() for f in files}
per_file_exec = {f: collections.Counter() for f in files}
stalls = {f: collections.defaultdict(collections.Counter) for f in files}
for r, ch in zip(body[:n], chains[:n]):
    s = int(r[isamp])
    e = int(r[iexec])
    if ch:
        inner[ch[0]] += s
    for f in files:
        for fr in reversed(ch):       # outermost frame in that file
            if fr[0] == f:
                per_file[f][fr[1]] += s
                per_file_exec[f][fr[1]] += e
                for i, h in stall_cols:
                    v = int(r[i] or 0)
                    if v:
                        stalls[f][fr[1]][h] += v
                break
print(f'total samples {tot}')
for f in files:
    print(f'--- by outermost line in {f}')
    for ln, s in sorted(per_file[f].items()):
        if s * 200 >= tot:
            top = ', '.join(f'{h[6:]}={v}' for h, v in stalls[f][ln].most_common(4))
            print(f'  line {ln:4d}: {100.0 * s / tot:5.1f}%  inst {per_file_exec[f][ln]:>12d}  {top}')
print('--- innermost top 25')
for (f, ln), s in inner.most_common(25):
    print(f'  {f}:{ln}  {100.0 * s / tot:5.1f}%')

import os
BOUNDS = [10 ** 9, 10 ** 9]
_ric = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'rotors_mpc_controller_b200', 'csrc', 'nmpc_riccati.cuh')
for _i, _l in enumerate(open(_ric), 1):
    if 'void forward_sweep' in _l:
        BOUNDS[0] = _i - 8
    if 'void delta_backward' in _l:
        BOUNDS[1] = _i - 4
# coarse buckets: which function of nmpc_riccati.cuh / nmpc_solver.cuh the samples fall in
def bucket(ch):
    for fr in reversed(ch):
        if fr[0] == 'nmpc_riccati.cuh':
            ln = fr[1]
            return 'factor_sweep' if ln < BOUNDS[0] else ('forward_sweep' if ln < BOUNDS[1] else 'delta_backward')
    for fr in reversed(ch):
        if fr[0] == 'nmpc_solver.cuh' and fr[1] < 225:
            return 'linearize/gradients/helpers'
    for fr in reversed(ch):
        if fr[0] == 'nmpc_model.cuh':
            return 'model'
    return 'ipm/sqp elementwise + other'
bs, be, bn = collections.Counter(), collections.Counter(), collections.Counter()
for r, ch in zip(body[:n], chains[:n]):
    b = bucket(ch)
    bs[b] += int(r[isamp])
    be[b] += int(r[iexec])
    bn[b] += 1
print('--- buckets (samples %, executed warp-instructions %, static SASS instructions)')
te = sum(be.values())
for b, s in bs.most_common():
    print(f'  {b:32s} {100.0 * s / tot:5.1f}%  {100.0 * be[b] / te:5.1f}%  {bn[b]:6d}')
