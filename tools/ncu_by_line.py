#!/usr/bin/env python
"""Join an `ncu --page source --csv` export with `nvdisasm -gi` output: executed instructions, thread
instructions and stall samples per source line (innermost line and outermost kernel-file line)."""
import collections
import csv
import re
import sys

src_csv, dis_txt, func_pat = sys.argv[1:4]
top_n = int(sys.argv[4]) if len(sys.argv) > 4 else 40
KERNEL_FILE = 'avr_volpath.cu'  # phase ranges refer to the innermost frame inside this file
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
inst = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    inst.append((r[ix['Source']].strip(), int(r[ix['Instructions Executed']] or 0),
                 int(r[ix['Thread Instructions Executed']] or 0), int(r[ix['# Samples']] or 0)))
# parse disassembly
lines = open(dis_txt).read().split('\n')
start = next(i for i, l in enumerate(lines) if l.startswith('.text.') and func_pat in l)
cur_inner = cur_outer = None
locs = []
chain = []
prev_was_loc = False
for l in lines[start + 1:]:
    if l.startswith('.text.') or l.startswith('.section'):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        if not prev_was_loc:
            chain = []
        chain.append(f'{m.group(1).split("/")[-1]}:{m.group(2)}')
        if m.group(3):
            chain.append(f'{m.group(3).split("/")[-1]}:{m.group(4)}')
        cur_inner = chain[0]
        kf = [c for c in chain if c.startswith(KERNEL_FILE + ':')]
        cur_outer = kf[0] if kf else chain[-1]
        prev_was_loc = True
        continue
    prev_was_loc = False
    m = re.search(r'/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        locs.append((cur_inner, cur_outer, m.group(2)))
print(f'# {len(inst)} ncu rows, {len(locs)} disassembled instructions', file=sys.stderr)
n = min(len(inst), len(locs))
for key_i, title in ((1, 'kernel-file line (innermost frame inside the kernel file)'), (0, 'innermost line')):
    ex = collections.Counter(); th = collections.Counter(); sm = collections.Counter()
    for (s, e, t, p), loc in zip(inst[:n], locs[:n]):
        ex[loc[key_i]] += e; th[loc[key_i]] += t; sm[loc[key_i]] += p
    T = sum(ex.values()); S = sum(sm.values())
    print(f'## by {title}: total warp-inst {T}, samples {S}')
    for k, v in ex.most_common(top_n):
        print(f'{k:34s} inst {v:11d} {100*v/T:5.1f}%  lanes {th[k]/max(v,1):5.1f}  samples {100*sm[k]/S:5.1f}%')

# optional: phase ranges "name:lo-hi,name:lo-hi" on the kernel file's lines (argv[5])
if len(sys.argv) > 5:
    ex = collections.Counter(); th = collections.Counter(); sm = collections.Counter()
    ranges = []
    for part in sys.argv[5].split(','):
        name, r = part.split(':')
        lo, hi = r.split('-')
        ranges.append((name, int(lo), int(hi)))
    for (s, e, t, p), loc in zip(inst[:n], locs[:n]):
        f, ln_ = loc[1].rsplit(':', 1)
        name = 'other'
        for nm, lo, hi in ranges:
            if lo <= int(ln_) <= hi:
                name = nm
        ex[name] += e; th[name] += t; sm[name] += p
    T = sum(ex.values()); TT = sum(th.values()); S = sum(sm.values())
    print('## by phase')
    for k, v in ex.most_common():
        print(f'{k:10s} warp-inst {v:11d} {100*v/T:5.1f}%  thread-inst {th[k]:12d} {100*th[k]/TT:5.1f}%  lanes {th[k]/max(v,1):5.1f}  samples {100*sm[k]/S:5.1f}%')
    # static code size per phase (SASS instructions)
    st = collections.Counter()
    for loc in locs[:n]:
        f, ln_ = loc[1].rsplit(':', 1)
        name = 'other'
        for nm, lo, hi in ranges:
            if lo <= int(ln_) <= hi:
                name = nm
        st[name] += 1
    print('## static SASS instructions per phase:', dict(st.most_common()))
