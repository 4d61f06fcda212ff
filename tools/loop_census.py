#!/usr/bin/env python
"""Opcode census of the hot loop (between the target of the last backward branch and the branch) of
one kernel in the shipped library.  python tools/loop_census.py <mangled-substring> [unroll]"""
import collections, re, subprocess, sys, os
lib = os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
if len(sys.argv) > 3: lib = sys.argv[3]
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
name, unroll = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 3
ins, on = [], False
for l in out.split('\n'):
    if 'Function :' in l:
        on = name in l
        continue
    if on:
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
        if m:
            t = re.sub(r'^@!?U?P\w+\s+', '', m.group(2).strip())
            ins.append((int(m.group(1), 16), t))
best = None
for a, t in ins:
    if t.startswith('BRA'):
        m = re.search(r'0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) < a and best is None:   # first backward branch = the hot loop (cold shuffle paths follow EXIT)
            best = (int(m.group(1), 16), a)
ops = collections.Counter(t.split()[0].split('.')[0] for a, t in ins if best[0] <= a <= best[1])
n = sum(ops.values())
print('loop %x..%x: %d instructions, %.1f per step' % (best[0], best[1], n, n / unroll))
print(' '.join('%s:%.1f' % (k, v / unroll) for k, v in ops.most_common(40)))
