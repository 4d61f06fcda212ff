#!/usr/bin/env python
"""Decode the scoreboard control bits of one kernel's SASS and list, inside the hot loop, every
instruction that waits on a scoreboard a global load (LDG) writes -- to spot false dependencies
(e.g. a write-after-write on a dead component of a vector load).
    python tools/sb_waits.py <mangled-substring> [lib.so]"""
import os, re, subprocess, sys
lib = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout.split('\n')
ins, on, i = [], False, 0
while i < len(out):
    l = out[i]
    if 'Function :' in l:
        on = sys.argv[1] in l
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/', l) if on else None
    if m:
        m2 = re.match(r'\s+/\* (0x[0-9a-f]+) \*/', out[i + 1])
        c = (int(m2.group(1), 16) if m2 else 0) >> 41
        ins.append((int(m.group(1), 16), m.group(2), c & 0xf, (c >> 5) & 7, (c >> 8) & 7, (c >> 11) & 0x3f))
        i += 1
    i += 1
loop = None
for a, t, *_ in ins:
    if 'BRA' in t and not 'DIV' in t:
        m = re.search(r'BRA 0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) < a and loop is None:
            loop = (int(m.group(1), 16), a)
print('loop %x..%x' % loop)
ldg_sb = set(wr for a, t, st, wr, rd, w in ins if loop[0] <= a <= loop[1] and 'LDG' in t and wr != 7)
print('scoreboards written by LDG:', sorted(ldg_sb))
for a, t, st, wr, rd, w in ins:
    if loop[0] <= a <= loop[1]:
        if 'LDG' in t:
            print('  %04x  %-70s wr %d' % (a, t[:70], wr))
        elif any(w >> b & 1 for b in ldg_sb):
            print('  %04x  %-70s WAITS %s' % (a, t[:70], format(w, '06b')))
