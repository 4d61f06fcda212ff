#!/usr/bin/env python
"""Opcode histogram of the biggest loop (largest backward-branch span that holds a BAR or a barrier-free body) of a kernel.
   python tools/loop_hist.py <mangled-substring> [lib.so] [--dump file]"""
import collections, os, re, subprocess, sys
name = sys.argv[1]
lib = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith('--') else os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
ins, on = [], False
for l in out.split('\n'):
    if 'Function :' in l:
        on = name in l
        continue
    if on:
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
cands = []
for a, t in ins:
    m = re.search(r'BRA(?:\.\w+)* 0x([0-9a-f]+)', t)
    if m and int(m.group(1), 16) < a and 'U.ANY' not in t:
        cands.append((int(m.group(1), 16), a))
# the hot loop: the span with the most SHFL.UP/DOWN per instruction among spans > 300 instructions
def body(c): return [(a, t) for a, t in ins if c[0] <= a <= c[1]]
best = max((c for c in cands if len(body(c)) > 300), key=lambda c: sum('SHFL.UP' in t or 'SHFL.DOWN' in t for a, t in body(c)) / len(body(c)))
b = body(best)
ops = collections.Counter(re.sub(r'^@!?U?P\w+\s+', '', t).split()[0] for a, t in b)
print('loop %x..%x: %d instructions' % (best[0], best[1], len(b)))
print(' '.join('%s:%d' % kv for kv in ops.most_common(80)))
if '--dump' in sys.argv:
    open(sys.argv[sys.argv.index('--dump') + 1], 'w').write('\n'.join('%x  %s' % x for x in b))
