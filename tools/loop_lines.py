#!/usr/bin/env python
"""Per-source-line opcode census of the HOT LOOP (last backward branch) of one kernel in the shipped library.
   python tools/loop_lines.py <mangled-substring> [lib.so] [top]"""
import collections, os, re, subprocess, sys, tempfile
kname = sys.argv[1]
lib = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
best = None
with tempfile.TemporaryDirectory() as td:
    subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.abspath(lib)], cwd=td, stdout=subprocess.DEVNULL)
    for f in sorted(os.listdir(td)):
        out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(td, f)], capture_output=True, text=True).stdout
        for fn in re.split(r'\n\s*\.section\s+\.text\.', out)[1:]:
            if kname in fn.split(',')[0].split()[0]:
                best = fn
if best is None:
    sys.exit('kernel not found')
ins, line = [], None
labels = {}
for l in best.split('\n'):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        line = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r'^(\.L_x_\d+):', l.strip())
    if m:
        labels[m.group(1)] = len(ins)
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        t = re.sub(r'^@!?U?P\w+\s+', '', m.group(2).strip())
        ins.append((int(m.group(1), 16), t, line))
# hot loop: the backward branch spanning the most instructions
span = None
for i, (a, t, ln) in enumerate(ins):
    if t.startswith('BRA'):
        m = re.search(r'(\.L_x_\d+)', t)
        if m and m.group(1) in labels and labels[m.group(1)] < i:
            if span is None or i - labels[m.group(1)] > span[1] - span[0]:
                span = (labels[m.group(1)], i)
body = ins[span[0]:span[1] + 1]
print('loop: %d instructions' % len(body))
ops = collections.Counter(t.split()[0].split('.')[0] for a, t, ln in body)
print(' '.join('%s:%d' % kv for kv in ops.most_common(40)))
agg, per = collections.Counter(), collections.defaultdict(collections.Counter)
for a, t, ln in body:
    agg[ln] += 1
    per[ln][t.split()[0].split('.')[0]] += 1
srcs = {}
for ln, n in agg.most_common(top):
    text = ''
    if ln:
        path = os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'csrc', ln[0])
        if ln[0] not in srcs and os.path.exists(path):
            srcs[ln[0]] = open(path).read().split('\n')
        if ln[0] in srcs and ln[1] <= len(srcs[ln[0]]):
            text = srcs[ln[0]][ln[1] - 1].strip()[:64]
    print('%4d %s:%s %-64s %s' % (n, ln[0] if ln else '?', ln[1] if ln else '?', text, ' '.join('%s%d' % kv for kv in per[ln].most_common(6))))
