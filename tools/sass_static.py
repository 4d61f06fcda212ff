#!/usr/bin/env python
"""Static SASS census of one kernel of the shipped library: instructions per CUDA source line and
per opcode (no profile needed).   python tools/sass_static.py <kernel-substring> [lib.so] [top]"""
import collections, os, re, subprocess, sys, tempfile

def main():
    kname = sys.argv[1]
    lib = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
    best = None
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.abspath(lib)], cwd=td, stdout=subprocess.DEVNULL)
        for f in sorted(os.listdir(td)):
            out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(td, f)], capture_output=True, text=True).stdout
            for fn in re.split(r'\n\s*\.section\s+\.text\.', out)[1:]:
                name = fn.split(',')[0].split()[0]
                if kname in name:
                    best = fn
    if best is None:
        sys.exit('kernel not found')
    line, agg, ops = None, collections.Counter(), collections.defaultdict(collections.Counter)
    srcs = {}
    for l in best.split('\n'):
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
        if m:
            op = re.sub(r'^@!?U?P\w+\s+', '', m.group(2).strip()).split()[0].split('.')[0]
            agg[line] += 1
            ops[line][op] += 1
    tot = sum(agg.values())
    print('total %d instructions' % tot)
    for ln, n in agg.most_common(top):
        text = ''
        if ln:
            path = os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'csrc', ln[0])
            if ln[0] not in srcs and os.path.exists(path):
                srcs[ln[0]] = open(path).read().split('\n')
            if ln[0] in srcs and ln[1] <= len(srcs[ln[0]]):
                text = srcs[ln[0]][ln[1] - 1].strip()[:70]
        print('%5d  %s:%s  %-70s %s' % (n, ln[0] if ln else '?', ln[1] if ln else '?', text,
                                       ' '.join('%s%d' % kv for kv in ops[ln].most_common(5))))

if __name__ == '__main__':
    main()
