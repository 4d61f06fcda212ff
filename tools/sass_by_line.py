#!/usr/bin/env python
"""Join an `ncu --page source --csv` dump with nvdisasm line info of the shipped library:
per CUDA source line, executed warp-instructions and stall samples of one kernel.

    python tools/sass_by_line.py <ncu_source.csv> <kernel-substring> [lib.so] [top]
"""
import collections, csv, os, re, subprocess, sys, tempfile

def main():
    src_csv, kname = sys.argv[1], sys.argv[2]
    lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'lib', 'libpsfm_b200.so')
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    rows = list(csv.reader(open(src_csv)))
    secs = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
    sec = [i for i in secs if kname in rows[i][1]][0]
    end = min([i for i in secs if i > sec] + [len(rows)])
    hdr = rows[sec + 1]
    ie, ismp, isrc = hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Source')
    prof = [(r[isrc].strip(), int(r[ie]), int(r[ismp])) for r in rows[sec + 2:end] if len(r) > ie]
    mangled = None
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.abspath(lib)], cwd=td, stdout=subprocess.DEVNULL)
        best = None
        for f in sorted(os.listdir(td)):
            out = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(td, f)], capture_output=True, text=True).stdout
            # split per function
            funcs = re.split(r'\n\s*\.section\s+\.text\.', out)
            for fn in funcs[1:]:
                name = fn.split(',')[0].split()[0]
                demangled = subprocess.run(['cu++filt', name], capture_output=True, text=True).stdout.strip()
                if kname_match(demangled, rows[sec][1]):
                    best = fn
                    mangled = name
        if best is None:
            sys.exit('kernel not found in ' + lib)
    line, seq = None, []
    for l in best.split('\n'):
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
        if m:
            seq.append((line, m.group(2).strip()))
    if len(seq) != len(prof):
        print('warning: %d SASS instructions in the library vs %d in the profile' % (len(seq), len(prof)), file=sys.stderr)
    agg = collections.defaultdict(lambda: [0, 0])
    tot = 0
    for (ln, _), (_, n, s) in zip(seq, prof):
        agg[ln][0] += n
        agg[ln][1] += s
        tot += n
    srcs = {}
    print('kernel %s: %d warp-instructions' % (mangled, tot))
    for ln, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        text = ''
        if ln:
            path = os.path.join(os.path.dirname(__file__), '..', 'packnet-sfm-resnet-san_b200', 'csrc', ln[0])
            if ln[0] not in srcs and os.path.exists(path):
                srcs[ln[0]] = open(path).read().split('\n')
            if ln[0] in srcs and ln[1] <= len(srcs[ln[0]]):
                text = srcs[ln[0]][ln[1] - 1].strip()[:90]
        print('%5.1f%% %9d smp %6d  %s:%s  %s' % (100.0 * n / tot, n, s, ln[0] if ln else '?', ln[1] if ln else '?', text))

def kname_match(demangled, profiled):
    norm = lambda x: re.sub(r'\(int\)|\(bool\)|\s+|void', '', x).replace('psfm::', '').replace('false', '0').replace('true', '1')
    return norm(demangled).split('(')[0] == norm(profiled).split('(')[0]

if __name__ == '__main__':
    main()
