"""Map local-memory (spill) instructions and key ops to regions of a kernel's SASS.  Usage: sass_map.py <lib.so> <name-substring>"""
import re, subprocess, sys
out = subprocess.run(['cuobjdump', '-sass', sys.argv[1]], capture_output=True, text=True).stdout
for f in out.split('Function : ')[1:]:
    name = f.split('\n')[0]
    if sys.argv[2] not in name: continue
    lines = [l for l in f.split('\n') if re.match(r'\s+/\*[0-9a-f]{4,5}\*/', l)]
    marks = []
    for i, l in enumerate(lines):
        for k in ('LDL', 'STL', 'MUFU.EX2', 'UTCHMMA', 'UTMALDG', 'LDTM', 'STTM', 'BAR.SYNC', 'USETMAXREG', 'SYNCS.PHASECHK'):
            if k in l: marks.append((i, k))
    prev = None; cnt = 0; start = 0; last = 0
    for i, k in marks + [(None, None)]:
        if k != prev:
            if prev: print(f"{start:6d}-{last:6d} {prev} x{cnt}")
            prev = k; cnt = 0; start = i
        cnt += 1; last = i
    print(len(lines), 'instructions')
