"""Compact opcode stream of the softmax step loop of a forward-kernel instance (reads cuobjdump -sass output):
M MUFU.EX2, f FFMA2, a FADD2, p F2FP pack, S STTM, L LDTM, x FMNMX3, W mbarrier try_wait, R mbarrier arrive, B branch.
  python tools/sass_stream.py libfa100.so [mangled-substring, default the bf16 d128 instance]"""
import re, subprocess, sys
lib = sys.argv[1]
sub = sys.argv[2] if len(sys.argv) > 2 else "ILi128ELi2ELb0ELb0ELb0E"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
blocks = out.split("Function : ")
blk = [b for b in blocks if sub in b.split("\n")[0] and "fa100_fwd_kernel" in b.split("\n")[0]][0]
code = {'MUFU.EX2': 'M', 'FFMA2': 'f', 'FADD2': 'a', 'F2FP': 'p', 'STTM': 'S', 'FMNMX3': 'x', 'FMNMX': 'x', 'LDTM': 'L', 'BRA': 'B',
        'SYNCS.PHASECHK': 'W', 'SYNCS.ARRIVE': 'R', 'VOTE': 'V', 'WARPSYNC': 'w', 'FFMA': 'F', 'FADD': 'A', 'FMUL': 'A', 'UTCHMMA': 'T', 'UTCQMMA': 'T'}
s = ''
for l in blk.splitlines():
    m = re.match(r'\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);', l)
    if not m: continue
    t = m.group(2).strip().split()
    op = t[1] if t[0].startswith('@') else t[0]
    ch = '.'
    for k, v in code.items():
        if op.startswith(k): ch = v; break
    s += ch
# print from the first LDTM burst to the last STTM of the main loop region that contains >= 100 MUFUs
i0 = s.find('L')
print(len(s), "instructions; stream from first LDTM:")
seg = s[i0:i0 + 4200]
for i in range(0, len(seg), 120): print(seg[i:i + 120])
