"""Lists the loops (backward branches) of one kernel in a built library with their instruction count and local-memory traffic.
usage: python tools/sass_loops.py <lib.so> <kernel-name-substring>"""
import re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
names = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
cur, ins = None, {}
for l in names.split('\n'):
    m = re.search(r'Function : (\S+)', l)
    if m:
        cur = m.group(1); ins[cur] = []; continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', l)
    if m and cur:
        ins[cur].append((int(m.group(1), 16), m.group(2).strip()))
for fn, body in ins.items():
    if pat not in fn:
        continue
    print(fn, len(body), 'instructions')
    for a, t in body:
        m = re.search(r'BRA\S*\s+(?:\S+,\s*)?(0x[0-9a-f]+)', t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt < a and a - tgt < 0x3000:
                inner = [x for x in body if tgt <= x[0] <= a]
                loc = sum(1 for x in inner if re.search(r'\b(LDL|STL)', x[1]))
                print('  loop %#x..%#x  %4d instr  local ld/st %d' % (tgt, a, len(inner), loc))
