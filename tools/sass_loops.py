#!/usr/bin/env python
"""List the backward branches (loops) of one kernel in a cuobjdump -sass listing with their opcode mix:
usage: cuobjdump -sass -fun <mangled> lib.so | tools/sass_loops.py"""
import collections
import re
import sys

s = sys.stdin.read()
ins = [(int(a, 16), t) for a, t in re.findall(r"/\*([0-9a-f]{4})\*/\s+(.*?);", s)]
print(len(ins), "instructions")
for a, txt in ins:
    m = re.search(r"BRA(?:\.\w+)*\s+(?:!?UP\d+,\s*)?0x([0-9a-f]+)", txt)  # also BRA.U UP0, 0x... (uniform loop counter)
    if not m or int(m.group(1), 16) >= a:
        continue
    t = int(m.group(1), 16)
    c = collections.Counter()
    for b, x in ins:
        if t <= b <= a:
            parts = x.split()
            op = parts[1] if parts[0].startswith("@") else parts[0]
            c[op.split(".")[0]] += 1
    print(f"loop {t:#06x}..{a:#06x} ({(a - t) // 16 + 1} instr, start % 64 = {t % 64}): "
          + " ".join(f"{k}={v}" for k, v in c.most_common(12)))
