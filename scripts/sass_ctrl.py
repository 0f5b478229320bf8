#!/usr/bin/env python
"""Decode the scheduling control fields (stall, write/read barrier, wait mask) next to each SASS
instruction of one kernel: cuobjdump -sass <obj> | sass_ctrl.py <kernel-substring>"""
import re, sys
want = sys.argv[1]
on = False
pend = None
for line in sys.stdin:
    if "Function :" in line:
        on = want in line
        if on: print(line.strip())
        continue
    if not on: continue
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* (0x[0-9a-f]{16}) \*/", line)
    if m:
        pend = (m.group(1), m.group(2).strip())
        continue
    m = re.match(r"\s+/\* (0x[0-9a-f]{16}) \*/", line)
    if m and pend:
        hi = int(m.group(1), 16)
        stall = (hi >> 41) & 0xF; yld = (hi >> 45) & 1; wbar = (hi >> 46) & 7; rbar = (hi >> 49) & 7; wait = (hi >> 52) & 0x3F
        w = "".join(str(i) if wait >> i & 1 else "-" for i in range(6))
        print(f"{pend[0]} st{stall:2d} {'Y' if yld==0 else ' '} W{wbar if wbar<7 else '-'} R{rbar if rbar<7 else '-'} wait[{w}]  {pend[1]}")
        pend = None
