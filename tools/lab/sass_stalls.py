#!/usr/bin/env python
"""Stall counts (control bits 105-108 of each 128-bit SASS instruction) of a kernel's hottest loop, summed.

    python tools/lab/sass_stalls.py lib.so k_force_warpILi2ELi1ELi4ELi3ELb0 [--list]
"""
import re
import subprocess
import sys

lib, key = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
ins, cur, on = [], None, False
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        on = key in m.group(1)
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/", line)
    if m:
        cur = [int(m.group(1), 16), m.group(2).strip(), None]
        ins.append(cur)
        continue
    m = re.match(r"\s+/\* 0x([0-9a-f]{16}) \*/", line)
    if m and cur is not None and cur[2] is None:
        hi = int(m.group(1), 16)
        cur[2] = ((hi >> 41) & 0xf, (hi >> 45) & 1, (hi >> 46) & 7, (hi >> 49) & 7, (hi >> 52) & 0x3f)
# the loop with the most FP64 instructions
best = None
loops = []
for a, t, c in ins:
    m = re.search(r"BRA\S*\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a:
        loops.append((int(m.group(1), 16), a))
for tgt, a in loops:
    if any((t2, a2) != (tgt, a) and tgt <= t2 and a2 <= a for t2, a2 in loops):
        continue
    if True:
        seg = [x for x in ins if tgt <= x[0] <= a]
        n = sum(1 for x in seg if re.sub(r"^@!?U?P\d\s+", "", x[1]).startswith(("DFMA", "DADD", "DMUL")))
        if best is None or n > best[0]:
            best = (n, seg)
seg = best[1]
tot = sum(c[0] for _, _, c in seg)
print(f"loop of {len(seg)} instr, {best[0]} FP64, sum of stall counts {tot}")
if "--list" in sys.argv:
    for a, t, c in seg:
        print(f"  /*{a:05x}*/ stall {c[0]:2d} y{c[1]} wb{c[2]} rb{c[3]} wait{c[4]:02x}  {t}")
