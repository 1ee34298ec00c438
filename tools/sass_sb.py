"""Scoreboard view of a kernel's SASS: for every instruction matching a pattern print its write-barrier index, and for every
instruction that waits on barriers the wait mask (decoded from the control bits of `cuobjdump -sass` hex words).
    cuobjdump -sass -fun <mangled> file.o | python tools/sass_sb.py [pattern]"""
import re, sys
pat = sys.argv[1] if len(sys.argv) > 1 else "LDG.E.NA.128"
lines = sys.stdin.read().split("\n")
ins = []
i = 0
while i < len(lines):
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/", lines[i])
    if m and i + 1 < len(lines):
        m2 = re.match(r"\s+/\* 0x([0-9a-f]{16}) \*/", lines[i + 1])
        if m2:
            hi = int(m2.group(1), 16)
            ctrl = (hi >> 41) & 0x1FFFFF
            ins.append((m.group(1), m.group(2).strip(), ctrl & 0xF, (ctrl >> 5) & 7, (ctrl >> 8) & 7, (ctrl >> 11) & 0x3F))
            i += 2
            continue
    i += 1
for k, (addr, txt, stall, wr, rd, wait) in enumerate(ins):
    if pat in txt:
        print(f"{addr} wr=SB{wr} {txt[:70]}")
    elif wait and "--waits" in sys.argv:
        print(f"{addr}        wait={wait:06b} {txt[:60]}")
