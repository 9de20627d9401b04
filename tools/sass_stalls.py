"""Static schedule estimate from SASS control codes (Volta+ encoding: bits 105..108 of the 128-bit
instruction word = stall cycles the issuing warp waits before its next instruction).
usage: tools/sass_stalls.py file.sass [start_addr end_addr]   (hex addresses; default whole kernel)"""
import collections
import re
import sys

lines = open(sys.argv[1], errors="replace").read().splitlines()
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 60
ins = []
i = 0
while i < len(lines):
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", lines[i])
    if m and i + 1 < len(lines):
        m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", lines[i + 1])
        if m2:
            ins.append((int(m.group(1), 16), m.group(2).strip(), int(m2.group(1), 16)))
            i += 2
            continue
    i += 1
sel = [x for x in ins if lo <= x[0] < hi]
tot = collections.Counter()
cnt = collections.Counter()
for addr, text, hiword in sel:
    stall = (hiword >> 41) & 0xF
    op = re.sub(r"^@!?U?P\d+\s+", "", text).split()[0].split(".")[0]
    tot[op] += stall
    cnt[op] += 1
n = len(sel)
s = sum(tot.values())
print(f"{n} instructions, sum of stall counts {s} cycles ({s / max(n, 1):.2f} per instruction)")
for op, c in cnt.most_common(14):
    print(f"  {op:10s} n={c:5d}  stall cycles={tot[op]:6d}  avg={tot[op] / c:.2f}")
