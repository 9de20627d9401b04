"""Newton-loop breakdown of a generated kernel's SASS: per opcode count, stall cycles and 'excess' stall
(beyond 2 cycles for FP64-pipe instructions, 1 for the rest).  usage: tools/sass_loop.py file.sass"""
import collections, re, sys
lines = open(sys.argv[1], errors="replace").read().splitlines()
ins = []
i = 0
while i < len(lines):
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", lines[i])
    if m and i + 1 < len(lines):
        m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", lines[i + 1])
        if m2:
            ins.append((int(m.group(1), 16), m.group(2).strip(), int(m2.group(1), 16))); i += 2; continue
    i += 1
spans = []
for a, t, h in ins:
    mm = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", t) if "BRA" in t else None
    if mm and int(mm.group(1), 16) < a:
        spans.append((a - int(mm.group(1), 16), int(mm.group(1), 16), a))
is64 = lambda t: re.match(r"(@!?U?P\d+ )?D(ADD|MUL|FMA|SETP)", t) is not None
tot64 = sum(1 for a, t, h in ins if is64(t))
newton = min((sp for sp in spans if sum(1 for a, t, h in ins if sp[1] <= a <= sp[2] and is64(t)) >= 0.5 * tot64), default=max(spans))
sel = [x for x in ins if newton[1] <= x[0] <= newton[2]]
cnt = collections.Counter(); st = collections.Counter(); ex = collections.Counter()
for a, t, h in sel:
    s = (h >> 41) & 0xF
    op = re.sub(r"^@!?U?P\d+\s+", "", t).split()[0]
    op = op.split(".")[0] + (".MOV" if ".MOV" in op else "")
    cnt[op] += 1; st[op] += s; ex[op] += max(0, s - (2 if is64(t) else 1))
print(f"loop {newton[1]:x}..{newton[2]:x}: {len(sel)} instr, stall sum {sum(st.values())}, excess {sum(ex.values())}")
for op, c in cnt.most_common(25):
    print(f"  {op:12s} n={c:5d} stall={st[op]:6d} excess={ex[op]:6d}")
