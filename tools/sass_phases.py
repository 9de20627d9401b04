"""Where one warp's cycles go inside the Newton loop of a generated thread-per-instance kernel: static stall counts
(SASS control codes) summed per PHASE of the generated source — MOSFET linearisation, LU rows (assembly +
elimination + parking), back substitution, redo check, damping/norm — using nvdisasm's inlining chain (-gi) to map
every instruction to the line of the generated source it comes from.  Optional third argument: `rows` also prints
the LU phase per matrix row.
usage: tools/sass_phases.py build/tpi_<name>_<mask>.cubin build/tpi_<name>_<mask>.cu [rows]"""
import collections, re, subprocess, sys
cubin, src = sys.argv[1], sys.argv[2]
lines = open(src).read().splitlines()
phase_of_line, row_of_line = {}, {}
cur, row = "prologue", -1
for n, l in enumerate(lines, 1):
    if "for (int iter = 0" in l: cur = "newton-head"
    elif "mos_eval" in l and cur in ("newton-head", "mos"): cur = "mos"
    elif "// position " in l:
        cur = "lu"
        row = int(re.search(r"// position (\d+)", l).group(1))
    elif re.match(r"^\s*double q0(_\d+)?;", l) and cur == "lu": cur = "backsub"
    elif "const bool redo" in l: cur = "redo-check"
    elif re.match(r"^\s*const double (xo0|n0)(_\d+)? =", l) and cur == "redo-check": cur = "damp+norm"
    elif "+= count" in l and "total" in l: cur = "step-tail"
    phase_of_line[n] = cur
    m = re.search(r"\b(?:r|rb|f)(\d+)(?:_\d+)*_\d+ =", l)  # statements of the interleaved rows name their row
    row_of_line[n] = int(m.group(1)) if (m and cur == "lu") else row
out = subprocess.run(["nvdisasm", "-gi", "-c", "-hex", cubin], capture_output=True, text=True).stdout.splitlines()
is64 = lambda t: re.match(r"(@!?U?P\d+ )?D(ADD|MUL|FMA|SETP)", t) is not None
tot, cnt, f64 = collections.Counter(), collections.Counter(), collections.Counter()
rtot, rcnt = collections.Counter(), collections.Counter()
file_, line_ = None, None
i = 0
while i < len(out):
    l = out[i]
    m = re.search(r'//## File "([^"]+)", line (\d+)\s*$', l) or re.search(r'//## File "([^"]+)", line (\d+) inlined', l)
    if m:
        if l.rstrip().endswith(m.group(2)) or file_ is None or "inlined at" not in l:
            file_, line_ = m.group(1), int(m.group(2))
        i += 1
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", l)
    if m and i + 1 < len(out):
        m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", out[i + 1])
        if m2:
            t, h = m.group(2).strip(), int(m2.group(1), 16)
            ph = phase_of_line.get(line_, "?") if (file_ or "").endswith("spec.cu") else "other-file"
            st = (h >> 41) & 0xF
            tot[ph] += st; cnt[ph] += 1; f64[ph] += 1 if is64(t) else 0
            if ph == "lu":
                rtot[row_of_line.get(line_, -1)] += st; rcnt[row_of_line.get(line_, -1)] += 1
            i += 2
            continue
    i += 1
for ph in ("newton-head", "mos", "lu", "backsub", "redo-check", "damp+norm", "step-tail", "prologue", "other-file", "?"):
    if cnt[ph]:
        print(f"{ph:12s} instr {cnt[ph]:5d}  fp64 {f64[ph]:5d}  stall sum {tot[ph]:6d}  (fp64 pipe floor {2*f64[ph]:5d})")
if len(sys.argv) > 3:
    for r in sorted(rcnt):
        print(f"  row {r:3d}: instr {rcnt[r]:4d} stall {rtot[r]:5d}")
