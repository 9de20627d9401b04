"""Where one warp's cycles go inside the Newton loop of a generated thread-per-instance kernel: static stall counts
(SASS control codes) summed per PHASE of the generated source (MOSFET linearisation, LU rows, back substitution,
damping/norm, redo check), using nvdisasm's per-instruction line info.
usage: tools/sass_phases.py build/tpi_<name>_<mask>.cubin build/tpi_<name>_<mask>.cu"""
import collections, re, subprocess, sys
cubin, src = sys.argv[1], sys.argv[2]
lines = open(src).read().splitlines()
phase_of_line = {}
cur = "prologue"
for n, l in enumerate(lines, 1):
    if "for (int iter = 0" in l: cur = "newton-head"
    elif "mos_eval" in l and cur in ("newton-head", "mos"): cur = "mos"
    elif "// position " in l: cur = "lu"
    elif re.search(r"^\s*double q0_?\d*;", l) or ("double q" in l and cur == "lu" and "qq" not in l and re.search(r"double q\d+_0;", l)): cur = "backsub"
    elif "const double n0" in l: cur = "damp+norm"
    elif "const bool redo" in l: cur = "redo-check"
    elif "total" in l and "+= count" in l: cur = "step-tail"
    phase_of_line[n] = cur
out = subprocess.run(["nvdisasm", "-g", "-c", "-hex", cubin], capture_output=True, text=True).stdout.splitlines()
ins = []
file_, line_ = None, None
i = 0
while i < len(out):
    l = out[i]
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        file_, line_ = m.group(1), int(m.group(2)); i += 1; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", l)
    if m and i + 1 < len(out):
        m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", out[i + 1])
        if m2:
            ins.append((int(m.group(1), 16), m.group(2).strip(), int(m2.group(1), 16), file_, line_)); i += 2; continue
    i += 1
spans = []
for a, t, h, f, ln in ins:
    mm = re.search(r"BRA\S*\s+(?:\S+,\s*)?`?\(?\.?(L_x_\d+)", t) if "BRA" in t else None
is64 = lambda t: re.match(r"(@!?U?P\d+ )?D(ADD|MUL|FMA|SETP)", t) is not None
# Newton loop = instructions whose spec.cu line falls in the newton phases
tot = collections.Counter(); cnt = collections.Counter(); f64 = collections.Counter()
for a, t, h, f, ln in ins:
    if f is None: continue
    if f.endswith("spec.cu"):
        ph = phase_of_line.get(ln, "?")
    else:
        ph = None
    if ph is None:
        # inlined runtime helper: attribute to the phase of the last spec.cu line seen
        ph = last if 'last' in dir() else "?"
    else:
        last = ph
    st = (h >> 41) & 0xF
    tot[ph] += st; cnt[ph] += 1; f64[ph] += 1 if is64(t) else 0
for ph in tot:
    print(f"{ph:12s} instr {cnt[ph]:5d}  fp64 {f64[ph]:5d}  stall sum {tot[ph]:6d}  (fp64 pipe floor {2*f64[ph]:5d})")
