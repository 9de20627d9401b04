"""Static stall sums (SASS control codes) of the two instruction streams of a warp-pair kernel, between consecutive
named-barrier instructions of the generated source: where each warp's cycles go and what it waits for.
usage: tools/sass_pair.py build/tpi_<name>_<mask>.cubin build/tpi_<name>_<mask>.cu"""
import collections, re, subprocess, sys
cubin, src = sys.argv[1], sys.argv[2]
lines = open(src).read().splitlines()
seg_of_line = {}
role, seg, inloop = "pre", 0, False
names = {}
for n, l in enumerate(lines, 1):
    if "if (tid & 32) {" in l: role, seg = "helper", 0
    elif role == "helper" and l.strip() == "return;": role = "master-pre"
    elif "for (int iter = 0" in l: role, seg = "master", 0
    elif role == "master" and "+= count" in l and "total" in l: role = "master-tail"
    m = re.search(r"PAIR_(SYNC|ARRIVE)\((\d+)\)", l)
    if m and role in ("helper", "master") and "#define" not in l:
        seg += 1
        names[(role, seg)] = f"{m.group(1).lower()} {m.group(2)}"
    if role == "master" and "double q0_0;" in l: seg += 1; names[(role, seg)] = "backsub.."
    if role == "master" and "const bool redo" in l: seg += 1; names[(role, seg)] = "redo/damp/norm"
    seg_of_line[n] = (role, seg)
out = subprocess.run(["nvdisasm", "-gi", "-c", "-hex", cubin], capture_output=True, text=True).stdout.splitlines()
tot, cnt = collections.Counter(), collections.Counter()
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
            h = int(m2.group(1), 16)
            key = seg_of_line.get(line_, ("?", 0)) if (file_ or "").endswith("spec.cu") else ("other-file", 0)
            tot[key] += (h >> 41) & 0xF; cnt[key] += 1
            i += 2
            continue
    i += 1
for role in ("helper", "master"):
    acc = 0
    print(role)
    for (r, s) in sorted(k for k in tot if k[0] == role):
        acc += tot[(r, s)]
        print(f"  after {names.get((r, s), 'start'):16s} instr {cnt[(r, s)]:5d} stall {tot[(r, s)]:5d}  cumulative {acc:6d}")

# ---- steady-state timeline: replay both streams' segments against each other (arrive -> released sync: 40 cycles) ----
def stream(role):
    segs = sorted(k for k in tot if k[0] == role)
    ops = []
    for (r, s) in segs:
        nm = names.get((r, s), "start")
        ops.append((nm, tot[(r, s)]))
    return ops
L = 40
H, M = stream("helper"), stream("master")
# helper: "start" segment + first "sync 3" block run once (prologue); its loop body starts at "sync 1"
def body(ops, first):
    idx = [i for i, (n, c) in enumerate(ops) if n == first]
    return ops[idx[0]:] if idx else ops
Hb, Mb = body(H, "sync 1"), body(M, "sync 1")
arr = {}      # (barrier id, iteration) -> time of arrive / of each side reaching a both-sides sync
def run(iters=4):
    th = tm = 0.0
    # event-driven: iterate segments alternately until both finish `iters` iterations
    hi = mi = 0; hit = mit = 0
    waiting = {}
    times = []
    while hit < iters or mit < iters:
        progressed = False
        for who in ("m", "h"):
            ops, i, it, t = (Mb, mi, mit, tm) if who == "m" else (Hb, hi, hit, th)
            if it >= iters: continue
            name, cost = ops[i]
            kind, bid = (name.split() + ["0"])[:2] if name.split()[0] in ("sync", "arrive") else ("none", "0")
            key = (bid, it if not (kind == "sync" and bid == "3" and who == "h") else it)  # helper's trailing sync 3 pairs with the master's arrive 3 of the same iteration
            if kind == "arrive":
                arr[(who, key)] = t
            elif kind == "sync":
                other = "h" if who == "m" else "m"
                if bid == "1":      # both sides sync
                    arr[(who, key)] = t
                    if (other, key) not in arr: continue
                    t = max(t, arr[(other, key)]) + 17
                else:
                    if (other, key) not in arr: continue
                    t = max(t, arr[(other, key)] + L)
            t += cost
            i += 1
            if i == len(ops):
                i = 0; it += 1
                if who == "m": times.append(t)
            if who == "m": mi, mit, tm = i, it, t
            else: hi, hit, th = i, it, t
            progressed = True
        if not progressed: print("timeline: stuck"); break
    return times
ts = run()
if len(ts) >= 3: print(f"timeline: steady-state Newton iteration {ts[-1] - ts[-2]:.0f} cycles (static), master alone {sum(c for n, c in Mb)}, helper alone {sum(c for n, c in Hb)}")
