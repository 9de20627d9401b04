"""Container-side (no GPU) schedule estimate of the generated thread-per-instance kernel: emit it for a netlist
under the CSIM_TPI_OPTS masks given on the command line, compile with NVRTC for sm_100a and, inside the Newton
loop (the largest backward-branch span), report instruction count, FP64-pipe instructions and the sum of the
SASS stall counts = cycles one warp alone needs per Newton iteration before any scoreboard wait.
usage: tools/tpi_static.py netlist mask [mask ...]     (development aid; oracle only used for the pivot order)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
from cuda.bindings import nvrtc

name = sys.argv[1]
g, topo = support.load_golden(name)
plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes, topo.elem_branch,
                 topo.elem_param, topo.n_params, device=-1)
dt = float(g["dt"])
m, grp, tab = O.make_source_tables(topo, topo.params, dt, 4)
x = g["x_dc"]
G, I = O.stamp(topo, 1, topo.params, x, x, O.src_tran(topo, topo.params, m, grp, tab, 0, dt), dt=dt, gmin=1e-6)
_, perm, _ = O.lu_solve(G, I)
shim = b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n"
os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)


def parse(sass):
    lines = sass.splitlines()
    ins = []
    i = 0
    while i < len(lines):
        mm = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", lines[i])
        if mm and i + 1 < len(lines):
            m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", lines[i + 1])
            if m2:
                ins.append((int(mm.group(1), 16), mm.group(2).strip(), int(m2.group(1), 16)))
                i += 2
                continue
        i += 1
    return ins


for mask in sys.argv[2:]:
    if mask != "default":
        os.environ["CSIM_TPI_OPTS"] = mask
    src = plan.generate_source(perm, 3).encode()
    out = os.path.join(ROOT, "build", f"tpi_{name}_{mask}.cu")
    open(out, "wb").write(src)
    err, prog = nvrtc.nvrtcCreateProgram(src, b"spec.cu", 1, [shim], [b"stdint.h"])
    opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false", b"-lineinfo", b"--ptxas-options=-v",
            ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
    opts += [o.encode() for o in os.environ.get("TPI_NVRTC_EXTRA", "").split()]
    err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    _, ls = nvrtc.nvrtcGetProgramLogSize(prog); log = b" " * ls; nvrtc.nvrtcGetProgramLog(prog, log)
    log = log.decode(errors="replace")
    _, cs = nvrtc.nvrtcGetCUBINSize(prog)
    if not cs:
        print(mask, "compile failed", log[-800:]); continue
    cub = b" " * cs; nvrtc.nvrtcGetCUBIN(prog, cub); open(out + "bin", "wb").write(cub)
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", "csim_tran_tpi", out + "bin"], capture_output=True, text=True).stdout
    open(out[:-3] + ".sass", "w").write(sass)
    ins = parse(sass)
    spans = []
    for a, t, h in ins:
        mm = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", t) if "BRA" in t else None
        if mm and int(mm.group(1), 16) < a:
            spans.append((a - int(mm.group(1), 16), int(mm.group(1), 16), a))
    # Newton loop: the smallest backward-branch span that holds most of the kernel's FP64 arithmetic of ONE variant
    # (the time loop and the kernel as a whole are larger spans around it)
    is64 = lambda t: re.match(r"(@!?U?P\d+ )?D(ADD|MUL|FMA)", t) is not None
    tot64 = sum(1 for a, t, h in ins if is64(t))
    newton = min((sp for sp in spans if sum(1 for a, t, h in ins if sp[1] <= a <= sp[2] and is64(t)) >= 0.5 * tot64),
                 default=max(spans))
    sel = [x for x in ins if newton[1] <= x[0] <= newton[2]]
    st = sum((h >> 41) & 0xF for a, t, h in sel)
    f64 = sum(1 for a, t, h in sel if re.match(r"(@!?U?P\d+ )?D(ADD|MUL|FMA|SETP)", t))
    lm = sum(1 for a, t, h in sel if re.search(r"\b(LDL|STL)", t))
    regs = re.search(r"Used (\d+) registers", log[log.find("csim_tran_tpi"):])
    spill = re.search(r"(\d+) bytes spill stores", log[log.find("csim_tran_tpi"):])
    head = src.decode().splitlines()[1:3]
    print(f"opts {mask}: newton loop {len(sel)} instr, fp64 {f64}, local ld/st {lm}, static stall sum {st} cycles; "
          f"regs {regs.group(1) if regs else '?'}, spill stores {spill.group(1) if spill else '?'} B; {head[0][3:]} {head[1][3:]}")
