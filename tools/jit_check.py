"""Container-side check of the generator: emit the specialised kernel for a netlist, compile it with NVRTC for
sm_100a (no GPU needed) and report registers / spills / instruction counts.  Uses the oracle only to obtain a
plausible pivot permutation (development aid)."""
import ctypes as C, os, sys, subprocess
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import support
from oracle import oraclelib as O
from circuitsimulator_b200 import capi
from cuda.bindings import nvrtc
name = sys.argv[1] if len(sys.argv) > 1 else "dbmixer"
g, topo = support.load_golden(name)
plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes, topo.elem_branch,
                 topo.elem_param, topo.n_params, device=-1)
dt = float(g["dt"])
m, grp, tab = O.make_source_tables(topo, topo.params, dt, 4)
x = g["x_dc"]
G, I = O.stamp(topo, 1, topo.params, x, x, O.src_tran(topo, topo.params, m, grp, tab, 0, dt), dt=dt, gmin=1e-6)
_, perm, _ = O.lu_solve(G, I)
kind = int(sys.argv[2]) if len(sys.argv) > 2 else 2
src = plan.generate_source(perm, kind).encode()
os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
out = os.path.join(ROOT, "build", f"spec_{name}_{kind}.cu"); open(out, "wb").write(src)
print("source:", out, len(src.splitlines()), "lines; perm", perm.tolist())
shim = b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n"
err, prog = nvrtc.nvrtcCreateProgram(src, b"spec.cu", 1, [shim], [b"stdint.h"])
opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false", b"-lineinfo", b"--ptxas-options=-v",
        ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
_, ls = nvrtc.nvrtcGetProgramLogSize(prog); log = b" " * ls; nvrtc.nvrtcGetProgramLog(prog, log)
print("nvrtc:", err, log.decode()[-1500:])
_, cs = nvrtc.nvrtcGetCUBINSize(prog)
if cs:
    cub = b" " * cs; nvrtc.nvrtcGetCUBIN(prog, cub); open(out + "bin", "wb").write(cub)
    sass = subprocess.run(["cuobjdump", "-sass", out + "bin"], capture_output=True, text=True).stdout
    open(out[:-3] + ".sass", "w").write(sass)
    import re, collections
    ops = collections.Counter()
    for line in sass.splitlines():
        mm = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if mm: ops[mm.group(2).split(".")[0]] += 1
    print("static SASS instructions:", sum(ops.values()), dict(ops.most_common(14)))
