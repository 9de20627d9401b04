"""Container-side check of the generated thread-per-instance DC kernel: emit it for a netlist and the pivot order of
the first DC solve, compile with NVRTC for sm_100a, report registers / spills.  usage: tools/dc_static.py netlist"""
import os, re, subprocess, sys
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
od = O.dc(topo, topo.params, trace_cap=2000)
perm = od["trace"].perm[-1]
src = plan.generate_source(perm, 4).encode()
out = os.path.join(ROOT, "build", f"dc_tpi_{name}.cu")
os.makedirs(os.path.dirname(out), exist_ok=True)
open(out, "wb").write(src)
shim = b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n"
err, prog = nvrtc.nvrtcCreateProgram(src, b"spec.cu", 1, [shim], [b"stdint.h"])
opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false", b"-lineinfo", b"--ptxas-options=-v",
        ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
_, ls = nvrtc.nvrtcGetProgramLogSize(prog); log = b" " * ls; nvrtc.nvrtcGetProgramLog(prog, log)
log = log.decode(errors="replace")
print("nvrtc:", err, "\n".join(l for l in log.splitlines() if "error" in l or "Used" in l or "spill" in l)[-1500:])
print(len(src.splitlines()), "lines;", [l for l in src.decode().splitlines() if l.startswith("//")][:2])
