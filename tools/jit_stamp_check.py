"""Container-side check of the assembly-kernel generator: emit, compile with NVRTC for sm_100a (no GPU
needed) and report registers / spills / SASS mix.  usage: tools/jit_stamp_check.py [netlist] [mode]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import support  # noqa: E402
from circuitsimulator_b200 import capi  # noqa: E402
from cuda.bindings import nvrtc  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "dbmixer"
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
g, topo = support.load_golden(name)
plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes, topo.elem_branch,
                 topo.elem_param, topo.n_params, device=-1)
src = plan.generate_stamp_source(mode).encode()
os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
out = os.path.join(ROOT, "build", f"stamp_{name}_{mode}")
open(out + ".cu", "wb").write(src)
print("source:", out + ".cu", len(src.splitlines()), "lines;", plan.stamp_info(mode, True))
shim = (b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef "
        b"unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; "
        b"typedef unsigned long long uint64_t;\n")
err, prog = nvrtc.nvrtcCreateProgram(src, b"stamp.cu", 1, [shim], [b"stdint.h"])
opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false", b"-lineinfo", b"--ptxas-options=-v",
        ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
_, ls = nvrtc.nvrtcGetProgramLogSize(prog)
log = b" " * ls
nvrtc.nvrtcGetProgramLog(prog, log)
print("nvrtc:", err, log.decode()[-1200:])
_, cs = nvrtc.nvrtcGetCUBINSize(prog)
if cs:
    cub = b" " * cs
    nvrtc.nvrtcGetCUBIN(prog, cub)
    open(out + ".cubin", "wb").write(cub)
    sass = subprocess.run(["cuobjdump", "-sass", out + ".cubin"], capture_output=True, text=True).stdout
    open(out + ".sass", "w").write(sass)
    ops = collections.Counter()
    for line in sass.splitlines():
        mm = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if mm:
            ops[mm.group(2).split(".")[0]] += 1
    print("static SASS instructions:", sum(ops.values()), dict(ops.most_common(16)))
