"""CPU: the kernel generator emits source for a plan without any device, and NVRTC (no GPU needed to compile)
accepts it for sm_100a."""
import os

import numpy as np
import pytest

from oracle import oraclelib as O
from support import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _perm(topo, g):
    dt = float(g["dt"])
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, 2)
    x = g["x_dc"]
    G, I = O.stamp(topo, 1, topo.params, x, x, O.src_tran(topo, topo.params, m, grp, tab, 0, dt), dt=dt, gmin=1e-6)
    return O.lu_solve(G, I)[1]


@pytest.mark.parametrize("kind", [2, 3])
@pytest.mark.parametrize("name", ["buffer", "dbmixer"])
def test_generated_source_compiles_for_sm100a(name, kind):
    from circuitsimulator_b200 import capi
    nvrtc = pytest.importorskip("cuda.bindings.nvrtc")
    g, topo = load_golden(name)
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    src = plan.generate_source(_perm(topo, g), kind)
    assert {2: "csim_tran_spec", 3: "csim_tran_tpi"}[kind] in src and f"#define SPEC_N {topo.n_unknowns}" in src
    shim = (b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; "
            b"typedef unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; "
            b"typedef long long int64_t; typedef unsigned long long uint64_t;\n")
    err, prog = nvrtc.nvrtcCreateProgram(src.encode(), b"spec.cu", 1, [shim], [b"stdint.h"])
    opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false",
            ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
    err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    _, n = nvrtc.nvrtcGetProgramLogSize(prog)
    log = b" " * n
    nvrtc.nvrtcGetProgramLog(prog, log)
    assert int(err) == 0, log.decode()
    _, cs = nvrtc.nvrtcGetCUBINSize(prog)
    assert cs > 10000


def test_multi_variant_source_compiles_for_sm100a():
    """Two predicted permutations: both Newton bodies in one kernel, selected per lane."""
    from circuitsimulator_b200 import capi
    nvrtc = pytest.importorskip("cuda.bindings.nvrtc")
    g, topo = load_golden("buffer")
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    p0 = np.asarray(_perm(topo, g), np.int32)
    p1 = p0.copy()
    p1[[2, 5]] = p1[[5, 2]]
    src = plan.generate_source_variants(np.stack([p0, p1]))
    assert "(2 variant(s))" in src and "var_0 == 1" in src
    assert _nvrtc_compile(nvrtc, src, b"tpi2.cu") > 10000


def _nvrtc_compile(nvrtc, src, name):
    shim = (b"#pragma once\ntypedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; "
            b"typedef unsigned short uint16_t;\ntypedef int int32_t; typedef unsigned int uint32_t; "
            b"typedef long long int64_t; typedef unsigned long long uint64_t;\n")
    err, prog = nvrtc.nvrtcCreateProgram(src.encode(), name, 1, [shim], [b"stdint.h"])
    opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"--fmad=false",
            ("-I" + os.path.join(ROOT, "circuitsimulator_b200", "csrc")).encode()]
    err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    _, n = nvrtc.nvrtcGetProgramLogSize(prog)
    log = b" " * n
    nvrtc.nvrtcGetProgramLog(prog, log)
    assert int(err) == 0, log.decode()
    _, cs = nvrtc.nvrtcGetCUBINSize(prog)
    return cs


@pytest.mark.parametrize("name,mode", [("buffer", 0), ("buffer", 1), ("dbmixer", 0), ("dbmixer", 1)])
def test_generated_stamp_source_compiles_and_counts_traffic(name, mode):
    from circuitsimulator_b200 import capi
    nvrtc = pytest.importorskip("cuda.bindings.nvrtc")
    g, topo = load_golden(name)
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    src = plan.generate_stamp_source(mode)
    assert "csim_stamp_spec" in src
    assert _nvrtc_compile(nvrtc, src, b"stamp.cu") > 5000
    info = plan.stamp_info(mode, True)
    nnz = plan.pattern(mode)[1].size
    assert info["bytes_written"] == 8 * (nnz + topo.n_unknowns)
    # every distinct input once: never more than SURVEY.md §8(d)'s P + N + N (+ sources)
    n_src = len(topo.source_elems())
    assert 0 < info["bytes_read"] <= 8 * (topo.n_params + 2 * topo.n_unknowns + n_src)
    assert src.count("LDP(") - 1 <= topo.n_params + 2 * n_src  # "-1": the macro definition


def test_generator_rejects_non_permutations():
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    with pytest.raises(capi.CsimError, match="permutation"):
        plan.generate_source(np.zeros(topo.n_unknowns, np.int32))
