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


@pytest.mark.parametrize("name", ["dbmixer", "buffer"])
def test_warp_pair_source_hand_overs_are_consistent_and_compile(name, monkeypatch):
    """The warp-pair form (CSIM_TPI_OPTS bit 20): every matrix row is emitted exactly once, in one of the two streams;
    every hand-over barrier has exactly one arrive in one stream and one sync in the other, and a sync never precedes
    (in the order of the master's stream interleaved by barrier) a use of a row the other warp has not parked; barrier
    ids stay below 16; NVRTC accepts the source for sm_100a."""
    import re
    from circuitsimulator_b200 import capi
    nvrtc = pytest.importorskip("cuda.bindings.nvrtc")
    g, topo = load_golden(name)
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    monkeypatch.setenv("CSIM_TPI_OPTS", str(1 + 2 + 32 + (2 << 12) + (1 << 20)))  # defaults + two pairs per CTA + pair
    src = plan.generate_source(_perm(topo, g), 3)
    monkeypatch.delenv("CSIM_TPI_OPTS")
    N = topo.n_unknowns
    assert "#define PAIRS 2" in src and "if (tid & 32) {  // helper warp" in src
    lines = src.splitlines()
    h0 = next(i for i, l in enumerate(lines) if "if (tid & 32) {" in l)
    h1 = next(i for i in range(h0, len(lines)) if lines[i].strip() == "return;")
    helper, master = lines[h0:h1], lines[h1:]
    rows_h = [int(m.group(1)) for l in helper for m in [re.search(r"// position (\d+) = row", l)] if m]
    rows_m = [int(m.group(1)) for l in master for m in [re.search(r"// position (\d+) = row", l)] if m]
    assert sorted(rows_h + rows_m) == list(range(N)) and rows_h and rows_m
    assert rows_h == sorted(rows_h) and rows_m == sorted(rows_m)  # each warp takes its rows in ascending order

    def ops(block, loop_only):
        out = []
        for l in block:
            m = re.search(r"PAIR_(SYNC|ARRIVE)\((\d+)\)", l)
            if m and "#define" not in l:
                out.append((m.group(1), int(m.group(2))))
        return out
    oh, om = ops(helper, True), ops(master, True)
    ids = {i for _, i in oh + om}
    assert max(ids) <= 15 and {1, 2, 3} <= ids
    for i in sorted(ids - {1, 2, 3}):  # hand-overs: one arrive on one side, one sync on the other, once per iteration
        a = [(w, k) for w, o in (("h", oh), ("m", om)) for k, j in o if j == i]
        assert sorted(a) in ([("h", "ARRIVE"), ("m", "SYNC")], [("h", "SYNC"), ("m", "ARRIVE")]), (i, a)
    assert [k for k, j in oh if j == 2] == ["ARRIVE"] and [k for k, j in om if j == 2] == ["SYNC"]
    # barrier 3 (next iterate at the MOSFET terminals): master arrives at the start and once per iteration, helper syncs likewise
    assert [k for k, j in om if j == 3] == ["ARRIVE", "ARRIVE"] and [k for k, j in oh if j == 3] == ["SYNC", "SYNC"]
    # barrier 1: helper once per round; master once per iteration and once to dismiss the helpers
    assert [k for k, j in oh if j == 1] == ["SYNC"] and [k for k, j in om if j == 1] == ["SYNC", "SYNC"]
    assert _nvrtc_compile(nvrtc, src, b"tpi_pair.cu") > 10000
