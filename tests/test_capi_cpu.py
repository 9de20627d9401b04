"""CPU: the C-ABI library loads, exports every symbol include/csim_b200.h declares, builds plans without a
device, and refuses to compute without CUDA (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from support import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from circuitsimulator_b200 import capi
    L = capi.lib()
    hdr = open(os.path.join(ROOT, "include", "csim_b200.h")).read()
    declared = set(re.findall(r"\b(csim_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/csim_b200.h but not exported"
    assert declared == set(capi.EXPORTS)
    assert L.csim_abi_version() == 1


def test_struct_layouts_match_header():
    from circuitsimulator_b200 import capi
    assert C.sizeof(capi.Constants) == 8 + 9 * 8 + 8 + 5 * 8
    c = capi.default_constants()
    assert (c.dc_ramp_steps, c.dc_max_iters, c.tran_max_iters) == (10, 50, 50)
    assert (c.dc_tol, c.dc_alpha, c.tran_tol, c.tran_gmin, c.tran_alpha) == (1e-9, 0.35, 1e-6, 1e-6, 0.45)
    assert (c.gmin_high_base, c.gmin_low_base, c.gmin_abs_max, c.lu_eps, c.mos_off_gds) == (1e-6, 3.35e-7, 1e-4, 1e-15, 1e-12)


@pytest.mark.parametrize("name,nnz_dc,nnz_tran", [("buffer", 34, 39), ("dbmixer", 91, 106)])
def test_plan_pattern_known_answers(name, nnz_dc, nnz_tran):
    """Structural nnz of the DC / TRAN assemblies from a device-less plan."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden(name)
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    for mode, want in ((0, nnz_dc), (1, nnz_tran)):
        rowptr, colidx = plan.pattern(mode)
        assert colidx.size == want and rowptr[-1] == want
        for i in range(topo.n_unknowns):
            seg = colidx[rowptr[i]:rowptr[i + 1]]
            assert np.all(np.diff(seg) > 0)
    assert (plan.n_unknowns, plan.n_params) == (topo.n_unknowns, topo.n_params)


def test_plan_rejects_bad_topology():
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    bad = topo.elem_branch.copy()
    bad[0] = -1  # VDD without a branch row: assignEquationIndices not called
    with pytest.raises(capi.CsimError, match="branch"):
        capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes, bad,
                  topo.elem_param, topo.n_params, device=-1)


def test_pivot_hash_definition_shared_with_oracle():
    from circuitsimulator_b200 import capi
    from oracle import oraclelib as O
    rng = np.random.default_rng(0)
    h1 = h2 = capi.PIVOT_HASH_INIT
    for _ in range(20):
        perm = rng.permutation(31).astype(np.int32)
        h1 = capi.pivot_hash_update(h1, perm)
        h2 = O.pivot_hash_update(h2, perm)
    assert h1 == h2 != capi.PIVOT_HASH_INIT


def test_no_cpu_fallback_without_device():
    """On a machine without CUDA every compute entry point must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    plan = capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=-1)
    with pytest.raises(capi.CsimError):
        capi.Batch(plan, 4)
    with pytest.raises(capi.CsimError):
        capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                  topo.elem_branch, topo.elem_param, topo.n_params, device=0)
