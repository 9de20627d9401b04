"""GPU: the large-matrix path (N > 32, matrix in HBM/L2) against the oracle — both kernel families (dense:
dense_kernels.cu, one CTA per instance, blocked LU; warp: big_kernels.cu, one warp per instance, zero
skipping) forced onto the two reference netlists, on generated inverter/RC ladders (BASELINE config 5's
topology) and on dense random systems.  Same bar as everywhere: bit-identical doubles, Newton counts and
pivot sequences."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, make_plan, n_steps_of, perturb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _topo_of(nl):
    return O.Topology(nl.n_unknowns, nl.n_node_eq, nl.elem_type, nl.elem_flags, nl.elem_nodes, nl.elem_branch,
                      nl.elem_param, nl.n_params, nl.params, nl.node_names, nl.elem_names)


def _run_and_compare(torch, capi, plan, topo, P, dt, ns, what, insts=None):
    B = P.shape[0]
    params = dev(torch, P.T)
    d = capi.dc(plan, params)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.arange(topo.n_unknowns))
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone(),
                  want_iters_step=True)
    torch.cuda.synchronize()
    wave = t["wave"].cpu().numpy()
    for b in (range(B) if insts is None else insts):
        od = O.dc(topo, P[b])
        assert_same_bits(d["x"][:, b].cpu().numpy(), od["x"], f"{what} DC inst {b}")
        assert np.array_equal(d["iters"][:, b].cpu().numpy(), od["iters"])
        ot = O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(wave[b].T, ot["wave"], f"{what} inst {b}")
        assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), ot["iters_step"])
        assert np.uint64(t["hash"][b].item() & 0xFFFFFFFFFFFFFFFF) == np.uint64(ot["hash"])
        assert int(t["status"][b]) == (od["status"] | ot["status"])
    return t


DENSE, WARP = 1, 2  # csim_plan_force_large_path


@pytest.mark.parametrize("kind", [DENSE, WARP])
@pytest.mark.parametrize("name,B,ns", [("buffer", 10, None), ("dbmixer", 6, 150)])
def test_large_path_forced_on_reference_netlists(torch, name, B, ns, kind):
    from circuitsimulator_b200 import capi
    g, topo = load_golden(name)
    ns = ns or n_steps_of(g)
    P = perturb(topo.params, topo, B, 99)
    plan = make_plan(topo)
    plan.force_large_path(kind)
    _run_and_compare(torch, capi, plan, topo, P, float(g["dt"]), ns, name + f" (large path, kind {kind})")
    assert not plan.spec_info()["active"]  # the on-chip kernels were never built


@pytest.mark.parametrize("kind", [DENSE, WARP])
@pytest.mark.parametrize("stages", [12, 20, 40])
def test_inverter_ladder(torch, stages, kind):
    """12 stages: N = 28, the on-chip path and the large path must agree with the oracle and with each other;
    20 / 40 stages: N = 44 / 84 > 32, only the large path applies."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist, ladder_netlist
    nl = Netlist(ladder_netlist(stages))
    topo = _topo_of(nl)
    assert nl.n_unknowns == 2 * stages + 4
    B, ns = 5, 60
    P = np.ascontiguousarray(nl.monte_carlo(B, 5).T)
    plan = nl.plan()
    if stages == 12:
        a = _run_and_compare(torch, capi, plan, topo, P, nl.tran["tstep"], ns, "ladder-12 on chip")
        big = nl.plan()
        big.force_large_path(kind)
        b = _run_and_compare(torch, capi, big, topo, P, nl.tran["tstep"], ns, "ladder-12 large path")
        assert_same_bits(a["wave"].cpu().numpy(), b["wave"].cpu().numpy(), "on-chip vs large path")
    else:
        plan.force_large_path(kind)
        _run_and_compare(torch, capi, plan, topo, P, nl.tran["tstep"], ns, f"ladder-{stages}", insts=(0, B - 1))


@pytest.mark.parametrize("kind", ["dense", "warp"])
@pytest.mark.parametrize("n", [33, 64, 100, 177])
def test_large_lu_solve_bit_exact(torch, n, kind, monkeypatch):
    from circuitsimulator_b200 import capi
    monkeypatch.setenv("CSIM_BIG_KIND", kind)
    rng = np.random.default_rng(n)
    B = 7
    A = rng.normal(size=(B, n, n))
    A[rng.random(size=A.shape) < 0.7] = 0.0  # sparse-ish: exercises zero skipping and fill-in logging
    for b in range(B):
        A[b] += np.diag(rng.normal(size=n))
    A[5] = rng.normal(size=(n, n))  # fully dense
    A[6, :, :] = np.round(A[6] * 2) / 2  # many exact ties in the pivot search: first maximum wins
    A[1, :, 3] = 0.0  # singular: zero vector (include/solver.hpp:94-97)
    A[2, 5, :] = A[2, 9, :]  # dependent rows: a pivot falls below eps (or rounds to something tiny)
    A[3, 0, 0] = 0.0  # forces a row exchange at the first step
    rhs = rng.normal(size=(B, n))
    x, perm, sing = capi.lu_solve(dev(torch, A.reshape(B, n * n).T), dev(torch, rhs.T))
    torch.cuda.synchronize()
    x, perm, sing = x.cpu().numpy(), perm.cpu().numpy(), sing.cpu().numpy()
    for b in range(B):
        ox, operm, osing = O.lu_solve(A[b], rhs[b])
        assert bool(sing[b]) == bool(osing), f"system {b} singular flag"
        assert_same_bits(x[:, b], ox, f"n={n} system {b}")
        if not osing:
            assert np.array_equal(perm[:, b], operm)
    assert sing[1] == 1
