"""GPU: periodic steady state (csim_dev_sstate) against the CPU restatement of the same definition.
The reference's src/sstateanalysis.cpp is empty — parity with the reference is unpinned by construction;
what is checked is that the device driver composes the (pinned) transient step exactly as defined in
include/csim_b200.h: same period of convergence per instance, same frozen state bit for bit, same
Newton totals and pivot hashes."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, make_plan, perturb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("jit", [1, 0])
def test_buffer_steady_state_matches_definition(torch, jit):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    dt, spp, K, tol = float(g["dt"]), 100, 5, 2e-8  # Vin is a 10 MHz sine: T = 100 steps of 1 ns
    B = 9
    P = perturb(topo.params, topo, B, 77)
    plan = make_plan(topo)
    plan.jit_mode(jit)
    params = dev(torch, P.T)
    d = capi.dc(plan, params)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, spp * K)
    setup = capi.TranSetup(dt, spp * K, m, grp, tab, probes=np.zeros(0, np.int32))
    r = capi.sstate(plan, params, d["x"].clone(), setup, spp, K, tol, pivot_hash=d["hash"].clone(),
                    status=d["status"].clone())
    torch.cuda.synchronize()
    periods = r["periods"].cpu().numpy()
    assert set(periods.tolist()) == {2, 3}, periods  # the tolerance splits the batch: frozen instances must stay frozen
    for b in range(B):
        od = O.dc(topo, P[b])
        o = O.sstate(topo, P[b], dt, spp, K, tol, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert periods[b] == o["period"]
        assert_same_bits(r["x"][:, b].cpu().numpy(), o["x"], f"inst {b} state")
        assert_same_bits(r["delta"][b].cpu().numpy(), np.float64(o["delta"]), f"inst {b} delta")
        assert int(r["iters_total"][b]) == o["total"]
        assert np.uint64(r["hash"][b].item() & 0xFFFFFFFFFFFFFFFF) == np.uint64(o["hash"])
        assert int(r["status"][b]) == (od["status"] | o["status"])


def test_steady_state_gives_up_after_max_periods(torch):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    dt, spp, K = float(g["dt"]), 100, 2
    P = perturb(topo.params, topo, 3, 5)
    plan = make_plan(topo)
    params = dev(torch, P.T)
    d = capi.dc(plan, params)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, spp * K)
    setup = capi.TranSetup(dt, spp * K, m, grp, tab, probes=np.zeros(0, np.int32))
    r = capi.sstate(plan, params, d["x"].clone(), setup, spp, K, 1e-30)
    assert (r["periods"].cpu().numpy() == -1).all()
    o = O.sstate(topo, P[1], dt, spp, K, 1e-30, m, grp, tab, O.dc(topo, P[1])["x"])
    assert o["period"] == -1
    assert_same_bits(r["x"][:, 1].cpu().numpy(), o["x"], "state after max_periods")
    with pytest.raises(capi.CsimError):
        capi.sstate(plan, params, d["x"].clone(), setup, 0, K, 1e-6)


def test_host_layer_driver_equals_device_api(torch):
    """runSteadyStateAnalysisBatch (C++ host layer, host buffers, .hb/.TRAN cards) == csim_dev_sstate."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    B, K, tol = 5, 4, 2e-8
    P = nl.monte_carlo(B, 77)
    h = nl.steady_state(P, f0=10e6, max_periods=K, tol=tol)  # buffer.sp's own .hb card says 1e-2 Hz
    assert h["steps_per_period"] == 100
    plan = nl.plan()
    params = dev(torch, P)
    d = capi.dc(plan, params)
    m, grp, tab = nl.source_tables(n_steps=100 * K)
    setup = capi.TranSetup(nl.tran["tstep"], 100 * K, m, grp, tab, probes=np.zeros(0, np.int32))
    r = capi.sstate(plan, params, d["x"].clone(), setup, 100, K, tol)
    assert np.array_equal(h["periods"], r["periods"].cpu().numpy()) and (h["periods"] > 0).all()
    assert_same_bits(h["x"], r["x"].cpu().numpy(), "host-layer steady state")
    assert_same_bits(h["x_dc"], d["x"].cpu().numpy(), "host-layer DC")
    assert np.array_equal(h["iters_total"], r["iters_total"].cpu().numpy())


def test_shooting_finds_the_periodic_state(torch):
    """Newton shooting with the perturbation columns as batch instances (runSteadyStateShootingBatch): the returned
    state is periodic (one more period of the pinned transient returns to it) and agrees with period-by-period
    integration.  Floating-point algorithm without a reference counterpart: tolerances stated here."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    C_, spp = 4, 100  # Vin is a 10 MHz sine: T = 100 steps of 1 ns
    P = nl.monte_carlo(C_, 21)
    r = nl.shooting(P, f0=10e6, max_iters=8, tol=1e-9)
    assert r["steps_per_period"] == spp
    assert (r["iterations"] >= 0).all() and (r["iterations"] <= 8).all(), r["iterations"]
    assert (r["residual"] < 1e-9).all()
    assert r["transient_instances"] == (int(r["iterations"].max()) + 1) * C_ * (nl.n_unknowns + 1)
    # independent check of periodicity with the device API
    plan = nl.plan()
    params = dev(torch, P)
    m, grp, tab = nl.source_tables(nl.tran["tstep"], spp)
    setup = capi.TranSetup(nl.tran["tstep"], spp, m, grp, tab, probes=np.zeros(0, np.int32))
    x = dev(torch, r["x"])
    t = capi.tran(plan, params, x.clone(), setup)
    assert np.linalg.norm(t["x"].cpu().numpy() - r["x"], axis=0).max() < 1e-9
    # brute force: many periods of plain integration end at the same orbit
    K = 12
    m, grp, tab = nl.source_tables(nl.tran["tstep"], spp * K)
    setup = capi.TranSetup(nl.tran["tstep"], spp * K, m, grp, tab, probes=np.zeros(0, np.int32))
    d = capi.dc(plan, params)
    s = capi.sstate(plan, params, d["x"].clone(), setup, spp, K, 1e-12)
    assert np.abs(s["x"].cpu().numpy() - r["x"]).max() < 1e-6
