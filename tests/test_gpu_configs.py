"""GPU: BASELINE.json's configurations at (or near) full size — parity on subsets against per-instance
oracle runs plus size-independent properties (a shard of a batch gives the bits the whole batch gives;
instance 0 equals the golden nominal run; no status bits)."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, n_steps_of

pytestmark = pytest.mark.gpu
SEED = 20261104


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def _topo_of(nl):
    return O.Topology(nl.n_unknowns, nl.n_node_eq, nl.elem_type, nl.elem_flags, nl.elem_nodes, nl.elem_branch,
                      nl.elem_param, nl.n_params, nl.params, nl.node_names, nl.elem_names)


def test_config3_buffer_monte_carlo_65536(torch):
    """tests/buffer.sp, 65,536 Monte-Carlo instances, full .TRAN."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    topo = _topo_of(nl)
    B, ns = 65536, n_steps_of(g)
    assert ns == nl.n_steps == 300
    P = nl.monte_carlo(B, SEED + 3)  # [P, B]
    plan = nl.plan()
    probes = np.arange(nl.n_unknowns, dtype=np.int32)
    setup = nl.tran_setup(probes=probes, out_stride=10)  # every 10th point of every unknown: 13 x 30 doubles/instance
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone())
    torch.cuda.synchronize()
    st = t["status"].cpu().numpy().astype(np.uint32)
    assert not (st & (capi.ST_LU_SINGULAR | capi.ST_NONFINITE)).any()  # (DC_UNCONVERGED is the reference's own behaviour here)
    wave = t["wave"].cpu().numpy()
    assert np.isfinite(wave).all()
    # instance 0 is the nominal circuit: equals the golden (reference-pinned) run
    assert_same_bits(wave[0].T, g["wave"][9::10], "instance 0 vs golden")
    # subset parity: first 128 + 128 random instances against the oracle, one instance at a time
    rng = np.random.default_rng(1)
    idx = np.concatenate([np.arange(128), rng.choice(np.arange(128, B), 128, replace=False)])
    m, grp, tab = O.make_source_tables(topo, nl.params, setup.dt, ns)
    hashes = t["hash"].cpu().numpy().view(np.uint64)
    totals = t["iters_total"].cpu().numpy()
    for b in idx:
        od = O.dc(topo, P[:, b])
        ot = O.tran(topo, P[:, b], setup.dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(wave[b].T, ot["wave"][9::10], f"instance {b}")
        assert hashes[b] == np.uint64(ot["hash"]) and totals[b] == ot["total"]
    # a shard run alone gives the bits the whole batch gave (no cross-instance coupling, any batch size)
    lo, n = 40000, 1000
    ps = params[:, lo:lo + n].contiguous()
    d2 = capi.dc(plan, ps)
    t2 = capi.tran(plan, ps, d2["x"].clone(), setup, pivot_hash=d2["hash"].clone(), status=d2["status"].clone())
    assert_same_bits(t2["wave"].cpu().numpy(), wave[lo:lo + n], "shard vs whole batch")
    assert np.array_equal(t2["hash"].cpu().numpy(), t["hash"][lo:lo + n].cpu().numpy())


def test_dbmixer_batch_beyond_one_wave_uses_wide_ctas(torch):
    """dbmixer.sp, 16,384 Monte-Carlo instances (config 4's batch size): more warps than one wave of one-warp CTAs
    holds (3 per SM), so the launch uses the 2-warp-CTA variant of the thread-per-instance kernel (4 per SM).
    A shard of the batch run alone (one-warp CTAs) gives the same bits; a subset equals the oracle."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("dbmixer")
    nl = Netlist(str(g["netlist"]))
    topo = _topo_of(nl)
    B, ns = 16384, 24
    P = nl.monte_carlo(B, SEED + 5)
    plan = nl.plan()
    setup = nl.tran_setup(probes=np.arange(nl.n_unknowns, dtype=np.int32), n_steps=ns)
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone())
    torch.cuda.synchronize()
    info = plan.spec_info()
    assert info["kind"] == 3, info["log"]
    if torch.cuda.get_device_properties(0).multi_processor_count * 3 * 32 < B:
        assert "wide variant" in info["log"], info["log"]
    wave = t["wave"].cpu().numpy()
    assert int(t["status"].max()) == 0 and np.isfinite(wave).all()
    m, grp, tab = O.make_source_tables(topo, nl.params, setup.dt, ns)
    hashes = t["hash"].cpu().numpy().view(np.uint64)
    for b in (0, 31, 32, 63, 8191, 8192, 14207, 14208, 16383):
        od = O.dc(topo, P[:, b])
        ot = O.tran(topo, P[:, b], setup.dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(wave[b].T, ot["wave"], f"instance {b}")
        assert hashes[b] == np.uint64(ot["hash"])
    lo, n = 9000, 700
    ps = params[:, lo:lo + n].contiguous()
    d2 = capi.dc(plan, ps)
    t2 = capi.tran(plan, ps, d2["x"].clone(), setup, pivot_hash=d2["hash"].clone(), status=d2["status"].clone())
    assert_same_bits(t2["wave"].cpu().numpy(), wave[lo:lo + n], "shard (one-warp CTAs) vs whole batch (wide CTAs)")
    assert np.array_equal(t2["hash"].cpu().numpy(), t["hash"][lo:lo + n].cpu().numpy())


def test_config4_amplitude_frequency_sweep_per_instance_sources(torch):
    """dbmixer.sp LO sweep (config 4's batch layout at 6 x 5 points): every instance has its own LO amplitude
    (a parameter) and frequency (its own host-tabulated sin row, selected through src_group_b).  Transient
    prefix and the steady-state driver against per-instance oracle runs."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("dbmixer")
    nl = Netlist(str(g["netlist"]))
    topo = _topo_of(nl)
    va = np.linspace(0.1, 1.0, 6)
    fr = np.linspace(0.8e9, 1.0e9, 5)
    ns = 240
    probes = np.arange(nl.n_unknowns, dtype=np.int32)
    P, setup = nl.sin_source_sweep("Vlo+", va, fr, ns, probes=probes)
    B = P.shape[1]
    assert B == 30 and setup.src_group_b.shape == (nl.n_sources, B)
    plan = nl.plan()
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone(),
                  want_iters_step=True)
    torch.cuda.synchronize()
    wave = t["wave"].cpu().numpy()
    for b in range(B):
        m, grp, tab = O.make_source_tables(topo, P[:, b], setup.dt, ns)  # the oracle tabulates this instance's own sine
        od = O.dc(topo, P[:, b])
        ot = O.tran(topo, P[:, b], setup.dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(wave[b].T, ot["wave"], f"sweep point {b}")
        assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), ot["iters_step"])
    assert not np.array_equal(wave[0], wave[1]) and not np.array_equal(wave[0], wave[5])  # both axes matter
    # steady-state driver on the same batch (short "period" so the test stays small): composition check
    spp, K = 60, 3
    r = capi.sstate(plan, params, d["x"].clone(), setup, spp, K, 1e-3)
    for b in (0, 7, 29):
        m, grp, tab = O.make_source_tables(topo, P[:, b], setup.dt, ns)
        o = O.sstate(topo, P[:, b], setup.dt, spp, K, 1e-3, m, grp, tab, O.dc(topo, P[:, b])["x"])
        assert int(r["periods"][b]) == o["period"]
        assert_same_bits(r["x"][:, b].cpu().numpy(), o["x"], f"sweep point {b} steady state")


def test_config5_ladder200_subset_parity(torch):
    """BASELINE config 5's netlist at full width (200 stages, N = 404, dense large-matrix path): DC operating
    point (490 Newton solves of a 404 x 404 system per instance) and the first 20 transient steps of four
    Monte-Carlo instances, bit for bit against the oracle (waveform of all 404 unknowns, per-step Newton counts,
    pivot hash, status); the whole 4,096-instance x 200-step run is tools/run_config5.py."""
    from concurrent.futures import ThreadPoolExecutor
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist, ladder_netlist
    nl = Netlist(ladder_netlist(200))
    topo = _topo_of(nl)
    assert nl.n_unknowns == 404
    B, ns = 4, 20
    P = nl.monte_carlo(B, SEED + 5)
    plan = nl.plan()
    probes = np.arange(nl.n_unknowns, dtype=np.int32)
    setup = nl.tran_setup(probes=probes, n_steps=ns)
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    x_dc = d["x"].clone()
    t = capi.tran(plan, params, d["x"], setup, pivot_hash=d["hash"].clone(), status=d["status"].clone(),
                  want_iters_step=True)
    torch.cuda.synchronize()
    wave = t["wave"].cpu().numpy()
    m, grp, tab = O.make_source_tables(topo, nl.params, setup.dt, ns)

    def one(b):
        od = O.dc(topo, P[:, b])
        return b, od, O.tran(topo, P[:, b], setup.dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])

    with ThreadPoolExecutor(4) as ex:
        runs = list(ex.map(one, range(B)))
    for b, od, ot in runs:
        assert_same_bits(x_dc[:, b].cpu().numpy(), od["x"], f"ladder-200 DC inst {b}")
        assert np.array_equal(d["iters"][:, b].cpu().numpy(), od["iters"])
        assert_same_bits(wave[b].T, ot["wave"], f"ladder-200 transient inst {b}")
        assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), ot["iters_step"])
        assert int(t["iters_total"][b]) == ot["total"]
        assert np.uint64(t["hash"][b].item() & 0xFFFFFFFFFFFFFFFF) == np.uint64(ot["hash"])
        assert int(t["status"][b]) == (od["status"] | ot["status"])
