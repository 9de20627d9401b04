"""GPU: the "next" rows of SURVEY §8(f) built on the batched path — the reference's CSV contract for a batch
(byte-for-byte against the unmodified reference's own CSV of the same instance) and the `.DC` sweep card
(every point bit-identical to the oracle's DC operating point with that source value)."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def _topo_of(nl):
    return O.Topology(nl.n_unknowns, nl.n_node_eq, nl.elem_type, nl.elem_flags, nl.elem_nodes, nl.elem_branch,
                      nl.elem_param, nl.n_params, nl.params, nl.node_names, nl.elem_names)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref (the unmodified reference build) is not present")
def test_batched_csv_equals_reference_csv(torch, tmp_path):
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    topo = _topo_of(nl)
    B = 4
    P = nl.monte_carlo(B, 11)
    setup = nl.tran_setup(probes=nl.csv_columns())
    plan = nl.plan()
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone())
    torch.cuda.synchronize()
    wave, x_dc, st = t["wave"].cpu().numpy(), d["x"].cpu().numpy(), t["status"].cpu().numpy().astype(np.uint32)
    for b in (0, 3):
        ours, theirs = tmp_path / f"ours{b}.csv", tmp_path / f"ref{b}.csv"
        nl.write_instance_csv(str(ours), b, setup, wave, x_dc=x_dc, status=st)
        ref = O.Ref(O.netlist_from_params(topo, P[:, b], tran=nl.tran))
        ref.tran(str(theirs))
        ref.close()
        assert ours.read_bytes() == theirs.read_bytes(), f"instance {b}: CSV differs from the reference's"


def test_dc_sweep_points_equal_oracle(torch):
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]) + "\n.DC Vin 0 3 0.25\n")
    topo = _topo_of(nl)
    r = nl.dc_sweep(0)
    K = r["values"].size
    assert K == 13 and r["values"][0] == 0.0 and r["values"][-1] == 3.0
    src = [e for e in range(nl.n_elems) if nl.elem_names[e] == "Vin"][0]
    for k in range(K):
        p = nl.params.copy()
        p[nl.elem_param[src]] = r["values"][k]
        od = O.dc(topo, p)
        assert_same_bits(r["x"][:, k], od["x"], f"sweep point {k}")
        assert np.array_equal(r["iters"][:, k], od["iters"]) and int(r["status"][k]) == od["status"]
    assert not np.array_equal(r["x"][:, 0], r["x"][:, K - 1])


def test_harmonics_match_definition_and_dft(torch):
    """csim_dev_harmonics against the CPU restatement of its definition (bit for bit) and against numpy's FFT
    (to rounding), on the last period of buffer.sp's probes (Vin is a 10 MHz sine: T = 100 steps of 1 ns)."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    B, spp, H = 5, 100, 4
    P = nl.monte_carlo(B, 3)
    setup = nl.tran_setup(probes=nl.csv_columns()[:6])
    plan = nl.plan()
    params = torch.from_numpy(P).cuda()
    d = capi.dc(plan, params)
    t = capi.tran(plan, params, d["x"].clone(), setup)
    tables = capi.twiddle_tables(spp)
    assert np.allclose(tables[0], np.cos(2 * np.pi * np.arange(spp) / spp), atol=1e-15)
    re, im = capi.harmonics(t["wave"], spp, H, tables)
    torch.cuda.synchronize()
    wave = t["wave"].cpu().numpy()
    re, im = re.cpu().numpy(), im.cpu().numpy()
    assert re.shape == (H + 1, B, 6)
    for b in (0, 4):
        for p in range(6):
            x = wave[b, p, -spp:]
            ore, oim = O.harmonics(x, H, *tables)
            assert_same_bits(re[:, b, p], ore, f"re inst {b} probe {p}")
            assert_same_bits(im[:, b, p], oim, f"im inst {b} probe {p}")
            F = np.fft.fft(x)[: H + 1] / spp
            assert np.allclose(re[:, b, p], F.real, atol=1e-12) and np.allclose(im[:, b, p], F.imag, atol=1e-12)
    # ragged length (not a multiple of 32, shorter than one unrolled pass) and the n_harm limit
    re2, im2 = capi.harmonics(t["wave"], 37, 2)
    ore, oim = O.harmonics(wave[1, 2, -37:], 2, *capi.twiddle_tables(37))
    assert_same_bits(re2[:, 1, 2].cpu().numpy(), ore, "ragged re")
    assert_same_bits(im2[:, 1, 2].cpu().numpy(), oim, "ragged im")
    with pytest.raises(capi.CsimError):
        capi.harmonics(t["wave"], spp, 65)
    # the reference's own card: `.hb 100e6 50` (tests/dbmixer.sp:45) -> 50 harmonics, 13 passes of four
    re50, im50 = capi.harmonics(t["wave"], spp, 50, tables)
    torch.cuda.synchronize()
    re50, im50 = re50.cpu().numpy(), im50.cpu().numpy()
    assert re50.shape == (51, B, 6)
    for b, p in ((0, 0), (B - 1, 1)):
        ore, oim = O.harmonics(wave[b, p, -spp:], 50, *tables)
        assert_same_bits(re50[:, b, p], ore, "50 harmonics, real parts")
        assert_same_bits(im50[:, b, p], oim, "50 harmonics, imaginary parts")
    assert_same_bits(re50[: H + 1], re, "the first harmonics do not depend on how many are asked for")


def test_gauss_seidel_batch_bit_exact(torch):
    """csim_dev_gauss_seidel against the oracle's restatement (itself pinned to the unmodified reference in
    tests/test_gauss_seidel_cpu.py): iterates and sweep counts, converging, non-converging and regularised."""
    from circuitsimulator_b200 import capi
    rng = np.random.default_rng(9)
    for n in (1, 13, 31, 40, 77):
        B = 70  # more than one CTA of 64 threads
        A = rng.normal(size=(B, n, n))
        for b in range(B):
            A[b] += np.diag(np.abs(A[b]).sum(1) + 1.0)
        rhs = rng.normal(size=(B, n))
        x0 = rng.normal(size=(B, n))
        if n >= 5:
            A[2, 1, 1], A[2, 3, 3] = 0.0, -1e-13  # regularised diagonals: diverges, every sweep still defined
            A[3] = rng.normal(size=(n, n))  # not dominant: runs to max_iters
        dA = torch.from_numpy(np.ascontiguousarray(A.reshape(B, n * n).T)).cuda()
        db = torch.from_numpy(np.ascontiguousarray(rhs.T)).cuda()
        dx0 = torch.from_numpy(np.ascontiguousarray(x0.T)).cuda()
        for kwargs, start in ((dict(max_iters=200, tol=1e-10), None), (dict(max_iters=4, tol=1e-30), dx0)):
            x, sweeps = capi.gauss_seidel(dA, db, start, **kwargs)
            torch.cuda.synchronize()
            x, sweeps = x.cpu().numpy(), sweeps.cpu().numpy()
            for b in (0, 1, 2, 3, 63, 64, B - 1):
                with np.errstate(all="ignore"):
                    ox, osw = O.gauss_seidel(A[b], rhs[b], None if start is None else x0[b], **kwargs)
                assert sweeps[b] == osw, f"n={n} system {b}: sweeps {sweeps[b]} vs {osw}"
                assert_same_bits(x[:, b], ox, f"n={n} system {b}")
    with pytest.raises(capi.CsimError):
        capi.gauss_seidel(torch.zeros((201 * 201, 1), dtype=torch.float64, device="cuda"),
                          torch.zeros((201, 1), dtype=torch.float64, device="cuda"))


def test_gauss_seidel_host_buffers(torch):
    """csim_host_gauss_seidel (what the host layer's Solver::solveLinearSystemGaussSeidel calls) == the oracle."""
    import ctypes as C
    from circuitsimulator_b200 import capi
    rng = np.random.default_rng(2)
    n = 13
    A = rng.normal(size=(n, n)) + np.diag(np.full(n, 20.0))
    b = rng.normal(size=n)
    x = np.zeros(n)
    sweeps = np.zeros(1, np.int32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    L = capi.lib()
    L.csim_host_gauss_seidel.argtypes = [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                                         C.c_void_p]
    assert L.csim_host_gauss_seidel(n, 1, vp(np.ascontiguousarray(A)), vp(b), vp(x), 1000, 1e-10, vp(sweeps)) == 0
    ox, osw = O.gauss_seidel(A, b)
    assert sweeps[0] == osw
    assert_same_bits(x, ox, "host-buffer Gauss-Seidel")


@pytest.mark.parametrize("jit", [3, 2, 0])
@pytest.mark.parametrize("case", ["const", "pulse_periodic_pwl", "pulse_single"])
def test_current_source_and_pulse_pwl_waveforms_in_transient(torch, case, jit):
    """Rows a3 / a9: an I source in the transient (src/element.cpp:34-66) and PULSE / PWL sources through
    CSIM_SRC_VAL_TABLE (include/sim.hpp:80-138) — against tests/golden/sources.npz, generated from the unmodified
    reference's own SourceSpec code (make_golden.py: pin_sources), for every kernel family and a Monte-Carlo batch
    against the oracle."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    from test_sources_cpu import CASES, load_sources
    g, topo, dt, ns = load_sources()
    nl = Netlist(str(g["netlist"]))
    for idx, (kind, spec) in CASES[case].items():
        nl.set_source_waveform(idx, "pwl" if kind == O.WAVE_PWL else "pulse", spec)
    plan = nl.plan()
    plan.jit_mode(jit)
    setup = nl.tran_setup(probes=np.arange(nl.n_unknowns, dtype=np.int32))
    assert setup.n_steps == ns
    B = 9
    P = O.monte_carlo(topo, B, 4242)
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    d = capi.dc(plan, params)
    x_dc = d["x"].clone()
    t = capi.tran(plan, params, d["x"], setup, pivot_hash=d["hash"], status=d["status"], want_iters_step=True)
    torch.cuda.synchronize()
    assert plan.spec_info()["kind"] == jit
    wave = t["wave"].cpu().numpy()
    assert_same_bits(x_dc[:, 0].cpu().numpy(), g[case + "_x_dc"], "DC vs reference-pinned fixture")
    assert_same_bits(wave[0].T, g[case + "_wave"], "waveform vs reference-pinned fixture")
    assert np.array_equal(t["iters_step"][:, 0].cpu().numpy(), g[case + "_iters_step"])
    assert np.uint64(t["hash"][0].item() & 0xFFFFFFFFFFFFFFFF) == g[case + "_hash"]
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns, waveforms=CASES[case])
    for b in range(1, B):
        od = O.dc(topo, P[b])
        ot = O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(wave[b].T, ot["wave"], f"instance {b}")
        assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), ot["iters_step"])
        assert np.uint64(t["hash"][b].item() & 0xFFFFFFFFFFFFFFFF) == np.uint64(ot["hash"])


def test_monte_carlo_card_runs_the_batch_the_recipe_defines(torch):
    """Row (f)3: `.MC instances seed rel` through the C++ host layer (csim::runMonteCarloTransientBatch): parameters from
    the C++ restatement of numpy's PCG64, DC + full .TRAN on the device; every instance against the oracle run with
    the Python recipe's parameters."""
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("buffer")
    nl = Netlist(str(g["netlist"]) + "\n.MC 7 20261107 0.05\n")
    r = nl.run_mc_card(probes=np.arange(nl.n_unknowns, dtype=np.int32))
    P = O.monte_carlo(topo, 7, 20261107)
    dt, ns = float(g["dt"]), nl.n_steps
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    for b in range(7):
        od = O.dc(topo, P[b])
        ot = O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(r["wave"][b].T, ot["wave"], f".MC instance {b}")
        assert r["hash"][b] == np.uint64(ot["hash"]) and r["iters_total"][b] == ot["total"]
    assert_same_bits(r["wave"][0].T, g["wave"], "instance 0 is the nominal circuit (golden)")
