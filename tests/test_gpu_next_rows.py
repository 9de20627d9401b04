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
        capi.harmonics(t["wave"], spp, 17)
