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
