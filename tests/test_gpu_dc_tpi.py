"""GPU: the generated thread-per-instance DC kernel (csim_dc_tpi; dcSolveNewtonLU + ConvController, src/dcanalysis.cpp:95-163,
264-307) against the oracle and the generic warp-per-instance dc_kernel: operating point bits, Newton iterations per
ramp step, pivot hash, status — on every launch while the kernel is still learning permutation variants, with a
deliberately wrong prediction, and on Monte-Carlo batches of all three circuits."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, make_plan

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def u64(t):
    return t.cpu().numpy().view(np.uint64)


def _topo(name):
    if name == "sources":
        from test_sources_cpu import load_sources
        g, topo, _, _ = load_sources()
        return g, topo
    return load_golden(name)


def _same(a, b, what):
    assert_same_bits(a["x"].cpu().numpy(), b["x"].cpu().numpy(), what + ": x")
    assert np.array_equal(a["iters"].cpu().numpy(), b["iters"].cpu().numpy()), what + ": iterations per ramp step"
    assert np.array_equal(u64(a["hash"]), u64(b["hash"])), what + ": pivot hash"
    assert np.array_equal(a["status"].cpu().numpy(), b["status"].cpu().numpy()), what + ": status"


@pytest.mark.parametrize("name", ["buffer", "dbmixer", "sources"])
def test_dc_tpi_equals_generic_kernel_and_oracle_while_learning(torch, name):
    from circuitsimulator_b200 import capi
    g, topo = _topo(name)
    B = 3000
    P = O.monte_carlo(topo, B, 20261104 + 3)
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    gen = make_plan(topo)
    gen.jit_mode(0)
    want = capi.dc(gen, params)
    assert not gen.dc_spec_info()["active"]
    plan = make_plan(topo)
    seen = []
    for launch in range(6):  # variants are added between launches; results never depend on them
        got = capi.dc(plan, params)
        torch.cuda.synchronize()
        info = plan.dc_spec_info()
        assert info["active"], info["log"]
        seen.append((info["n_variants"], info["fallback_solves"], info["solves"]))
        _same(got, want, f"{name} launch {launch}")
    assert seen[-1][2] > 0
    if name == "buffer":  # the ramp crosses operating regions: more than one pivot order
        assert seen[-1][0] > 1, seen
    assert_same_bits(want["x"][:, 0].cpu().numpy(), g["x_dc" if name != "sources" else "const_x_dc"], "nominal vs golden")
    for b in list(range(12)) + [B - 1]:
        od = O.dc(topo, P[b])
        assert_same_bits(got["x"][:, b].cpu().numpy(), od["x"], f"{name} instance {b} vs oracle")
        assert np.array_equal(got["iters"][:, b].cpu().numpy(), od["iters"])
        assert u64(got["hash"])[b] == np.uint64(od["hash"]) and int(got["status"][b]) == od["status"]


def test_dc_tpi_wrong_prediction_is_redone_generically(torch):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    N = topo.n_unknowns
    B = 70
    P = O.monte_carlo(topo, B, 5)
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    plan = make_plan(topo)
    plan.specialize_dc_variants(np.arange(N)[::-1].copy())  # nobody's matrix takes this order
    got = capi.dc(plan, params)
    torch.cuda.synchronize()
    info = plan.dc_spec_info()
    assert info["active"] and info["fallback_solves"] == info["solves"] > 0
    for b in range(0, B, 7):
        od = O.dc(topo, P[b])
        assert_same_bits(got["x"][:, b].cpu().numpy(), od["x"], f"instance {b}")
        assert np.array_equal(got["iters"][:, b].cpu().numpy(), od["iters"]) and u64(got["hash"])[b] == np.uint64(od["hash"])


def test_dc_tpi_degenerate_and_nonfinite_paths(torch):
    """R = 0 (stamp skipped), L <= 0 (singular -> zero vector) and a huge transconductance (non-finite solve -> gmin x 10):
    the controller's branches, lane by lane."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    P = np.repeat(topo.params[None], 6, 0)
    e_l = [e for e in range(topo.n_elems) if topo.elem_type[e] == O.ELEM_L]
    e_r = [e for e in range(topo.n_elems) if topo.elem_type[e] == O.ELEM_R]
    e_m = [e for e in range(topo.n_elems) if topo.elem_type[e] == O.ELEM_MOS]
    P[1, topo.elem_param[e_l[0]]] = 0.0
    P[2, topo.elem_param[e_r[1]]] = 0.0
    P[3, topo.elem_param[e_m[0]] + 1] = 1e300     # K: overflows the linearisation
    P[4, topo.elem_param[e_m[1]] + 1] = 1e-300
    P[5, topo.elem_param[e_m[2]]] = 1e3           # Vth: never on
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    plan = make_plan(topo)
    for launch in range(2):
        got = capi.dc(plan, params)
        torch.cuda.synchronize()
        assert plan.dc_spec_info()["active"]
        for b in range(6):
            od = O.dc(topo, P[b])
            assert_same_bits(got["x"][:, b].cpu().numpy(), od["x"], f"instance {b}")
            assert np.array_equal(got["iters"][:, b].cpu().numpy(), od["iters"]), b
            assert u64(got["hash"])[b] == np.uint64(od["hash"]) and int(got["status"][b]) == od["status"], b


def test_dc_tpi_candidate_retry_recovers_from_a_wrong_variant_without_the_generic_redo(torch):
    """Two predicted permutations, the first wrong at elimination step 0: every lane's first solve fails there and the
    candidate table (same pivot prefix, another pivot at the failing step) sends it to the other variant."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    N, B = topo.n_unknowns, 64
    P = O.monte_carlo(topo, B, 9)
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    first = O.dc(topo, topo.params, trace_cap=4)["trace"].perm[0]
    wrong = first.copy()
    wrong[0], wrong[-1] = first[-1], first[0]  # another pivot at step 0
    only_wrong, only_right, both = make_plan(topo), make_plan(topo), make_plan(topo)
    only_wrong.specialize_dc_variants(wrong)
    only_right.specialize_dc_variants(first)
    both.specialize_dc_variants(np.stack([wrong, first]))
    a, r, b = capi.dc(only_wrong, params), capi.dc(only_right, params), capi.dc(both, params)
    torch.cuda.synchronize()
    _same(a, b, "wrong only vs wrong + right")
    _same(r, b, "right only vs wrong + right")
    ia, ir, ib = only_wrong.dc_spec_info(), only_right.dc_spec_info(), both.dc_spec_info()
    assert ia["fallback_solves"] == ia["solves"] == ib["solves"] == ir["solves"]
    # starting every lane on the wrong variant costs at most ONE extra generic solve per lane (none when the wrong pivot is
    # merely smaller: the failure at step 0 then leads to the right variant through the candidate table; one when the wrong
    # pivot is a structural zero, which is a range failure, not an order failure); the first solve's order holds for part
    # of the ramp only, hence the redos both plans share
    assert ir["fallback_solves"] < ia["fallback_solves"]
    assert ib["fallback_solves"] <= ir["fallback_solves"] + B, (ia, ir, ib)
    od = O.dc(topo, P[5])
    assert_same_bits(b["x"][:, 5].cpu().numpy(), od["x"], "instance 5 vs oracle")


def test_plan_cache_hands_learned_variants_to_the_next_plan(torch, tmp_path, monkeypatch):
    """CSIM_PLAN_CACHE=<dir>: the pivot-order variants a plan learned are stored per topology hash; a NEW plan of the same
    topology starts with them (no learning launches, few redos on its first launch).  Results are unaffected."""
    from circuitsimulator_b200 import capi
    monkeypatch.setenv("CSIM_PLAN_CACHE", str(tmp_path))
    g, topo = load_golden("buffer")
    B, dt, ns = 3000, float(g["dt"]), 120
    P = O.monte_carlo(topo, B, 20261104 + 3)
    params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([8], np.int32))
    first = make_plan(topo)
    for _ in range(5):
        d1 = capi.dc(first, params)
        t1 = capi.tran(first, params, d1["x"].clone(), setup, pivot_hash=d1["hash"].clone(), status=d1["status"].clone())
    torch.cuda.synchronize()
    i_dc, i_tr = first.dc_spec_info(), first.spec_info()
    assert i_dc["n_variants"] > 1 and "predicted permutation(s)" in i_tr["log"]
    assert len(list(tmp_path.glob("csim_*_dc.perms"))) == 1 and len(list(tmp_path.glob("csim_*_tran.perms"))) == 1
    second = make_plan(topo)
    d2 = capi.dc(second, params)
    t2 = capi.tran(second, params, d2["x"].clone(), setup, pivot_hash=d2["hash"].clone(), status=d2["status"].clone())
    torch.cuda.synchronize()
    j_dc, j_tr = second.dc_spec_info(), second.spec_info()
    assert j_dc["n_variants"] == i_dc["n_variants"], (i_dc, j_dc)
    assert j_dc["fallback_solves"] * 10 < j_dc["solves"]                      # its FIRST launch already runs on the variants
    assert j_tr["log"].split(" predicted")[0].split()[-1] == i_tr["log"].split(" predicted")[0].split()[-1]
    _same(d1, d2, "cached variants")
    assert torch.equal(t1["hash"], t2["hash"]) and torch.equal(t1["x"].view(torch.int64), t2["x"].view(torch.int64))
    # a corrupt file is ignored
    for f in tmp_path.glob("*.perms"):
        f.write_text("13 2\n0 0 0\n")
    third = make_plan(topo)
    d3 = capi.dc(third, params)
    torch.cuda.synchronize()
    _same(d1, d3, "corrupt cache ignored")
