"""GPU: the run-time specialised transient kernel (NVRTC) against the generic kernel and the oracle.
The pivot permutation it is generated for is only a prediction; results must not depend on it."""
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


def u64(t):
    return t.cpu().numpy().view(np.uint64)


def run(capi, torch, plan, P, topo, dt, ns):
    params = dev(torch, P.T)
    d = capi.dc(plan, params)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.arange(topo.n_unknowns))
    t = capi.tran(plan, params, d["x"].clone(), setup, pivot_hash=d["hash"].clone(), status=d["status"].clone(),
                  want_iters_step=True)
    torch.cuda.synchronize()
    return dict(x=t["x"].cpu().numpy(), wave=t["wave"].cpu().numpy(), its=t["iters_step"].cpu().numpy(),
                hash=u64(t["hash"]), status=t["status"].cpu().numpy(), total=t["iters_total"].cpu().numpy())


def same(a, b, what):
    assert_same_bits(a["wave"], b["wave"], what + " waveforms")
    assert_same_bits(a["x"], b["x"], what + " final state")
    assert np.array_equal(a["its"], b["its"]) and np.array_equal(a["hash"], b["hash"]), what
    assert np.array_equal(a["status"], b["status"]) and np.array_equal(a["total"], b["total"]), what


def test_split_division_is_bit_identical_to_ieee_division(torch):
    """rcp_seq + fdiv (spec_runtime.cuh) == a / b for random, huge, tiny, zero, inf and NaN operands."""
    from circuitsimulator_b200 import capi
    rng = np.random.default_rng(0)
    n = 1 << 22
    a = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    b = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    # adversarial exponents / mantissas
    k = 1 << 18
    a[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1074, 1023, k))
    b[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1074, 1023, k))
    b[k:2 * k] = np.nextafter(np.ldexp(1.0, rng.integers(-500, 500, k)), 0)        # mantissa all ones
    a[2 * k:3 * k] = rng.integers(1, 1 << 20, k).astype(float)                        # exact small quotients
    b[2 * k:3 * k] = rng.integers(1, 1 << 20, k).astype(float)
    special = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1.0, -1.0, 5e-324, 1.7976931348623157e308, 2.2250738585072014e-308])
    aa, bb = np.meshgrid(special, special)
    a[-aa.size:], b[-bb.size:] = aa.ravel(), bb.ravel()
    qs, qr = capi.selftest_div(dev(torch, a), dev(torch, b))
    torch.cuda.synchronize()
    qs, qr = qs.cpu().numpy(), qr.cpu().numpy()
    nan = np.isnan(qr)
    assert np.array_equal(np.isnan(qs), nan)
    assert np.array_equal(qs[~nan].view(np.uint64), qr[~nan].view(np.uint64))
    with np.errstate(all="ignore"):
        host = a / b  # x86 IEEE division == what the reference computes
    assert np.array_equal(qr[~nan].view(np.uint64), host[~nan].view(np.uint64))


def test_guarded_division_of_the_thread_per_instance_kernel(torch):
    """tdiv (spec_runtime.cuh): wherever its own range guards accept a division the quotient is IEEE a / b bit for
    bit — signed zeros included (the sign comes from a*rcp(b)) — and ordinary magnitudes are always accepted, so the
    redo path is not taken for them.  Divisors span 2^-1021 .. 2^1021 (the kernel's pivot test; outside it the
    solve is redone generically)."""
    from circuitsimulator_b200 import capi
    rng = np.random.default_rng(1)
    n = 1 << 22
    a = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    b = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    k = 1 << 18
    a[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1074, 1023, k)) * rng.choice([-1.0, 1.0], k)
    b[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1021, 1021, k)) * rng.choice([-1.0, 1.0], k)
    b[k:2 * k] = np.nextafter(np.ldexp(1.0, rng.integers(-500, 500, k)), 0)        # mantissa all ones
    a[2 * k:3 * k] = rng.integers(1, 1 << 20, k).astype(float)                        # exact small quotients
    b[2 * k:3 * k] = rng.integers(1, 1 << 20, k).astype(float)
    a[3 * k:4 * k] = rng.choice([0.0, -0.0], k)                                       # exact zeros of either sign
    a[4 * k:5 * k] = rng.integers(1, 1 << 30, k) * 5e-324 * rng.choice([-1.0, 1.0], k)  # denormal numerators
    b[4 * k:5 * k] = rng.integers(1, 64, k).astype(float)                             # ... with halfway quotients
    b[b == 0.0] = 1.0
    q, qr, ok = capi.selftest_tdiv(dev(torch, a), dev(torch, b))
    torch.cuda.synchronize()
    q, qr, ok = q.cpu().numpy(), qr.cpu().numpy(), ok.cpu().numpy().astype(bool)
    with np.errstate(all="ignore"):
        host = a / b
    assert np.array_equal(qr.view(np.uint64), host.view(np.uint64))
    diff = ok & (q.view(np.uint64) != qr.view(np.uint64))
    assert not diff.any(), (int(diff.sum()), [(float.hex(x), float.hex(y), float.hex(u), float.hex(v))
                                              for x, y, u, v in zip(a[diff][:6], b[diff][:6], q[diff][:6], qr[diff][:6])])
    assert ok[3 * k:4 * k].all() and np.array_equal(np.signbit(q[3 * k:4 * k]), np.signbit(host[3 * k:4 * k]))
    assert not ok[4 * k:5 * k].any()                       # denormal numerators are never trusted
    plain = (np.abs(a) > 1e-200) & (np.abs(a) < 1e200) & (np.abs(host) > 1e-200) & (np.abs(host) < 1e200)
    assert ok[plain].all() and plain.sum() > n // 2


def test_multiplier_division_with_folded_guards(torch):
    """tdiv_mul / tdiv_nz (spec_runtime.cuh): one unsigned accumulator instead of exponent min/max bookkeeping.
    (1) wherever the range guard accepts, the quotient is IEEE a / b bit for bit; (2) ordinary magnitudes are
    accepted; exact-zero, denormal and non-finite numerators are not (the solve is redone); (3) the pivot guard fires
    for EVERY |a| > |b| (a row the reference's pivot search would have preferred, include/solver.hpp:46-56) and only
    for |a| >= |b|; pairs one ulp apart and exact ties included."""
    from circuitsimulator_b200 import capi
    rng = np.random.default_rng(2)
    n = 1 << 22
    a = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    b = rng.normal(size=n) * 10.0 ** rng.uniform(-30, 30, n)
    k = 1 << 18
    a[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1074, 1023, k)) * rng.choice([-1.0, 1.0], k)
    b[:k] = np.ldexp(rng.uniform(1, 2, k), rng.integers(-1021, 1021, k)) * rng.choice([-1.0, 1.0], k)
    b[k:2 * k] = np.nextafter(np.ldexp(1.0, rng.integers(-500, 500, k)), 0)        # mantissa all ones
    a[k:2 * k] = b[k:2 * k] * rng.choice([-1.0, 1.0], k)                              # exact ties
    a[2 * k:3 * k] = rng.choice([0.0, -0.0], k)                                       # exact zeros
    a[3 * k:4 * k] = rng.integers(1, 1 << 30, k) * 5e-324 * rng.choice([-1.0, 1.0], k)  # denormal numerators
    b[3 * k:4 * k] = rng.integers(1, 64, k).astype(float)
    # neighbours: |a| one ulp above / below |b|, every sign combination
    bb = np.ldexp(rng.uniform(1, 2, k), rng.integers(-300, 300, k))
    up = np.nextafter(bb, np.inf)
    dn = np.nextafter(bb, 0)
    b[4 * k:5 * k] = bb * rng.choice([-1.0, 1.0], k)
    a[4 * k:5 * k] = np.where(rng.uniform(size=k) < 0.5, up, dn) * rng.choice([-1.0, 1.0], k)
    a[5 * k:6 * k] = np.array([np.inf, -np.inf, np.nan])[rng.integers(0, 3, k)]
    b[b == 0.0] = 1.0
    q, qr, fl = capi.selftest_tdiv_mul(dev(torch, a), dev(torch, b))
    torch.cuda.synchronize()
    q, qr, fl = q.cpu().numpy(), qr.cpu().numpy(), fl.cpu().numpy()
    ok, piv = (fl & 1).astype(bool), (fl & 2).astype(bool)
    with np.errstate(all="ignore"):
        host = a / b
    nan = np.isnan(host)
    assert np.array_equal(qr[~nan].view(np.uint64), host[~nan].view(np.uint64))
    diff = ok & (q.view(np.uint64) != qr.view(np.uint64))
    assert not diff.any(), (int(diff.sum()), [(float.hex(x), float.hex(y)) for x, y in zip(a[diff][:6], b[diff][:6])])
    assert not ok[2 * k:4 * k].any() and not ok[5 * k:6 * k].any()
    plain = (np.abs(a) > 1e-200) & (np.abs(a) < 1e200) & (np.abs(host) > 1e-200) & (np.abs(host) < 1e200)
    assert ok[plain].all() and plain.sum() > n // 2
    bigger, smaller = ok & (np.abs(a) > np.abs(b)), ok & (np.abs(a) < np.abs(b))
    assert piv[bigger].all() and bigger[4 * k:5 * k].sum() > k // 4
    assert not piv[smaller].any() and smaller[4 * k:5 * k].sum() > k // 4
    assert piv[k:2 * k][ok[k:2 * k]].all()                 # exact ties go to the redo path (conservative)


KINDS = pytest.mark.parametrize("kind", [2, 3])  # 2: warp per instance (csim_tran_spec), 3: thread per instance (csim_tran_tpi)
ENTRY = {2: "csim_tran_spec", 3: "csim_tran_tpi"}


@KINDS
def test_dbmixer_specialised_equals_generic_and_oracle(torch, kind):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("dbmixer")
    B, dt, ns = 12, float(g["dt"]), 600
    P = perturb(topo.params, topo, B, 20261104)
    generic = make_plan(topo)
    generic.jit_mode(0)
    a = run(capi, torch, generic, P, topo, dt, ns)
    assert not generic.spec_info()["active"]
    spec = make_plan(topo)
    spec.jit_mode(kind)
    b = run(capi, torch, spec, P, topo, dt, ns)
    info = spec.spec_info()
    assert info["kind"] == kind, info["log"]
    same(a, b, "dbmixer specialised vs generic")
    assert info["solves"] == int(b["total"].sum()) and info["fallback_solves"] == 0, info
    assert list(info["perm"]) == [21, 1, 2, 3, 4, 5, 6, 7, 22, 9, 23, 11, 12, 13, 24, 25, 26, 27, 28, 29, 30, 0, 8, 10,
                                  14, 15, 16, 17, 18, 19, 20]  # the reference's TRAN permutation
    assert_same_bits(b["wave"][0].T[:400], g["wave_first"], "nominal vs golden")
    assert ENTRY[kind] in spec.spec_source()
    for k in (0, 3, 7):
        od = O.dc(topo, P[k])
        m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
        ot = O.tran(topo, P[k], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(b["wave"][k].T, ot["wave"], f"inst {k} vs oracle")
        assert b["hash"][k] == np.uint64(ot["hash"])


@KINDS
def test_buffer_specialised_with_pivot_misses(torch, kind):
    """buffer.sp changes its pivot order ~12 % of the time: those solves must take the generic path and the
    result must still be bit-identical (full 300 steps, every unknown)."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    B, dt, ns = 16, float(g["dt"]), n_steps_of(g)
    P = perturb(topo.params, topo, B, 20261104 + 3)
    generic = make_plan(topo)
    generic.jit_mode(0)
    a = run(capi, torch, generic, P, topo, dt, ns)
    spec = make_plan(topo)
    spec.jit_mode(kind)
    b = run(capi, torch, spec, P, topo, dt, ns)
    info = spec.spec_info()
    assert info["kind"] == kind, info["log"]
    same(a, b, "buffer specialised vs generic")
    assert 0 < info["fallback_solves"] < info["solves"], info
    assert_same_bits(b["wave"][0].T, g["wave"], "nominal buffer vs golden")
    assert b["hash"][0] == np.uint64(g["hash"])


@KINDS
def test_wrong_prediction_only_costs_time(torch, kind):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    B, dt, ns = 6, float(g["dt"]), 60
    P = perturb(topo.params, topo, B, 5)
    generic = make_plan(topo)
    generic.jit_mode(0)
    a = run(capi, torch, generic, P, topo, dt, ns)
    wrong = make_plan(topo)
    wrong.jit_mode(kind)
    wrong.specialize(np.arange(topo.n_unknowns))  # identity: never what partial pivoting picks here
    b = run(capi, torch, wrong, P, topo, dt, ns)
    info = wrong.spec_info()
    assert info["kind"] == kind and info["fallback_solves"] == info["solves"] > 0
    same(a, b, "identity prediction")


@KINDS
def test_degenerate_and_chunked_through_specialised_kernel(torch, kind):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    plan = make_plan(topo)
    plan.jit_mode(kind)
    dt, ns = float(g["dt"]), 40
    P = np.repeat(topo.params[None], 3, 0)
    e_l = [e for e in range(topo.n_elems) if topo.elem_type[e] == O.ELEM_L]
    P[1, topo.elem_param[e_l[0]]] = 0.0  # empty branch row -> singular -> zero vector
    b = run(capi, torch, plan, P, topo, dt, ns)
    assert plan.spec_info()["kind"] == kind
    for k in range(3):
        od = O.dc(topo, P[k])
        m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
        ot = O.tran(topo, P[k], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(b["wave"][k].T, ot["wave"], f"inst {k}")
        assert int(b["status"][k]) == (od["status"] | ot["status"]) and b["hash"][k] == np.uint64(ot["hash"])
    assert int(b["status"][1]) & capi.ST_LU_SINGULAR


def test_variants_switch_per_lane_after_a_miss(torch):
    """Two predicted permutations, the FIRST one wrong (identity) and the second the one partial pivoting takes:
    every lane misses its first solve, learns from the redo which variant holds and runs that one from then on.
    Results equal the generic kernel's bit for bit; only the first solve of each lane (plus the real pivot-order
    changes of buffer.sp) goes through the redo path."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    B, dt, ns = 40, float(g["dt"]), 120
    P = perturb(topo.params, topo, B, 31)
    generic = make_plan(topo)
    generic.jit_mode(0)
    a = run(capi, torch, generic, P, topo, dt, ns)
    one = make_plan(topo)
    one.jit_mode(3)
    b1 = run(capi, torch, one, P, topo, dt, ns)
    true_perm = one.spec_info()["perm"]
    same(a, b1, "one variant")
    two = make_plan(topo)
    two.jit_mode(3)
    two.specialize_variants(np.stack([np.arange(topo.n_unknowns, dtype=np.int32), true_perm]))
    b2 = run(capi, torch, two, P, topo, dt, ns)
    same(a, b2, "wrong variant first")
    i1, i2 = one.spec_info(), two.spec_info()
    assert i2["kind"] == 3 and "2 predicted permutation(s)" in i2["log"], i2["log"]
    assert i1["solves"] == i2["solves"]
    # B lanes miss once more each (their first solve), nothing else changes
    assert i1["fallback_solves"] < i2["fallback_solves"] <= i1["fallback_solves"] + 2 * B, (i1, i2)


def test_automatic_mode_adds_variants_then_picks_the_kernel_by_miss_rate(torch):
    """Default mode: thread-per-instance first.  buffer.sp changes its pivot order while its stages switch (about
    12 % of the solves): the following launches add the permutations the redo path met as further variants and the
    miss rate drops; only if four variants do not bring it under 2 % does the warp-per-instance kernel take over.
    dbmixer.sp stays on its single permutation.  Every launch is bit-identical to the generic kernel."""
    from circuitsimulator_b200 import capi
    for name, ns in (("buffer", 150), ("dbmixer", 120)):
        g, topo = load_golden(name)
        B, dt = 64, float(g["dt"])
        P = perturb(topo.params, topo, B, 11)
        generic = make_plan(topo)
        generic.jit_mode(0)
        a = run(capi, torch, generic, P, topo, dt, ns)
        auto = make_plan(topo)
        rates = []
        for launch in range(5):
            before = auto.spec_info() if launch else dict(fallback_solves=0, solves=0, log="")
            b = run(capi, torch, auto, P, topo, dt, ns)
            same(a, b, f"{name} auto launch {launch}")
            info = auto.spec_info()
            if info["kind"] != 3:
                break
            rebuilt = info["solves"] < before["solves"] + 1 or info["log"].count("added") != before["log"].count("added")
            f0, s0 = (0, 0) if rebuilt else (before["fallback_solves"], before["solves"])
            rates.append((info["fallback_solves"] - f0) / max(info["solves"] - s0, 1))
        info = auto.spec_info()
        if name == "dbmixer":
            assert info["kind"] == 3 and max(rates) < 0.02, (rates, info["log"])
        else:
            assert rates[0] > 0.02, rates
            assert "variant 1 added" in info["log"] or info["kind"] == 2, info["log"]
            assert info["kind"] == 2 or rates[-1] < rates[0] / 3, (rates, info["log"])


TIE_NETLIST = """* two voltage sources on one node: an exact tie (|+1| = |+1|) in the first pivot search
V1 a 0 DC 1
V2 a b DC 0.5
R1 b 0 1k
R2 a c 2k
C1 c 0 1e-12
M1 c b 0 n 10e-6 0.35e-6 1
.MODEL 1 VT 0.5 MU 3e-2 COX 6e-3 LAMBDA 0.05 CJ0 4e-14
.TRAN 1e-11 1e-9
"""


@pytest.mark.parametrize("jit,opts", [(3, None), (3, "0"), (2, None)])
def test_exact_tie_goes_to_the_row_the_reference_scans_first(torch, monkeypatch, jit, opts):
    """include/solver.hpp:46-56 keeps the FIRST maximum of the pivot search.  Rows V1 and V2 tie in column 0 and the
    reference takes V1's; a kernel specialised for the other choice ([4,3,...]) must notice (the tied row is scanned
    before its predicted pivot) and redo every solve generically instead of silently factorising in another order."""
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    if opts is not None:
        monkeypatch.setenv("CSIM_TPI_OPTS", opts)  # 0: no constant propagation -> the tie is tested numerically
    nl = Netlist(TIE_NETLIST)
    topo = O.Topology(nl.n_unknowns, nl.n_node_eq, nl.elem_type, nl.elem_flags, nl.elem_nodes, nl.elem_branch,
                      nl.elem_param, nl.n_params, nl.params, nl.node_names, nl.elem_names)
    B, dt, ns = 7, nl.tran["tstep"], nl.n_steps
    P = O.monte_carlo(topo, B, 31)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    want = []
    for b in range(B):
        od = O.dc(topo, P[b])
        want.append(O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"], trace_cap=4))
    assert want[0]["trace"].perm[0].tolist() == [3, 4, 2, 0, 1]  # the reference's order: V1's row first
    for perm, all_redone in (([3, 4, 2, 0, 1], False), ([4, 3, 2, 0, 1], True)):
        plan = nl.plan()
        plan.jit_mode(jit)
        plan.specialize(np.array(perm, np.int32))
        params = torch.from_numpy(np.ascontiguousarray(P.T)).cuda()
        d = capi.dc(plan, params)
        setup = nl.tran_setup(probes=np.arange(nl.n_unknowns, dtype=np.int32))
        t = capi.tran(plan, params, d["x"], setup, pivot_hash=d["hash"], status=d["status"], want_iters_step=True)
        torch.cuda.synchronize()
        info = plan.spec_info()
        assert info["kind"] == jit
        assert (info["fallback_solves"] == info["solves"]) if all_redone else (info["fallback_solves"] == 0), info
        wave = t["wave"].cpu().numpy()
        for b in range(B):
            assert_same_bits(wave[b].T, want[b]["wave"], f"perm {perm} inst {b}")
            assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), want[b]["iters_step"])
            assert np.uint64(t["hash"][b].item() & 0xFFFFFFFFFFFFFFFF) == np.uint64(want[b]["hash"])


def test_variants_are_added_between_the_chunks_of_a_resumed_transient(torch, monkeypatch):
    """csim_batch_tran cuts a run into chunks of timepoints (CSIM_WAVE_CHUNKS); every chunk is a launch, so in automatic
    mode the thread-per-instance kernel is re-evaluated — and rebuilt with one more predicted permutation — BETWEEN the
    chunks of one run, with every lane's state (x, variant, hash, totals) carried across.  buffer.sp changes its pivot
    order while its stages switch: the chunked run must equal the single-launch run and the oracle bit for bit."""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("buffer")
    B, dt, ns = 600, float(g["dt"]), n_steps_of(g)
    P = perturb(topo.params, topo, B, 11)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([8, 1], np.int32))
    outs, infos = [], []
    for chunks in ("1", "6"):
        monkeypatch.setenv("CSIM_WAVE_CHUNKS", chunks)
        plan = make_plan(topo)  # automatic mode, fresh plan: starts with one variant
        batch = capi.Batch(plan, B)
        batch.set_params(P.T)
        outs.append(batch.tran(setup, want_iters_step=True))
        infos.append(plan.spec_info())
    one, many = outs
    for k in ("x_dc", "x_final", "wave"):
        assert_same_bits(many[k], one[k], k)
    for k in ("iters_total", "iters_step", "hash", "status"):
        assert np.array_equal(many[k], one[k]), k
    assert "1 predicted permutation(s)" in infos[0]["log"]           # one launch: no chance to adapt
    assert "variant" in infos[1]["log"] and "added" in infos[1]["log"], infos[1]["log"]  # rebuilt between chunks
    for b in (0, 7, B - 1):
        od = O.dc(topo, P[b])
        ot = O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(many["wave"][b].T, ot["wave"][:, [8, 1]], f"instance {b} vs oracle")
        assert many["hash"][b] == np.uint64(ot["hash"]) and many["iters_total"][b] == ot["total"]
