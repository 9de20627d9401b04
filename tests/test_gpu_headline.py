"""GPU: the HEADLINE workload itself, end to end — tests/dbmixer.sp, DC operating point + the full `.TRAN 1e-13 5e-9`
(50,000 backward-Euler steps, src/tanalisis.cpp:238,255-420) — against the reference-pinned golden of the nominal
circuit (tests/golden/dbmixer_nominal.npz: every 100th row of the CSV, the final state, the pivot hash and all
50,000 per-step Newton counts) and against full-length oracle runs of Monte-Carlo instances of bench.py's batch."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, make_plan, n_steps_of, perturb

pytestmark = pytest.mark.gpu
SEED = 20261104                           # bench.py's and config 2's Monte-Carlo batch
SUBSET = (0, 1, 2, 3, 17, 63, 511, 1023)  # instances with a full-length oracle run (row k of the draw does not depend on B)
NEWTON_TOTAL_NOMINAL = 492304             # Newton iterations of the nominal 50,000-step run


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def u64(t):
    return t.cpu().numpy().view(np.uint64)


@pytest.fixture(scope="module")
def full_runs():
    """Full-length oracle runs (DC + 50,000 steps, ~6 s each on one core; ctypes releases the GIL)."""
    g, topo = load_golden("dbmixer")
    dt, ns = float(g["dt"]), n_steps_of(g)
    assert ns == 50000
    P = perturb(topo.params, topo, 1024, SEED)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)

    def one(b):
        od = O.dc(topo, P[b])
        ot = O.tran(topo, P[b], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        return b, dict(x_dc=od["x"], wave100=ot["wave"][99::100].copy(), x=ot["x"], iters_step=ot["iters_step"],
                       total=ot["total"], hash=np.uint64(ot["hash"]), status=od["status"] | ot["status"])

    with ThreadPoolExecutor(8) as ex:
        runs = dict(ex.map(one, SUBSET))
    # the restatement reproduces the reference-pinned golden over the whole run
    r0 = runs[0]
    assert_same_bits(r0["wave100"], g["wave_dec100"], "oracle vs golden, every 100th row")
    assert_same_bits(r0["x"], g["x_final"], "oracle vs golden, final state")
    assert r0["hash"] == np.uint64(g["hash"]) and r0["total"] == NEWTON_TOTAL_NOMINAL
    assert np.array_equal(r0["iters_step"], g["iters_step"])
    return dict(g=g, topo=topo, P=P, tables=(m, grp, tab), dt=dt, ns=ns, runs=runs)


def _run(torch, fr, B, jit, want_iters_step=True):
    from circuitsimulator_b200 import capi
    topo = fr["topo"]
    plan = make_plan(topo)
    if jit is not None:
        plan.jit_mode(jit)
    m, grp, tab = fr["tables"]
    setup = capi.TranSetup(fr["dt"], fr["ns"], m, grp, tab, probes=np.arange(topo.n_unknowns), out_stride=100)
    params = torch.from_numpy(np.ascontiguousarray(fr["P"][:B].T)).cuda()
    d = capi.dc(plan, params)
    x_dc = d["x"].clone()
    t = capi.tran(plan, params, d["x"], setup, pivot_hash=d["hash"], status=d["status"], want_iters_step=want_iters_step)
    torch.cuda.synchronize()
    return plan, x_dc, t


def _check(fr, x_dc, t, instances):
    wave = t["wave"].cpu().numpy()  # [B, N, 500]
    g = fr["g"]
    # instance 0 = the nominal circuit: the reference's own run (golden fixture)
    assert_same_bits(wave[0].T, g["wave_dec100"], "nominal dbmixer, every 100th timepoint of all 50,000 vs golden")
    assert_same_bits(t["x"][:, 0].cpu().numpy(), g["x_final"], "nominal dbmixer final state vs golden")
    assert u64(t["hash"])[0] == np.uint64(g["hash"])
    assert int(t["iters_total"][0]) == NEWTON_TOTAL_NOMINAL
    if t["iters_step"] is not None:
        assert np.array_equal(t["iters_step"][:, 0].cpu().numpy(), g["iters_step"]), "all 50,000 Newton counts vs golden"
    for b in instances:
        r = fr["runs"][b]
        assert_same_bits(x_dc[:, b].cpu().numpy(), r["x_dc"], f"DC inst {b}")
        assert_same_bits(wave[b].T, r["wave100"], f"waveform inst {b}")
        assert_same_bits(t["x"][:, b].cpu().numpy(), r["x"], f"final state inst {b}")
        assert u64(t["hash"])[b] == r["hash"], f"pivot hash inst {b}"
        assert int(t["iters_total"][b]) == r["total"]
        assert int(t["status"][b]) == r["status"]
        if t["iters_step"] is not None:
            assert np.array_equal(t["iters_step"][:, b].cpu().numpy(), r["iters_step"]), f"Newton counts inst {b}"


@pytest.mark.parametrize("jit", [3, 2, 0])
def test_dbmixer_full_transient_all_kernel_families(torch, full_runs, jit):
    """64 instances x 50,000 steps on the thread-per-instance (3), warp-per-instance (2) and generic (0) kernels:
    instance 0 vs the golden, five Monte-Carlo instances vs full-length oracle runs — waveform, final state,
    per-step Newton counts, pivot hash, status."""
    plan, x_dc, t = _run(torch, full_runs, 64, jit)
    info = plan.spec_info()
    assert info["kind"] == jit
    _check(full_runs, x_dc, t, [b for b in SUBSET if b < 64])


def test_config2_dbmixer_1024_instances_full_length(torch, full_runs):
    """BASELINE config 2: tests/dbmixer.sp DC + transient, 1,024-instance batch (default kernel choice), full
    length; an 8-instance subset against full-length oracle runs, every instance finite and converged."""
    from circuitsimulator_b200 import capi
    plan, x_dc, t = _run(torch, full_runs, 1024, None, want_iters_step=False)
    assert plan.spec_info()["kind"] == 3
    _check(full_runs, x_dc, t, SUBSET)
    st = t["status"].cpu().numpy().astype(np.uint32)
    assert not (st & (capi.ST_LU_SINGULAR | capi.ST_NONFINITE | capi.ST_TRAN_UNCONVERGED)).any()
    assert np.isfinite(t["wave"].cpu().numpy()).all()
    # a shard of the batch alone gives the bits the whole batch gave
    lo, hi = 500, 532
    from circuitsimulator_b200 import capi as C2
    topo = full_runs["topo"]
    m, grp, tab = full_runs["tables"]
    setup = C2.TranSetup(full_runs["dt"], 3000, m, grp, tab, probes=np.arange(topo.n_unknowns), out_stride=100)
    ps = torch.from_numpy(np.ascontiguousarray(full_runs["P"][lo:hi].T)).cuda()
    ds = C2.dc(plan, ps)
    assert_same_bits(ds["x"].cpu().numpy(), x_dc[:, lo:hi].cpu().numpy(), "shard DC == batch DC")
    ts = C2.tran(plan, ps, ds["x"], setup)
    torch.cuda.synchronize()
    assert_same_bits(ts["wave"].cpu().numpy()[:, :, :30], t["wave"][lo:hi, :, :30].cpu().numpy(), "shard == batch")
