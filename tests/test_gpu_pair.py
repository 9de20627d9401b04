"""GPU: the warp-PAIR variant of the thread-per-instance transient kernel (two warps per 32 instances: the master runs
the Newton loop, the helper computes part of the matrix rows and linearises the MOSFETs for the next iteration; the
library picks it for batches that fit one wave of its CTAs).  Which warp computes a row changes no operand and no
operation order, so everything must be bit-identical to the one-warp kernel, the generic kernel and the oracle."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, make_plan, perturb
from test_gpu_specialized import run, same

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def fresh(monkeypatch, topo, pair):
    """a plan whose first transient launch may (pair) or may not pick the warp-pair variant"""
    monkeypatch.delenv("CSIM_TPI_OPTS", raising=False)
    if pair:
        monkeypatch.delenv("CSIM_TPI_PAIR", raising=False)
    else:
        monkeypatch.setenv("CSIM_TPI_PAIR", "0")
    plan = make_plan(topo)
    plan.jit_mode(3)
    return plan


def test_pair_equals_one_warp_kernel_generic_kernel_and_oracle(torch, monkeypatch):
    from circuitsimulator_b200 import capi
    g, topo = load_golden("dbmixer")
    B, dt, ns = 80, float(g["dt"]), 1500  # 80 instances: one full CTA of two pairs and a CTA with one half-empty pair
    P = perturb(topo.params, topo, B, 20261105)
    pair = fresh(monkeypatch, topo, True)
    a = run(capi, torch, pair, P, topo, dt, ns)
    info = pair.spec_info()
    assert info["kind"] == 3 and "warp-pair variant" in info["log"], info["log"]
    assert info["fallback_solves"] == 0 and info["solves"] == int(a["total"].sum())
    single = fresh(monkeypatch, topo, False)
    b = run(capi, torch, single, P, topo, dt, ns)
    assert single.spec_info()["kind"] == 3 and "warp-pair variant" not in single.spec_info()["log"]
    same(a, b, "pair vs one warp")
    generic = make_plan(topo)
    generic.jit_mode(0)
    c = run(capi, torch, generic, P[:8], topo, dt, ns)
    assert_same_bits(a["wave"][:8], c["wave"], "pair vs generic waveforms")
    assert_same_bits(a["x"][:, :8], c["x"], "pair vs generic final state")
    assert np.array_equal(a["its"][:, :8], c["its"]) and np.array_equal(a["hash"][:8], c["hash"])
    assert np.array_equal(a["status"][:8], c["status"]) and np.array_equal(a["total"][:8], c["total"])
    assert_same_bits(a["wave"][0].T[:400], g["wave_first"], "nominal vs golden")
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    for k in (0, 33, 79):
        od = O.dc(topo, P[k])
        ot = O.tran(topo, P[k], dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(a["wave"][k].T, ot["wave"], f"inst {k} vs oracle")
        assert a["hash"][k] == np.uint64(ot["hash"]) and int(a["total"][k]) == int(ot["total"])


def test_pair_with_every_solve_redone_generically(torch, monkeypatch):
    """identity prediction: every lane of every master takes the generic redo while its helper waits at the barrier"""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("dbmixer")
    B, dt, ns = 40, float(g["dt"]), 120
    P = perturb(topo.params, topo, B, 7)
    pair = fresh(monkeypatch, topo, True)
    pair.specialize(np.arange(topo.n_unknowns))  # never what partial pivoting picks
    a = run(capi, torch, pair, P, topo, dt, ns)
    info = pair.spec_info()
    assert "warp-pair variant" in info["log"] and info["fallback_solves"] == info["solves"] > 0, info["log"]
    generic = make_plan(topo)
    generic.jit_mode(0)
    c = run(capi, torch, generic, P, topo, dt, ns)
    same(a, c, "pair, identity prediction, vs generic")


def test_pair_resumed_in_chunks_through_the_batch_interface(torch, monkeypatch):
    """csim_batch_tran in chunks of timepoints: every chunk is a launch of the pair kernel that resumes from the state the
    previous one left (iterate, hash, totals); MOSFET linearisations are rebuilt from the iterate at every launch"""
    from circuitsimulator_b200 import capi
    g, topo = load_golden("dbmixer")
    B, dt, ns = 70, float(g["dt"]), 400
    P = perturb(topo.params, topo, B, 3)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    setup = capi.TranSetup(dt, ns, m, grp, tab, probes=np.array([1, 2], np.int32))
    outs = []
    for chunks, pair in (("1", False), ("5", True)):
        monkeypatch.setenv("CSIM_WAVE_CHUNKS", chunks)
        plan = fresh(monkeypatch, topo, pair)
        batch = capi.Batch(plan, B)
        batch.set_params(P.T)
        outs.append(batch.tran(setup, want_iters_step=True))
        assert ("warp-pair variant" in plan.spec_info()["log"]) == pair
    one, many = outs
    for k in ("x_dc", "x_final", "wave"):
        assert_same_bits(many[k], one[k], k)
    for k in ("iters_total", "iters_step", "hash", "status"):
        assert np.array_equal(many[k], one[k]), k
