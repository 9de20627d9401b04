#!/usr/bin/env python
"""Pins the oracle and writes the golden fixtures.  RUN IN THE BUILD CONTAINER ONLY
(needs /root/reference):   python tests/golden/make_golden.py

1. Compiles the UNMODIFIED reference (oracle/_ref) and the C restatement (oracle/).
2. Pins the restatement against the unmodified reference drivers, bit for bit:
   raw LU solution + pivot permutation of EVERY Newton iteration (DC ramp and
   transient), every Newton step norm, the CSV the reference writes, plus
   Element::stamp and Solver::solveLinearSystemLU on random inputs, for the
   shipped netlists and for parameter-perturbed instances.
3. Writes small raw-double fixtures (tests/golden/*.npz) that the CPU and GPU
   test-suites check against on machines where /root/reference does not exist.
The reference has NO golden vectors of its own; these are generated
from its own code.
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oraclelib as O  # noqa: E402

REF_TESTS = "/root/reference/tests"
OUT = os.path.dirname(os.path.abspath(__file__))


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def same_bits(a, b):
    return a.shape == b.shape and np.array_equal(bits(a), bits(b))


def perturb(nominal, topo, n, seed):
    """SURVEY.md §8(d) recipe: instance 0 nominal, others nominal*(1+0.05u), u~U(-1,1),
    PCG64(seed); applied to R, C, L and the MOS Vth, K, lambda, Cj0; sources nominal."""
    rng = np.random.Generator(np.random.PCG64(seed))
    P = nominal.size
    mask = np.zeros(P, bool)
    for e in range(topo.n_elems):
        ty = int(topo.elem_type[e])
        if ty in (O.ELEM_R, O.ELEM_C, O.ELEM_L, O.ELEM_MOS):
            mask[topo.elem_param[e]: topo.elem_param[e] + O.PARAM_COUNT[ty]] = True
    u = rng.uniform(-1.0, 1.0, size=(n, P))
    out = np.repeat(nominal[None, :], n, 0)
    out[1:, mask] = nominal[mask] * (1.0 + 0.05 * u[1:, mask])
    return out


def csv_from_wave(topo, x_dc, wave, dt):
    """Re-creates the reference's CSV text (src/tanalisis.cpp:183-231) from raw doubles."""
    hdr = ["time"] + [f"V({n})" for n in topo.node_names]
    for e in range(topo.n_elems):
        if topo.elem_type[e] in (O.ELEM_V, O.ELEM_L):
            hdr.append(f"I({topo.elem_names[e]})")
    cols = list(range(topo.n_node_eq)) + [int(topo.elem_branch[e]) for e in range(topo.n_elems)
                                            if topo.elem_type[e] in (O.ELEM_V, O.ELEM_L)]
    lines = [",".join(hdr)]

    def row(t, x):
        return ",".join(["%.9e" % t] + ["%.9e" % x[c] for c in cols])

    lines.append(row(0.0, x_dc))
    for s in range(wave.shape[0]):
        lines.append(row((s + 1) * dt, wave[s]))
    return "\n".join(lines) + "\n"


def pin_netlist(name, text, expect_sha16=None, cap=600_000):
    print(f"== {name}")
    ref = O.Ref(text)
    topo = ref.topology()
    cfg = ref.tran_cfg()
    dt = cfg["tstep"]
    n_steps = int(np.floor(cfg["tstop"] / dt + 1e-12))
    N = topo.n_unknowns
    print(f"   N={N} elems={topo.n_elems} P={topo.n_params} steps={n_steps}")

    # --- unmodified reference: DC alone (raw doubles) and the full transient driver with trace
    x_ref_dc, tr_dc = ref.dc(trace_cap=2000)
    with tempfile.TemporaryDirectory() as td:
        csv = os.path.join(td, "ref.csv")
        rc, tr = ref.tran(csv, trace_cap=cap)
        assert rc == 0
        ref_csv = open(csv).read()
    sha = hashlib.sha256(ref_csv.encode()).hexdigest()
    print(f"   reference CSV sha256 {sha[:16]}  newton solves (DC+TRAN) {tr.n_iters}")
    if expect_sha16:
        assert sha.startswith(expect_sha16), "reference CSV differs from SURVEY.md §4 known answer"

    # --- restatement
    o_dc = O.dc(topo, topo.params, trace_cap=2000)
    mode, group, table = O.make_source_tables(topo, topo.params, dt, n_steps)
    for s in range(len(mode)):  # table convention reproduces SourceSpec::evalTran bit for bit
        for st in (0, 1, n_steps // 3, n_steps - 1):
            v = O.src_tran(topo, topo.params, mode, group, table, st, dt)[s]
            assert bits(np.array(v)) == bits(np.array(ref.source_tran(s, (st + 1) * dt))), (s, st)
    o_tr = O.tran(topo, topo.params, dt, 0, n_steps, mode, group, table, o_dc["x"], trace_cap=cap,
                  pivot_hash=o_dc["hash"])

    # DC: raw solution, every iteration
    assert same_bits(o_dc["x"], x_ref_dc), "DC operating point differs"
    assert o_dc["trace"].n_iters == tr_dc.n_iters
    assert same_bits(o_dc["trace"].xraw, tr_dc.xraw) and np.array_equal(o_dc["trace"].perm, tr_dc.perm)
    assert same_bits(o_dc["trace"].err, tr_dc.err)
    # transient driver = DC (again) followed by the step loop
    nd = o_dc["trace"].n_iters
    assert tr.n_iters == nd + o_tr["trace"].n_iters == nd + o_tr["total"], (tr.n_iters, nd, o_tr["total"])
    assert same_bits(tr.xraw[:nd], o_dc["trace"].xraw)
    assert same_bits(tr.xraw[nd:], o_tr["trace"].xraw), "raw LU solutions differ in the transient"
    assert np.array_equal(tr.perm[nd:], o_tr["trace"].perm), "pivot permutations differ"
    assert same_bits(tr.err[o_dc["trace"].n_err:], o_tr["trace"].err), "Newton step norms differ"
    my_csv = csv_from_wave(topo, o_dc["x"], o_tr["wave"], dt)
    assert my_csv == ref_csv, "CSV text differs"
    margin = min(o_dc["trace"].min_margin, o_tr["trace"].min_margin)
    print(f"   PINNED: {tr.n_iters} raw solutions, permutations and norms bit-identical; CSV identical;"
          f" min threshold margin {margin:.3e}")

    # distinct permutations
    allperm = np.concatenate([o_dc["trace"].perm, o_tr["trace"].perm])
    uniq = np.unique(allperm, axis=0)
    print(f"   distinct pivot permutations: {len(uniq)}")

    # Element::stamp / solveLinearSystemLU on random iterates
    rng = np.random.default_rng(7)
    for k in range(25):
        x = rng.uniform(-1, 4, N)
        scale = (k % 10 + 1) / 10.0
        Gr, Ir = ref.stamp(False, x, scale=scale)
        Go, Io = O.stamp(topo, O.MODE_DC, topo.params, x, None, O.src_dc(topo, topo.params, scale), gmin=0.0)
        assert same_bits(Gr, Go) and same_bits(Ir, Io), "DC stamp differs"
        A = Go + np.diag(np.r_[np.full(topo.n_node_eq, 1e-6), np.zeros(N - topo.n_node_eq)])
        xr, pr = O.ref_lu_solve(A, Io)
        xo, po, failed = O.lu_solve(A, Io)
        assert same_bits(xr, xo) and np.array_equal(pr, po) and not failed
    return topo, cfg, o_dc, o_tr, dict(sha=sha, margin=margin, n_perm=len(uniq), n_steps=n_steps)


SOURCES_NETLIST = """* current source + PULSE / PWL waveforms (fixture; reference grammar)
Vdd vdd 0 DC 3
I1 0 n1 DC 1e-3
R1 n1 0 1k
C1 n1 0 1e-12
Vg g 0 DC 0.2
M1 vdd g n2 n 10e-6 0.35e-6 1
R2 n2 0 10k
C2 n2 0 5e-13
L1 n1 n3 1e-9
R3 n3 0 2k
.MODEL 1 VT 0.5 MU 3e-2 COX 6e-3 LAMBDA 0.05 CJ0 4e-14
.TRAN 1e-11 2e-9
.end
"""
# source index (V and I in netlist order: Vdd, I1, Vg) -> (kind, spec); PULSE: v1 v2 td tr tf ton per
SOURCES_CASES = {
    "const": {},                                                                   # I source, evalTran = dcValue
    "pulse_periodic_pwl": {1: (O.WAVE_PULSE, [0.0, 2e-3, 1e-10, 5e-11, 1e-10, 3e-10, 8e-10]),
                           2: (O.WAVE_PWL, ([0.0, 2e-10, 6e-10, 1.2e-9, 2e-9], [0.0, 1.0, 2.5, 0.3, 1.5]))},
    "pulse_single": {1: (O.WAVE_PULSE, [1e-3, -1e-3, 3e-10, 2e-10, 4e-10, 5e-10, 0.0]),
                     2: (O.WAVE_PULSE, [0.0, 2.0, 0.0, 1e-10, 1e-10, 2e-10, 5e-10])},
}


def pin_sources():
    """a3 / a9: an I source in the transient and PULSE / PWL waveforms (include/sim.hpp:80-138,
    src/element.cpp:34-66).  The reference's parser never produces them, its classes implement them: the
    harness sets the reference's own SourceSpec objects and its unmodified drivers run."""
    out = {"netlist": SOURCES_NETLIST}
    for name, wf in SOURCES_CASES.items():
        ref = O.Ref(SOURCES_NETLIST)
        topo = ref.topology()
        cfg = ref.tran_cfg()
        dt = cfg["tstep"]
        n_steps = int(np.floor(cfg["tstop"] / dt + 1e-12))
        for idx, (kind, spec) in wf.items():
            ref.set_source_waveform(idx, kind, spec)
        mode, group, table = O.make_source_tables(topo, topo.params, dt, n_steps, waveforms=wf)
        for s in range(len(mode)):  # the tabulated rows reproduce SourceSpec::evalTran at EVERY step, bit for bit
            for st in range(n_steps):
                v = O.src_tran(topo, topo.params, mode, group, table, st, dt)[s]
                assert bits(np.array(v)) == bits(np.array(ref.source_tran(s, (st + 1) * dt))), (name, s, st)
        with tempfile.TemporaryDirectory() as td:
            csv = os.path.join(td, "ref.csv")
            rc, tr = ref.tran(csv, trace_cap=50000)
            assert rc == 0
            ref_csv = open(csv).read()
        o_dc = O.dc(topo, topo.params, trace_cap=2000)
        o_tr = O.tran(topo, topo.params, dt, 0, n_steps, mode, group, table, o_dc["x"], trace_cap=50000,
                      pivot_hash=o_dc["hash"])
        nd = o_dc["trace"].n_iters
        assert tr.n_iters == nd + o_tr["total"], (name, tr.n_iters, nd, o_tr["total"])
        assert same_bits(tr.xraw[:nd], o_dc["trace"].xraw)
        assert same_bits(tr.xraw[nd:], o_tr["trace"].xraw), f"{name}: raw LU solutions differ"
        assert np.array_equal(tr.perm[nd:], o_tr["trace"].perm)
        assert csv_from_wave(topo, o_dc["x"], o_tr["wave"], dt) == ref_csv, f"{name}: CSV differs"
        print(f"   sources/{name}: PINNED ({tr.n_iters} Newton solves, CSV sha256 "
              f"{hashlib.sha256(ref_csv.encode()).hexdigest()[:16]})")
        out[name + "_x_dc"] = o_dc["x"]
        out[name + "_wave"] = o_tr["wave"]
        out[name + "_iters_step"] = o_tr["iters_step"]
        out[name + "_hash"] = np.uint64(o_tr["hash"])
        out[name + "_table"] = table
        out[name + "_mode"] = mode
        out[name + "_group"] = group
    out.update(params=topo.params, elem_type=topo.elem_type, elem_flags=topo.elem_flags, elem_nodes=topo.elem_nodes,
               elem_branch=topo.elem_branch, elem_param=topo.elem_param, n_node_eq=topo.n_node_eq, dt=dt,
               tstop=cfg["tstop"])
    np.savez_compressed(os.path.join(OUT, "sources.npz"), **out)


def pin_boundary():
    """(b): Solver::luDecompose's packed factors and the Gauss-Seidel DC driver of the unmodified reference."""
    rng = np.random.default_rng(11)
    n_lu = 0
    for n in list(range(1, 34)) + [40, 77]:
        for k in range(4):
            A = rng.uniform(-2, 2, (n, n))
            if k == 1 and n > 2:
                A[rng.integers(0, n)] = 0.0          # singular
            if k == 2 and n > 1:
                A[:, 0] = np.where(rng.uniform(size=n) < 0.5, 1.5, -1.5)  # ties in the first pivot search
            if k == 3:
                A[np.abs(A) < 1.2] = 0.0             # sparse, exact zeros
            Lr, pr, okr = O.ref_lu_decompose(A)
            Lo, po, failed = O.lu_factor(A)
            assert okr == (not failed), (n, k)
            assert np.array_equal(pr, po) and same_bits(Lr, Lo), f"luDecompose n={n} case {k}"
            n_lu += 1
    print(f"   Solver::luDecompose: {n_lu} matrices, packed factors and permutations bit-identical (incl. failures)")
    out = {}
    texts = {"buffer": open(os.path.join(REF_TESTS, "buffer.sp")).read(),
             "dbmixer": open(os.path.join(REF_TESTS, "dbmixer.sp")).read(),
             "sources": SOURCES_NETLIST,
             "linear": "* linear divider with a current source\nI1 0 a DC 2e-3\nR1 a b 1k\nR2 b 0 3k\nR3 a 0 10k\n.end\n",
             "mosdiode": "* current-driven MOS stack: no voltage source, so Gauss-Seidel has non-zero diagonals\n"
                         "I1 0 d DC 1e-4\nR1 d 0 10k\nM1 d d s n 10e-6 0.35e-6 1\nR2 s 0 500\nI2 0 s DC 2e-5\n"
                         ".MODEL 1 VT 0.5 MU 3e-2 COX 6e-3 LAMBDA 0.05 CJ0 4e-14\n.end\n"}
    for name, text in texts.items():
        x, nw = O.Ref(text).dc_gauss_seidel()
        out[name + "_netlist"] = text
        out[name + "_x"] = x
        out[name + "_warnings"] = nw
        print(f"   dcSolveGaussSeidel {name}: {nw} warning lines, max|x| = {np.abs(x).max():.3e}")
    np.savez_compressed(os.path.join(OUT, "dc_gauss_seidel.npz"), **out)


def main():
    O.build(force=True)
    summary = {}
    if "--sources-only" in sys.argv:
        pin_sources()
        return
    if "--boundary-only" in sys.argv:
        pin_boundary()
        return
    pin_sources()
    pin_boundary()

    # ---- buffer.sp (config 1): full raw waveform
    text = open(os.path.join(REF_TESTS, "buffer.sp")).read()
    topo, cfg, dc, tr, info = pin_netlist("buffer.sp", text, "32dfdc4364637c7a")
    summary["buffer"] = info
    np.savez_compressed(os.path.join(OUT, "buffer_nominal.npz"), netlist=text, x_dc=dc["x"], dc_iters=dc["iters"],
                        dc_hash=np.uint64(dc["hash"]), wave=tr["wave"], iters_step=tr["iters_step"],
                        hash=np.uint64(tr["hash"]), status=np.uint32(dc["status"] | tr["status"]),
                        params=topo.params, elem_type=topo.elem_type, elem_flags=topo.elem_flags,
                        elem_nodes=topo.elem_nodes, elem_branch=topo.elem_branch, elem_param=topo.elem_param,
                        n_node_eq=topo.n_node_eq, dt=cfg["tstep"], tstop=cfg["tstop"])

    # perturbed buffer instances through the reference's own parser + drivers
    pert = perturb(topo.params, topo, 6, 20261104 + 3)
    n_steps = info["n_steps"]
    rows = []
    for b in range(1, 6):
        t2 = O.netlist_from_params(topo, pert[b], tran=cfg)
        r2 = O.Ref(t2)
        tp2 = r2.topology()
        assert same_bits(tp2.params, pert[b]), "netlist round trip is not bit exact"
        assert np.array_equal(tp2.elem_nodes, topo.elem_nodes) and np.array_equal(tp2.elem_type, topo.elem_type)
        _, trc = r2.tran("/dev/null", trace_cap=20000)
        d2 = O.dc(topo, pert[b], trace_cap=2000)
        m, g, tab = O.make_source_tables(topo, pert[b], cfg["tstep"], n_steps)
        o2 = O.tran(topo, pert[b], cfg["tstep"], 0, n_steps, m, g, tab, d2["x"], trace_cap=20000, pivot_hash=d2["hash"])
        nd = d2["trace"].n_iters
        assert trc.n_iters == nd + o2["total"]
        assert same_bits(trc.xraw[nd:], o2["trace"].xraw) and np.array_equal(trc.perm[nd:], o2["trace"].perm)
        rows.append(dict(params=pert[b], x_dc=d2["x"], x_final=o2["x"], iters=o2["iters_step"], hash=o2["hash"]))
    print(f"   perturbed buffer instances 1..5 pinned against the reference")
    np.savez_compressed(os.path.join(OUT, "buffer_perturbed.npz"), params=np.stack([r["params"] for r in rows]),
                        x_dc=np.stack([r["x_dc"] for r in rows]), x_final=np.stack([r["x_final"] for r in rows]),
                        iters_step=np.stack([r["iters"] for r in rows]),
                        hash=np.array([r["hash"] for r in rows], np.uint64))

    # ---- dbmixer.sp (config 2): DC point, per-step Newton counts, decimated waveform
    text = open(os.path.join(REF_TESTS, "dbmixer.sp")).read()
    topo, cfg, dc, tr, info = pin_netlist("dbmixer.sp", text, "46fcdf8fa4a14bf7")
    summary["dbmixer"] = info
    np.savez_compressed(os.path.join(OUT, "dbmixer_nominal.npz"), netlist=text, x_dc=dc["x"], dc_iters=dc["iters"],
                        dc_hash=np.uint64(dc["hash"]), wave_first=tr["wave"][:400], wave_dec100=tr["wave"][99::100],
                        x_final=tr["x"], iters_step=tr["iters_step"], hash=np.uint64(tr["hash"]),
                        status=np.uint32(dc["status"] | tr["status"]), params=topo.params,
                        elem_type=topo.elem_type, elem_flags=topo.elem_flags, elem_nodes=topo.elem_nodes,
                        elem_branch=topo.elem_branch, elem_param=topo.elem_param, n_node_eq=topo.n_node_eq,
                        dt=cfg["tstep"], tstop=cfg["tstop"])
    json.dump(summary, open(os.path.join(OUT, "pinning_summary.json"), "w"), indent=1)
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
