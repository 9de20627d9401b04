"""GPU: the drop-in boundary used the way a maintainer would —
  * INTEGRATION.md route A: the reference's src/main.cpp, UNCHANGED, compiled against include/csim and linked with
    libcsim_b200.so (oracle/_ref/mysim_routeA, built by `make -C oracle routeA` where /root/reference exists), run on
    both shipped netlists: the CSV it writes has the sha256 of the reference's own CSV;
  * Solver::luDecompose returns the reference's packed L\\U factors (include/solver.hpp:64-76);
  * dcSolveGaussSeidel runs the reference's Gauss-Seidel DC drivers (src/dcanalysis.cpp:71-92,166-237,251-257),
    warnings and near-zero result included."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from oracle import oraclelib as O
from support import GOLDEN, assert_same_bits, load_golden

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ROUTE_A = os.path.join(ROOT, "oracle", "_ref", "mysim_routeA")


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.mark.skipif(not os.path.exists(ROUTE_A), reason="oracle/_ref/mysim_routeA not built (needs /root/reference at build time)")
@pytest.mark.parametrize("name", ["buffer", "dbmixer"])
def test_route_a_unchanged_main_cpp_writes_the_reference_csv(torch, tmp_path, name):
    g, _ = load_golden(name)
    want = json.load(open(f"{GOLDEN}/pinning_summary.json"))[name]["sha"]
    sp = tmp_path / f"{name}.sp"
    sp.write_text(str(g["netlist"]))
    out = tmp_path / "tran_out.csv"
    r = subprocess.run([ROUTE_A, str(sp), str(out)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "DC analysis finished." in r.stdout and "Transient analysis (Backward Euler) finished" in r.stdout
    assert hashlib.sha256(out.read_bytes()).hexdigest() == want


def _lu_cases(ns):
    rng = np.random.default_rng(5)
    for n in ns:
        for k in range(4):
            A = rng.uniform(-2, 2, (n, n))
            if k == 1 and n > 2:
                A[rng.integers(0, n)] = 0.0
            if k == 2 and n > 1:
                A[:, 0] = np.where(rng.uniform(size=n) < 0.5, 1.5, -1.5)
            if k == 3:
                A[np.abs(A) < 1.2] = 0.0
            yield n, k, A


@pytest.mark.parametrize("ns", [(1, 2, 3, 7, 13, 31, 32), (33, 50, 97)])
def test_lu_factor_returns_the_reference_packed_factors(torch, ns):
    """csim_dev_lu_factor (on chip for n <= 32, dense large-matrix kernel above) vs the reference-pinned restatement
    of Solver::luDecompose: L\\U bits, permutation, failure flag; ties, singular and sparse matrices included."""
    from circuitsimulator_b200 import capi
    for n in ns:
        mats = [A for nn, k, A in _lu_cases([n])]
        B = len(mats)
        A = torch.from_numpy(np.ascontiguousarray(np.stack([m.reshape(-1) for m in mats], 1))).cuda()
        LU, perm, sing = capi.lu_factor(A, n)
        torch.cuda.synchronize()
        LU, perm, sing = LU.cpu().numpy(), perm.cpu().numpy(), sing.cpu().numpy()
        for b, m in enumerate(mats):
            Lo, po, failed = O.lu_factor(m)
            assert bool(sing[b]) == failed, (n, b)
            if failed and n > 32:
                continue  # contents unspecified for the dense path (csim_b200.h)
            assert np.array_equal(perm[:, b], po), (n, b)
            assert_same_bits(LU[:, b].reshape(n, n), Lo, f"packed factors n={n} case {b}")


def test_host_layer_luDecompose_is_drop_in(torch):
    from circuitsimulator_b200.netlist import lu_decompose
    for n, k, A in _lu_cases([5, 31, 40]):
        LU, perm, ok = lu_decompose(A)
        Lo, po, failed = O.lu_factor(A)
        assert ok == (not failed)
        if ok:
            assert np.array_equal(perm, po)
            assert_same_bits(LU, Lo, f"Solver::luDecompose n={n} case {k}")
            # a caller that back-substitutes with the factors gets the right answer
            L, U = np.tril(LU, -1) + np.eye(n), np.triu(LU)
            assert np.allclose(L @ U, A[perm], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("name", ["buffer", "sources", "linear", "mosdiode"])  # (dbmixer: same 660-warning path as buffer, 46 s)
def test_dc_gauss_seidel_drivers_equal_the_reference(torch, capfd, name):
    """dcSolveGaussSeidel of the host layer vs the unmodified reference's own result (tests/golden/dc_gauss_seidel.npz,
    make_golden.py: pin_boundary): bit for bit, same number of warning lines."""
    from circuitsimulator_b200.netlist import Netlist
    g = np.load(f"{GOLDEN}/dc_gauss_seidel.npz")
    nl = Netlist(str(g[name + "_netlist"]))
    x = nl.dc_gauss_seidel()
    err = capfd.readouterr().err
    assert_same_bits(x, g[name + "_x"], f"dcSolveGaussSeidel {name}")
    assert err.count("WARNING") == int(g[name + "_warnings"])
