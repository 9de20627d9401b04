"""Host-only: the oracle's Gauss-Seidel restatement against the UNMODIFIED reference's
Solver::solveLinearSystemGaussSeidel (include/solver.hpp:139-192, through oracle/_ref), bit for bit."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import load_golden


def _systems(n, seed):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(n, n))
    A += np.diag(np.abs(A).sum(1) + 1.0)  # diagonally dominant: converges
    return A, rng.normal(size=n)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref (the unmodified reference build) is not present")
@pytest.mark.parametrize("n", [1, 5, 13, 31, 40])
def test_restatement_equals_reference(n):
    A, b = _systems(n, n)
    x, sweeps = O.gauss_seidel(A, b)
    assert 1 <= sweeps < 1000 and np.abs(A @ x - b).max() < 1e-8
    assert np.array_equal(x.view(np.uint64), O.ref_gauss_seidel(A, b).view(np.uint64))
    x0 = np.linspace(-1, 1, n)  # warm start, few sweeps, no convergence
    x, sweeps = O.gauss_seidel(A, b, x0, max_iters=3, tol=1e-30)
    assert sweeps == (3 if n > 1 else 2)  # n = 1 repeats x = b/a exactly: the second sweep has err = 0
    assert np.array_equal(x.view(np.uint64), O.ref_gauss_seidel(A, b, x0, 3, 1e-30).view(np.uint64))
    if n >= 5:  # zero / tiny diagonals are replaced by +-1e-12 (solver.hpp:169-173)
        A[1, 1], A[3, 3] = 0.0, -1e-13
        x, _ = O.gauss_seidel(A, b, max_iters=20)
        assert np.array_equal(x.view(np.uint64), O.ref_gauss_seidel(A, b, max_iters=20).view(np.uint64))


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref (the unmodified reference build) is not present")
def test_mna_matrix_with_zero_diagonals_as_the_survey_probe_found():
    """buffer.sp's DC matrix has zero diagonals on its voltage-source rows: Gauss-Seidel does not solve it
    (SURVEY row a-); the restatement reproduces whatever the reference returns, bit for bit."""
    g, topo = load_golden("buffer")
    sv = O.src_dc(topo, topo.params, 1.0)
    n = topo.n_unknowns
    G, I = O.stamp(topo, 0, topo.params, np.zeros(n), np.zeros(n), sv, gmin=1e-6)
    with np.errstate(all="ignore"):
        x, sweeps = O.gauss_seidel(G, I, max_iters=60)
        r = O.ref_gauss_seidel(G, I, max_iters=60)
    assert np.array_equal(x.view(np.uint64), r.view(np.uint64)) and sweeps == 60
