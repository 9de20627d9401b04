"""CPU: boundary rows — the oracle's restatement of Solver::luDecompose against the unmodified reference
(when oracle/_ref is present) and the committed Gauss-Seidel DC fixture's sanity."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import GOLDEN, assert_same_bits


def lu_cases():
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 7, 13, 31, 32, 33, 50):
        for k in range(4):
            A = rng.uniform(-2, 2, (n, n))
            if k == 1 and n > 2:
                A[rng.integers(0, n)] = 0.0
            if k == 2 and n > 1:
                A[:, 0] = np.where(rng.uniform(size=n) < 0.5, 1.5, -1.5)
            if k == 3:
                A[np.abs(A) < 1.2] = 0.0
            yield n, k, A


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref (the unmodified reference build) is not present")
def test_lu_factor_restatement_equals_reference_luDecompose():
    for n, k, A in lu_cases():
        Lr, pr, okr = O.ref_lu_decompose(A)
        Lo, po, failed = O.lu_factor(A)
        assert okr == (not failed)
        assert np.array_equal(pr, po)
        assert_same_bits(Lr, Lo, f"packed factors n={n} case {k}")


def test_lu_factor_reconstructs_the_permuted_matrix():
    for n, k, A in lu_cases():
        LU, perm, failed = O.lu_factor(A)
        if failed:
            continue
        L = np.tril(LU, -1) + np.eye(n)
        U = np.triu(LU)
        assert np.allclose(L @ U, A[perm], rtol=1e-9, atol=1e-9)


def test_gauss_seidel_dc_fixture():
    g = np.load(f"{GOLDEN}/dc_gauss_seidel.npz")
    # SURVEY.md §8 row a-: with voltage-source rows the reference's own driver ends at x = 0 after 660 warnings
    for name in ("buffer", "dbmixer", "sources"):
        assert int(g[name + "_warnings"]) == 660 and not g[name + "_x"].any()
    assert int(g["linear_warnings"]) == 0 and np.isfinite(g["linear_x"]).all() and g["linear_x"].any()
    assert np.isfinite(g["mosdiode_x"]).all() and g["mosdiode_x"].any()
