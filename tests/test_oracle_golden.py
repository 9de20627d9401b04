"""CPU: the oracle restatement reproduces the committed golden vectors (which were generated
from, and pinned bit-for-bit against, the unmodified reference: tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden, n_steps_of


def test_norm2_definition():
    rng = np.random.default_rng(1)
    for n in (1, 13, 31, 32, 33, 404):
        v = rng.normal(size=n)
        acc = np.zeros(32)
        for i in range(n):
            acc[i & 31] = acc[i & 31] + v[i] * v[i]
        s = 16
        while s >= 1:
            for i in range(s):
                acc[i] = acc[i] + acc[i + s]
            s //= 2
        assert O.norm2(v) == np.sqrt(acc[0])


def test_buffer_full_waveform_matches_golden():
    g, topo = load_golden("buffer")
    dt, ns = float(g["dt"]), n_steps_of(g)
    assert ns == 300
    d = O.dc(topo, topo.params)
    assert_same_bits(d["x"], g["x_dc"], "buffer DC")
    assert np.array_equal(d["iters"], g["dc_iters"])
    assert list(d["iters"]) == [45, 45, 50, 47, 47, 50, 47, 45, 45, 45]  # SURVEY.md §4 known answers
    assert d["status"] & 4  # ramp steps 3 and 6 hit the 50-iteration cap
    assert d["x"][3] == 2.4900441327001546 and d["x"][2] == 1.4999949744451004  # V(107), V(102)
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    t = O.tran(topo, topo.params, dt, 0, ns, m, grp, tab, d["x"], pivot_hash=d["hash"])
    assert_same_bits(t["wave"], g["wave"], "buffer waveform")
    assert np.array_equal(t["iters_step"], g["iters_step"])
    assert t["total"] == 6035 and t["hash"] == int(g["hash"])
    hist = np.bincount(t["iters_step"])
    assert {k: int(hist[k]) for k in range(16, 24)} == {16: 6, 17: 21, 18: 22, 19: 47, 20: 91, 21: 39, 22: 48, 23: 26}


def test_buffer_perturbed_match_golden():
    g, topo = load_golden("buffer")
    p = np.load(__import__("os").path.join(__import__("support").GOLDEN, "buffer_perturbed.npz"))
    dt, ns = float(g["dt"]), n_steps_of(g)
    for k in range(p["params"].shape[0]):
        d = O.dc(topo, p["params"][k])
        assert_same_bits(d["x"], p["x_dc"][k], f"perturbed {k} DC")
        m, grp, tab = O.make_source_tables(topo, p["params"][k], dt, ns)
        t = O.tran(topo, p["params"][k], dt, 0, ns, m, grp, tab, d["x"], pivot_hash=d["hash"])
        assert_same_bits(t["x"], p["x_final"][k], f"perturbed {k} final state")
        assert np.array_equal(t["iters_step"], p["iters_step"][k]) and t["hash"] == int(p["hash"][k])


def test_dbmixer_dc_and_first_steps_match_golden():
    g, topo = load_golden("dbmixer")
    dt = float(g["dt"])
    d = O.dc(topo, topo.params)
    assert_same_bits(d["x"], g["x_dc"], "dbmixer DC")
    assert list(d["iters"]) == [46] * 10 and d["hash"] == int(g["dc_hash"])
    assert d["x"][1] == 2.8510884239329735 and d["x"][4] == 0.47206181986918888  # V(102), V(107)
    ns = 400
    m, grp, tab = O.make_source_tables(topo, topo.params, dt, ns)
    t = O.tran(topo, topo.params, dt, 0, ns, m, grp, tab, d["x"], pivot_hash=d["hash"])
    assert_same_bits(t["wave"], g["wave_first"], "dbmixer first 400 steps")
    assert np.array_equal(t["iters_step"], g["iters_step"][:ns])


def test_lu_matches_dense_textbook():
    """The restated LU is a plain partial-pivot Doolittle: check P*A = L*U and A x = b on random data,
    the zero-vector-on-singular rule and first-max-wins tie breaking."""
    rng = np.random.default_rng(3)
    for n in (1, 2, 5, 13, 31):
        A = rng.normal(size=(n, n))
        b = rng.normal(size=n)
        x, perm, failed = O.lu_solve(A, b)
        assert not failed and np.allclose(A @ x, b, atol=1e-9)
        assert sorted(perm) == list(range(n))
    A = np.array([[1.0, 2.0], [2.0, 4.0]])
    x, perm, failed = O.lu_solve(A, np.ones(2))
    assert failed and np.all(x == 0.0)
    A = np.array([[1.0, 0.0, 0.0], [-1.0, 1.0, 0.0], [1.0, 0.0, 1.0]])  # |.| ties in column 0: first wins
    x, perm, failed = O.lu_solve(A, np.ones(3))
    assert list(perm) == [0, 1, 2]


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
def test_restated_stamp_equals_reference_stamp():
    g, topo = load_golden("dbmixer")
    ref = O.Ref(str(g["netlist"]))
    rng = np.random.default_rng(11)
    for k in range(5):
        x = rng.uniform(-1, 4, topo.n_unknowns)
        Gr, Ir = ref.stamp(False, x, scale=0.3 * (k + 1))
        Go, Io = O.stamp(topo, O.MODE_DC, topo.params, x, None, O.src_dc(topo, topo.params, 0.3 * (k + 1)), gmin=0.0)
        assert_same_bits(Gr, Go, "G")
        assert_same_bits(Ir, Io, "I")
