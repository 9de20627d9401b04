"""Host-only: the batched CSV writer (row (f)1 of SURVEY §8) — column naming, number format and the
t = 0 row of the reference's CSV contract (src/tanalisis.cpp:183-231), on synthetic numbers (no GPU)."""
import numpy as np

from support import load_golden


def test_instance_csv_layout(tmp_path):
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    cols = nl.csv_columns()
    assert cols.size == nl.n_unknowns  # buffer.sp: every branch unknown is a V source or an inductor
    B, n_steps, stride = 3, 12, 4
    probes = cols[[0, 2, cols.size - 1]]
    m, grp, tab = nl.source_tables(nl.tran["tstep"], n_steps)
    setup = capi.TranSetup(nl.tran["tstep"], n_steps, m, grp, tab, probes, out_stride=stride)
    assert setup.n_out == 3
    rng = np.random.default_rng(0)
    wave = rng.normal(size=(B, probes.size, setup.n_out))
    x_dc = rng.normal(size=(nl.n_unknowns, B))
    path = tmp_path / "inst1.csv"
    nl.write_instance_csv(str(path), 1, setup, wave, x_dc=x_dc)
    lines = path.read_text().splitlines()
    names = nl.csv_header.split(",")
    assert lines[0] == "time," + ",".join(names[1 + int(np.where(cols == p)[0][0])] for p in probes)
    assert lines[0].endswith("I(L2)")
    assert lines[1] == "%.9e" % 0.0 + "".join(",%.9e" % x_dc[p, 1] for p in probes)
    for r in range(setup.n_out):
        want = "%.9e" % ((r + 1) * stride * nl.tran["tstep"]) + "".join(",%.9e" % wave[1, k, r] for k in range(probes.size))
        assert lines[2 + r] == want
    assert len(lines) == 2 + setup.n_out
    # a failed instance: rows from the first NaN on are never written (the reference throws there)
    wave[2, :, 1:] = np.nan
    nl.write_instance_csv(str(path), 2, setup, wave, status=np.array([0, 0, capi.ST_NONFINITE], np.uint32))
    assert len(path.read_text().splitlines()) == 1 + 1  # header + the one finite row (no x_dc given: no t = 0 row)


def test_dc_sweep_card_is_parsed():
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]) + "\n.DC Vin 0 3 0.25\n")
    assert nl.n_dc_sweeps == 1 and Netlist(str(g["netlist"])).n_dc_sweeps == 0


def test_oracle_harmonics_restatement_is_a_dft():
    """oracle.harmonics (the CPU statement of csim_dev_harmonics' summation order) equals numpy's FFT to rounding,
    for lengths that are and are not multiples of 32; the host twiddle tables are cos/sin(2*pi*n/M)."""
    from circuitsimulator_b200 import capi
    from oracle import oraclelib as O
    rng = np.random.default_rng(5)
    for M in (100, 37, 256):
        c, s = capi.twiddle_tables(M)
        assert np.allclose(c, np.cos(2 * np.pi * np.arange(M) / M), atol=1e-15)
        assert np.allclose(s, np.sin(2 * np.pi * np.arange(M) / M), atol=1e-15)
        x = rng.normal(size=M)
        re, im = O.harmonics(x, 5, c, s)
        F = np.fft.fft(x)[:6] / M
        assert np.allclose(re, F.real, atol=1e-13) and np.allclose(im, F.imag, atol=1e-13)
