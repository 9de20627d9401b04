"""CPU: independent-source rows a3 / a9 — an I source in the transient and PULSE / PWL waveforms
(src/element.cpp:34-66, include/sim.hpp:80-138).  tests/golden/sources.npz was generated from the UNMODIFIED
reference (make_golden.py: pin_sources); here the oracle restatement and the product's host-side tabulation are
checked against it without the reference."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import GOLDEN, assert_same_bits

CASES = {
    "const": {},
    "pulse_periodic_pwl": {1: (O.WAVE_PULSE, [0.0, 2e-3, 1e-10, 5e-11, 1e-10, 3e-10, 8e-10]),
                           2: (O.WAVE_PWL, ([0.0, 2e-10, 6e-10, 1.2e-9, 2e-9], [0.0, 1.0, 2.5, 0.3, 1.5]))},
    "pulse_single": {1: (O.WAVE_PULSE, [1e-3, -1e-3, 3e-10, 2e-10, 4e-10, 5e-10, 0.0]),
                     2: (O.WAVE_PULSE, [0.0, 2.0, 0.0, 1e-10, 1e-10, 2e-10, 5e-10])},
}


def load_sources():
    g = np.load(f"{GOLDEN}/sources.npz")
    topo = O.Topology(int(g["x_dc" if "x_dc" in g.files else "const_x_dc"].shape[0]), int(g["n_node_eq"]),
                      g["elem_type"].astype(np.int32), g["elem_flags"].astype(np.int32), g["elem_nodes"].astype(np.int32),
                      g["elem_branch"].astype(np.int32), g["elem_param"].astype(np.int32), int(g["params"].shape[0]),
                      g["params"].astype(np.float64))
    ns = int(np.floor(float(g["tstop"]) / float(g["dt"]) + 1e-12))
    return g, topo, float(g["dt"]), ns


@pytest.mark.parametrize("case", list(CASES))
def test_oracle_matches_reference_pinned_fixture(case):
    g, topo, dt, ns = load_sources()
    assert (topo.elem_type == O.ELEM_I).sum() == 1
    mode, group, table = O.make_source_tables(topo, topo.params, dt, ns, waveforms=CASES[case])
    assert_same_bits(table, g[case + "_table"], "tabulated waveforms")
    assert np.array_equal(mode, g[case + "_mode"]) and np.array_equal(group, g[case + "_group"])
    od = O.dc(topo, topo.params)
    ot = O.tran(topo, topo.params, dt, 0, ns, mode, group, table, od["x"], pivot_hash=od["hash"])
    assert_same_bits(od["x"], g[case + "_x_dc"], "DC")
    assert_same_bits(ot["wave"], g[case + "_wave"], "waveform")
    assert np.array_equal(ot["iters_step"], g[case + "_iters_step"]) and np.uint64(ot["hash"]) == g[case + "_hash"]
    if case != "const":  # the waveforms really move the circuit
        assert not np.array_equal(g[case + "_wave"], g["const_wave"])


@pytest.mark.parametrize("case", list(CASES))
def test_host_layer_tabulates_pulse_and_pwl_like_the_reference(case):
    """The product's own TranWaveform::eval (include/csim/sim.hpp) through csimh_tabulate_sources."""
    from circuitsimulator_b200.netlist import Netlist
    g, topo, dt, ns = load_sources()
    nl = Netlist(str(g["netlist"]))
    for idx, (kind, spec) in CASES[case].items():
        nl.set_source_waveform(idx, "pwl" if kind == O.WAVE_PWL else "pulse", spec)
    mode, group, table = nl.source_tables(dt, ns)
    assert np.array_equal(mode, g[case + "_mode"]) and np.array_equal(group, g[case + "_group"])
    assert_same_bits(table, g[case + "_table"], "host tabulation")
    assert_same_bits(nl.params, g["params"], "lowered parameters")
