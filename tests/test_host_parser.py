"""CPU: the product's own parser + lowering reproduce what the reference's parser + Circuit produce
(golden fixtures hold the topology exported from the UNMODIFIED reference objects)."""
import os

import numpy as np
import pytest

from oracle import oraclelib as O
from support import GOLDEN, assert_same_bits, load_golden, perturb


@pytest.mark.parametrize("name", ["buffer", "dbmixer"])
def test_lowering_matches_reference_objects(name):
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden(name)
    nl = Netlist(str(g["netlist"]))
    assert (nl.n_unknowns, nl.n_node_eq, nl.n_params) == (topo.n_unknowns, topo.n_node_eq, topo.n_params)
    for k in ("elem_type", "elem_flags", "elem_nodes", "elem_branch", "elem_param"):
        assert np.array_equal(getattr(nl, k), getattr(topo, k)), k
    assert_same_bits(nl.params, topo.params, "nominal parameters")
    assert nl.tran["tstep"] == float(g["dt"]) and nl.tran["tstop"] == float(g["tstop"]) and nl.tran["enabled"]
    assert nl.n_steps == {"buffer": 300, "dbmixer": 50000}[name]
    # known equation order and probes
    if name == "buffer":
        assert nl.node_names == ["103", "101", "102", "107", "104", "115", "116", "117", "118"]
        assert list(nl.plot_probes) == [1, 8] and nl.hb == dict(f0=1e-2, n_harm=3, enabled=True)
        assert nl.csv_header.startswith("time,V(103),V(101)") and nl.csv_header.endswith("I(VDD),I(Vin),I(L1),I(L2)")
    else:
        assert list(nl.plot_probes) == [1, 2] and nl.hb["f0"] == 100e6 and nl.hb["n_harm"] == 50


def test_source_tables_match_oracle_and_quirks():
    """SIN v0 va freq td phi: the 5th number is a delay in SECONDS (dbmixer's '180' = 180 s -> constant)."""
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("dbmixer")
    nl = Netlist(str(g["netlist"]))
    m, grp, tab = nl.source_tables(n_steps=700)
    mo, go, to = O.make_source_tables(topo, topo.params, float(g["dt"]), 700)
    assert np.array_equal(m, mo) and np.array_equal(grp, go)
    assert_same_bits(tab, to, "sin tables (host libm) vs oracle")
    src = nl.source_elems()
    td = [nl.params[nl.elem_param[e] + 4] for e in src]
    assert td == [0.0, 0.0, 180.0, 180.0, 180.0, 0.0, 0.0]  # Vdd, Vlo+, Vlo-, Vrf1+, Vrf2+, Vrf1-, Vrf2-


def test_monte_carlo_recipe_and_sharding():
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    P = nl.monte_carlo(64, 20261104 + 3)
    assert_same_bits(P.T, perturb(topo.params, topo, 64, 20261104 + 3), "recipe == tests/support.perturb")
    pg = np.load(os.path.join(GOLDEN, "buffer_perturbed.npz"))
    assert_same_bits(P.T[1:6], pg["params"], "recipe == the instances pinned against the reference")
    shard = nl.monte_carlo(16, 20261104 + 3, first=32)
    assert_same_bits(shard, P[:, 32:48], "shard [32,48) of the global batch")


def test_parser_grammar_edge_cases():
    from circuitsimulator_b200.netlist import Netlist
    text = """a title line that is reported and skipped
* comment
; another comment
R1 in out 10k $ inline comment
C1 out 0 1p
+
L1 out n2
+ 3.3meg
Vs in GND DC 2 SIN 0.5 1m 1k
I1 0 n2 dc 1u
M1 n2 in 0 mymod 2u 1u
M2 n2 in vdd p 4u 1u pm
Vd vdd 0 3
.model mymod vt 0.5 mu 2e-2 cox 1e-3 lambda 0.1 cjo 1f
.MODEL pm VT -0.6 MU 1e-2 COX 1e-3
.tran 1n 10n 2n
.print tran V(out) V(in,out) I(Vs)
.plotnc M1(d) R1
.bogus 1 2
"""
    nl = Netlist(text)
    assert nl.node_names == ["in", "out", "n2", "vdd"] and nl.n_unknowns == 4 + 3  # + L1, Vs, Vd branches
    assert list(nl.elem_type) == [0, 1, 2, 3, 4, 5, 5, 3]
    p = nl.params
    assert p[0] == 10e3 and p[1] == 1e-12 and p[2] == 3.3e6
    assert list(p[3:9]) == [2.0, 0.5, 1e-3, 1e3, 0.0, 0.0] and nl.elem_flags[3] == 1
    assert p[9] == 1e-6
    assert list(p[15:19]) == [0.5, 2e-2 * 1e-3 * (2e-6 / 1e-6), 0.1, 1e-15] and nl.elem_flags[5] == 0
    assert list(p[19:23]) == [0.6, 1e-2 * 1e-3 * (4e-6 / 1e-6), 0.0, 0.0] and nl.elem_flags[6] == 1
    assert list(nl.elem_nodes[5]) == [2, 0, -1, -1] and list(nl.elem_branch) == [-1, -1, 4, 5, -1, -1, -1, 6]
    assert nl.tran == dict(tstep=1e-9, tstop=10e-9, tstart=2e-9, enabled=True)
    assert list(nl.plot_probes) == [1]  # V(out); the differential and current probes are not node probes
    setup = nl.tran_setup(probes=[1])
    assert setup.out_begin_step == 1 and setup.n_out == 9  # rows with t >= tstart: steps 1..9


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
def test_parser_against_reference_parser_on_edge_netlist():
    from circuitsimulator_b200.netlist import Netlist
    text = """* edge cases
V1 a 0 5
R1 a b 1k
R2 b gnd 2.2K
C1 b 0 10n
L1 b c 1u
R3 c 0 .5e3
Vx c d SIN 0 1 1meg 1u 0.5
R4 d 0 1e3
M1 d b 0 n 1e-6 .5e-6 2
.MODEL 2 VT 0.4 MU 3e-2 COX 6e-3 LAMBDA 0.05 CJ0 4e-14
.tran 1e-9 5e-8
"""
    nl = Netlist(text)
    t = O.Ref(text).topology()
    for k in ("elem_type", "elem_flags", "elem_nodes", "elem_branch", "elem_param"):
        assert np.array_equal(getattr(nl, k), getattr(t, k)), k
    assert_same_bits(nl.params, t.params, "params")
    assert nl.node_names == t.node_names and nl.elem_names == t.elem_names


def test_monte_carlo_card_and_cpp_generator_equal_numpy():
    """`.MC instances seed rel` (superset card, include/csim/montecarlo.hpp): the C++ host layer restates numpy's
    SeedSequence + PCG64 + uniform, so its batch equals the Python recipe's bit for bit — shards
    (bit_generator.advance) and 64-bit seeds included."""
    import numpy as np
    from circuitsimulator_b200.netlist import Netlist
    from support import load_golden
    g, _ = load_golden("dbmixer")
    text = str(g["netlist"]).replace(".end", ".MC 1024 20261104 0.05\n.end")
    nl = Netlist(text)
    assert nl.mc_card() == dict(instances=1024, seed=20261104, rel=0.05)
    assert Netlist(str(g["netlist"])).mc_card() is None
    for seed, B, first, rel in ((20261104, 300, 0, 0.05), (20261104, 77, 4096, 0.05), (0, 5, 0, 0.1),
                                (2**63 + 12345, 9, 3, 0.02), (7, 1, 0, 0.05)):
        a = nl.monte_carlo(B, seed, rel, first)
        b = nl.monte_carlo_host(B, seed, rel, first)
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), (seed, B, first)
    assert np.array_equal(nl.monte_carlo_host(4, 1)[:, 0], nl.params)  # global instance 0 is the nominal circuit
