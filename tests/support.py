"""Shared helpers of the test-suite (test infrastructure; may use oracle/)."""
import os

import numpy as np

from oracle import oraclelib as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def assert_same_bits(a, b, what=""):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    assert a.shape == b.shape, f"{what}: shape {a.shape} != {b.shape}"
    ne = bits(a) != bits(b)
    if ne.any():
        idx = np.argwhere(ne)[0]
        raise AssertionError(f"{what}: {int(ne.sum())} of {ne.size} doubles differ bitwise; first at {tuple(idx)}: "
                             f"{a[tuple(idx)]!r} vs {b[tuple(idx)]!r}")


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, f"{name}_nominal.npz"))
    n_node_eq = int(g["n_node_eq"])
    n_unknowns = int(g["x_dc"].shape[0])
    topo = O.Topology(n_unknowns, n_node_eq, g["elem_type"].astype(np.int32), g["elem_flags"].astype(np.int32),
                      g["elem_nodes"].astype(np.int32), g["elem_branch"].astype(np.int32),
                      g["elem_param"].astype(np.int32), int(g["params"].shape[0]), g["params"].astype(np.float64))
    return g, topo


def perturb(nominal, topo, n, seed, rel=0.05):
    """instance 0 nominal; others nominal*(1+rel*u), u~U(-1,1), PCG64(seed); R, C, L and MOS
    Vth/K/lambda/Cj0 only (sources nominal).  Returns [n, P]."""
    assert nominal is topo.params or np.array_equal(nominal, topo.params)
    return O.monte_carlo(topo, n, seed, rel)


def make_plan(topo, device=0):
    from circuitsimulator_b200 import capi
    return capi.Plan(topo.n_unknowns, topo.n_node_eq, topo.elem_type, topo.elem_flags, topo.elem_nodes,
                     topo.elem_branch, topo.elem_param, topo.n_params, device=device)


def n_steps_of(g):
    return int(np.floor(float(g["tstop"]) / float(g["dt"]) + 1e-12))
