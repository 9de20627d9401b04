"""GPU: csim_multi_* — one batch sharded over several devices below the C ABI (SURVEY.md §8(b) "Threading", §8(e)):
one host thread, plan, batch and stream per device, contiguous shards, results gathered into the caller's global host
arrays.  On a one-GPU box the two shards share device 0 (same code path: two threads, two plans, two streams); with
two or more GPUs visible (gpurun --gpus 2) they run on devices 0 and 1."""
import numpy as np
import pytest

from oracle import oraclelib as O
from support import assert_same_bits, load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.mark.parametrize("n_shards", [2, 3])
def test_multi_device_batch_equals_single_device_batch_and_oracle(torch, n_shards):
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, topo = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    B, ns = 37, nl.n_steps  # ragged: 37 = 19 + 18 or 13 + 12 + 12
    P = nl.monte_carlo(B, 77)
    ndev = torch.cuda.device_count()
    devices = [r % ndev for r in range(n_shards)]
    mg = nl.multi(B, devices)
    sh = [mg.shard(r) for r in range(n_shards)]
    assert [s[0] for s in sh] == devices and sh[0][1] == 0 and sum(s[2] for s in sh) == B
    assert all(sh[r][1] + sh[r][2] == sh[r + 1][1] for r in range(n_shards - 1)) and sh[0][2] - sh[-1][2] in (0, 1)
    mg.set_params(P)
    setup = nl.tran_setup(probes=np.arange(nl.n_unknowns, dtype=np.int32))
    d = mg.dc()
    out = mg.tran(setup)
    assert mg.last_kernel_ms > 0
    # the same batch on one device through csim_batch_*
    batch = capi.Batch(nl.plan(), B)
    batch.set_params(P)
    one = batch.tran(setup)
    for k in ("x_dc", "x_final", "wave"):
        assert_same_bits(out[k], one[k], f"multi vs single: {k}")
    for k in ("iters_total", "hash", "status"):
        assert np.array_equal(out[k], one[k]), k
    assert_same_bits(d["x"], one["x_dc"], "csim_multi_dc")
    # and against the oracle, instances on either side of every shard boundary
    m, grp, tab = O.make_source_tables(topo, topo.params, setup.dt, ns)
    for b in sorted({0, B - 1} | {s[1] for s in sh} | {s[1] + s[2] - 1 for s in sh}):
        od = O.dc(topo, P[:, b])
        ot = O.tran(topo, P[:, b], setup.dt, 0, ns, m, grp, tab, od["x"], pivot_hash=od["hash"])
        assert_same_bits(out["wave"][b].T, ot["wave"], f"instance {b}")
        assert out["hash"][b] == np.uint64(ot["hash"]) and out["iters_total"][b] == ot["total"]
    assert_same_bits(out["wave"][0].T, g["wave"], "instance 0 vs golden")
    mg.close()


def test_cpp_host_layer_multi_gpu_driver(torch):
    """runTransientAnalysisBackwardEulerMultiGpu (through the `.MC` card's facade) == the single-device driver."""
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]) + "\n.MC 11 5 0.05\n")
    probes = np.arange(nl.n_unknowns, dtype=np.int32)
    ndev = torch.cuda.device_count()
    one = nl.run_mc_card(probes=probes)
    two = nl.run_mc_card(probes=probes, devices=[0, 1 % ndev])
    for k in ("x_final", "wave"):
        assert_same_bits(two[k], one[k], k)
    for k in ("iters_total", "hash", "status"):
        assert np.array_equal(two[k], one[k]), k


def test_multi_rejects_bad_devices(torch):
    from circuitsimulator_b200 import capi
    from circuitsimulator_b200.netlist import Netlist
    g, _ = load_golden("buffer")
    nl = Netlist(str(g["netlist"]))
    with pytest.raises(capi.CsimError):
        nl.multi(8, [0, 99])
