"""GPU: csim_dev_lu_solve_noalloc — the allocation-free variant of the dense batched LU solve (no counterpart in the
reference; added for callers that cannot call cudaMalloc).  Its scratch is a
static 32 MiB device array inside the library, shared by every call.  Checked through the C ABI against the existing
entry point csim_dev_lu_solve, bit for bit (the tolerance of every integer/FP64 parity test here), on calls in a row
whose workspaces overlap in the shared array (stale contents of the previous call must not matter).

Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch
    assert torch.cuda.is_available()
    return torch


def _systems(rng, n, B):
    """B systems of order n, instance-minor: a few singular, a few with ties in the first pivot search, some sparse."""
    A = rng.uniform(-2, 2, (B, n, n))
    A[1 % B, rng.integers(0, n)] = 0.0                                  # singular -> zero vector
    A[2 % B, :, 0] = np.where(rng.uniform(size=n) < 0.5, 1.5, -1.5)     # ties: first maximum wins
    A[3 % B][np.abs(A[3 % B]) < 1.2] = 0.0                              # exact zeros
    b = rng.uniform(-1, 1, (B, n))
    return np.ascontiguousarray(A.reshape(B, n * n).T), np.ascontiguousarray(b.T)


def _both(torch, capi, n, B, seed, out):
    rng = np.random.default_rng(seed)
    Ah, bh = _systems(rng, n, B)
    A, b = torch.from_numpy(Ah).cuda(), torch.from_numpy(bh).cuda()
    x0, p0, s0 = capi.lu_solve(A, b)                                    # existing entry point (allocates its scratch)
    x, perm, sing = out
    x = x[: n * B].view(n, B)
    perm = perm[: n * B].view(n, B)
    sing = sing[:B]
    x.fill_(float("nan"))
    perm.fill_(-7)
    sing.fill_(-7)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    capi.lu_solve_noalloc(A, b, x, perm, sing)                          # new entry point: caller's buffers only
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert np.array_equal(x.cpu().numpy().view(np.uint64), x0.cpu().numpy().view(np.uint64)), f"x differs, n={n}"
    assert np.array_equal(sing.cpu().numpy(), s0.cpu().numpy())
    ok = s0.cpu().numpy() == 0
    assert np.array_equal(perm.cpu().numpy()[:, ok], p0.cpu().numpy()[:, ok]), f"perm differs, n={n}"
    assert ok.sum() >= B - 2 and (~ok).any()
    return free0 - free1


def test_noalloc_lu_solve_equals_lu_solve_on_calls_in_a_row(torch):
    from circuitsimulator_b200 import capi
    cap = 404 * 600
    out = (torch.empty(cap, dtype=torch.float64, device="cuda"), torch.empty(cap, dtype=torch.int32, device="cuda"),
           torch.empty(600, dtype=torch.int32, device="cuda"))
    # first call: more systems than workspaces fit the static array (404 x 408 doubles each -> 25), so CTAs loop
    _both(torch, capi, 404, 40, 1, out)
    # second and third call in a row: other orders, workspaces overlap the first call's leftovers at other strides
    drop = _both(torch, capi, 97, 600, 2, out)
    assert drop < (2 << 20), f"device memory shrank by {drop} bytes across an allocation-free call"
    drop = _both(torch, capi, 404, 40, 3, out)
    assert drop < (2 << 20), f"device memory shrank by {drop} bytes across an allocation-free call"
    # small systems take the same kernel (csim_dev_lu_solve uses the on-chip kernel there: same bits by parity)
    _both(torch, capi, 31, 64, 4, out)


def test_noalloc_lu_solve_rejects_sizes_beyond_the_static_array(torch):
    from circuitsimulator_b200 import capi
    A = torch.zeros((4, 1), dtype=torch.float64, device="cuda")
    b = torch.zeros((2, 1), dtype=torch.float64, device="cuda")
    L = capi.lib()
    import ctypes as C
    L.csim_dev_lu_solve_noalloc.argtypes = [C.c_int32, C.c_int64] + [C.c_void_p] * 5 + [C.c_double, C.c_void_p]
    rc = L.csim_dev_lu_solve_noalloc(1281, 1, A.data_ptr(), b.data_ptr(), b.data_ptr(), None, None, 1e-15, None)
    assert rc == -2  # CSIM_E_UNSUPPORTED
    msg = L.csim_last_error().decode()
    assert "1281" in msg and "1280" in msg and "32 MiB" in msg
