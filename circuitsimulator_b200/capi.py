"""ctypes binding of the C ABI in include/csim_b200.h (libcsim_b200.so, built in-tree).

PyTorch is plumbing only: device buffers (torch tensors), streams, and
torch.distributed in bench.py.  Every compute entry point runs hand-written
sm_100a kernels from csrc/; there is no CPU or eager fallback — without the
library or without a CUDA device the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcsim_b200.so")

MODE_DC, MODE_TRAN = 0, 1
ELEM_R, ELEM_C, ELEM_L, ELEM_V, ELEM_I, ELEM_MOS = range(6)
FLAG_PMOS = 1
FLAG_SRC_SIN = 1
SRC_CONST, SRC_SIN_TABLE, SRC_VAL_TABLE = 0, 1, 2
PARAM_COUNT = {ELEM_R: 1, ELEM_C: 1, ELEM_L: 1, ELEM_V: 6, ELEM_I: 6, ELEM_MOS: 4}
ST_LU_SINGULAR, ST_NONFINITE, ST_DC_UNCONVERGED, ST_TRAN_UNCONVERGED = 1, 2, 4, 8
PIVOT_HASH_INIT = 0xCBF29CE484222325

EXPORTS = [
    "csim_last_error", "csim_abi_version", "csim_device_count", "csim_default_constants",
    "csim_plan_create", "csim_plan_destroy", "csim_plan_pattern", "csim_plan_dims",
    "csim_dev_stamp_csr", "csim_dev_lu_solve", "csim_dev_dc", "csim_dev_tran",
    "csim_batch_create", "csim_batch_destroy", "csim_batch_set_params", "csim_batch_dc",
    "csim_batch_tran", "csim_batch_tran_from", "csim_batch_last_kernel_ms", "csim_pivot_hash_update",
    "csim_host_stamp_csr", "csim_host_lu_solve",
    "csim_plan_jit_mode", "csim_plan_specialize", "csim_plan_spec_info", "csim_plan_spec_source",
    "csim_plan_generate_source", "csim_dev_selftest_div",
    "csim_plan_stamp_info", "csim_plan_generate_stamp_source", "csim_plan_generate_source_kind",
    "csim_plan_specialize_variants", "csim_plan_generate_source_variants", "csim_dev_selftest_tdiv",
    "csim_dev_sstate", "csim_batch_sstate", "csim_plan_force_large_path", "csim_dev_harmonics", "csim_dev_gauss_seidel", "csim_host_gauss_seidel",
    "csim_dev_fp64_peak", "csim_dev_lu_factor", "csim_host_lu_factor", "csim_dev_lu_solve_noalloc", "csim_dev_selftest_tdiv_mul",
    "csim_multi_create", "csim_multi_destroy", "csim_multi_shard", "csim_multi_set_params", "csim_multi_dc", "csim_multi_tran",
    "csim_multi_last_kernel_ms", "csim_set_last_error", "csim_plan_specialize_dc_variants", "csim_plan_dc_spec_info", "csim_dev_margin_probe",
]


class CsimError(RuntimeError):
    pass


class Constants(C.Structure):
    _fields_ = [
        ("dc_ramp_steps", C.c_int32), ("dc_max_iters", C.c_int32), ("dc_tol", C.c_double),
        ("dc_alpha", C.c_double), ("gmin_high_base", C.c_double), ("gmin_low_base", C.c_double),
        ("gmin_abs_max", C.c_double), ("fast_conv_ratio", C.c_double), ("slow_conv_ratio", C.c_double),
        ("dc_nonfinite_mul", C.c_double), ("dc_nonfinite_cap", C.c_double),
        ("tran_max_iters", C.c_int32), ("_pad0", C.c_int32), ("tran_tol", C.c_double),
        ("tran_gmin", C.c_double), ("tran_alpha", C.c_double), ("lu_eps", C.c_double),
        ("mos_off_gds", C.c_double),
    ]


class TopologyC(C.Structure):
    _fields_ = [
        ("n_unknowns", C.c_int32), ("n_node_eq", C.c_int32), ("n_elems", C.c_int32), ("n_params", C.c_int32),
        ("elem_type", C.POINTER(C.c_int32)), ("elem_flags", C.POINTER(C.c_int32)),
        ("elem_nodes", C.POINTER(C.c_int32)), ("elem_branch", C.POINTER(C.c_int32)),
        ("elem_param", C.POINTER(C.c_int32)),
    ]


class SstateArgsC(C.Structure):
    _fields_ = [("steps_per_period", C.c_int32), ("max_periods", C.c_int32), ("tol", C.c_double)]


class TranArgsC(C.Structure):
    _fields_ = [
        ("dt", C.c_double), ("step_begin", C.c_int32), ("step_end", C.c_int32),
        ("n_groups", C.c_int32), ("table_steps", C.c_int32),
        ("src_table", C.c_void_p), ("src_mode", C.POINTER(C.c_int32)), ("src_group", C.POINTER(C.c_int32)),
        ("src_group_b", C.c_void_p),
        ("n_probes", C.c_int32), ("probes", C.POINTER(C.c_int32)),
        ("out_stride", C.c_int32), ("out_begin_step", C.c_int32), ("wave_ld", C.c_int64),
    ]


_lib = None


def lib():
    """Loads libcsim_b200.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CsimError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(LIB_PATH)
        L.csim_last_error.restype = C.c_char_p
        L.csim_batch_last_kernel_ms.restype = C.c_float
        L.csim_batch_last_kernel_ms.argtypes = [C.c_void_p]
        L.csim_pivot_hash_update.restype = C.c_uint64
        L.csim_pivot_hash_update.argtypes = [C.c_uint64, C.POINTER(C.c_int32), C.c_int32]
        L.csim_plan_create.argtypes = [C.POINTER(TopologyC), C.POINTER(Constants), C.c_int, C.POINTER(C.c_void_p)]
        L.csim_plan_destroy.argtypes = [C.c_void_p]
        L.csim_plan_pattern.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                        C.POINTER(C.c_int32)]
        L.csim_plan_dims.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.csim_dev_stamp_csr.argtypes = [C.c_void_p, C.c_int, C.c_int64] + [C.c_void_p] * 4 + [C.c_double] * 3 + \
                                        [C.c_void_p] * 3
        L.csim_dev_lu_solve.argtypes = [C.c_int32, C.c_int64] + [C.c_void_p] * 5 + [C.c_double, C.c_void_p]
        L.csim_dev_dc.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 6
        L.csim_dev_tran.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(TranArgsC)] + [C.c_void_p] * 7
        L.csim_dev_sstate.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(TranArgsC), C.POINTER(SstateArgsC)] + \
            [C.c_void_p] * 7
        L.csim_plan_jit_mode.argtypes = [C.c_void_p, C.c_int]
        L.csim_plan_specialize.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        L.csim_plan_specialize_variants.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]
        L.csim_plan_generate_source_variants.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_char_p, C.c_int32]
        L.csim_plan_spec_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_uint64),
                                          C.POINTER(C.c_uint64), C.c_char_p, C.c_int32]
        L.csim_plan_spec_source.argtypes = [C.c_void_p, C.c_char_p, C.c_int32]
        L.csim_plan_generate_source.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_char_p, C.c_int32]
        L.csim_dev_selftest_div.argtypes = [C.c_int64] + [C.c_void_p] * 5
        L.csim_dev_selftest_tdiv.argtypes = [C.c_int64] + [C.c_void_p] * 6
        L.csim_plan_stamp_info.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int64),
                                           C.POINTER(C.c_int64), C.c_char_p, C.c_int32]
        L.csim_plan_generate_stamp_source.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int32]
        L.csim_plan_generate_source_kind.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.c_char_p, C.c_int32]
        L.csim_batch_create.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]
        L.csim_batch_destroy.argtypes = [C.c_void_p]
        L.csim_batch_set_params.argtypes = [C.c_void_p, C.c_void_p]
        L.csim_batch_dc.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.csim_batch_tran.argtypes = [C.c_void_p, C.POINTER(TranArgsC)] + [C.c_void_p] * 7
        _lib = L
    return _lib


def _check(rc: int, what: str):
    if rc != 0:
        raise CsimError(f"{what} failed ({rc}): {lib().csim_last_error().decode(errors='replace')}")


def default_constants() -> Constants:
    c = Constants()
    lib().csim_default_constants(C.byref(c))
    return c


def pivot_hash_update(h: int, perm) -> int:
    perm = np.ascontiguousarray(perm, np.int32)
    return int(lib().csim_pivot_hash_update(C.c_uint64(h), perm.ctypes.data_as(C.POINTER(C.c_int32)), perm.size))


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class Plan:
    """csim_plan: a netlist topology lowered to device-resident stamp programs."""

    def __init__(self, n_unknowns, n_node_eq, elem_type, elem_flags, elem_nodes, elem_branch, elem_param, n_params,
                 device: int = 0, consts: Constants | None = None):
        self._keep = [_i32(elem_type), _i32(elem_flags), _i32(np.asarray(elem_nodes).reshape(-1, 4)), _i32(elem_branch),
                      _i32(elem_param)]
        ty, fl, nd, br, po = self._keep
        topo = TopologyC(int(n_unknowns), int(n_node_eq), int(ty.shape[0]), int(n_params), _ip(ty), _ip(fl), _ip(nd),
                         _ip(br), _ip(po))
        self.consts = consts or default_constants()
        self.device = device
        h = C.c_void_p()
        _check(lib().csim_plan_create(C.byref(topo), C.byref(self.consts), device, C.byref(h)), "csim_plan_create")
        self.handle = h
        n, p, s = C.c_int32(), C.c_int32(), C.c_int32()
        _check(lib().csim_plan_dims(h, C.byref(n), C.byref(p), C.byref(s)), "csim_plan_dims")
        self.n_unknowns, self.n_params, self.n_sources = n.value, p.value, s.value
        self.n_node_eq = int(n_node_eq)

    def close(self):
        if getattr(self, "handle", None):
            lib().csim_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def jit_mode(self, mode: int):
        """0: generic kernels only; 1: automatic (default); 2: warp-per-instance; 3: thread-per-instance."""
        _check(lib().csim_plan_jit_mode(self.handle, int(mode)), "csim_plan_jit_mode")

    def force_large_path(self, on: bool = True):
        """Run this plan on the large-matrix (N > 32) kernels whatever its size."""
        lib().csim_plan_force_large_path.argtypes = [C.c_void_p, C.c_int]
        _check(lib().csim_plan_force_large_path(self.handle, int(on)), "csim_plan_force_large_path")

    def specialize(self, perm):
        perm = _i32(perm)
        assert perm.size == self.n_unknowns
        _check(lib().csim_plan_specialize(self.handle, _ip(perm)), "csim_plan_specialize")

    def specialize_variants(self, perms):
        """Thread-per-instance kernel with several predicted permutations, perms[n_variants][N]."""
        perms = np.ascontiguousarray(perms, np.int32)
        assert perms.ndim == 2 and perms.shape[1] == self.n_unknowns
        _check(lib().csim_plan_specialize_variants(self.handle, perms.shape[0], _ip(perms)), "csim_plan_specialize_variants")

    def generate_source_variants(self, perms) -> str:
        perms = np.ascontiguousarray(perms, np.int32)
        n = lib().csim_plan_generate_source_variants(self.handle, perms.shape[0], _ip(perms), None, 0)
        if n < 0:
            _check(n, "csim_plan_generate_source_variants")
        buf = C.create_string_buffer(n + 1)
        lib().csim_plan_generate_source_variants(self.handle, perms.shape[0], _ip(perms), buf, n + 1)
        return buf.value.decode()

    def spec_info(self):
        act, fb, sv = C.c_int32(), C.c_uint64(), C.c_uint64()
        perm = np.zeros(self.n_unknowns, np.int32)
        log = C.create_string_buffer(1 << 14)
        _check(lib().csim_plan_spec_info(self.handle, C.byref(act), _ip(perm), C.byref(fb), C.byref(sv), log, len(log)),
               "csim_plan_spec_info")
        return dict(active=bool(act.value), kind=int(act.value), perm=perm, fallback_solves=int(fb.value), solves=int(sv.value),
                    log=log.value.decode(errors="replace"))

    def specialize_dc_variants(self, perms):
        """Thread-per-instance DC kernel with the given predicted permutations, perms[n_variants][N]."""
        perms = np.ascontiguousarray(perms, np.int32).reshape(-1, self.n_unknowns)
        lib().csim_plan_specialize_dc_variants.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]
        _check(lib().csim_plan_specialize_dc_variants(self.handle, perms.shape[0], _ip(perms)), "csim_plan_specialize_dc_variants")

    def dc_spec_info(self):
        act, nv, fb, sv = C.c_int32(), C.c_int32(), C.c_uint64(), C.c_uint64()
        log = C.create_string_buffer(1 << 14)
        lib().csim_plan_dc_spec_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_uint64),
                                                 C.POINTER(C.c_uint64), C.c_char_p, C.c_int32]
        _check(lib().csim_plan_dc_spec_info(self.handle, C.byref(act), C.byref(nv), C.byref(fb), C.byref(sv), log, len(log)),
               "csim_plan_dc_spec_info")
        return dict(active=bool(act.value), n_variants=int(nv.value), fallback_solves=int(fb.value), solves=int(sv.value),
                    log=log.value.decode(errors="replace"))

    def spec_source(self) -> str:
        n = lib().csim_plan_spec_source(self.handle, None, 0)
        buf = C.create_string_buffer(n + 1)
        lib().csim_plan_spec_source(self.handle, buf, n + 1)
        return buf.value.decode()

    def generate_source(self, perm, kind: int = 2) -> str:
        """Source of a specialised transient kernel: kind 2 = warp per instance, 3 = thread per instance."""
        perm = _i32(perm)
        n = lib().csim_plan_generate_source_kind(self.handle, kind, _ip(perm), None, 0)
        if n < 0:
            _check(n, "csim_plan_generate_source_kind")
        buf = C.create_string_buffer(n + 1)
        lib().csim_plan_generate_source_kind(self.handle, kind, _ip(perm), buf, n + 1)
        return buf.value.decode()

    def stamp_info(self, mode: int, has_src_val: bool = True):
        """Specialisation state of the standalone assembly kernel and its algorithmic bytes per instance."""
        act, br, bw = C.c_int32(), C.c_int64(), C.c_int64()
        log = C.create_string_buffer(1 << 14)
        _check(lib().csim_plan_stamp_info(self.handle, mode, int(has_src_val), C.byref(act), C.byref(br), C.byref(bw),
                                          log, len(log)), "csim_plan_stamp_info")
        return dict(active=bool(act.value), bytes_read=int(br.value), bytes_written=int(bw.value),
                    log=log.value.decode(errors="replace"))

    def generate_stamp_source(self, mode: int) -> str:
        n = lib().csim_plan_generate_stamp_source(self.handle, mode, None, 0)
        if n < 0:
            _check(n, "csim_plan_generate_stamp_source")
        buf = C.create_string_buffer(n + 1)
        lib().csim_plan_generate_stamp_source(self.handle, mode, buf, n + 1)
        return buf.value.decode()

    def pattern(self, mode: int):
        nnz = C.c_int32()
        _check(lib().csim_plan_pattern(self.handle, mode, C.byref(nnz), None, None), "csim_plan_pattern")
        rowptr = np.zeros(self.n_unknowns + 1, np.int32)
        colidx = np.zeros(nnz.value, np.int32)
        _check(lib().csim_plan_pattern(self.handle, mode, C.byref(nnz), _ip(rowptr), _ip(colidx)), "csim_plan_pattern")
        return rowptr, colidx


def _torch():
    import torch
    return torch


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(_torch().cuda.current_stream().cuda_stream)


def _dev_f64(t, shape=None):
    torch = _torch()
    assert t.is_cuda and t.dtype == torch.float64 and t.is_contiguous(), "expected a contiguous CUDA float64 tensor"
    if shape is not None:
        assert tuple(t.shape) == tuple(shape), f"shape {tuple(t.shape)} != {tuple(shape)}"
    return t


def stamp_csr(plan: Plan, mode: int, params, x, x_prev=None, src_val=None, dt=0.0, scale=1.0, gmin=0.0,
              out=None):
    """csim_dev_stamp_csr on torch CUDA tensors: params [P,B], x [N,B] -> (vals [nnz,B], rhs [N,B])."""
    torch = _torch()
    B = params.shape[1]
    _dev_f64(params, (plan.n_params, B))
    _dev_f64(x, (plan.n_unknowns, B))
    nnz = plan.pattern(mode)[1].size
    if out is None:
        vals = torch.empty((nnz, B), dtype=torch.float64, device=params.device)
        rhs = torch.empty((plan.n_unknowns, B), dtype=torch.float64, device=params.device)
    else:
        vals, rhs = out
    _check(lib().csim_dev_stamp_csr(plan.handle, mode, B, _ptr(params), _ptr(x), _ptr(x_prev), _ptr(src_val),
                                    float(dt), float(scale), float(gmin), _ptr(vals), _ptr(rhs), _stream()),
           "csim_dev_stamp_csr")
    return vals, rhs


def lu_solve(A, b, lu_eps=1e-15):
    """csim_dev_lu_solve: A [N*N,B] (row-major (i,j) major, instance minor), b [N,B]."""
    torch = _torch()
    n, B = b.shape
    _dev_f64(A, (n * n, B))
    _dev_f64(b)
    x = torch.empty_like(b)
    perm = torch.empty((n, B), dtype=torch.int32, device=b.device)
    sing = torch.empty((B,), dtype=torch.int32, device=b.device)
    _check(lib().csim_dev_lu_solve(n, B, _ptr(A), _ptr(b), _ptr(x), _ptr(perm), _ptr(sing), float(lu_eps), _stream()),
           "csim_dev_lu_solve")
    return x, perm, sing


def lu_solve_noalloc(A, b, x, perm, sing, lu_eps=1e-15):
    """csim_dev_lu_solve_noalloc: like lu_solve but every buffer (outputs included) is the caller's, and the library
    allocates nothing (static device scratch).  A [N*N,B], b [N,B] -> x [N,B], perm [N,B] int32, sing [B] int32."""
    n, B = b.shape
    _dev_f64(A, (n * n, B))
    _dev_f64(b)
    _dev_f64(x, (n, B))
    lib().csim_dev_lu_solve_noalloc.argtypes = [C.c_int32, C.c_int64] + [C.c_void_p] * 5 + [C.c_double, C.c_void_p]
    _check(lib().csim_dev_lu_solve_noalloc(n, B, _ptr(A), _ptr(b), _ptr(x), _ptr(perm), _ptr(sing), float(lu_eps),
                                           _stream()), "csim_dev_lu_solve_noalloc")


def lu_factor(A, n, lu_eps=1e-15):
    """csim_dev_lu_factor: A [N*N,B] -> (LU [N*N,B] packed L\\U as Solver::luDecompose leaves it, perm [N,B], singular [B])."""
    torch = _torch()
    B = A.shape[1]
    _dev_f64(A, (n * n, B))
    LU = torch.empty_like(A)
    perm = torch.empty((n, B), dtype=torch.int32, device=A.device)
    sing = torch.empty((B,), dtype=torch.int32, device=A.device)
    lib().csim_dev_lu_factor.argtypes = [C.c_int32, C.c_int64] + [C.c_void_p] * 4 + [C.c_double, C.c_void_p]
    _check(lib().csim_dev_lu_factor(n, B, _ptr(A), _ptr(LU), _ptr(perm), _ptr(sing), float(lu_eps), _stream()),
           "csim_dev_lu_factor")
    return LU, perm, sing


def new_diag(B, device):
    """(pivot_hash int64-as-uint64 [B], status int32-as-uint32 [B]) initialised for a fresh analysis."""
    torch = _torch()
    h = torch.full((B,), np.uint64(PIVOT_HASH_INIT).astype(np.int64).item(), dtype=torch.int64, device=device)
    st = torch.zeros((B,), dtype=torch.int32, device=device)
    return h, st


def dc(plan: Plan, params, pivot_hash=None, status=None):
    """csim_dev_dc -> dict(x [N,B], iters [ramp,B], hash, status) (device tensors)."""
    torch = _torch()
    B = params.shape[1]
    _dev_f64(params, (plan.n_params, B))
    if pivot_hash is None:
        pivot_hash, status = new_diag(B, params.device)
    x = torch.empty((plan.n_unknowns, B), dtype=torch.float64, device=params.device)
    iters = torch.zeros((plan.consts.dc_ramp_steps, B), dtype=torch.int32, device=params.device)
    _check(lib().csim_dev_dc(plan.handle, B, _ptr(params), _ptr(x), _ptr(iters), _ptr(pivot_hash), _ptr(status),
                             _stream()), "csim_dev_dc")
    return dict(x=x, iters=iters, hash=pivot_hash, status=status)


class TranSetup:
    """Host-side description of a transient run (sources, probes, output decimation)."""

    def __init__(self, dt, n_steps, src_mode, src_group, table, probes, out_stride=1, out_begin_step=0,
                 src_group_b=None):
        self.dt = float(dt)
        self.n_steps = int(n_steps)
        self.src_mode = _i32(src_mode)
        self.src_group = _i32(src_group)
        self.table = np.ascontiguousarray(table, np.float64)  # [n_groups, table_steps]
        self.probes = _i32(probes)
        self.out_stride = int(out_stride)
        self.out_begin_step = int(out_begin_step)
        # per-instance source groups [n_sources, B] (numpy int32, or None: every instance uses src_group)
        self.src_group_b = None if src_group_b is None else np.ascontiguousarray(src_group_b, np.int32)
        self.n_out = self.n_steps // self.out_stride - self.out_begin_step // self.out_stride
        self._dev_table = {}

    def device_table(self, device):
        torch = _torch()
        key = str(device)
        if key not in self._dev_table:
            self._dev_table[key] = torch.from_numpy(self.table).to(device)
        return self._dev_table[key]

    def device_group_b(self, device):
        if self.src_group_b is None:
            return None
        key = "gb" + str(device)
        if key not in self._dev_table:
            self._dev_table[key] = _torch().from_numpy(self.src_group_b).to(device)
        return self._dev_table[key]

    def args(self, step_begin, step_end, table_ptr, group_b_ptr, wave_ld):
        return TranArgsC(self.dt, step_begin, step_end, int(self.table.shape[0]), int(self.table.shape[1]),
                         table_ptr, _ip(self.src_mode), _ip(self.src_group), group_b_ptr, int(self.probes.size),
                         _ip(self.probes), self.out_stride, self.out_begin_step, int(wave_ld))


def tran(plan: Plan, params, x, setup: TranSetup, step_begin=0, step_end=None, wave=None, iters_total=None,
         iters_step=None, pivot_hash=None, status=None, want_iters_step=False):
    """csim_dev_tran on device tensors; x [N,B] is updated in place.  Returns a dict of device tensors."""
    torch = _torch()
    B = params.shape[1]
    step_end = setup.n_steps if step_end is None else step_end
    _dev_f64(params, (plan.n_params, B))
    _dev_f64(x, (plan.n_unknowns, B))
    dev = params.device
    if pivot_hash is None:
        pivot_hash, status = new_diag(B, dev)
    npr = int(setup.probes.size)
    if wave is None and npr:
        wave = torch.full((B, npr, max(setup.n_out, 1)), float("nan"), dtype=torch.float64, device=dev)
    if iters_total is None:
        iters_total = torch.zeros((B,), dtype=torch.int32, device=dev)
    if iters_step is None and want_iters_step:
        iters_step = torch.zeros((setup.n_steps, B), dtype=torch.uint8, device=dev)
    table = setup.device_table(dev)
    gb = setup.device_group_b(dev)
    a = setup.args(step_begin, step_end, table.data_ptr(), gb.data_ptr() if gb is not None else None,
                   wave.shape[2] if wave is not None else 0)
    _check(lib().csim_dev_tran(plan.handle, B, _ptr(params), C.byref(a), _ptr(x), _ptr(wave), _ptr(iters_total),
                               _ptr(iters_step), _ptr(pivot_hash), _ptr(status), _stream()), "csim_dev_tran")
    return dict(x=x, wave=wave, iters_total=iters_total, iters_step=iters_step, hash=pivot_hash, status=status)


def sstate(plan: Plan, params, x, setup: TranSetup, steps_per_period: int, max_periods: int, tol: float,
           pivot_hash=None, status=None, iters_total=None):
    """csim_dev_sstate: periodic steady state by period-to-period integration (definition: csim_b200.h).
    x [N,B] is updated in place; setup.table must cover max_periods*steps_per_period steps."""
    torch = _torch()
    B = params.shape[1]
    _dev_f64(params, (plan.n_params, B))
    _dev_f64(x, (plan.n_unknowns, B))
    dev = params.device
    if pivot_hash is None:
        pivot_hash, status = new_diag(B, dev)
    if iters_total is None:
        iters_total = torch.zeros((B,), dtype=torch.int32, device=dev)
    periods = torch.empty((B,), dtype=torch.int32, device=dev)
    delta = torch.empty((B,), dtype=torch.float64, device=dev)
    table = setup.device_table(dev)
    gb = setup.device_group_b(dev)
    a = setup.args(0, 0, table.data_ptr(), gb.data_ptr() if gb is not None else None, 0)
    ss = SstateArgsC(int(steps_per_period), int(max_periods), float(tol))
    _check(lib().csim_dev_sstate(plan.handle, B, _ptr(params), C.byref(a), C.byref(ss), _ptr(x), _ptr(periods),
                                 _ptr(delta), _ptr(iters_total), _ptr(pivot_hash), _ptr(status), _stream()),
           "csim_dev_sstate")
    return dict(x=x, periods=periods, delta=delta, iters_total=iters_total, hash=pivot_hash, status=status)


def gauss_seidel(A, b, x0=None, max_iters: int = 1000, tol: float = 1e-10):
    """csim_dev_gauss_seidel: A [n*n, B], b [n, B] (device, instance-minor).  Returns (x [n, B], sweeps [B])."""
    torch = _torch()
    n, B = b.shape
    _dev_f64(A, (n * n, B))
    _dev_f64(b, (n, B))
    x = torch.zeros_like(b) if x0 is None else x0.clone()
    sweeps = torch.empty((B,), dtype=torch.int32, device=b.device)
    L = lib()
    L.csim_dev_gauss_seidel.argtypes = [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                                        C.c_void_p, C.c_void_p]
    _check(L.csim_dev_gauss_seidel(int(n), int(B), _ptr(A), _ptr(b), _ptr(x), int(max_iters), float(tol), _ptr(sweeps),
                                   _stream()), "csim_dev_gauss_seidel")
    return x, sweeps


def twiddle_tables(n_samples: int):
    """cos/sin(2*pi*n/M) tabulated by the host layer with the host libm (csimh_twiddle_table)."""
    c, s = np.zeros(n_samples, np.float64), np.zeros(n_samples, np.float64)
    L = lib()
    L.csimh_twiddle_table.argtypes = [C.c_int32, C.c_void_p, C.c_void_p]
    if L.csimh_twiddle_table(int(n_samples), c.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p)) != 0:
        raise CsimError("csimh_twiddle_table failed")
    return c, s


def harmonics(wave, n_samples: int, n_harm: int, tables=None):
    """csim_dev_harmonics over the LAST n_samples columns of every row of `wave` (device, [..., n_out], e.g. the
    wave of csim_dev_tran with the last period recorded).  Returns (re, im), each [n_harm+1, *wave.shape[:-1]]."""
    torch = _torch()
    assert wave.is_cuda and wave.dtype == torch.float64 and wave.is_contiguous()
    n_out = wave.shape[-1]
    assert 0 < n_samples <= n_out
    n_sig = wave.numel() // n_out
    c, s = tables if tables is not None else twiddle_tables(n_samples)
    ct, st = torch.from_numpy(c).to(wave.device), torch.from_numpy(s).to(wave.device)
    re = torch.empty((n_harm + 1, n_sig), dtype=torch.float64, device=wave.device)
    im = torch.empty_like(re)
    L = lib()
    L.csim_dev_harmonics.argtypes = [C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p]
    first = wave.data_ptr() + 8 * (n_out - n_samples)
    _check(L.csim_dev_harmonics(n_sig, int(n_samples), n_out, C.c_void_p(first), int(n_harm), _ptr(ct), _ptr(st),
                                _ptr(re), _ptr(im), _stream()), "csim_dev_harmonics")
    shape = (n_harm + 1,) + tuple(wave.shape[:-1])
    return re.reshape(shape), im.reshape(shape)


class Batch:
    """csim_batch: the host-buffer API (numpy in, numpy out; copies inside)."""

    def __init__(self, plan: Plan, B: int):
        self.plan, self.B = plan, int(B)
        h = C.c_void_p()
        _check(lib().csim_batch_create(plan.handle, self.B, C.byref(h)), "csim_batch_create")
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            lib().csim_batch_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_params(self, params):
        params = np.ascontiguousarray(params, np.float64)
        assert params.shape == (self.plan.n_params, self.B)
        _check(lib().csim_batch_set_params(self.handle, params.ctypes.data_as(C.c_void_p)), "csim_batch_set_params")

    def dc(self):
        N, B = self.plan.n_unknowns, self.B
        x = np.empty((N, B), np.float64)
        iters = np.empty((self.plan.consts.dc_ramp_steps, B), np.int32)
        h = np.empty(B, np.uint64)
        st = np.empty(B, np.uint32)
        _check(lib().csim_batch_dc(self.handle, x.ctypes.data_as(C.c_void_p), iters.ctypes.data_as(C.c_void_p),
                                   h.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p)), "csim_batch_dc")
        return dict(x=x, iters=iters, hash=h, status=st)

    def tran(self, setup: TranSetup, want_iters_step=False, out=None):
        """DC operating point + whole transient (the reference's runTransientAnalysisBackwardEuler)."""
        N, B = self.plan.n_unknowns, self.B
        npr = int(setup.probes.size)
        if out is None:
            out = dict(x_dc=np.empty((N, B), np.float64), x_final=np.empty((N, B), np.float64),
                       wave=np.empty((B, npr, max(setup.n_out, 1)), np.float64) if npr else None,
                       iters_total=np.empty(B, np.int32),
                       iters_step=np.empty((setup.n_steps, B), np.uint8) if want_iters_step else None,
                       hash=np.empty(B, np.uint64), status=np.empty(B, np.uint32))
        gb = setup.src_group_b
        a = setup.args(0, setup.n_steps, setup.table.ctypes.data, gb.ctypes.data if gb is not None else None,
                       out["wave"].shape[2] if out["wave"] is not None else 0)

        def p(k):
            return out[k].ctypes.data_as(C.c_void_p) if out.get(k) is not None else None

        _check(lib().csim_batch_tran(self.handle, C.byref(a), p("x_dc"), p("x_final"), p("wave"), p("iters_total"),
                                     p("iters_step"), p("hash"), p("status")), "csim_batch_tran")
        return out

    @property
    def last_kernel_ms(self) -> float:
        return float(lib().csim_batch_last_kernel_ms(self.handle))


class Multi:
    """csim_multi: one batch of B instances sharded over several GPUs below the C ABI (one host thread, plan, batch
    and stream per device; contiguous shards).  Host numpy arrays in, host numpy arrays out."""

    def __init__(self, topo_arrays, B: int, devices, consts: Constants | None = None):
        n_unknowns, n_node_eq, elem_type, elem_flags, elem_nodes, elem_branch, elem_param, n_params = topo_arrays
        self._keep = [_i32(elem_type), _i32(elem_flags), _i32(np.asarray(elem_nodes).reshape(-1, 4)), _i32(elem_branch),
                      _i32(elem_param)]
        ty, fl, nd, br, po = self._keep
        topo = TopologyC(int(n_unknowns), int(n_node_eq), int(ty.shape[0]), int(n_params), _ip(ty), _ip(fl), _ip(nd),
                         _ip(br), _ip(po))
        self.consts = consts or default_constants()
        self.N, self.P, self.B = int(n_unknowns), int(n_params), int(B)
        dev = _i32(devices)
        L = lib()
        L.csim_multi_create.argtypes = [C.POINTER(TopologyC), C.POINTER(Constants), C.c_int32, C.POINTER(C.c_int32), C.c_int64,
                                        C.POINTER(C.c_void_p)]
        L.csim_multi_destroy.argtypes = [C.c_void_p]
        L.csim_multi_shard.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.csim_multi_set_params.argtypes = [C.c_void_p, C.c_void_p]
        L.csim_multi_dc.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.csim_multi_tran.argtypes = [C.c_void_p, C.POINTER(TranArgsC)] + [C.c_void_p] * 6
        L.csim_multi_last_kernel_ms.restype = C.c_float
        L.csim_multi_last_kernel_ms.argtypes = [C.c_void_p]
        h = C.c_void_p()
        _check(L.csim_multi_create(C.byref(topo), C.byref(self.consts), int(dev.size), _ip(dev), int(B), C.byref(h)),
               "csim_multi_create")
        self.handle, self.n_shards = h, int(dev.size)

    def close(self):
        if getattr(self, "handle", None):
            lib().csim_multi_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def shard(self, rank: int):
        d, f, c = C.c_int32(), C.c_int64(), C.c_int64()
        _check(lib().csim_multi_shard(self.handle, rank, C.byref(d), C.byref(f), C.byref(c)), "csim_multi_shard")
        return int(d.value), int(f.value), int(c.value)

    def set_params(self, params):
        params = np.ascontiguousarray(params, np.float64)
        assert params.shape == (self.P, self.B)
        _check(lib().csim_multi_set_params(self.handle, params.ctypes.data_as(C.c_void_p)), "csim_multi_set_params")

    def dc(self):
        x = np.empty((self.N, self.B), np.float64)
        it = np.empty((self.consts.dc_ramp_steps, self.B), np.int32)
        h = np.empty(self.B, np.uint64)
        st = np.empty(self.B, np.uint32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        _check(lib().csim_multi_dc(self.handle, vp(x), vp(it), vp(h), vp(st)), "csim_multi_dc")
        return dict(x=x, iters=it, hash=h, status=st)

    def tran(self, setup: "TranSetup"):
        npr, ld = int(setup.probes.size), max(setup.n_out, 1)
        out = dict(x_dc=np.empty((self.N, self.B)), x_final=np.empty((self.N, self.B)),
                   wave=np.empty((self.B, npr, ld)), iters_total=np.empty(self.B, np.int32),
                   hash=np.empty(self.B, np.uint64), status=np.empty(self.B, np.uint32))
        gb = setup.src_group_b
        a = setup.args(0, setup.n_steps, setup.table.ctypes.data, gb.ctypes.data if gb is not None else None, ld)
        vp = lambda a_: a_.ctypes.data_as(C.c_void_p)  # noqa: E731
        _check(lib().csim_multi_tran(self.handle, C.byref(a), vp(out["x_dc"]), vp(out["x_final"]), vp(out["wave"]),
                                     vp(out["iters_total"]), vp(out["hash"]), vp(out["status"])), "csim_multi_tran")
        return out

    @property
    def last_kernel_ms(self) -> float:
        return float(lib().csim_multi_last_kernel_ms(self.handle))


def margin_probe(margin):
    """csim_dev_margin_probe: margin = CUDA float64 tensor [B] initialised to +inf (or None to switch the probe off)."""
    lib().csim_dev_margin_probe.argtypes = [C.c_void_p]
    _check(lib().csim_dev_margin_probe(_ptr(margin)), "csim_dev_margin_probe")


def fp64_peak(device: int = 0) -> float:
    """csim_dev_fp64_peak: sustained separate DMUL+DADD Tflop/s of the chip (the fused kernels' roofline peak)."""
    v = C.c_double()
    lib().csim_dev_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    _check(lib().csim_dev_fp64_peak(int(device), C.byref(v)), "csim_dev_fp64_peak")
    return float(v.value)


def selftest_tdiv(a, b):
    """(q, q_ref, in_range) on the device for float64 CUDA tensors a, b (see csim_dev_selftest_tdiv)."""
    torch = _torch()
    _dev_f64(a)
    _dev_f64(b, a.shape)
    q, qr = torch.empty_like(a), torch.empty_like(a)
    ok = torch.empty(a.shape, dtype=torch.int32, device=a.device)
    _check(lib().csim_dev_selftest_tdiv(a.numel(), _ptr(a), _ptr(b), _ptr(q), _ptr(qr), _ptr(ok), _stream()),
           "csim_dev_selftest_tdiv")
    return q, qr, ok


def selftest_tdiv_mul(a, b):
    """csim_dev_selftest_tdiv_mul -> (q, q_ref, flags): bit 0 range guard accepts, bit 1 pivot guard fires."""
    torch = _torch()
    q = torch.empty_like(a)
    qr = torch.empty_like(a)
    fl = torch.empty(a.shape, dtype=torch.int32, device=a.device)
    lib().csim_dev_selftest_tdiv_mul.argtypes = [C.c_int64] + [C.c_void_p] * 6
    _check(lib().csim_dev_selftest_tdiv_mul(a.numel(), _ptr(a), _ptr(b), _ptr(q), _ptr(qr), _ptr(fl), _stream()),
           "csim_dev_selftest_tdiv_mul")
    return q, qr, fl


def selftest_div(a, b):
    """(q_split, q_ref) on the device for float64 CUDA tensors a, b (see csim_dev_selftest_div)."""
    torch = _torch()
    _dev_f64(a)
    _dev_f64(b, a.shape)
    qs, qr = torch.empty_like(a), torch.empty_like(a)
    _check(lib().csim_dev_selftest_div(a.numel(), _ptr(a), _ptr(b), _ptr(qs), _ptr(qr), _stream()),
           "csim_dev_selftest_div")
    return qs, qr
