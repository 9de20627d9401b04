"""ctypes bindings for oracle/libcsim_oracle.so and oracle/_ref/libcsim_ref.so.

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.  The product
package (circuitsimulator_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libcsim_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libcsim_ref.so")

ELEM_R, ELEM_C, ELEM_L, ELEM_V, ELEM_I, ELEM_MOS = range(6)
FLAG_PMOS = 1
FLAG_SRC_SIN = 1
SRC_CONST, SRC_SIN_TABLE, SRC_VAL_TABLE = 0, 1, 2
MODE_DC, MODE_TRAN = 0, 1
PIVOT_HASH_INIT = 0xCBF29CE484222325
PARAM_COUNT = {ELEM_R: 1, ELEM_C: 1, ELEM_L: 1, ELEM_V: 6, ELEM_I: 6, ELEM_MOS: 4}


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


def default_constants() -> Constants:
    """The reference's literals (see include/csim_b200.h: csim_constants)."""
    return Constants(10, 50, 1e-9, 0.35, 1e-6, 3.35e-7, 1e-4, 0.7, 1.05, 10.0, 1e-2,
                     50, 0, 1e-6, 1e-6, 0.45, 1e-15, 1e-12)


class TopologyC(C.Structure):
    _fields_ = [
        ("n_unknowns", C.c_int32), ("n_node_eq", C.c_int32), ("n_elems", C.c_int32), ("n_params", C.c_int32),
        ("elem_type", C.POINTER(C.c_int32)), ("elem_flags", C.POINTER(C.c_int32)),
        ("elem_nodes", C.POINTER(C.c_int32)), ("elem_branch", C.POINTER(C.c_int32)),
        ("elem_param", C.POINTER(C.c_int32)),
    ]


class TraceC(C.Structure):
    _fields_ = [
        ("xraw", C.POINTER(C.c_double)), ("err", C.POINTER(C.c_double)), ("perm", C.POINTER(C.c_int32)),
        ("cap_iters", C.c_int64), ("n_iters", C.c_int64), ("n_err", C.c_int64), ("min_margin", C.c_double),
    ]


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


@dataclass
class Topology:
    """Flat topology (numpy mirror of csim_topology) plus nominal parameters and names."""
    n_unknowns: int
    n_node_eq: int
    elem_type: np.ndarray
    elem_flags: np.ndarray
    elem_nodes: np.ndarray      # [n_elems, 4]
    elem_branch: np.ndarray
    elem_param: np.ndarray
    n_params: int
    params: np.ndarray          # nominal [P]
    node_names: list = field(default_factory=list)   # by equation index (node rows only)
    elem_names: list = field(default_factory=list)

    @property
    def n_elems(self):
        return int(self.elem_type.shape[0])

    def as_c(self) -> TopologyC:
        for name in ("elem_type", "elem_flags", "elem_nodes", "elem_branch", "elem_param"):
            a = np.ascontiguousarray(getattr(self, name), dtype=np.int32)
            setattr(self, name, a)
        return TopologyC(self.n_unknowns, self.n_node_eq, self.n_elems, self.n_params,
                         _p(self.elem_type, C.c_int32), _p(self.elem_flags, C.c_int32),
                         _p(self.elem_nodes, C.c_int32), _p(self.elem_branch, C.c_int32),
                         _p(self.elem_param, C.c_int32))

    def source_elems(self):
        return [e for e in range(self.n_elems) if self.elem_type[e] in (ELEM_V, ELEM_I)]


@dataclass
class Trace:
    xraw: np.ndarray
    perm: np.ndarray
    err: np.ndarray
    n_iters: int
    n_err: int
    min_margin: float = float("inf")


def build(force: bool = False) -> None:
    """(Re)build the oracle; builds oracle/_ref too when /root/reference is present."""
    if force or not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if os.path.isdir("/root/reference/src") and (force or not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        build()
        L = C.CDLL(ORACLE_SO)
        L.oracle_norm2.restype = C.c_double
        L.oracle_norm2.argtypes = [C.POINTER(C.c_double), C.c_int64]
        L.oracle_pivot_hash_update.restype = C.c_uint64
        L.oracle_pivot_hash_update.argtypes = [C.c_uint64, C.POINTER(C.c_int32), C.c_int32]
        L.oracle_tran.restype = C.c_int64
        L.oracle_count_sources.restype = C.c_int32
        L.oracle_lu_solve.restype = C.c_int
        _oracle = L
    return _oracle


def ref_available() -> bool:
    return os.path.exists(REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        build()
        if not os.path.exists(REF_SO):
            raise RuntimeError("oracle/_ref/libcsim_ref.so missing (built only where /root/reference exists)")
        L = C.CDLL(REF_SO)
        L.ref_open.restype = C.c_void_p
        L.ref_open.argtypes = [C.c_char_p]
        L.ref_close.argtypes = [C.c_void_p]
        L.ref_source_tran.restype = C.c_double
        L.ref_source_tran.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.ref_run_batch.restype = C.c_double
        L.ref_hardware_threads.restype = C.c_int
        _ref = L
    return _ref


# ----------------------------------------------------------------- oracle API

def norm2(v) -> float:
    v = np.ascontiguousarray(v, dtype=np.float64)
    return oracle_lib().oracle_norm2(_p(v, C.c_double), v.size)


def pivot_hash_update(h: int, perm) -> int:
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    return int(oracle_lib().oracle_pivot_hash_update(C.c_uint64(h), _p(perm, C.c_int32), perm.size))


WAVE_PULSE, WAVE_PWL = 1, 3  # the reference's WaveformType numbering (include/sim.hpp:25-30)


def waveform_table(kind: int, spec, dt: float, n_steps: int):
    """One CSIM_SRC_VAL_TABLE row: TranWaveform::eval((s+1)*dt) of a PULSE (spec: v1 v2 td tr tf ton per) or PWL
    (spec: (times, values)) waveform, by the oracle's restatement of include/sim.hpp:80-138."""
    L = oracle_lib()
    if kind == WAVE_PWL:
        t, v = np.asarray(spec[0], np.float64), np.asarray(spec[1], np.float64)
        flat, n = np.ascontiguousarray(np.concatenate([t, v])), t.size
    else:
        flat, n = np.ascontiguousarray(spec, np.float64), 0
    row = np.empty(n_steps, np.float64)
    L.oracle_waveform_table(C.c_int32(kind), _p(flat, C.c_double), C.c_int32(n), C.c_double(dt), C.c_int32(n_steps),
                            _p(row, C.c_double))
    return row


def make_source_tables(topo: Topology, params, dt: float, n_steps: int, waveforms=None):
    """Table convention of csim_src_mode for netlist-level sources: one SIN group per
    SIN source, CONST otherwise; waveforms = {source index: (WAVE_PULSE | WAVE_PWL, spec)} adds a
    CSIM_SRC_VAL_TABLE group for each.  Returns (src_mode, src_group, table[n_groups, n_steps])."""
    L = oracle_lib()
    params = np.asarray(params, dtype=np.float64)
    src = topo.source_elems()
    mode = np.zeros(len(src), np.int32)
    group = np.zeros(len(src), np.int32)
    rows = []
    for s, e in enumerate(src):
        if waveforms and s in waveforms:
            mode[s] = SRC_VAL_TABLE
            group[s] = len(rows)
            rows.append(waveform_table(waveforms[s][0], waveforms[s][1], dt, n_steps))
            continue
        if topo.elem_flags[e] & FLAG_SRC_SIN:
            p = params[topo.elem_param[e]: topo.elem_param[e] + 6]
            row = np.empty(n_steps, np.float64)
            L.oracle_sin_table(C.c_double(p[3]), C.c_double(p[4]), C.c_double(p[5]), C.c_double(dt),
                               C.c_int32(n_steps), _p(row, C.c_double))
            mode[s] = SRC_SIN_TABLE
            group[s] = len(rows)
            rows.append(row)
    table = np.stack(rows) if rows else np.zeros((1, max(n_steps, 1)), np.float64)
    return mode, group, np.ascontiguousarray(table)


def _trace(cap, n):
    if not cap:
        return None, None
    xraw = np.zeros((cap, n), np.float64)
    perm = np.zeros((cap, n), np.int32)
    err = np.zeros(cap, np.float64)
    t = TraceC(_p(xraw, C.c_double), _p(err, C.c_double), _p(perm, C.c_int32), cap, 0, 0, float("inf"))
    return t, (xraw, perm, err)


def _trace_out(t, bufs):
    if t is None:
        return None
    xraw, perm, err = bufs
    return Trace(xraw[: t.n_iters], perm[: t.n_iters], err[: t.n_err], int(t.n_iters), int(t.n_err),
                 float(t.min_margin))


def stamp(topo: Topology, mode: int, params, x, x_prev, src_val, dt=0.0, gmin=0.0, consts=None):
    L = oracle_lib()
    n = topo.n_unknowns
    consts = consts or default_constants()
    tc = topo.as_c()
    params = np.ascontiguousarray(params, np.float64)
    x = np.ascontiguousarray(x, np.float64)
    xp = np.ascontiguousarray(x_prev, np.float64) if x_prev is not None else None
    sv = np.ascontiguousarray(src_val, np.float64)
    G = np.zeros((n, n), np.float64)
    I = np.zeros(n, np.float64)
    L.oracle_stamp(C.byref(tc), C.byref(consts), C.c_int(mode), _p(params, C.c_double), _p(x, C.c_double),
                   _p(xp, C.c_double), _p(sv, C.c_double), C.c_double(dt), C.c_double(gmin),
                   _p(G, C.c_double), _p(I, C.c_double))
    return G, I


def src_dc(topo: Topology, params, scale: float):
    L = oracle_lib()
    tc = topo.as_c()
    params = np.ascontiguousarray(params, np.float64)
    sv = np.zeros(max(len(topo.source_elems()), 1), np.float64)
    L.oracle_src_dc(C.byref(tc), _p(params, C.c_double), C.c_double(scale), _p(sv, C.c_double))
    return sv[: len(topo.source_elems())]


def src_tran(topo: Topology, params, mode, group, table, step: int, dt: float):
    L = oracle_lib()
    tc = topo.as_c()
    params = np.ascontiguousarray(params, np.float64)
    sv = np.zeros(max(len(topo.source_elems()), 1), np.float64)
    L.oracle_src_tran(C.byref(tc), _p(params, C.c_double), _p(mode, C.c_int32), _p(group, C.c_int32),
                      _p(table, C.c_double), C.c_int32(table.shape[1]), C.c_int32(step), C.c_double(dt),
                      _p(sv, C.c_double))
    return sv[: len(topo.source_elems())]


def lu_solve(A, b, lu_eps=1e-15):
    L = oracle_lib()
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    n = b.size
    x = np.zeros(n, np.float64)
    perm = np.zeros(n, np.int32)
    failed = L.oracle_lu_solve(C.c_int32(n), _p(A, C.c_double), _p(b, C.c_double), _p(x, C.c_double),
                               _p(perm, C.c_int32), C.c_double(lu_eps))
    return x, perm, int(failed)


def dc(topo: Topology, params, consts=None, trace_cap: int = 0):
    """oracle_dc -> dict(x, iters, hash, status, trace)."""
    L = oracle_lib()
    consts = consts or default_constants()
    tc = topo.as_c()
    params = np.ascontiguousarray(params, np.float64)
    n = topo.n_unknowns
    x = np.zeros(n, np.float64)
    iters = np.zeros(consts.dc_ramp_steps, np.int32)
    h = C.c_uint64(PIVOT_HASH_INIT)
    st = C.c_uint32(0)
    t, bufs = _trace(trace_cap, n)
    L.oracle_dc(C.byref(tc), C.byref(consts), _p(params, C.c_double), _p(x, C.c_double), _p(iters, C.c_int32),
                C.byref(h), C.byref(st), C.byref(t) if t is not None else None)
    return dict(x=x, iters=iters, hash=int(h.value), status=int(st.value), trace=_trace_out(t, bufs))


def tran(topo: Topology, params, dt, step_begin, step_end, src_mode, src_group, table, x0, consts=None,
         want_wave=True, pivot_hash=PIVOT_HASH_INIT, trace_cap: int = 0):
    """oracle_tran -> dict(x, wave[nsteps,N], iters_step, total, hash, status, trace)."""
    L = oracle_lib()
    consts = consts or default_constants()
    tc = topo.as_c()
    params = np.ascontiguousarray(params, np.float64)
    n = topo.n_unknowns
    ns = step_end - step_begin
    x = np.array(x0, np.float64, copy=True)
    wave = np.full((ns, n), np.nan, np.float64) if want_wave else None
    its = np.zeros(ns, np.uint8)
    h = C.c_uint64(pivot_hash)
    st = C.c_uint32(0)
    src_mode = np.ascontiguousarray(src_mode, np.int32)
    src_group = np.ascontiguousarray(src_group, np.int32)
    table = np.ascontiguousarray(table, np.float64)
    t, bufs = _trace(trace_cap, n)
    total = L.oracle_tran(C.byref(tc), C.byref(consts), _p(params, C.c_double), C.c_double(dt),
                          C.c_int32(step_begin), C.c_int32(step_end), _p(src_mode, C.c_int32),
                          _p(src_group, C.c_int32), _p(table, C.c_double), C.c_int32(table.shape[1]),
                          _p(x, C.c_double), _p(wave, C.c_double), _p(its, C.c_uint8), C.byref(h), C.byref(st),
                          C.byref(t) if t is not None else None)
    return dict(x=x, wave=wave, iters_step=its, total=int(total), hash=int(h.value), status=int(st.value),
                trace=_trace_out(t, bufs))


def sstate(topo: Topology, params, dt, steps_per_period, max_periods, tol, src_mode, src_group, table, x0,
           consts=None, pivot_hash=PIVOT_HASH_INIT):
    """Periodic steady state, CPU restatement of the DEFINITION in include/csim_b200.h (the reference's
    src/sstateanalysis.cpp is empty: parity against the reference is unpinned; this composes the pinned
    transient oracle).  Returns dict(x, period (-1: not reached), delta, total, hash, status)."""
    x = np.array(x0, np.float64, copy=True)
    h, st, total, d = pivot_hash, 0, 0, 0.0
    for k in range(1, max_periods + 1):
        r = tran(topo, params, dt, (k - 1) * steps_per_period, k * steps_per_period, src_mode, src_group, table, x,
                 consts=consts, want_wave=False, pivot_hash=h)
        d = norm2(r["x"] - x)
        x, h, st, total = r["x"], r["hash"], st | r["status"], total + r["total"]
        if d < tol:
            return dict(x=x, period=k, delta=d, total=total, hash=h, status=st)
    return dict(x=x, period=-1, delta=d, total=total, hash=h, status=st)


# -------------------------------------------------------------- reference API

def harmonics(x, n_harm: int, cos_tab, sin_tab):
    """CPU restatement of csim_dev_harmonics' definition (include/csim_b200.h) for ONE signal x[M]: 32 lane-strided
    partial sums in ascending n with separately rounded products and sums, xor butterfly 16..1, division by M.
    The reference has no harmonic extraction (its .hb card is parsed and unused): parity unpinned."""
    x = np.asarray(x, np.float64)
    M = x.size
    rows = (M + 31) // 32
    xp = np.zeros(rows * 32, np.float64)
    xp[:M] = x
    valid = (np.arange(rows * 32) < M).reshape(rows, 32)
    xp = xp.reshape(rows, 32)
    n = np.arange(rows * 32, dtype=np.int64).reshape(rows, 32)
    re, im = np.zeros(n_harm + 1), np.zeros(n_harm + 1)
    for h in range(n_harm + 1):
        idx = (h * n) % M
        pc, ps = xp * cos_tab[idx], xp * sin_tab[idx]
        ar, ai = np.zeros(32), np.zeros(32)
        for r in range(rows):  # lanes past the end of the signal add nothing (the kernel skips them)
            ar = np.where(valid[r], ar + pc[r], ar)
            ai = np.where(valid[r], ai + ps[r], ai)
        for s in (16, 8, 4, 2, 1):
            ar = ar + ar[np.arange(32) ^ s]
            ai = ai + ai[np.arange(32) ^ s]
        re[h], im[h] = ar[0] / M, -(ai[0] / M)
    return re, im


class Ref:
    """The unmodified reference (oracle/_ref) on one netlist text."""

    def __init__(self, netlist_text: str):
        self.L = ref_lib()
        self.h = self.L.ref_open(netlist_text.encode())
        d = [C.c_int32() for _ in range(4)]
        self.L.ref_dims(C.c_void_p(self.h), *[C.byref(v) for v in d])
        self.n_unknowns, self.n_node_eq, self.n_elems, self.n_params = [int(v.value) for v in d]

    def close(self):
        if self.h:
            self.L.ref_close(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def topology(self) -> Topology:
        ne = self.n_elems
        ty = np.zeros(ne, np.int32)
        fl = np.zeros(ne, np.int32)
        nd = np.zeros((ne, 4), np.int32)
        br = np.zeros(ne, np.int32)
        po = np.zeros(ne, np.int32)
        pr = np.zeros(max(self.n_params, 1), np.float64)
        rc = self.L.ref_topology(C.c_void_p(self.h), _p(ty, C.c_int32), _p(fl, C.c_int32), _p(nd, C.c_int32),
                                 _p(br, C.c_int32), _p(po, C.c_int32), _p(pr, C.c_double))
        assert rc == 0
        buf = C.create_string_buffer(256)
        nn = []
        for eq in range(self.n_node_eq):
            self.L.ref_node_name(C.c_void_p(self.h), C.c_int(eq), buf, 256)
            nn.append(buf.value.decode())
        en = []
        for e in range(ne):
            self.L.ref_elem_name(C.c_void_p(self.h), C.c_int(e), buf, 256)
            en.append(buf.value.decode())
        return Topology(self.n_unknowns, self.n_node_eq, ty, fl, nd, br, po, self.n_params,
                        pr[: self.n_params].copy(), nn, en)

    def tran_cfg(self):
        a, b, c_, e = C.c_double(), C.c_double(), C.c_double(), C.c_int32()
        self.L.ref_tran_cfg(C.c_void_p(self.h), C.byref(a), C.byref(b), C.byref(c_), C.byref(e))
        return dict(tstep=a.value, tstop=b.value, tstart=c_.value, enabled=bool(e.value))

    def hb_cfg(self):
        f0, nh, e = C.c_double(), C.c_int32(), C.c_int32()
        self.L.ref_hb_cfg(C.c_void_p(self.h), C.byref(f0), C.byref(nh), C.byref(e))
        return dict(f0=f0.value, n_harm=int(nh.value), enabled=bool(e.value))

    def _run(self, fn, cap, *pre):
        n = self.n_unknowns
        xraw = np.zeros((max(cap, 1), n), np.float64)
        perm = np.zeros((max(cap, 1), n), np.int32)
        err = np.zeros(max(cap, 1), np.float64)
        ni, ne = C.c_int64(), C.c_int64()
        rc = fn(C.c_void_p(self.h), *pre, _p(xraw, C.c_double), _p(perm, C.c_int32), _p(err, C.c_double),
                C.c_int64(cap), C.byref(ni), C.byref(ne))
        k = min(int(ni.value), cap)
        return rc, Trace(xraw[:k], perm[:k], err[: min(int(ne.value), cap)], int(ni.value), int(ne.value))

    def dc(self, trace_cap: int = 0):
        x = np.zeros(self.n_unknowns, np.float64)
        rc, tr = self._run(self.L.ref_dc, trace_cap, _p(x, C.c_double))
        return x, tr

    def tran(self, csv_path: str, trace_cap: int = 0):
        rc, tr = self._run(self.L.ref_tran, trace_cap, csv_path.encode())
        return rc, tr

    def stamp(self, tran: bool, x, scale=1.0, time=0.0):
        n = self.n_unknowns
        x = np.ascontiguousarray(x, np.float64)
        G = np.zeros((n, n), np.float64)
        I = np.zeros(n, np.float64)
        self.L.ref_stamp(C.c_void_p(self.h), C.c_int(1 if tran else 0), _p(x, C.c_double), C.c_double(scale),
                         C.c_double(time), _p(G, C.c_double), _p(I, C.c_double))
        return G, I

    def dc_gauss_seidel(self):
        """The reference's dcSolveGaussSeidel (src/dcanalysis.cpp:251-257): (x, number of warning lines)."""
        x = np.zeros(self.n_unknowns, np.float64)
        nw = C.c_int64()
        self.L.ref_dc_gauss_seidel(C.c_void_p(self.h), _p(x, C.c_double), C.byref(nw))
        return x, int(nw.value)

    def source_tran(self, idx: int, t: float) -> float:
        return float(self.L.ref_source_tran(C.c_void_p(self.h), idx, t))

    def set_source_waveform(self, idx: int, kind: int, spec):
        """PULSE / PWL on the reference's own SourceSpec of source idx (its parser only produces SIN)."""
        if kind == WAVE_PWL:
            t, v = np.asarray(spec[0], np.float64), np.asarray(spec[1], np.float64)
            flat, n = np.ascontiguousarray(np.concatenate([t, v])), t.size
        else:
            flat, n = np.ascontiguousarray(spec, np.float64), 0
        rc = self.L.ref_set_source_waveform(C.c_void_p(self.h), C.c_int(idx), C.c_int(kind), _p(flat, C.c_double), C.c_int(n))
        assert rc == 0, rc


def ref_lu_solve(A, b):
    L = ref_lib()
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    n = b.size
    x = np.zeros(n, np.float64)
    perm = np.zeros(n, np.int32)
    L.ref_lu_solve(C.c_int(n), _p(A, C.c_double), _p(b, C.c_double), _p(x, C.c_double), _p(perm, C.c_int32))
    return x, perm


def lu_factor(A, lu_eps=1e-15):
    """oracle_lu_factor (restatement of Solver::luDecompose): (LU packed, perm, failed)."""
    L = oracle_lib()
    A = np.ascontiguousarray(A, np.float64)
    n = A.shape[0]
    LU = np.zeros((n, n), np.float64)
    perm = np.zeros(n, np.int32)
    L.oracle_lu_factor.restype = C.c_int
    failed = L.oracle_lu_factor(C.c_int32(n), _p(A, C.c_double), _p(LU, C.c_double), _p(perm, C.c_int32), C.c_double(lu_eps))
    return LU, perm, bool(failed)


def ref_lu_decompose(A):
    """The unmodified reference's Solver::luDecompose: (LU, perm, ok)."""
    L = ref_lib()
    A = np.ascontiguousarray(A, np.float64)
    n = A.shape[0]
    LU = np.zeros((n, n), np.float64)
    perm = np.zeros(n, np.int32)
    ok = L.ref_lu_decompose(C.c_int(n), _p(A, C.c_double), _p(LU, C.c_double), _p(perm, C.c_int32))
    return LU, perm, bool(ok)


def gauss_seidel(A, b, x0=None, max_iters=1000, tol=1e-10):
    """oracle_gauss_seidel (restatement of include/solver.hpp:139-192): returns (x, sweeps done)."""
    L = oracle_lib()
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    n = b.size
    x = np.zeros(n, np.float64) if x0 is None else np.array(x0, np.float64)
    L.oracle_gauss_seidel.restype = C.c_int32
    it = L.oracle_gauss_seidel(C.c_int32(n), _p(A, C.c_double), _p(b, C.c_double), _p(x, C.c_double),
                               C.c_int32(int(max_iters)), C.c_double(float(tol)))
    return x, int(it)


def ref_gauss_seidel(A, b, x0=None, max_iters=1000, tol=1e-10):
    """The unmodified reference's Solver::solveLinearSystemGaussSeidel through oracle/_ref."""
    L = ref_lib()
    A = np.ascontiguousarray(A, np.float64)
    b = np.ascontiguousarray(b, np.float64)
    n = b.size
    x = np.zeros(n, np.float64) if x0 is None else np.array(x0, np.float64)
    L.ref_gauss_seidel(C.c_int(n), _p(A, C.c_double), _p(b, C.c_double), _p(x, C.c_double), C.c_int(int(max_iters)),
                       C.c_double(float(tol)))
    return x


def ref_run_batch(netlists, n_threads: int):
    """Timed CPU baseline: returns (wall_seconds, total_timepoints)."""
    L = ref_lib()
    arr = (C.c_char_p * len(netlists))(*[s.encode() for s in netlists])
    steps = C.c_int64()
    wall = L.ref_run_batch(arr, C.c_int(len(netlists)), C.c_int(n_threads), C.byref(steps))
    return float(wall), int(steps.value)


def ref_hardware_threads() -> int:
    return int(ref_lib().ref_hardware_threads())


def monte_carlo(topo: Topology, n: int, seed: int, rel: float = 0.05):
    """SURVEY.md §8(d) recipe, [n, P]: instance 0 nominal; instance i > 0 nominal*(1+rel*u_i), u~U(-1,1) drawn
    row-wise from PCG64(seed); R, C, L and MOS Vth/K/lambda/Cj0 only (sources nominal)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    P = topo.params.size
    mask = np.zeros(P, bool)
    for e in range(topo.n_elems):
        ty = int(topo.elem_type[e])
        if ty in (ELEM_R, ELEM_C, ELEM_L, ELEM_MOS):
            mask[topo.elem_param[e]: topo.elem_param[e] + PARAM_COUNT[ty]] = True
    u = rng.uniform(-1.0, 1.0, size=(n, P))
    out = np.repeat(topo.params[None, :], n, 0)
    out[1:, mask] = topo.params[mask] * (1.0 + rel * u[1:, mask])
    return out


# ------------------------------------------------------------ netlist writer

def netlist_from_params(topo: Topology, params, tran=None, title="generated instance") -> str:
    """Netlist text in the reference grammar that reproduces `params` BIT-EXACTLY in the
    reference's own Circuit: values are printed with %.17g (round-trips through stod);
    every MOSFET gets its own .MODEL with COX=1, W=L=1 so that the reference's
    K = MU*COX*(W/L) (src/circuit.cpp:144) equals the requested K exactly."""
    params = np.asarray(params, np.float64)

    def nm(eq):
        return "0" if eq < 0 else topo.node_names[eq]

    def f(v):
        return "%.17g" % v

    lines = ["* " + title]
    models = []
    for e in range(topo.n_elems):
        ty = int(topo.elem_type[e])
        nd = topo.elem_nodes[e]
        p = params[topo.elem_param[e]: topo.elem_param[e] + PARAM_COUNT[ty]]
        name = topo.elem_names[e]
        if ty in (ELEM_R, ELEM_C, ELEM_L):
            lines.append(f"{name} {nm(nd[0])} {nm(nd[1])} {f(p[0])}")
        elif ty in (ELEM_V, ELEM_I):
            s = f"{name} {nm(nd[0])} {nm(nd[1])} DC {f(p[0])}"
            if topo.elem_flags[e] & FLAG_SRC_SIN:
                assert ty == ELEM_V, "the reference parser accepts SIN on V sources only"
                s += f" SIN {f(p[1])} {f(p[2])} {f(p[3])} {f(p[4])} {f(p[5])}"
            lines.append(s)
        elif ty == ELEM_MOS:
            isp = bool(topo.elem_flags[e] & FLAG_PMOS)
            mid = f"mdl{e}"
            vt = -p[0] if isp else p[0]
            models.append(f".MODEL {mid} VT {f(vt)} MU {f(p[1])} COX 1 LAMBDA {f(p[2])} CJ0 {f(p[3])}")
            lines.append(f"{name} {nm(nd[0])} {nm(nd[1])} {nm(nd[2])} {'p' if isp else 'n'} 1 1 {mid}")
    lines += models
    if tran:
        lines.append(f".TRAN {f(tran['tstep'])} {f(tran['tstop'])}" + (f" {f(tran['tstart'])}" if tran.get('tstart') else ""))
    return "\n".join(lines) + "\n"
