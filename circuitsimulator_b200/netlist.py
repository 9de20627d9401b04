"""Host side above the C ABI: the product's own .sp parser + lowering (C++: include/csim/*.hpp,
circuitsimulator_b200/host/*.cpp) seen from Python, plus batch parameter generation."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi


def _hlib():
    L = capi.lib()
    if not getattr(L, "_csimh_ready", False):
        L.csimh_last_error.restype = C.c_char_p
        L.csimh_parse.restype = C.c_void_p
        L.csimh_parse.argtypes = [C.c_char_p]
        L.csimh_free.argtypes = [C.c_void_p]
        for name in ("csimh_dims", "csimh_topology", "csimh_node_name", "csimh_elem_name", "csimh_csv_header",
                     "csimh_tran_cfg", "csimh_hb_cfg", "csimh_plot_probes", "csimh_run_tran_csv",
                     "csimh_write_instance_csv", "csimh_n_dc_sweeps", "csimh_run_dc_sweep"):
            getattr(L, name).restype = C.c_int
        L.csimh_tabulate_sources.argtypes = [C.c_void_p, C.c_double, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p]
        L.csimh_sin_table.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, C.c_void_p]
        L.csimh_write_instance_csv.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p,
                                               C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.csimh_set_source_waveform.restype = C.c_int
        L.csimh_set_source_waveform.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32]
        L.csimh_n_dc_sweeps.argtypes = [C.c_void_p]
        L.csimh_run_dc_sweep.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p]
        L._csimh_ready = True
    return L


class Netlist:
    """A parsed netlist lowered to the flat topology of include/csim_b200.h."""

    def __init__(self, text: str):
        L = _hlib()
        self.text = text
        self._h = L.csimh_parse(text.encode())
        if not self._h:
            raise capi.CsimError("csimh_parse: " + L.csimh_last_error().decode(errors="replace"))
        h = C.c_void_p(self._h)
        d = [C.c_int32() for _ in range(5)]
        L.csimh_dims(h, *[C.byref(v) for v in d])
        self.n_unknowns, self.n_node_eq, self.n_elems, self.n_params, self.n_sources = [int(v.value) for v in d]
        ne = self.n_elems
        self.elem_type = np.zeros(ne, np.int32)
        self.elem_flags = np.zeros(ne, np.int32)
        self.elem_nodes = np.zeros((ne, 4), np.int32)
        self.elem_branch = np.zeros(ne, np.int32)
        self.elem_param = np.zeros(ne, np.int32)
        self.params = np.zeros(max(self.n_params, 1), np.float64)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        L.csimh_topology(h, vp(self.elem_type), vp(self.elem_flags), vp(self.elem_nodes), vp(self.elem_branch),
                         vp(self.elem_param), vp(self.params))
        self.params = self.params[: self.n_params].copy()
        buf = C.create_string_buffer(1 << 16)
        self.node_names, self.elem_names = [], []
        for eq in range(self.n_node_eq):
            L.csimh_node_name(h, eq, buf, len(buf))
            self.node_names.append(buf.value.decode())
        for e in range(ne):
            L.csimh_elem_name(h, e, buf, len(buf))
            self.elem_names.append(buf.value.decode())
        L.csimh_csv_header(h, buf, len(buf))
        self.csv_header = buf.value.decode()
        a, b, c_, en = C.c_double(), C.c_double(), C.c_double(), C.c_int32()
        L.csimh_tran_cfg(h, C.byref(a), C.byref(b), C.byref(c_), C.byref(en))
        self.tran = dict(tstep=a.value, tstop=b.value, tstart=c_.value, enabled=bool(en.value))
        f0, nh = C.c_double(), C.c_int32()
        L.csimh_hb_cfg(h, C.byref(f0), C.byref(nh), C.byref(en))
        self.hb = dict(f0=f0.value, n_harm=int(nh.value), enabled=bool(en.value))
        n = L.csimh_plot_probes(h, None, 0)
        pr = np.zeros(max(n, 1), np.int32)
        L.csimh_plot_probes(h, vp(pr), n)
        self.plot_probes = pr[:n].copy()

    @classmethod
    def from_file(cls, path: str) -> "Netlist":
        with open(path) as f:
            return cls(f.read())

    def close(self):
        if getattr(self, "_h", None):
            _hlib().csimh_free(C.c_void_p(self._h))
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_steps(self) -> int:
        """nSteps = floor(tstop/dt + 1e-12) (reference: src/tanalisis.cpp:238)."""
        return int(np.floor(self.tran["tstop"] / self.tran["tstep"] + 1e-12))

    def source_elems(self):
        return [e for e in range(self.n_elems) if self.elem_type[e] in (capi.ELEM_V, capi.ELEM_I)]

    def plan(self, device: int = 0, consts=None) -> capi.Plan:
        return capi.Plan(self.n_unknowns, self.n_node_eq, self.elem_type, self.elem_flags, self.elem_nodes,
                         self.elem_branch, self.elem_param, self.n_params, device=device, consts=consts)

    def set_source_waveform(self, source, kind: str, spec):
        """PULSE (spec = v1 v2 td tr tf ton per) or PWL (spec = (times, values)) waveform on an independent source
        (name or index among the V/I sources in netlist order) — include/sim.hpp:80-138."""
        src = self.source_elems()
        idx = source if isinstance(source, int) else [self.elem_names[e].lower() for e in src].index(source.lower())
        if kind.lower() == "pwl":
            t, v = np.asarray(spec[0], np.float64), np.asarray(spec[1], np.float64)
            flat, n, ty = np.ascontiguousarray(np.concatenate([t, v])), t.size, 3
        else:
            flat, n, ty = np.ascontiguousarray(spec, np.float64), 0, 1
            assert flat.size == 7
        rc = _hlib().csimh_set_source_waveform(C.c_void_p(self._h), int(idx), ty, flat.ctypes.data_as(C.c_void_p), int(n))
        if rc:
            raise capi.CsimError("csimh_set_source_waveform failed")

    def multi(self, B: int, devices) -> "capi.Multi":
        """csim_multi over `devices` for B instances of this netlist."""
        return capi.Multi((self.n_unknowns, self.n_node_eq, self.elem_type, self.elem_flags, self.elem_nodes,
                           self.elem_branch, self.elem_param, self.n_params), B, devices)

    def source_tables(self, dt: float | None = None, n_steps: int | None = None):
        """(src_mode, src_group, table[n_groups, n_steps]) tabulated with the host libm."""
        dt = self.tran["tstep"] if dt is None else dt
        n_steps = self.n_steps if n_steps is None else n_steps
        ns = max(self.n_sources, 1)
        mode = np.zeros(ns, np.int32)
        group = np.zeros(ns, np.int32)
        table = np.zeros((ns, max(n_steps, 1)), np.float64)
        ng = C.c_int32()
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        _hlib().csimh_tabulate_sources(C.c_void_p(self._h), float(dt), int(n_steps), vp(mode), vp(group), vp(table),
                                       C.byref(ng))
        return mode[: self.n_sources], group[: self.n_sources], np.ascontiguousarray(table[: max(ng.value, 1)])

    def tran_setup(self, probes=None, out_stride: int = 1, n_steps: int | None = None) -> capi.TranSetup:
        n_steps = self.n_steps if n_steps is None else n_steps
        dt = self.tran["tstep"]
        mode, group, table = self.source_tables(dt, n_steps)
        first = 0
        while first < n_steps and (first + 1) * dt < self.tran["tstart"]:
            first += 1
        probes = self.plot_probes if probes is None else probes
        return capi.TranSetup(dt, n_steps, mode, group, table, probes, out_stride=out_stride, out_begin_step=first)

    def monte_carlo(self, B: int, seed: int, rel: float = 0.05, first: int = 0) -> np.ndarray:
        """Instance parameters [P, B] (instance-minor): global instance 0 is nominal, instance i>0 is
        nominal*(1+rel*u_i), u_i~U(-1,1) drawn row-wise from PCG64(seed); applied to R, C, L and the MOS
        Vth/K/lambda/Cj0; sources stay nominal (SURVEY.md §8(d) recipe).  `first` selects a shard
        [first, first+B) of one global batch, so ranks draw disjoint, reproducible slices."""
        P = self.n_params
        mask = np.zeros(P, bool)
        for e in range(self.n_elems):
            ty = int(self.elem_type[e])
            if ty in (capi.ELEM_R, capi.ELEM_C, capi.ELEM_L, capi.ELEM_MOS):
                mask[self.elem_param[e]: self.elem_param[e] + capi.PARAM_COUNT[ty]] = True
        rng = np.random.Generator(np.random.PCG64(seed))
        if first:
            rng.bit_generator.advance(first * P)
        u = rng.uniform(-1.0, 1.0, size=(B, P))
        out = np.repeat(self.params[None, :], B, 0)
        lo = 1 if first == 0 else 0
        out[lo:, mask] = self.params[mask] * (1.0 + rel * u[lo:, mask])
        return np.ascontiguousarray(out.T)

    def monte_carlo_host(self, B: int, seed: int, rel: float = 0.05, first: int = 0) -> np.ndarray:
        """The same batch as monte_carlo(), drawn by the C++ host layer (csim::monteCarloParams: numpy's SeedSequence +
        PCG64 restated) — what a C++ caller of the `.MC` card gets."""
        out = np.empty((self.n_params, B), np.float64)
        L = _hlib()
        L.csimh_monte_carlo.argtypes = [C.c_void_p, C.c_int64, C.c_uint64, C.c_double, C.c_int64, C.c_void_p]
        if L.csimh_monte_carlo(C.c_void_p(self._h), int(B), int(seed), float(rel), int(first), out.ctypes.data_as(C.c_void_p)):
            raise capi.CsimError("csimh_monte_carlo: " + L.csimh_last_error().decode(errors="replace"))
        return out

    def mc_card(self):
        """The netlist's `.MC instances [seed [rel]]` card, or None."""
        n, sd, rel = C.c_int64(), C.c_uint64(), C.c_double()
        L = _hlib()
        L.csimh_mc_cfg.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        if L.csimh_mc_cfg(C.c_void_p(self._h), C.byref(n), C.byref(sd), C.byref(rel)) != 1:
            return None
        return dict(instances=int(n.value), seed=int(sd.value), rel=float(rel.value))

    def run_mc_card(self, probes=None, out_stride: int = 1, devices=None) -> dict:
        """csim::runMonteCarloTransientBatch: the netlist's `.MC` card as one batched DC + full .TRAN (C++ host layer)."""
        card = self.mc_card()
        if card is None:
            raise capi.CsimError("the netlist has no .MC card")
        B, N = card["instances"], self.n_unknowns
        probes = np.ascontiguousarray(self.plot_probes if probes is None else probes, np.int32)
        n_out = self.n_steps // out_stride
        out = dict(x_final=np.empty((N, B)), wave=np.empty((B, probes.size, max(n_out, 1))), iters_total=np.empty(B, np.int32),
                   hash=np.empty(B, np.uint64), status=np.empty(B, np.uint32))
        L = _hlib()
        L.csimh_run_mc_tran_devices.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32] + [C.c_void_p] * 5
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        dev = np.ascontiguousarray(devices if devices is not None else [], np.int32)
        rc = L.csimh_run_mc_tran_devices(C.c_void_p(self._h), int(dev.size), vp(dev) if dev.size else None, int(probes.size),
                                         vp(probes), int(out_stride), vp(out["x_final"]), vp(out["wave"]),
                                         vp(out["iters_total"]), vp(out["hash"]), vp(out["status"]))
        if rc < 0:
            raise capi.CsimError("csimh_run_mc_tran: " + L.csimh_last_error().decode(errors="replace"))
        assert rc == n_out or self.tran["tstart"] > 0
        out["wave"] = out["wave"].reshape(-1)[: B * probes.size * rc].reshape(B, probes.size, rc)
        return out

    def sin_source_sweep(self, source_name: str, va_values, freq_values, n_steps: int, dt: float | None = None,
                         out_stride: int = 1, probes=None):
        """Sweep of one SIN source over amplitude x frequency (BASELINE config 4): instance b = ia*len(freq)+ifr
        gets va = va_values[ia], freq = freq_values[ifr]; everything else nominal.  Frequencies need their own
        host-tabulated sin rows, selected per instance through src_group_b.  Returns (params [P,B], TranSetup)."""
        dt = self.tran["tstep"] if dt is None else dt
        src = self.source_elems()
        e = [k for k in src if self.elem_names[k].lower() == source_name.lower()]
        if not e or not (self.elem_flags[e[0]] & capi.FLAG_SRC_SIN):
            raise ValueError(f"{source_name}: not a SIN source of this netlist")
        e = e[0]
        s_idx = src.index(e)
        po = int(self.elem_param[e])
        va_values = np.asarray(va_values, np.float64)
        freq_values = np.asarray(freq_values, np.float64)
        B = va_values.size * freq_values.size
        P = np.repeat(self.params[:, None], B, 1)
        P[po + 2] = np.repeat(va_values, freq_values.size)
        P[po + 3] = np.tile(freq_values, va_values.size)
        mode, group, table = self.source_tables(dt, n_steps)
        td, phi = self.params[po + 4], self.params[po + 5]
        extra = np.stack([sin_table(f, td, phi, dt, n_steps) for f in freq_values])
        g0 = table.shape[0]
        table = np.ascontiguousarray(np.concatenate([table, extra], 0))
        group_b = np.repeat(group[:, None].astype(np.int32), B, 1)
        group_b[s_idx] = g0 + np.tile(np.arange(freq_values.size, dtype=np.int32), va_values.size)
        probes = self.plot_probes if probes is None else probes
        setup = capi.TranSetup(dt, n_steps, mode, group, table, probes, out_stride=out_stride,
                               src_group_b=np.ascontiguousarray(group_b))
        return np.ascontiguousarray(P), setup

    def steady_state(self, params, f0: float = 0.0, tstep: float = 0.0, max_periods: int = 20, tol: float = 1e-6):
        """runSteadyStateAnalysisBatch of the C++ host layer: params [P, B] -> dict of numpy arrays."""
        L = _hlib()
        params = np.ascontiguousarray(params, np.float64)
        B = params.shape[1]
        N = self.n_unknowns
        out = dict(x_dc=np.empty((N, B)), x=np.empty((N, B)), periods=np.empty(B, np.int32), delta=np.empty(B),
                   iters_total=np.empty(B, np.int32))
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        L.csimh_run_sstate.restype = C.c_int
        L.csimh_run_sstate.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_double, C.c_double, C.c_int32,
                                       C.c_double] + [C.c_void_p] * 5
        spp = L.csimh_run_sstate(C.c_void_p(self._h), B, vp(params), float(f0), float(tstep), int(max_periods),
                                 float(tol), vp(out["x_dc"]), vp(out["x"]), vp(out["periods"]), vp(out["delta"]),
                                 vp(out["iters_total"]))
        if spp < 0:
            raise capi.CsimError(L.csimh_last_error().decode(errors="replace"))
        out["steps_per_period"] = spp
        return out

    def csv_columns(self) -> np.ndarray:
        """Equation indices in the reference's CSV column order (node rows, then V / L branch currents)."""
        cols = list(range(self.n_node_eq))
        cols += [int(self.elem_branch[e]) for e in range(self.n_elems)
                 if int(self.elem_type[e]) in (capi.ELEM_V, capi.ELEM_L)]
        return np.asarray(cols, np.int32)

    def write_instance_csv(self, path: str, instance: int, setup: "capi.TranSetup", wave, x_dc=None, status=None):
        """One instance of a batched transient in the reference's CSV layout (src/tanalisis.cpp:183-231): `wave`
        [B][n_probes][n_out] and optionally `x_dc` [N][B] (the t = 0 row) and `status` [B], host arrays (numpy).
        With setup.probes == csv_columns() and out_stride 1 the file is what the reference writes for that instance."""
        wave = np.ascontiguousarray(wave, np.float64)
        B, n_probes, n_out = wave.shape
        probes = np.ascontiguousarray(setup.probes, np.int32)
        assert probes.size == n_probes
        vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
        xd = None if x_dc is None else np.ascontiguousarray(x_dc, np.float64)
        stt = None if status is None else np.ascontiguousarray(status, np.uint32)
        L = _hlib()
        if L.csimh_write_instance_csv(C.c_void_p(self._h), path.encode(), B, int(instance), n_probes, vp(probes),
                                      int(setup.n_steps), n_out, int(setup.out_stride), vp(xd), vp(wave), vp(stt)) != 0:
            raise capi.CsimError("csimh_write_instance_csv: " + L.csimh_last_error().decode(errors="replace"))

    @property
    def n_dc_sweeps(self) -> int:
        return int(_hlib().csimh_n_dc_sweeps(C.c_void_p(self._h)))

    def dc_sweep(self, index: int = 0, base=None) -> dict:
        """The netlist's index-th `.DC src start stop step` card as one batch of DC operating points on the GPU
        (runDcSweepBatch, include/csim/batchio.hpp): values [K], x [N][K], iters [ramp][K], status [K]."""
        L = _hlib()
        h = C.c_void_p(self._h)
        K = L.csimh_run_dc_sweep(h, int(index), None, 0, None, None, None, None)
        if K < 0:
            raise capi.CsimError("csimh_run_dc_sweep: " + L.csimh_last_error().decode(errors="replace"))
        ramp = capi.default_constants().dc_ramp_steps
        out = {"values": np.zeros(K, np.float64), "x": np.zeros((self.n_unknowns, K), np.float64),
               "iters": np.zeros((ramp, K), np.int32), "status": np.zeros(K, np.uint32)}
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        b = None if base is None else np.ascontiguousarray(base, np.float64)
        if L.csimh_run_dc_sweep(h, int(index), None if b is None else vp(b), K, vp(out["values"]), vp(out["x"]),
                                vp(out["iters"]), vp(out["status"])) != K:
            raise capi.CsimError("csimh_run_dc_sweep: " + L.csimh_last_error().decode(errors="replace"))
        return out

    def shooting(self, params, f0=None, tstep=None, max_iters: int = 10, tol: float = 1e-9, rel_step: float = 1e-6):
        """runSteadyStateShootingBatch of the C++ host layer: Newton shooting on Phi_T(x) = x for C circuits
        (params [P, C]); every iteration integrates C * (N + 1) instances for one period on the GPU."""
        L = _hlib()
        L.csimh_run_shooting.restype = C.c_int
        L.csimh_run_shooting.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_double, C.c_double, C.c_int32, C.c_double,
                                         C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        params = np.ascontiguousarray(params, np.float64)
        Cn = params.shape[1]
        out = {"x_dc": np.zeros((self.n_unknowns, Cn)), "x": np.zeros((self.n_unknowns, Cn)),
               "iterations": np.zeros(Cn, np.int32), "residual": np.zeros(Cn)}
        n_inst = C.c_int64(0)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        spp = L.csimh_run_shooting(C.c_void_p(self._h), Cn, vp(params), float(f0 or 0.0), float(tstep or 0.0),
                                   int(max_iters), float(tol), float(rel_step), vp(out["x_dc"]), vp(out["x"]),
                                   vp(out["iterations"]), vp(out["residual"]), C.byref(n_inst))
        if spp < 0:
            raise capi.CsimError(L.csimh_last_error().decode(errors="replace"))
        out["steps_per_period"] = spp
        out["transient_instances"] = int(n_inst.value)
        return out

    def dc_gauss_seidel(self) -> np.ndarray:
        """dcSolveGaussSeidel of the C++ host layer (the reference's Gauss-Seidel DC driver) on the nominal circuit."""
        x = np.zeros(self.n_unknowns, np.float64)
        L = _hlib()
        L.csimh_dc_gauss_seidel.argtypes = [C.c_void_p, C.c_void_p]
        if L.csimh_dc_gauss_seidel(C.c_void_p(self._h), x.ctypes.data_as(C.c_void_p)):
            raise capi.CsimError("csimh_dc_gauss_seidel: " + L.csimh_last_error().decode(errors="replace"))
        return x

    def run_reference_style_csv(self, path: str):
        """runTransientAnalysisBackwardEuler(ckt, sim, path) of the C++ host layer (B = 1 on the GPU)."""
        L = _hlib()
        if L.csimh_run_tran_csv(C.c_void_p(self._h), path.encode()) != 0:
            raise capi.CsimError(L.csimh_last_error().decode(errors="replace"))


def lu_decompose(A):
    """Solver::luDecompose of the C++ host layer: (LU packed as the reference leaves it, perm, ok)."""
    A = np.ascontiguousarray(A, np.float64)
    n = A.shape[0]
    LU = np.zeros((n, n), np.float64)
    perm = np.zeros(n, np.int32)
    L = _hlib()
    L.csimh_lu_decompose.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    rc = L.csimh_lu_decompose(n, A.ctypes.data_as(C.c_void_p), LU.ctypes.data_as(C.c_void_p), perm.ctypes.data_as(C.c_void_p))
    if rc < 0:
        raise capi.CsimError("csimh_lu_decompose: " + L.csimh_last_error().decode(errors="replace"))
    return LU, perm, bool(rc)


def ladder_netlist(stages: int = 200, tstep: float = 1e-11, tstop: float = 2e-9, vin_offset: float = 1.5,
                   vin_ampl: float = 1.5) -> str:
    """BASELINE config 5: a chain of `stages` CMOS inverters, each driving an R-C section,
    written in the reference's netlist grammar.  Nodes: vdd, n0 (input), then o_i / n_{i+1} per stage;
    N = 2*stages + 2 node equations + 2 branch equations (404 for 200 stages).
    vin_offset = 1.5 is the survey's netlist: the input sits at mid-rail at DC, every stage is at its high-gain
    point and the reference's own fixed-damping DC ramp does not converge on it (status bits DC_UNCONVERGED,
    LU_SINGULAR; reproduced by the oracle).  vin_offset = 0 biases the input at a rail: the stages alternate between the
    rails and the same ramp converges — the variant reported next to it."""
    lines = ["* generated inverter / RC ladder", "VDD vdd 0 DC 3", f"Vin n0 0 SIN {vin_offset:g} {vin_ampl:g} 10e6 0"]
    for i in range(stages):
        lines += [f"MP{i} o{i} n{i} vdd p 30e-6 0.35e-6 1", f"MN{i} o{i} n{i} 0 n 10e-6 0.35e-6 2",
                  f"R{i} o{i} n{i + 1} 100", f"C{i} n{i + 1} 0 10e-15"]
    lines += [".MODEL 1 VT -0.75 MU 5e-2 COX 0.3e-4 LAMBDA 0.05 CJ0 4.0e-14",
              ".MODEL 2 VT 0.83 MU 1.5e-1 COX 0.3e-4 LAMBDA 0.05 CJ0 4.0e-14",
              f".PLOTNV n{stages}", f".TRAN {tstep:g} {tstop:g}"]
    return "\n".join(lines) + "\n"


def sin_table(freq, td, phi, dt, n_steps) -> np.ndarray:
    out = np.zeros(n_steps, np.float64)
    _hlib().csimh_sin_table(float(freq), float(td), float(phi), float(dt), int(n_steps), out.ctypes.data_as(C.c_void_p))
    return out
