// facade.cpp — extern "C" window onto the C++ host layer for Python (ctypes): parse a netlist
// with the product's own parser, lower it, tabulate sources.  Declared in include/csim_host.h.
#include <cmath>
#include <cstring>
#include <memory>
#include <sstream>
#include <string>

#include "../../include/csim_host.h"
#include <algorithm>

#include "csim/batchio.hpp"
#include "csim/dcanalysis.hpp"
#include "csim/solver.hpp"
#include "csim/lower.hpp"
#include "csim/montecarlo.hpp"
#include "csim/sstateanalysis.hpp"
#include "csim/parser.hpp"
#include "csim/tanalisis.hpp"

struct csimh_netlist {
    Circuit ckt;
    SimulationConfig sim;
    csim::Lowered low;
    std::string header;
};

namespace {
thread_local std::string g_herr;
void copy_str(const std::string& s, char* out, int cap) {
    if (!out || cap <= 0) return;
    std::snprintf(out, static_cast<std::size_t>(cap), "%s", s.c_str());
}
}  // namespace

extern "C" {

const char* csimh_last_error(void) { return g_herr.c_str(); }

csimh_netlist* csimh_parse(const char* text) {
    if (!text) return nullptr;
    try {
        std::unique_ptr<csimh_netlist> h(new csimh_netlist());
        std::istringstream in(text);
        NetlistParser parser(h->ckt, h->sim);
        parser.parseStream(in, "<memory>");
        h->sim.ensureDefaultOp();
        h->ckt.assignEquationIndices();
        h->low = csim::lower(h->ckt);
        h->header = h->low.csv_header();
        return h.release();
    } catch (const std::exception& e) {
        g_herr = e.what();
        return nullptr;
    }
}

void csimh_free(csimh_netlist* h) { delete h; }

// PULSE (type 1: spec = v1 v2 td tr tf ton per) or PWL (type 3: spec = n times then n values) waveform on the
// idx-th independent source (V and I in netlist order) — TranWaveform's other two kinds, which the reference
// implements (include/sim.hpp:80-138) but its parser never produces (src/parser.cpp:346-351).  Re-lowers.
int csimh_set_source_waveform(csimh_netlist* h, int32_t idx, int32_t type, const double* spec, int32_t n) {
    if (!h || !spec || (type != 1 && type != 3) || (type == 3 && n <= 0)) return -1;
    int s = 0;
    for (const auto& e : h->ckt.elements) {
        auto* v = dynamic_cast<VoltageSource*>(e.get());
        auto* c = dynamic_cast<CurrentSource*>(e.get());
        if (!v && !c) continue;
        if (s++ != idx) continue;
        SourceSpec sp = v ? v->getSpec() : c->getSpec();
        if (type == 1) {
            sp.tran.type = WaveformType::PULSE;
            sp.tran.pulse = PulseSpec{spec[0], spec[1], spec[2], spec[3], spec[4], spec[5], spec[6]};
        } else {
            sp.tran.type = WaveformType::PWL;
            sp.tran.pwl.t.assign(spec, spec + n);
            sp.tran.pwl.v.assign(spec + n, spec + 2 * n);
        }
        if (v) v->replaceSpec(sp);
        else c->replaceSpec(sp);
        h->low = csim::lower(h->ckt);
        return 0;
    }
    g_herr = "csimh_set_source_waveform: no such source";
    return -1;
}

int csimh_dims(const csimh_netlist* h, int32_t* n_unknowns, int32_t* n_node_eq, int32_t* n_elems, int32_t* n_params,
               int32_t* n_sources) {
    if (!h) return -1;
    if (n_unknowns) *n_unknowns = h->low.n_unknowns;
    if (n_node_eq) *n_node_eq = h->low.n_node_eq;
    if (n_elems) *n_elems = static_cast<int32_t>(h->low.elem_type.size());
    if (n_params) *n_params = h->low.n_params;
    if (n_sources) *n_sources = static_cast<int32_t>(h->low.source_elems.size());
    return 0;
}

int csimh_topology(const csimh_netlist* h, int32_t* type, int32_t* flags, int32_t* nodes, int32_t* branch,
                   int32_t* param_off, double* nominal) {
    if (!h) return -1;
    const csim::Lowered& L = h->low;
    const std::size_t ne = L.elem_type.size();
    std::memcpy(type, L.elem_type.data(), ne * 4);
    std::memcpy(flags, L.elem_flags.data(), ne * 4);
    std::memcpy(nodes, L.elem_nodes.data(), ne * 16);
    std::memcpy(branch, L.elem_branch.data(), ne * 4);
    std::memcpy(param_off, L.elem_param.data(), ne * 4);
    std::memcpy(nominal, L.nominal.data(), L.nominal.size() * 8);
    return 0;
}

int csimh_node_name(const csimh_netlist* h, int32_t eq, char* out, int32_t cap) {
    if (!h || eq < 0 || eq >= h->low.n_node_eq) return -1;
    copy_str(h->low.node_names[eq], out, cap);
    return 0;
}
int csimh_elem_name(const csimh_netlist* h, int32_t idx, char* out, int32_t cap) {
    if (!h || idx < 0 || idx >= static_cast<int32_t>(h->low.elem_names.size())) return -1;
    copy_str(h->low.elem_names[idx], out, cap);
    return 0;
}
int csimh_csv_header(const csimh_netlist* h, char* out, int32_t cap) {
    if (!h) return -1;
    copy_str(h->header, out, cap);
    return static_cast<int>(h->header.size());
}

int csimh_tran_cfg(const csimh_netlist* h, double* tstep, double* tstop, double* tstart, int32_t* enabled) {
    if (!h) return -1;
    *tstep = h->sim.tran.tstep;
    *tstop = h->sim.tran.tstop;
    *tstart = h->sim.tran.tstart;
    *enabled = h->sim.tran.enabled ? 1 : 0;
    return 0;
}
int csimh_hb_cfg(const csimh_netlist* h, double* f0, int32_t* n_harm, int32_t* enabled) {
    if (!h) return -1;
    *f0 = h->sim.hb.f0;
    *n_harm = h->sim.hb.nHarm;
    *enabled = h->sim.hb.enabled ? 1 : 0;
    return 0;
}

// equation indices of the .PLOTNV / .PRINT V(node) probes, in card order; returns the count
int csimh_plot_probes(const csimh_netlist* h, int32_t* out, int32_t cap) {
    if (!h) return -1;
    int n = 0;
    for (const PrintCommand& pc : h->sim.printCommands)
        for (const ProbeSpec& p : pc.probes) {
            if (p.kind != ProbeKind::NodeVoltage) continue;
            auto it = h->ckt.nodeNameToId.find(p.node1);
            if (it == h->ckt.nodeNameToId.end()) continue;
            const int eq = h->ckt.nodes[it->second].eqIndex;
            if (eq < 0) continue;
            if (out && n < cap) out[n] = eq;
            ++n;
        }
    return n;
}

// Host tabulation of the independent sources (host libm), csim_src_mode convention.
// mode/group: [n_sources]; table: [n_sources][n_steps] capacity (n_groups <= n_sources rows used).
int csimh_tabulate_sources(const csimh_netlist* h, double dt, int32_t n_steps, int32_t* mode, int32_t* group,
                           double* table, int32_t* n_groups) {
    if (!h) return -1;
    csim::SourceTables T = csim::tabulate_sources(h->ckt, dt, n_steps);
    std::memcpy(mode, T.mode.data(), T.mode.size() * 4);
    std::memcpy(group, T.group.data(), T.group.size() * 4);
    std::memcpy(table, T.table.data(), T.table.size() * 8);
    *n_groups = T.n_groups;
    return 0;
}

// S[s] = sin(2*pi*freq*(t-td)+phi), t=(s+1)*dt — one table row for a swept source
int csimh_sin_table(double freq, double td, double phi, double dt, int32_t n_steps, double* out) {
    for (int32_t s = 0; s < n_steps; ++s) {
        const double t = (s + 1) * dt;
        const double tau = t - td;
        const double w = 2.0 * M_PI * freq;
        out[s] = std::sin(w * tau + phi);
    }
    return 0;
}

// `.MC instances [seed [rel]]` card (include/csim/montecarlo.hpp); returns 1 when present
int csimh_mc_cfg(const csimh_netlist* h, int64_t* instances, uint64_t* seed, double* rel) {
    if (!h) return -1;
    if (instances) *instances = h->sim.mc.instances;
    if (seed) *seed = h->sim.mc.seed;
    if (rel) *rel = h->sim.mc.rel;
    return h->sim.mc.enabled ? 1 : 0;
}

// csim::monteCarloParams: out [P][B] instance-minor, instances [first, first + B) of the batch drawn from `seed`
int csimh_monte_carlo(const csimh_netlist* h, int64_t B, uint64_t seed, double rel, int64_t first, double* out) {
    if (!h || !out) return -1;
    try {
        const std::vector<double> p = csim::monteCarloParams(h->low, B, seed, rel, first);
        std::memcpy(out, p.data(), p.size() * sizeof(double));
        return 0;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// csim::runMonteCarloTransientBatch: the `.MC` card as one batched DC + transient.  Outputs sized by the caller from
// csimh_mc_cfg (B) and the .TRAN card: x_final [N][B], wave [B][n_probes][n_out], iters_total / hash / status [B].
// Returns n_out (recorded points per probe) or -1.
int csimh_run_mc_tran(const csimh_netlist* h, int32_t n_probes, const int32_t* probes, int32_t out_stride,
                      double* x_final, double* wave, int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status) {
    return csimh_run_mc_tran_devices(h, 0, nullptr, n_probes, probes, out_stride, x_final, wave, iters_total, pivot_hash,
                                     status);
}

// ... n_devices > 0: sharded over those CUDA devices (csim::runMonteCarloTransientMultiGpu)
int csimh_run_mc_tran_devices(const csimh_netlist* h, int32_t n_devices, const int32_t* devices, int32_t n_probes,
                              const int32_t* probes, int32_t out_stride, double* x_final, double* wave,
                              int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status) {
    if (!h || (n_probes > 0 && !probes) || (n_devices > 0 && !devices)) return -1;
    try {
        const std::vector<int32_t> pr(probes, probes + n_probes);
        const csim::BatchResult R =
            n_devices > 0 ? csim::runMonteCarloTransientMultiGpu(h->ckt, h->sim, pr, out_stride,
                                                                 std::vector<int32_t>(devices, devices + n_devices))
                          : csim::runMonteCarloTransientBatch(h->ckt, h->sim, pr, out_stride);
        if (x_final) std::memcpy(x_final, R.x_final.data(), R.x_final.size() * sizeof(double));
        if (wave) std::memcpy(wave, R.wave.data(), R.wave.size() * sizeof(double));
        if (iters_total) std::memcpy(iters_total, R.iters_total.data(), R.iters_total.size() * sizeof(int32_t));
        if (pivot_hash) std::memcpy(pivot_hash, R.pivot_hash.data(), R.pivot_hash.size() * sizeof(uint64_t));
        if (status) std::memcpy(status, R.status.data(), R.status.size() * sizeof(uint32_t));
        return R.n_out;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// dcSolveGaussSeidel of the host layer (src/dcanalysis.cpp:251-257) on the nominal circuit; x: [N].
int csimh_dc_gauss_seidel(const csimh_netlist* h, double* x) {
    if (!h || !x) return -1;
    try {
        const Eigen::VectorXd sol = dcSolveGaussSeidel(h->ckt);
        for (int i = 0; i < h->low.n_unknowns; ++i) x[i] = sol(i);
        return 0;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// Solver::luDecompose of the host layer on one row-major n x n matrix: LU [n*n], perm [n]; returns 1 (true),
// 0 (false: singular) or -1.
int csimh_lu_decompose(int32_t n, const double* A, double* LU, int32_t* perm) {
    if (n <= 0 || !A || !LU || !perm) return -1;
    try {
        Eigen::MatrixXd Am(n, n), F;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) Am(i, j) = A[static_cast<std::size_t>(i) * n + j];
        std::vector<int> p;
        const bool ok = Solver::luDecompose(Am, F, p);
        for (int i = 0; i < n && i < (int)p.size(); ++i) perm[i] = p[i];
        if (F.rows() == n)
            for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j) LU[static_cast<std::size_t>(i) * n + j] = F(i, j);
        return ok ? 1 : 0;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// The reference's own driver signature, for drop-in checks from Python: DC + transient of the
// nominal circuit, CSV written to `path`.  Returns 0, or -1 with csimh_last_error().
int csimh_run_tran_csv(const csimh_netlist* h, const char* path) {
    if (!h || !path) return -1;
    try {
        runTransientAnalysisBackwardEuler(h->ckt, h->sim, path);
        return 0;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// runSteadyStateAnalysisBatch of the C++ host layer (include/csim/sstateanalysis.hpp); f0 / tstep > 0
// override the netlist's `.hb` / `.TRAN` values.  Arrays as documented there; returns steps per period
// (> 0) or -1 with csimh_last_error().
int csimh_run_sstate(const csimh_netlist* h, int64_t B, const double* params, double f0, double tstep,
                     int32_t max_periods, double tol, double* x_dc, double* x, int32_t* periods, double* delta,
                     int32_t* iters_total) {
    if (!h || !params || B <= 0) return -1;
    try {
        SimulationConfig sim = h->sim;
        if (f0 > 0.0) {
            sim.hb.enabled = true;
            sim.hb.f0 = f0;
        }
        if (tstep > 0.0) sim.tran.tstep = tstep;
        const std::vector<double> p(params, params + (size_t)h->low.n_params * B);
        csim::SteadyStateResult R = csim::runSteadyStateAnalysisBatch(h->ckt, sim, p, B, max_periods, tol);
        auto put = [](auto& v, auto* out) {
            if (out) std::copy(v.begin(), v.end(), out);
        };
        put(R.x_dc, x_dc);
        put(R.x, x);
        put(R.periods, periods);
        put(R.delta, delta);
        put(R.iters_total, iters_total);
        return R.steps_per_period;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// runSteadyStateShootingBatch (include/csim/sstateanalysis.hpp); f0 / tstep > 0 override `.hb` / `.TRAN`.
// x_dc, x: [N][C]; iterations [C]; residual [C].  Returns steps per period (> 0) or -1 with csimh_last_error().
int csimh_run_shooting(const csimh_netlist* h, int64_t C, const double* params, double f0, double tstep,
                       int32_t max_iters, double tol, double rel_step, double* x_dc, double* x, int32_t* iterations,
                       double* residual, int64_t* transient_instances) {
    if (!h || !params || C <= 0) return -1;
    try {
        SimulationConfig sim = h->sim;
        if (f0 > 0.0) {
            sim.hb.enabled = true;
            sim.hb.f0 = f0;
        }
        if (tstep > 0.0) sim.tran.tstep = tstep;
        std::vector<double> p(params, params + static_cast<std::size_t>(h->low.n_params) * C);
        csim::ShootingResult R = csim::runSteadyStateShootingBatch(h->ckt, sim, p, C, max_iters, tol, rel_step);
        if (x_dc) std::memcpy(x_dc, R.x_dc.data(), R.x_dc.size() * 8);
        if (x) std::memcpy(x, R.x.data(), R.x.size() * 8);
        if (iterations) std::memcpy(iterations, R.iterations.data(), R.iterations.size() * 4);
        if (residual) std::memcpy(residual, R.residual.data(), R.residual.size() * 8);
        if (transient_instances) *transient_instances = R.transient_instances;
        return R.steps_per_period;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

int csimh_write_instance_csv(const csimh_netlist* h, const char* path, int64_t B, int64_t instance, int32_t n_probes,
                             const int32_t* probes, int32_t n_steps, int32_t n_out, int32_t out_stride,
                             const double* x_dc, const double* wave, const uint32_t* status) {
    if (!h || !path || !wave || (n_probes > 0 && !probes) || instance < 0 || instance >= B) return -1;
    try {
        csim::BatchResult R;  // a one-instance view of the caller's batch
        R.B = 1;
        R.n_steps = n_steps;
        R.n_probes = n_probes;
        R.n_out = n_out;
        if (x_dc) {
            R.x_dc.resize(static_cast<std::size_t>(h->low.n_unknowns));
            for (int32_t i = 0; i < h->low.n_unknowns; ++i)
                R.x_dc[static_cast<std::size_t>(i)] = x_dc[static_cast<int64_t>(i) * B + instance];
        }
        const double* w = wave + static_cast<std::size_t>(instance) * n_probes * n_out;
        R.wave.assign(w, w + static_cast<std::size_t>(n_probes) * n_out);
        R.status.assign(1, status ? status[instance] : 0u);
        csim::writeInstanceCsv(h->low, h->sim, R, std::vector<int32_t>(probes, probes + n_probes), out_stride, 0, path);
        return 0;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

// cos/sin(2*pi*n/M), n = 0..M-1, host libm (csim_dev_harmonics' tables)
int csimh_twiddle_table(int32_t M, double* cos_out, double* sin_out) {
    if (M <= 0 || !cos_out || !sin_out) return -1;
    for (int32_t n = 0; n < M; ++n) {
        const double a = 2.0 * M_PI * static_cast<double>(n) / static_cast<double>(M);
        cos_out[n] = std::cos(a);
        sin_out[n] = std::sin(a);
    }
    return 0;
}

int csimh_n_dc_sweeps(const csimh_netlist* h) { return h ? static_cast<int>(h->sim.dcSweeps.size()) : -1; }

int csimh_run_dc_sweep(const csimh_netlist* h, int32_t index, const double* base, int32_t cap_points, double* values,
                       double* x, int32_t* iters, uint32_t* status) {
    if (!h || index < 0 || index >= static_cast<int32_t>(h->sim.dcSweeps.size())) {
        g_herr = "csimh_run_dc_sweep: no such .DC card";
        return -1;
    }
    try {
        const DCSweepConfig& sw = h->sim.dcSweeps[static_cast<std::size_t>(index)];
        if (!values) {  // size query
            if (!(sw.step != 0.0) || (sw.stop - sw.start) / sw.step < 0.0) throw std::runtime_error("invalid .DC range");
            return static_cast<int>(std::floor((sw.stop - sw.start) / sw.step + 1e-9)) + 1;
        }
        std::vector<double> b = base ? std::vector<double>(base, base + h->low.n_params) : h->low.nominal;
        csim::DcSweepResult R = csim::runDcSweepBatch(h->ckt, sw, b);
        const int K = static_cast<int>(R.values.size());
        if (K > cap_points) throw std::runtime_error("csimh_run_dc_sweep: output arrays too small");
        std::memcpy(values, R.values.data(), R.values.size() * 8);
        if (x) std::memcpy(x, R.x.data(), R.x.size() * 8);
        if (iters) std::memcpy(iters, R.iters.data(), R.iters.size() * 4);
        if (status) std::memcpy(status, R.status.data(), R.status.size() * 4);
        return K;
    } catch (const std::exception& e) {
        g_herr = e.what();
        return -1;
    }
}

}  // extern "C"
