// lower.cpp — Circuit -> csim_topology lowering, host source tabulation, batched drivers,
// and the device-backed Element::stamp.
#include "csim/lower.hpp"
#include "csim/sstateanalysis.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <stdexcept>

namespace {
[[noreturn]] void raise(const char* what) {
    throw std::runtime_error(std::string(what) + ": " + csim_last_error());
}
struct PlanGuard {
    csim_plan* p = nullptr;
    ~PlanGuard() { csim_plan_destroy(p); }
};
struct BatchGuard {
    csim_batch* b = nullptr;
    ~BatchGuard() { csim_batch_destroy(b); }
};
struct MultiGuard {
    csim_multi* m = nullptr;
    ~MultiGuard() { csim_multi_destroy(m); }
};
void putSource(const SourceSpec& s, std::vector<double>& out) {
    out.insert(out.end(), {s.dcValue, s.tran.sine.v0, s.tran.sine.va, s.tran.sine.freq, s.tran.sine.td,
                           s.tran.sine.phi});
}
}  // namespace

// ---- C-ABI view of each element class ---------------------------------------------------
int Resistor::abiType() const { return CSIM_ELEM_R; }
int CapacitorElement::abiType() const { return CSIM_ELEM_C; }
int Inductor::abiType() const { return CSIM_ELEM_L; }
int VoltageSource::abiType() const { return CSIM_ELEM_V; }
int CurrentSource::abiType() const { return CSIM_ELEM_I; }
int MosfetBase::abiType() const { return CSIM_ELEM_MOS; }
int VoltageSource::abiFlags() const { return spec.tran.type == WaveformType::SIN ? CSIM_FLAG_SRC_SIN : 0; }
int CurrentSource::abiFlags() const { return spec.tran.type == WaveformType::SIN ? CSIM_FLAG_SRC_SIN : 0; }
int MosfetBase::abiFlags() const { return isP ? CSIM_FLAG_PMOS : 0; }
void VoltageSource::abiParams(std::vector<double>& out) const { putSource(spec, out); }
void CurrentSource::abiParams(std::vector<double>& out) const { putSource(spec, out); }

namespace csim {

csim_topology Lowered::view() const {
    csim_topology t;
    t.n_unknowns = n_unknowns;
    t.n_node_eq = n_node_eq;
    t.n_elems = static_cast<int32_t>(elem_type.size());
    t.n_params = n_params;
    t.elem_type = elem_type.data();
    t.elem_flags = elem_flags.data();
    t.elem_nodes = elem_nodes.data();
    t.elem_branch = elem_branch.data();
    t.elem_param = elem_param.data();
    return t;
}

std::vector<int32_t> Lowered::csv_columns() const {
    std::vector<int32_t> cols;
    for (int32_t i = 0; i < n_node_eq; ++i) cols.push_back(i);
    for (std::size_t e = 0; e < elem_type.size(); ++e)
        if (elem_type[e] == CSIM_ELEM_V || elem_type[e] == CSIM_ELEM_L) cols.push_back(elem_branch[e]);
    return cols;
}

std::string Lowered::csv_header() const {
    std::string h = "time";
    for (const auto& n : node_names) h += ",V(" + n + ")";
    for (std::size_t e = 0; e < elem_type.size(); ++e)
        if (elem_type[e] == CSIM_ELEM_V || elem_type[e] == CSIM_ELEM_L) h += ",I(" + elem_names[e] + ")";
    return h;
}

Lowered lower(const Circuit& ckt) {
    Lowered L;
    L.n_node_eq = ckt.numNodeEquations();
    L.n_unknowns = ckt.numUnknowns();
    L.node_names.assign(static_cast<std::size_t>(L.n_node_eq), std::string());
    for (const Node& nd : ckt.nodes) {
        if (isGroundName(nd.name)) continue;
        if (nd.eqIndex < 0 || nd.eqIndex >= L.n_node_eq)
            throw std::runtime_error("lower: node '" + nd.name + "' has no equation index; call "
                                     "Circuit::assignEquationIndices() before any analysis");
        L.node_names[static_cast<std::size_t>(nd.eqIndex)] = nd.name;
    }
    for (std::size_t i = 0; i < ckt.elements.size(); ++i) {
        const Element& e = *ckt.elements[i];
        const int ty = e.abiType();
        L.elem_type.push_back(ty);
        L.elem_flags.push_back(e.abiFlags());
        L.elem_branch.push_back(e.abiBranch());
        L.elem_param.push_back(static_cast<int32_t>(L.nominal.size()));
        e.abiParams(L.nominal);
        const auto& ids = e.getNodeIds();
        for (int k = 0; k < 4; ++k)
            L.elem_nodes.push_back(k < static_cast<int>(ids.size()) ? ckt.nodes[ids[k]].eqIndex : -1);
        L.elem_names.push_back(e.getName());
        if (ty == CSIM_ELEM_V || ty == CSIM_ELEM_I) L.source_elems.push_back(static_cast<int32_t>(i));
        if ((ty == CSIM_ELEM_V || ty == CSIM_ELEM_L) && e.abiBranch() < 0)
            throw std::runtime_error("lower: element '" + e.getName() + "' has no branch equation; call "
                                     "Circuit::assignEquationIndices() before any analysis");
    }
    L.n_params = static_cast<int32_t>(L.nominal.size());
    return L;
}

SourceTables tabulate_sources(const Circuit& ckt, double dt, int32_t n_steps) {
    SourceTables T;
    T.n_steps = n_steps;
    for (const auto& e : ckt.elements) {
        const SourceSpec* sp = nullptr;
        if (auto* v = dynamic_cast<const VoltageSource*>(e.get())) sp = &v->getSpec();
        else if (auto* c = dynamic_cast<const CurrentSource*>(e.get())) sp = &c->getSpec();
        if (!sp) continue;
        const WaveformType wt = sp->tran.type;
        if (wt == WaveformType::NONE) {
            T.mode.push_back(CSIM_SRC_CONST);
            T.group.push_back(0);
            continue;
        }
        T.mode.push_back(wt == WaveformType::SIN ? CSIM_SRC_SIN_TABLE : CSIM_SRC_VAL_TABLE);
        T.group.push_back(T.n_groups++);
        for (int32_t s = 0; s < n_steps; ++s) {
            const double t = (s + 1) * dt;  // reference: src/tanalisis.cpp:256
            double val;
            if (wt == WaveformType::SIN) {  // reference: include/sim.hpp:117-122 (the sin() factor only)
                const double tau = t - sp->tran.sine.td;
                const double w = 2.0 * M_PI * sp->tran.sine.freq;
                val = std::sin(w * tau + sp->tran.sine.phi);
            } else {
                val = sp->tran.eval(t);
            }
            T.table.push_back(val);
        }
    }
    if (T.n_groups == 0) {  // keep the table non-empty for the ABI
        T.n_groups = 1;
        T.table.assign(static_cast<std::size_t>(n_steps > 0 ? n_steps : 1), 0.0);
    }
    return T;
}

std::vector<double> dcSolveBatch(const Circuit& ckt, const std::vector<double>& params, int64_t B, int device) {
    Lowered L = lower(ckt);
    if (static_cast<int64_t>(params.size()) != static_cast<int64_t>(L.n_params) * B)
        throw std::runtime_error("dcSolveBatch: params must hold P*B doubles");
    csim_topology t = L.view();
    PlanGuard pg;
    if (csim_plan_create(&t, nullptr, device, &pg.p)) raise("csim_plan_create");
    BatchGuard bg;
    if (csim_batch_create(pg.p, B, &bg.b)) raise("csim_batch_create");
    if (csim_batch_set_params(bg.b, params.data())) raise("csim_batch_set_params");
    std::vector<double> x(static_cast<std::size_t>(L.n_unknowns) * B);
    if (csim_batch_dc(bg.b, x.data(), nullptr, nullptr, nullptr)) raise("csim_batch_dc");
    return x;
}

BatchResult runTransientAnalysisBackwardEulerBatch(const Circuit& ckt, const SimulationConfig& sim,
                                                   const std::vector<double>& params, int64_t B,
                                                   const std::vector<int32_t>& probes, int out_stride, int device) {
    const TranConfig& cfg = sim.tran;
    if (!cfg.enabled) throw std::runtime_error("Transient analysis is not enabled (.TRAN missing).");
    if (cfg.tstep <= 0.0 || cfg.tstop <= 0.0)
        throw std::runtime_error("Invalid .TRAN card: tstep and tstop must be > 0.");
    Lowered L = lower(ckt);
    if (static_cast<int64_t>(params.size()) != static_cast<int64_t>(L.n_params) * B)
        throw std::runtime_error("runTransientAnalysisBackwardEulerBatch: params must hold P*B doubles");
    const int32_t n_steps = static_cast<int32_t>(std::floor(cfg.tstop / cfg.tstep + 1e-12));  // tanalisis.cpp:238
    SourceTables S = tabulate_sources(ckt, cfg.tstep, n_steps);
    // first step whose time passes the reference's "t < tstart -> skip" filter (tanalisis.cpp:209)
    int32_t first = 0;
    while (first < n_steps && (first + 1) * cfg.tstep < cfg.tstart) ++first;

    csim_topology t = L.view();
    PlanGuard pg;
    if (csim_plan_create(&t, nullptr, device, &pg.p)) raise("csim_plan_create");
    BatchGuard bg;
    if (csim_batch_create(pg.p, B, &bg.b)) raise("csim_batch_create");
    if (csim_batch_set_params(bg.b, params.data())) raise("csim_batch_set_params");

    BatchResult R;
    R.B = B;
    R.n_steps = n_steps;
    R.n_probes = static_cast<int32_t>(probes.size());
    if (out_stride < 1) out_stride = 1;
    R.n_out = n_steps / out_stride - first / out_stride;
    const std::size_t NB = static_cast<std::size_t>(L.n_unknowns) * B;
    R.x_dc.resize(NB);
    R.x_final.resize(NB);
    R.wave.assign(static_cast<std::size_t>(B) * R.n_probes * (R.n_out > 0 ? R.n_out : 1),
                  std::numeric_limits<double>::quiet_NaN());
    R.iters_total.resize(B);
    R.pivot_hash.resize(B);
    R.status.resize(B);

    csim_tran_args a;
    std::memset(&a, 0, sizeof(a));
    a.dt = cfg.tstep;
    a.step_begin = 0;
    a.step_end = n_steps;
    a.n_groups = S.n_groups;
    a.table_steps = n_steps;
    a.src_table = S.table.data();
    a.src_mode = S.mode.data();
    a.src_group = S.group.data();
    a.src_group_b = nullptr;
    a.n_probes = R.n_probes;
    a.probes = probes.data();
    a.out_stride = out_stride;
    a.out_begin_step = first;
    a.wave_ld = R.n_out > 0 ? R.n_out : 1;
    if (csim_batch_tran(bg.b, &a, R.x_dc.data(), R.x_final.data(), R.n_probes ? R.wave.data() : nullptr,
                        R.iters_total.data(), nullptr, R.pivot_hash.data(), R.status.data()))
        raise("csim_batch_tran");
    R.kernel_ms = csim_batch_last_kernel_ms(bg.b);
    return R;
}

// The same batch sharded over several GPUs (csim_multi_*: one host thread, plan, batch and stream per device,
// contiguous shards, every device copies its slice into R's global arrays); bit-identical to the single-device call.
BatchResult runTransientAnalysisBackwardEulerMultiGpu(const Circuit& ckt, const SimulationConfig& sim,
                                                   const std::vector<double>& params, int64_t B,
                                                   const std::vector<int32_t>& probes, int out_stride,
                                                       const std::vector<int32_t>& devices) {
    const TranConfig& cfg = sim.tran;
    if (!cfg.enabled) throw std::runtime_error("Transient analysis is not enabled (.TRAN missing).");
    if (cfg.tstep <= 0.0 || cfg.tstop <= 0.0)
        throw std::runtime_error("Invalid .TRAN card: tstep and tstop must be > 0.");
    Lowered L = lower(ckt);
    if (static_cast<int64_t>(params.size()) != static_cast<int64_t>(L.n_params) * B)
        throw std::runtime_error("runTransientAnalysisBackwardEulerBatch: params must hold P*B doubles");
    const int32_t n_steps = static_cast<int32_t>(std::floor(cfg.tstop / cfg.tstep + 1e-12));  // tanalisis.cpp:238
    SourceTables S = tabulate_sources(ckt, cfg.tstep, n_steps);
    // first step whose time passes the reference's "t < tstart -> skip" filter (tanalisis.cpp:209)
    int32_t first = 0;
    while (first < n_steps && (first + 1) * cfg.tstep < cfg.tstart) ++first;

    csim_topology t = L.view();
    if (devices.empty()) throw std::runtime_error("runTransientAnalysisBackwardEulerMultiGpu: no devices");
    MultiGuard mg;
    if (csim_multi_create(&t, nullptr, static_cast<int32_t>(devices.size()), devices.data(), B, &mg.m)) raise("csim_multi_create");
    if (csim_multi_set_params(mg.m, params.data())) raise("csim_multi_set_params");

    BatchResult R;
    R.B = B;
    R.n_steps = n_steps;
    R.n_probes = static_cast<int32_t>(probes.size());
    if (out_stride < 1) out_stride = 1;
    R.n_out = n_steps / out_stride - first / out_stride;
    const std::size_t NB = static_cast<std::size_t>(L.n_unknowns) * B;
    R.x_dc.resize(NB);
    R.x_final.resize(NB);
    R.wave.assign(static_cast<std::size_t>(B) * R.n_probes * (R.n_out > 0 ? R.n_out : 1),
                  std::numeric_limits<double>::quiet_NaN());
    R.iters_total.resize(B);
    R.pivot_hash.resize(B);
    R.status.resize(B);

    csim_tran_args a;
    std::memset(&a, 0, sizeof(a));
    a.dt = cfg.tstep;
    a.step_begin = 0;
    a.step_end = n_steps;
    a.n_groups = S.n_groups;
    a.table_steps = n_steps;
    a.src_table = S.table.data();
    a.src_mode = S.mode.data();
    a.src_group = S.group.data();
    a.src_group_b = nullptr;
    a.n_probes = R.n_probes;
    a.probes = probes.data();
    a.out_stride = out_stride;
    a.out_begin_step = first;
    a.wave_ld = R.n_out > 0 ? R.n_out : 1;
    if (csim_multi_tran(mg.m, &a, R.x_dc.data(), R.x_final.data(), R.n_probes ? R.wave.data() : nullptr,
                        R.iters_total.data(), R.pivot_hash.data(), R.status.data()))
        raise("csim_multi_tran");
    R.kernel_ms = csim_multi_last_kernel_ms(mg.m);
    return R;
}

SteadyStateResult runSteadyStateAnalysisBatch(const Circuit& ckt, const SimulationConfig& sim,
                                              const std::vector<double>& params, int64_t B, int max_periods,
                                              double tol, int device) {
    if (!sim.hb.enabled || !(sim.hb.f0 > 0.0)) throw std::runtime_error("Steady state: .hb card with f0 > 0 required.");
    if (!(sim.tran.tstep > 0.0)) throw std::runtime_error("Steady state: .TRAN card (tstep) required.");
    Lowered L = lower(ckt);
    if (static_cast<int64_t>(params.size()) != static_cast<int64_t>(L.n_params) * B)
        throw std::runtime_error("runSteadyStateAnalysisBatch: params must hold P*B doubles");
    SteadyStateResult R;
    R.B = B;
    R.steps_per_period = static_cast<int32_t>(std::llround(1.0 / sim.hb.f0 / sim.tran.tstep));
    if (R.steps_per_period < 1) throw std::runtime_error("Steady state: period shorter than one timestep.");
    const int64_t total = static_cast<int64_t>(R.steps_per_period) * max_periods;
    if (total > 2147483647LL) throw std::runtime_error("Steady state: too many steps.");
    SourceTables S = tabulate_sources(ckt, sim.tran.tstep, static_cast<int32_t>(total));
    csim_topology t = L.view();
    PlanGuard pg;
    if (csim_plan_create(&t, nullptr, device, &pg.p)) raise("csim_plan_create");
    BatchGuard bg;
    if (csim_batch_create(pg.p, B, &bg.b)) raise("csim_batch_create");
    if (csim_batch_set_params(bg.b, params.data())) raise("csim_batch_set_params");
    const std::size_t NB = static_cast<std::size_t>(L.n_unknowns) * B;
    R.x_dc.resize(NB);
    R.x.resize(NB);
    R.periods.resize(B);
    R.delta.resize(B);
    R.iters_total.resize(B);
    R.pivot_hash.resize(B);
    R.status.resize(B);
    csim_tran_args a;
    std::memset(&a, 0, sizeof(a));
    a.dt = sim.tran.tstep;
    a.n_groups = S.n_groups;
    a.table_steps = static_cast<int32_t>(total);
    a.src_table = S.table.data();
    a.src_mode = S.mode.data();
    a.src_group = S.group.data();
    csim_sstate_args ss{R.steps_per_period, max_periods, tol};
    if (csim_batch_sstate(bg.b, &a, &ss, R.x_dc.data(), R.x.data(), R.periods.data(), R.delta.data(),
                          R.iters_total.data(), R.pivot_hash.data(), R.status.data()))
        raise("csim_batch_sstate");
    R.kernel_ms = csim_batch_last_kernel_ms(bg.b);
    return R;
}

ShootingResult runSteadyStateShootingBatch(const Circuit& ckt, const SimulationConfig& sim,
                                           const std::vector<double>& params, int64_t C, int max_iters, double tol,
                                           double rel_step, int device) {
    if (!sim.hb.enabled || !(sim.hb.f0 > 0.0)) throw std::runtime_error("Steady state: .hb card with f0 > 0 required.");
    if (!(sim.tran.tstep > 0.0)) throw std::runtime_error("Steady state: .TRAN card (tstep) required.");
    Lowered L = lower(ckt);
    const int N = L.n_unknowns, P = L.n_params;
    if (static_cast<int64_t>(params.size()) != static_cast<int64_t>(P) * C)
        throw std::runtime_error("runSteadyStateShootingBatch: params must hold P*C doubles");
    ShootingResult R;
    R.C = C;
    R.steps_per_period = static_cast<int32_t>(std::llround(1.0 / sim.hb.f0 / sim.tran.tstep));
    if (R.steps_per_period < 1) throw std::runtime_error("Steady state: period shorter than one timestep.");
    SourceTables S = tabulate_sources(ckt, sim.tran.tstep, R.steps_per_period);
    csim_topology t = L.view();
    PlanGuard pg;
    if (csim_plan_create(&t, nullptr, device, &pg.p)) raise("csim_plan_create");
    // operating points of the C circuits
    {
        BatchGuard dc;
        if (csim_batch_create(pg.p, C, &dc.b)) raise("csim_batch_create");
        if (csim_batch_set_params(dc.b, params.data())) raise("csim_batch_set_params");
        R.x_dc.resize(static_cast<std::size_t>(N) * C);
        if (csim_batch_dc(dc.b, R.x_dc.data(), nullptr, nullptr, nullptr)) raise("csim_batch_dc");
    }
    R.x = R.x_dc;
    R.iterations.assign(static_cast<std::size_t>(C), -1);
    R.residual.assign(static_cast<std::size_t>(C), 0.0);
    // the big batch: circuit c owns instances c*(N+1) .. c*(N+1)+N (candidate first, then one column per unknown)
    const int64_t W = N + 1, Bt = C * W;
    std::vector<double> pb(static_cast<std::size_t>(P) * Bt);
    for (int p = 0; p < P; ++p)
        for (int64_t c = 0; c < C; ++c)
            for (int64_t k = 0; k < W; ++k) pb[static_cast<std::size_t>(p) * Bt + c * W + k] = params[static_cast<std::size_t>(p) * C + c];
    BatchGuard bg;
    if (csim_batch_create(pg.p, Bt, &bg.b)) raise("csim_batch_create");
    if (csim_batch_set_params(bg.b, pb.data())) raise("csim_batch_set_params");
    csim_tran_args a;
    std::memset(&a, 0, sizeof(a));
    a.dt = sim.tran.tstep;
    a.step_end = R.steps_per_period;
    a.n_groups = S.n_groups;
    a.table_steps = R.steps_per_period;
    a.src_table = S.table.data();
    a.src_mode = S.mode.data();
    a.src_group = S.group.data();
    a.out_stride = 1;
    std::vector<double> X(static_cast<std::size_t>(N) * Bt), h(static_cast<std::size_t>(N) * C);
    std::vector<double> J(static_cast<std::size_t>(N) * N * C), rhs(static_cast<std::size_t>(N) * C), delta(rhs.size());
    std::vector<uint8_t> open(static_cast<std::size_t>(C), 1);
    for (int it = 0; it <= max_iters; ++it) {
        for (int64_t c = 0; c < C; ++c)
            for (int i = 0; i < N; ++i) {
                const double xi = R.x[static_cast<std::size_t>(i) * C + c];
                h[static_cast<std::size_t>(i) * C + c] = rel_step * std::max(std::fabs(xi), 1.0);
                for (int64_t k = 0; k < W; ++k)
                    X[static_cast<std::size_t>(i) * Bt + c * W + k] = xi + (k == i + 1 ? h[static_cast<std::size_t>(i) * C + c] : 0.0);
            }
        if (csim_batch_tran_from(bg.b, &a, X.data(), nullptr, nullptr, nullptr, nullptr)) raise("csim_batch_tran_from");
        R.transient_instances += Bt;
        bool any_open = false;
        for (int64_t c = 0; c < C; ++c) {
            double ss = 0.0;
            for (int i = 0; i < N; ++i) {
                const double f = X[static_cast<std::size_t>(i) * Bt + c * W] - R.x[static_cast<std::size_t>(i) * C + c];
                rhs[static_cast<std::size_t>(i) * C + c] = -f;
                ss += f * f;
            }
            if (!open[static_cast<std::size_t>(c)]) continue;
            R.residual[static_cast<std::size_t>(c)] = std::sqrt(ss);
            if (R.residual[static_cast<std::size_t>(c)] < tol) {
                R.iterations[static_cast<std::size_t>(c)] = it;
                open[static_cast<std::size_t>(c)] = 0;
            } else {
                any_open = true;
            }
        }
        if (!any_open || it == max_iters) break;
        // (dPhi - I) delta = -F, one system per circuit, solved on the device (csim_host_lu_solve, A [N*N][C])
        for (int64_t c = 0; c < C; ++c)
            for (int i = 0; i < N; ++i)
                for (int j = 0; j < N; ++j) {
                    const double phi0 = X[static_cast<std::size_t>(i) * Bt + c * W];
                    const double phij = X[static_cast<std::size_t>(i) * Bt + c * W + 1 + j];
                    J[(static_cast<std::size_t>(i) * N + j) * C + c] = (phij - phi0) / h[static_cast<std::size_t>(j) * C + c] - (i == j ? 1.0 : 0.0);
                }
        if (csim_host_lu_solve(N, C, J.data(), rhs.data(), delta.data(), nullptr, nullptr, 1e-15)) raise("csim_host_lu_solve");
        for (int64_t c = 0; c < C; ++c)
            if (open[static_cast<std::size_t>(c)])
                for (int i = 0; i < N; ++i) R.x[static_cast<std::size_t>(i) * C + c] += delta[static_cast<std::size_t>(i) * C + c];
    }
    return R;
}

}  // namespace csim

// ---- Element::stamp through the device ---------------------------------------------------
// The element is lowered alone (same equation numbering as `ckt`, N = G.rows()), assembled by
// the CUDA stamp kernel for the analysis type in ctx, and the result ADDED into G / I.
// DC/OP: sources at evalDC(sourceScale); TRAN: sources at evalTran(time); AC/NONE/HB: nothing
// (reference: src/element.cpp:46-58).  Capacitors stamp nothing and inductors stamp their
// 0 V-source form in every analysis type (the reference puts the transient companion models
// in the driver, src/tanalisis.cpp:282-353, not in Element::stamp).
void Element::stamp(Eigen::MatrixXd& G, Eigen::VectorXd& I, const Circuit& ckt, const Eigen::VectorXd& x,
                    const AnalysisContext& ctx) const {
    const int ty = abiType();
    const bool is_src = ty == CSIM_ELEM_V || ty == CSIM_ELEM_I;
    if (is_src && ctx.type != AnalysisType::OP && ctx.type != AnalysisType::DC && ctx.type != AnalysisType::TRAN)
        return;
    const int N = static_cast<int>(G.rows());
    int32_t type = ty, flags = abiFlags(), branch = abiBranch(), poff = 0, nodes[4] = {-1, -1, -1, -1};
    for (std::size_t k = 0; k < nodeIds.size() && k < 4; ++k) nodes[k] = ckt.nodes[nodeIds[k]].eqIndex;
    if ((ty == CSIM_ELEM_V || ty == CSIM_ELEM_L) && (branch < 0 || branch >= N)) {
        std::fprintf(stderr, "Internal error: invalid branchEqIndex for %s\n", name.c_str());
        return;
    }
    std::vector<double> p;
    abiParams(p);
    csim_topology t;
    t.n_unknowns = N;
    t.n_node_eq = ckt.numNodeEquations();
    t.n_elems = 1;
    t.n_params = static_cast<int32_t>(p.size());
    t.elem_type = &type;
    t.elem_flags = &flags;
    t.elem_nodes = nodes;
    t.elem_branch = &branch;
    t.elem_param = &poff;
    PlanGuard pg;
    if (csim_plan_create(&t, nullptr, 0, &pg.p)) raise("Element::stamp: csim_plan_create");
    int32_t nnz = 0;
    csim_plan_pattern(pg.p, CSIM_MODE_DC, &nnz, nullptr, nullptr);
    std::vector<int32_t> rowptr(N + 1), col(nnz);
    csim_plan_pattern(pg.p, CSIM_MODE_DC, &nnz, rowptr.data(), col.data());
    std::vector<double> xv(N, 0.0), vals(nnz), rhs(N);
    for (int i = 0; i < N && i < x.size(); ++i) xv[i] = x(i);
    double sv = 0.0;
    if (is_src) {
        const SourceSpec& sp = ty == CSIM_ELEM_V ? static_cast<const VoltageSource*>(this)->getSpec()
                                                 : static_cast<const CurrentSource*>(this)->getSpec();
        sv = ctx.type == AnalysisType::TRAN ? sp.evalTran(ctx.time) : sp.evalDC(ctx.sourceScale);
    }
    if (csim_host_stamp_csr(pg.p, CSIM_MODE_DC, 1, p.data(), xv.data(), nullptr, is_src ? &sv : nullptr, 0.0, 1.0,
                            0.0, vals.data(), rhs.data()))
        raise("Element::stamp: csim_host_stamp_csr");
    for (int i = 0; i < N; ++i) {
        for (int k = rowptr[i]; k < rowptr[i + 1]; ++k) G(i, col[k]) += vals[k];
        I(i) += rhs[i];
    }
}
