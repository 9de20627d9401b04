// batchio.cpp — batched CSV writer and the .DC sweep front end (include/csim/batchio.hpp).
#include "../../include/csim/batchio.hpp"

#include <cmath>
#include <cstdio>
#include <stdexcept>

namespace csim {

std::string csv_column_name(const Lowered& L, int32_t eq) {
    if (eq >= 0 && eq < L.n_node_eq) return "V(" + L.node_names[static_cast<std::size_t>(eq)] + ")";
    for (std::size_t e = 0; e < L.elem_type.size(); ++e)
        if ((L.elem_type[e] == CSIM_ELEM_V || L.elem_type[e] == CSIM_ELEM_L) && L.elem_branch[e] == eq)
            return "I(" + L.elem_names[e] + ")";
    throw std::runtime_error("csv_column_name: no unknown with equation index " + std::to_string(eq));
}

void writeInstanceCsv(const Lowered& L, const SimulationConfig& sim, const BatchResult& R,
                      const std::vector<int32_t>& probes, int out_stride, int64_t b, const std::string& path) {
    if (b < 0 || b >= R.B) throw std::runtime_error("writeInstanceCsv: instance out of range");
    if (static_cast<int32_t>(probes.size()) != R.n_probes) throw std::runtime_error("writeInstanceCsv: probes do not match the result");
    if (out_stride < 1) out_stride = 1;
    const TranConfig& cfg = sim.tran;
    std::FILE* f = std::fopen(path.c_str(), "w");
    if (!f) throw std::runtime_error("Cannot open transient output file '" + path + "'.");
    std::string header = "time";
    for (int32_t p : probes) header += "," + csv_column_name(L, p);
    std::fprintf(f, "%s\n", header.c_str());
    if (!(0.0 < cfg.tstart) && !R.x_dc.empty()) {  // the reference's first dumpRow (src/tanalisis.cpp:233-236)
        std::fprintf(f, "%.9e", 0.0);
        for (int32_t p : probes) std::fprintf(f, ",%.9e", R.x_dc[static_cast<std::size_t>(p) * R.B + b]);
        std::fputc('\n', f);
    }
    const int first_row = R.n_steps / out_stride - R.n_out;  // recorded rows are the LAST n_out multiples of out_stride
    const bool dead = (R.status[static_cast<std::size_t>(b)] & CSIM_ST_NONFINITE) != 0;
    const double* w = R.wave.data() + static_cast<std::size_t>(b) * R.n_probes * R.n_out;
    for (int r = 0; r < R.n_out; ++r) {
        if (dead && R.n_probes > 0 && std::isnan(w[r])) break;
        const int step = (first_row + r + 1) * out_stride;  // 1-based step count at this row
        std::fprintf(f, "%.9e", step * cfg.tstep);
        for (int p = 0; p < R.n_probes; ++p) std::fprintf(f, ",%.9e", w[static_cast<std::size_t>(p) * R.n_out + r]);
        std::fputc('\n', f);
    }
    std::fclose(f);
}

DcSweepResult runDcSweepBatch(const Circuit& ckt, const DCSweepConfig& sw, const std::vector<double>& base, int device) {
    Lowered L = lower(ckt);
    if (static_cast<int32_t>(base.size()) != L.n_params) throw std::runtime_error("runDcSweepBatch: base must hold P doubles");
    if (!(sw.step != 0.0) || (sw.stop - sw.start) / sw.step < 0.0)
        throw std::runtime_error("runDcSweepBatch: invalid .DC range");
    int32_t param = -1;
    for (std::size_t e = 0; e < L.elem_type.size(); ++e)
        if ((L.elem_type[e] == CSIM_ELEM_V || L.elem_type[e] == CSIM_ELEM_I) && L.elem_names[e] == sw.sourceName)
            param = L.elem_param[e];  // first parameter of a source = its DC value (CSIM_PARAMS_SRC)
    if (param < 0) throw std::runtime_error("runDcSweepBatch: no independent source named '" + sw.sourceName + "'");
    const int64_t K = static_cast<int64_t>(std::floor((sw.stop - sw.start) / sw.step + 1e-9)) + 1;
    DcSweepResult R;
    R.values.resize(static_cast<std::size_t>(K));
    std::vector<double> params(static_cast<std::size_t>(L.n_params) * K);
    for (int64_t k = 0; k < K; ++k) {
        R.values[static_cast<std::size_t>(k)] = sw.start + static_cast<double>(k) * sw.step;
        for (int32_t p = 0; p < L.n_params; ++p)
            params[static_cast<std::size_t>(p) * K + k] = p == param ? R.values[static_cast<std::size_t>(k)] : base[static_cast<std::size_t>(p)];
    }
    csim_topology t = L.view();
    csim_plan* plan = nullptr;
    csim_batch* batch = nullptr;
    auto fail = [&](const char* what) {
        std::string msg = std::string(what) + ": " + csim_last_error();
        if (batch) csim_batch_destroy(batch);
        if (plan) csim_plan_destroy(plan);
        throw std::runtime_error(msg);
    };
    if (csim_plan_create(&t, nullptr, device, &plan)) fail("csim_plan_create");
    if (csim_batch_create(plan, K, &batch)) fail("csim_batch_create");
    if (csim_batch_set_params(batch, params.data())) fail("csim_batch_set_params");
    csim_constants c;
    csim_default_constants(&c);
    R.x.resize(static_cast<std::size_t>(L.n_unknowns) * K);
    R.iters.resize(static_cast<std::size_t>(c.dc_ramp_steps) * K);
    R.status.resize(static_cast<std::size_t>(K));
    if (csim_batch_dc(batch, R.x.data(), R.iters.data(), nullptr, R.status.data())) fail("csim_batch_dc");
    csim_batch_destroy(batch);
    csim_plan_destroy(plan);
    return R;
}

}  // namespace csim
