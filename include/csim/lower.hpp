// lower.hpp — Circuit -> flat csim_topology (+ nominal parameters, names, source tables) and
// the batched drivers that sit beside the reference's single-instance entry points.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../csim_b200.h"
#include "circuit.hpp"
#include "sim.hpp"

namespace csim {

struct Lowered {
    int32_t n_unknowns = 0, n_node_eq = 0, n_params = 0;
    std::vector<int32_t> elem_type, elem_flags, elem_nodes, elem_branch, elem_param;
    std::vector<double> nominal;          // [P]
    std::vector<std::string> node_names;  // by equation index (node rows)
    std::vector<std::string> elem_names;
    std::vector<int32_t> source_elems;    // element index of each independent source, in order
    csim_topology view() const;           // pointers into this object
    // CSV column order of the reference (src/tanalisis.cpp:191-205): node rows, then V/L branches
    std::vector<int32_t> csv_columns() const;
    std::string csv_header() const;
};

// Walks ckt.elements in order like the reference drivers do (src/tanalisis.cpp:128-136).
// assignEquationIndices() must have been called (src/main.cpp:34); throws otherwise.
Lowered lower(const Circuit& ckt);

struct SourceTables {
    std::vector<int32_t> mode, group;  // per source
    std::vector<double> table;         // [n_groups][n_steps]
    int32_t n_groups = 0, n_steps = 0;
};
// Tabulates every source of the circuit at t = (s+1)*dt, s in [0,n_steps), with the host libm
// (SIN: S = sin(2*pi*f*(t-td)+phi); PULSE/PWL: the waveform value; see csim_src_mode).
SourceTables tabulate_sources(const Circuit& ckt, double dt, int32_t n_steps);

struct BatchResult {
    int64_t B = 0;
    int32_t n_steps = 0, n_probes = 0, n_out = 0;
    std::vector<double> x_dc, x_final;  // [N][B]
    std::vector<double> wave;           // [B][n_probes][n_out]
    std::vector<int32_t> iters_total;   // [B]
    std::vector<uint64_t> pivot_hash;   // [B]
    std::vector<uint32_t> status;       // [B]
    float kernel_ms = 0.f;
};

// DC operating point + backward-Euler transient for B instances of `ckt` that differ only in
// parameter values: params [P][B] instance-minor (P, order and meaning: include/csim_b200.h).
// probes = equation indices to record; every out_stride-th step from tstart on is kept.
BatchResult runTransientAnalysisBackwardEulerBatch(const Circuit& ckt, const SimulationConfig& sim,
                                                   const std::vector<double>& params, int64_t B,
                                                   const std::vector<int32_t>& probes, int out_stride = 1,
                                                   int device = 0);
// The same over several GPUs of one box: contiguous shards of the batch, one host thread + stream per device
// (csim_multi_*, include/csim_b200.h); `devices` lists CUDA device ordinals.  Results equal the single-device call's.
BatchResult runTransientAnalysisBackwardEulerMultiGpu(const Circuit& ckt, const SimulationConfig& sim,
                                                      const std::vector<double>& params, int64_t B,
                                                      const std::vector<int32_t>& probes, int out_stride,
                                                      const std::vector<int32_t>& devices);
// DC operating points only: returns x [N][B].
std::vector<double> dcSolveBatch(const Circuit& ckt, const std::vector<double>& params, int64_t B, int device = 0);

}  // namespace csim
