// batchio.hpp — rows (f)1 and (f)3 of SURVEY §8: the reference's CSV contract for batched runs, and the
// `.DC src start stop step` card (parsed by the reference at src/parser.cpp:476-495 into DCSweepConfig,
// include/sim.hpp:165-170, and never executed) as one batch of DC operating points.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "lower.hpp"

namespace csim {

// Column name of one unknown in the reference's CSV (src/tanalisis.cpp:191-205): "V(<node>)" for a node row,
// "I(<element>)" for the branch current of a voltage source or inductor.
std::string csv_column_name(const Lowered& L, int32_t equation_index);

// One instance of a batched transient written in the reference's CSV layout (src/tanalisis.cpp:183-231):
// header "time,<columns>", "%.9e", a t = 0 row with the DC operating point when tstart <= 0, then one row per
// recorded point at t = (step + 1) * tstep.  `probes` are the recorded equation indices (BatchResult::wave rows);
// with probes == L.csv_columns() and out_stride == 1 the file is byte-identical to what the reference writes for
// that instance.  Rows after a non-finite failure (status & CSIM_ST_NONFINITE) are not written, as in the reference.
void writeInstanceCsv(const Lowered& L, const SimulationConfig& sim, const BatchResult& R,
                      const std::vector<int32_t>& probes, int out_stride, int64_t instance, const std::string& path);

struct DcSweepResult {
    std::vector<double> values;        // swept source value of each point
    std::vector<double> x;             // [N][K] operating points, instance-minor
    std::vector<int32_t> iters;        // [ramp][K] Newton iterations per ramp step
    std::vector<uint32_t> status;      // [K]
};
// `.DC`: point k sets the DC value of source `sweep.sourceName` to start + k * step, k = 0 .. floor((stop - start) /
// step + 1e-9); every point is an independent DC operating point (dcSolve, src/dcanalysis.cpp:242-262) from x = 0 —
// one batch of K instances on the device.  `base` = the other parameters ([P], e.g. Lowered::nominal).
DcSweepResult runDcSweepBatch(const Circuit& ckt, const DCSweepConfig& sweep, const std::vector<double>& base,
                              int device = 0);

}  // namespace csim
