// sstateanalysis.hpp — periodic steady-state driver.  The reference ships src/sstateanalysis.cpp as
// an EMPTY file and never reads the `.hb f0 nHarm` card it parses (src/parser.cpp:551-572,
// include/sim.hpp:187-191), so this driver is DEFINED here (see csim_sstate_args in csim_b200.h):
// period T = 1/f0 from `.hb`, integrate period after period with the reference's backward-Euler
// step (src/tanalisis.cpp:255-420, timestep from `.TRAN`) until || x(kT) - x((k-1)T) ||_2 < tol.
#pragma once
#include <cstdint>
#include <vector>

#include "circuit.hpp"
#include "sim.hpp"

namespace csim {

struct SteadyStateResult {
    int64_t B = 0;
    int32_t steps_per_period = 0;
    std::vector<double> x_dc, x;       // [N][B]: operating point, state at the period boundary reached
    std::vector<int32_t> periods;      // [B]: periods integrated until periodic, -1 = not within max_periods
    std::vector<double> delta;         // [B]: last period-to-period distance
    std::vector<int32_t> iters_total;  // [B]
    std::vector<uint64_t> pivot_hash;  // [B]
    std::vector<uint32_t> status;      // [B]
    float kernel_ms = 0.f;
};

// params [P][B] instance-minor as for runTransientAnalysisBackwardEulerBatch.  Needs `.hb` (f0 > 0) and
// `.TRAN` (tstep) cards; throws std::runtime_error otherwise.
SteadyStateResult runSteadyStateAnalysisBatch(const Circuit& ckt, const SimulationConfig& sim,
                                              const std::vector<double>& params, int64_t B, int max_periods = 20,
                                              double tol = 1e-6, int device = 0);

// Shooting (the method the north-star names "steady-state shooting candidates"): find x* with Phi_T(x*) = x*, where
// Phi_T integrates one period with the reference's backward-Euler transient, by Newton on F(x) = Phi_T(x) - x.
// The Jacobian dPhi_T/dx is taken by forward differences whose N perturbation columns are BATCH INSTANCES: every
// shooting iteration is one batched transient of C * (N + 1) instances (candidate + N perturbed copies per
// circuit), followed by one batched LU solve of the C systems (dPhi - I) delta = -F on the device.
//   x_{m+1} = x_m + delta,  column j: (Phi_T(x_m + h_j e_j) - Phi_T(x_m)) / h_j,  h_j = rel_step * max(|x_m,j|, 1).
// Requires sources that are T-periodic (time restarts at 0 every period).  Converged when ||F(x_m)||_2 < tol.
// No counterpart in the reference (empty src/sstateanalysis.cpp): parity unpinned; the tests check periodicity of
// the result and agreement with runSteadyStateAnalysisBatch.
struct ShootingResult {
    int64_t C = 0;
    int32_t steps_per_period = 0;
    std::vector<double> x_dc, x;      // [N][C]
    std::vector<int32_t> iterations;  // [C] shooting iterations used, -1 = not converged within max_iters
    std::vector<double> residual;     // [C] ||Phi_T(x) - x||_2 at the returned x
    int64_t transient_instances = 0;  // instances integrated for one period, summed over the iterations
};
ShootingResult runSteadyStateShootingBatch(const Circuit& ckt, const SimulationConfig& sim,
                                           const std::vector<double>& params, int64_t C, int max_iters = 10,
                                           double tol = 1e-9, double rel_step = 1e-6, int device = 0);

}  // namespace csim
