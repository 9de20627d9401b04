// montecarlo.hpp — Monte-Carlo parameter batches for the batched drivers ("plus Monte-Carlo
// spec cards").  The reference has no such card; its parser would print "unsupported control card" for it.  Card:
//
//     .MC <instances> [seed [rel]]          e.g.  .MC 65536 20261104 0.05
//
// Recipe (the one bench.py and the tests use): instance 0 is the nominal circuit; instance i > 0 is
// nominal * (1 + rel * u_i) with u_i ~ U(-1, 1) drawn ROW-wise (instance-major, parameter-minor) from
// numpy.random.Generator(PCG64(seed)).uniform(-1, 1, (instances, P)); only R, C, L and the MOSFET parameters
// Vth, K, lambda, Cj0 are perturbed, sources stay nominal.  The generator is restated here bit for bit (numpy's
// SeedSequence + PCG64 XSL-RR 128/64 + 53-bit doubles), so a C++ caller draws exactly the batch the Python layer draws.
#pragma once
#include <cstdint>
#include <vector>

#include "lower.hpp"

namespace csim {

// numpy.random.PCG64 seeded with one non-negative integer (SeedSequence(seed)), restated.
class Pcg64 {
public:
    explicit Pcg64(uint64_t seed);
    uint64_t next64();
    double nextDouble() { return static_cast<double>(next64() >> 11) * (1.0 / 9007199254740992.0); }
    void advance(uint64_t delta);  // bit_generator.advance(delta), delta < 2^64

private:
    unsigned __int128 state_, inc_;
};

// params [P][B] instance-minor for instances [first, first + B) of the global batch drawn from `seed`.
std::vector<double> monteCarloParams(const Lowered& L, int64_t B, uint64_t seed, double rel = 0.05, int64_t first = 0);

// The netlist's `.MC` card as one batched transient: parameters as above, then
// runTransientAnalysisBackwardEulerBatch.  Throws std::runtime_error when the card is absent.
BatchResult runMonteCarloTransientBatch(const Circuit& ckt, const SimulationConfig& sim, const std::vector<int32_t>& probes,
                                        int out_stride = 1, int device = 0);
// ... sharded over several GPUs (runTransientAnalysisBackwardEulerMultiGpu)
BatchResult runMonteCarloTransientMultiGpu(const Circuit& ckt, const SimulationConfig& sim, const std::vector<int32_t>& probes,
                                           int out_stride, const std::vector<int32_t>& devices);

}  // namespace csim
