// montecarlo.cpp — see include/csim/montecarlo.hpp.
#include "csim/montecarlo.hpp"

#include <stdexcept>

namespace csim {

namespace {
// numpy/random/bit_generator.pyx: SeedSequence (pool of four 32-bit words)
constexpr uint32_t kInitA = 0x43b0d7e5u, kMultA = 0x931e8875u, kInitB = 0x8b51f9ddu, kMultB = 0x58f38dedu;
constexpr uint32_t kMixL = 0xca01f9ddu, kMixR = 0x4973f715u;
constexpr int kShift = 16, kPool = 4;

uint32_t hashmix(uint32_t value, uint32_t& hash_const) {
    value ^= hash_const;
    hash_const *= kMultA;
    value *= hash_const;
    value ^= value >> kShift;
    return value;
}
uint32_t mix(uint32_t x, uint32_t y) {
    uint32_t r = kMixL * x - kMixR * y;
    r ^= r >> kShift;
    return r;
}
// SeedSequence(seed).generate_state(4, uint64)
void seed_words(uint64_t seed, uint64_t out[4]) {
    std::vector<uint32_t> entropy;  // the integer as little-endian 32-bit words, at least one
    entropy.push_back(static_cast<uint32_t>(seed));
    if (seed >> 32) entropy.push_back(static_cast<uint32_t>(seed >> 32));
    uint32_t pool[kPool], hc = kInitA;
    for (int i = 0; i < kPool; ++i) pool[i] = hashmix(i < (int)entropy.size() ? entropy[i] : 0u, hc);
    for (int s = 0; s < kPool; ++s)
        for (int d = 0; d < kPool; ++d)
            if (s != d) pool[d] = mix(pool[d], hashmix(pool[s], hc));
    for (size_t s = kPool; s < entropy.size(); ++s)
        for (int d = 0; d < kPool; ++d) pool[d] = mix(pool[d], hashmix(entropy[s], hc));
    uint32_t w[8], hb = kInitB;
    for (int i = 0; i < 8; ++i) {
        uint32_t v = pool[i % kPool];
        v ^= hb;
        hb *= kMultB;
        v *= hb;
        v ^= v >> kShift;
        w[i] = v;
    }
    for (int i = 0; i < 4; ++i) out[i] = static_cast<uint64_t>(w[2 * i]) | (static_cast<uint64_t>(w[2 * i + 1]) << 32);
}
const unsigned __int128 kPcgMult = (static_cast<unsigned __int128>(0x2360ED051FC65DA4ull) << 64) | 0x4385DF649FCCF645ull;
}  // namespace

// numpy/random/_pcg64.pyx: pcg64_set_seed(state = (w0 << 64 | w1), inc = (w2 << 64 | w3)), pcg64_srandom_r
Pcg64::Pcg64(uint64_t seed) {
    uint64_t w[4];
    seed_words(seed, w);
    const unsigned __int128 initstate = (static_cast<unsigned __int128>(w[0]) << 64) | w[1];
    const unsigned __int128 initseq = (static_cast<unsigned __int128>(w[2]) << 64) | w[3];
    state_ = 0;
    inc_ = (initseq << 1) | 1u;
    state_ = state_ * kPcgMult + inc_;
    state_ += initstate;
    state_ = state_ * kPcgMult + inc_;
}

uint64_t Pcg64::next64() {  // pcg64_random_r: step, then XSL-RR output
    state_ = state_ * kPcgMult + inc_;
    const uint64_t hi = static_cast<uint64_t>(state_ >> 64), lo = static_cast<uint64_t>(state_);
    const uint64_t x = hi ^ lo;
    const unsigned rot = static_cast<unsigned>(state_ >> 122);
    return (x >> rot) | (x << ((64u - rot) & 63u));
}

void Pcg64::advance(uint64_t delta) {  // pcg_advance_lcg_128
    unsigned __int128 acc_mult = 1, acc_plus = 0, cur_mult = kPcgMult, cur_plus = inc_;
    while (delta > 0) {
        if (delta & 1) {
            acc_mult *= cur_mult;
            acc_plus = acc_plus * cur_mult + cur_plus;
        }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    state_ = acc_mult * state_ + acc_plus;
}

std::vector<double> monteCarloParams(const Lowered& L, int64_t B, uint64_t seed, double rel, int64_t first) {
    if (B <= 0 || first < 0) throw std::runtime_error("monteCarloParams: bad batch range");
    const int P = L.n_params;
    std::vector<char> mask(P, 0);
    for (size_t e = 0; e < L.elem_type.size(); ++e) {
        const int ty = L.elem_type[e];
        const int cnt = ty == CSIM_ELEM_MOS ? CSIM_PARAMS_MOS : (ty == CSIM_ELEM_R || ty == CSIM_ELEM_C || ty == CSIM_ELEM_L) ? 1 : 0;
        for (int k = 0; k < cnt; ++k) mask[L.elem_param[e] + k] = 1;
    }
    Pcg64 rng(seed);
    if (first) rng.advance(static_cast<uint64_t>(first) * static_cast<uint64_t>(P));
    std::vector<double> out(static_cast<size_t>(P) * B);
    for (int64_t b = 0; b < B; ++b)
        for (int p = 0; p < P; ++p) {
            const double u = -1.0 + 2.0 * rng.nextDouble();  // Generator.uniform(-1, 1): low + (high - low) * next_double
            const bool nominal = (first + b == 0) || !mask[p];
            out[static_cast<size_t>(p) * B + b] = nominal ? L.nominal[p] : L.nominal[p] * (1.0 + rel * u);
        }
    return out;
}

BatchResult runMonteCarloTransientBatch(const Circuit& ckt, const SimulationConfig& sim, const std::vector<int32_t>& probes,
                                        int out_stride, int device) {
    if (!sim.mc.enabled) throw std::runtime_error("runMonteCarloTransientBatch: the netlist has no .MC card");
    const Lowered L = lower(ckt);
    const std::vector<double> params = monteCarloParams(L, sim.mc.instances, sim.mc.seed, sim.mc.rel);
    return runTransientAnalysisBackwardEulerBatch(ckt, sim, params, sim.mc.instances, probes, out_stride, device);
}

BatchResult runMonteCarloTransientMultiGpu(const Circuit& ckt, const SimulationConfig& sim, const std::vector<int32_t>& probes,
                                           int out_stride, const std::vector<int32_t>& devices) {
    if (!sim.mc.enabled) throw std::runtime_error("runMonteCarloTransientMultiGpu: the netlist has no .MC card");
    const Lowered L = lower(ckt);
    const std::vector<double> params = monteCarloParams(L, sim.mc.instances, sim.mc.seed, sim.mc.rel);
    return runTransientAnalysisBackwardEulerMultiGpu(ckt, sim, params, sim.mc.instances, probes, out_stride, devices);
}

}  // namespace csim
