// sim.hpp — analysis/source description types.  Same type and member names as the
// reference's include/sim.hpp so code written against it (src/main.cpp) compiles unchanged.
#pragma once
#include <cmath>
#include <cstddef>
#include <string>
#include <vector>

#include "utils.hpp"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

enum class AnalysisType { NONE, OP, DC, AC, TRAN, HB };
enum class AcSweepType { LIN, DEC, OCT };
enum class WaveformType { NONE, PULSE, SIN, PWL };
enum class ProbeKind { NodeVoltage, DiffVoltage, BranchCurrent };

struct AnalysisContext {
    AnalysisType type = AnalysisType::OP;
    double sourceScale = 1.0;
    double time = 0.0;
    double omega = 0.0;
};

struct PulseSpec {
    double v1 = 0.0, v2 = 0.0, td = 0.0, tr = 0.0, tf = 0.0, ton = 0.0, per = 0.0;
};
struct SinSpec {
    double v0 = 0.0, va = 0.0, freq = 0.0, td = 0.0, phi = 0.0;  // phi in radians, td in seconds
};
struct PwlSpec {
    std::vector<double> t, v;
};

// Time-domain waveform of an independent source (reference: include/sim.hpp:69-143).
// The batched device path never calls this on the GPU: the host tabulates it per
// timepoint with the host libm so that sin() is the reference's sin() (csim_src_mode).
struct TranWaveform {
    WaveformType type = WaveformType::NONE;
    PulseSpec pulse;
    SinSpec sine;
    PwlSpec pwl;

    double eval(double t) const {
        if (type == WaveformType::SIN) {
            if (t < sine.td) return sine.v0;
            const double tau = t - sine.td;
            const double w = 2.0 * M_PI * sine.freq;
            return sine.v0 + sine.va * std::sin(w * tau + sine.phi);
        }
        if (type == WaveformType::PULSE) return evalPulse(t);
        if (type == WaveformType::PWL) return evalPwl(t);
        return 0.0;
    }

private:
    double evalPulse(double t) const {
        const PulseSpec& p = pulse;
        double tau;
        if (p.per <= 0.0) {  // single shot
            tau = t - p.td;
            if (tau <= 0.0) return p.v1;
            if (tau < p.tr) return p.v1 + clamp01(tau / p.tr) * (p.v2 - p.v1);
            if (tau < p.tr + p.ton) return p.v2;
            const double k = clamp01((tau - (p.tr + p.ton)) / p.tf);
            return p.v2 + k * (p.v1 - p.v2);
        }
        if (t < p.td) return p.v1;
        tau = std::fmod(t - p.td, p.per);
        if (tau < 0.0) tau += p.per;
        if (tau < p.tr) return p.v1 + (p.v2 - p.v1) * clamp01(tau / p.tr);
        if (tau < p.tr + p.ton) return p.v2;
        if (tau < p.tr + p.ton + p.tf) return p.v2 + (p.v1 - p.v2) * clamp01((tau - (p.tr + p.ton)) / p.tf);
        return p.v1;
    }
    double evalPwl(double t) const {
        if (pwl.t.empty()) return 0.0;
        if (t <= pwl.t.front()) return pwl.v.front();
        if (t >= pwl.t.back()) return pwl.v.back();
        for (std::size_t i = 0; i + 1 < pwl.t.size(); ++i)
            if (t > pwl.t[i] && t <= pwl.t[i + 1]) {
                const double k = (t - pwl.t[i]) / (pwl.t[i + 1] - pwl.t[i]);
                return pwl.v[i] + (pwl.v[i + 1] - pwl.v[i]) * k;
            }
        return pwl.v.back();
    }
};

struct SourceSpec {
    double dcValue = 0.0;
    double acMag = 0.0;
    double acPhaseDeg = 0.0;
    TranWaveform tran;

    // reference: include/sim.hpp:152-158 — a SIN source contributes its offset v0 at DC
    double evalDC(double scale) const {
        double base = dcValue;
        if (tran.type == WaveformType::SIN) base += tran.sine.v0;
        return base * scale;
    }
    double evalTran(double t) const { return dcValue + tran.eval(t); }
};

struct DCSweepConfig {
    std::string sourceName;
    double start = 0.0, stop = 0.0, step = 0.0;
};
struct TranConfig {
    bool enabled = false;
    double tstep = 0.0, tstop = 0.0, tstart = 0.0;
};
struct AcConfig {
    bool enabled = false;
    AcSweepType sweepType = AcSweepType::DEC;
    int nPoints = 0;
    double fstart = 0.0, fstop = 0.0;
};
struct HbConfig {
    bool enabled = false;
    double f0 = 0.0;
    int nHarm = 0;
};
// `.MC instances [seed [rel]]` — Monte-Carlo batch card (superset; see montecarlo.hpp).  Not an analysis of its own:
// it says how many instances the batched transient / DC drivers run.
struct MonteCarloConfig {
    bool enabled = false;
    long long instances = 0;
    unsigned long long seed = 0;
    double rel = 0.05;
};
struct ProbeSpec {
    ProbeKind kind = ProbeKind::NodeVoltage;
    std::string expr, node1, node2, eleName, elePort;
};
struct PrintCommand {
    AnalysisType analysis = AnalysisType::NONE;
    std::vector<ProbeSpec> probes;
};

class SimulationConfig {
public:
    bool doOp = false;
    std::vector<DCSweepConfig> dcSweeps;
    TranConfig tran;
    AcConfig ac;
    HbConfig hb;
    MonteCarloConfig mc;
    std::vector<PrintCommand> printCommands;

    bool hasAnyAnalysis() const { return doOp || !dcSweeps.empty() || tran.enabled || ac.enabled || hb.enabled; }
    void ensureDefaultOp() { doOp = !hasAnyAnalysis(); }
};

inline bool matchAnalysis(const PrintCommand& pc, AnalysisType cur) {
    return pc.analysis == AnalysisType::NONE || pc.analysis == cur;
}
