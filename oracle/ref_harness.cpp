// ref_harness.cpp — TEST INFRASTRUCTURE.  extern "C" window onto the UNMODIFIED
// reference (ZyuRao/CircuitSimulator) compiled from /root/reference/src/*.cpp
// against oracle/eigen_shim.  Built only where /root/reference exists (the build
// container) into oracle/_ref/libcsim_ref.so, which travels to the GPU box.
//
// What it offers
//   * ref_open/ref_topology : parse a netlist with the reference's own parser and
//     export the flat topology + nominal parameters (to cross-check the product's
//     own parser and lowering).
//   * ref_dc / ref_tran     : run the reference's own dcSolveLU and
//     runTransientAnalysisBackwardEuler, optionally recording — through the
//     stand-in's trace hooks, i.e. without touching reference code — the raw LU
//     solution, pivot permutation and step norm of every Newton iteration.
//   * ref_stamp / ref_lu_solve : single calls of Element::stamp and
//     Solver::solveLinearSystemLU.
//   * ref_run_batch         : the timed CPU baseline: one instance per host
//     thread, each with its own Circuit built from its own netlist text
//     (BASELINE.md §3).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <streambuf>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include <pthread.h>
#include <sched.h>

#include <Eigen/Dense>  // the stand-in (include guards make later includes no-ops)

// The reference keeps R / K / Vth / lambda private or protected and offers no
// getters.  This TU — and only this TU — reads them directly;
// the reference's own translation units are compiled untouched.
#define private public
#define protected public
#include "circuit.hpp"
#include "element.hpp"
#undef private
#undef protected
#include "dcanalysis.hpp"
#include "parser.hpp"
#include "sim.hpp"
#include "solver.hpp"
#include "tanalisis.hpp"

#include "../include/csim_b200.h"

namespace {

struct NullBuf : std::streambuf {
    int overflow(int c) override { return c; }
};
NullBuf g_nullbuf;
struct Silence {  // the reference chats on cout/cerr; keep test logs readable
    std::streambuf *o, *e;
    Silence() : o(std::cout.rdbuf(&g_nullbuf)), e(std::cerr.rdbuf(&g_nullbuf)) {}
    ~Silence() {
        std::cout.rdbuf(o);
        std::cerr.rdbuf(e);
    }
};

struct Handle {
    Circuit ckt;
    SimulationConfig sim;
};

struct Recorder {
    int n = 0;
    std::vector<int32_t> perm_cur;
    double* xraw = nullptr;
    int32_t* perm = nullptr;
    double* err = nullptr;
    int64_t cap = 0, n_iters = 0, n_err = 0;
    void reset_perm() {
        perm_cur.resize(n);
        for (int i = 0; i < n; ++i) perm_cur[i] = i;
    }
};

void hook_all_finite(void* u, const double* v, Eigen::Index n) {
    auto* r = static_cast<Recorder*>(u);
    if ((int)n == r->n) {
        if (r->n_iters < r->cap) {
            if (r->xraw) std::memcpy(r->xraw + r->n_iters * n, v, sizeof(double) * n);
            if (r->perm) std::memcpy(r->perm + r->n_iters * n, r->perm_cur.data(), sizeof(int32_t) * n);
        }
        r->n_iters++;
    }
    r->reset_perm();
}
void hook_norm(void* u, const double*, Eigen::Index, double res) {
    auto* r = static_cast<Recorder*>(u);
    if (r->err && r->n_err < r->cap) r->err[r->n_err] = res;
    r->n_err++;
}
void hook_swap(void* u, Eigen::Index, Eigen::Index k, Eigen::Index p) {
    auto* r = static_cast<Recorder*>(u);
    std::swap(r->perm_cur[k], r->perm_cur[p]);
}

struct HookGuard {
    HookGuard(Recorder* r) {
        auto& h = Eigen::shim_trace::hooks();
        h.user = r;
        h.on_all_finite = hook_all_finite;
        h.on_norm = hook_norm;
        h.on_row_swap = hook_swap;
    }
    ~HookGuard() { Eigen::shim_trace::hooks() = Eigen::shim_trace::Hooks(); }
};

int param_count(const std::shared_ptr<Element>& e) {
    if (std::dynamic_pointer_cast<Resistor>(e)) return CSIM_PARAMS_R;
    if (std::dynamic_pointer_cast<CapacitorElement>(e)) return CSIM_PARAMS_C;
    if (std::dynamic_pointer_cast<Inductor>(e)) return CSIM_PARAMS_L;
    if (std::dynamic_pointer_cast<VoltageSource>(e)) return CSIM_PARAMS_SRC;
    if (std::dynamic_pointer_cast<CurrentSource>(e)) return CSIM_PARAMS_SRC;
    if (std::dynamic_pointer_cast<MosfetBase>(e)) return CSIM_PARAMS_MOS;
    return 0;
}

}  // namespace

extern "C" {

void* ref_open(const char* netlist_text) {
    Silence s;
    auto* h = new Handle();
    std::istringstream in(netlist_text);
    NetlistParser parser(h->ckt, h->sim);
    parser.parseStream(in, "<memory>");
    h->sim.ensureDefaultOp();
    h->ckt.assignEquationIndices();  // src/main.cpp:34
    return h;
}
void ref_close(void* hv) { delete static_cast<Handle*>(hv); }

int ref_dims(void* hv, int32_t* n_unknowns, int32_t* n_node_eq, int32_t* n_elems, int32_t* n_params) {
    auto* h = static_cast<Handle*>(hv);
    *n_unknowns = h->ckt.numUnknowns();
    *n_node_eq = h->ckt.numNodeEquations();
    *n_elems = (int32_t)h->ckt.elements.size();
    int p = 0;
    for (auto& e : h->ckt.elements) p += param_count(e);
    *n_params = p;
    return 0;
}

// arrays sized per ref_dims; nodes is [n_elems][4]
int ref_topology(void* hv, int32_t* type, int32_t* flags, int32_t* nodes, int32_t* branch, int32_t* param_off,
                 double* params) {
    auto* h = static_cast<Handle*>(hv);
    const Circuit& c = h->ckt;
    int off = 0;
    for (size_t i = 0; i < c.elements.size(); ++i) {
        const auto& e = c.elements[i];
        const auto& nid = e->getNodeIds();
        for (int k = 0; k < 4; ++k) nodes[4 * i + k] = k < (int)nid.size() ? c.nodes[nid[k]].eqIndex : -1;
        branch[i] = -1;
        flags[i] = 0;
        param_off[i] = off;
        auto put_src = [&](const SourceSpec& sp) {
            params[off + 0] = sp.dcValue;
            params[off + 1] = sp.tran.sine.v0;
            params[off + 2] = sp.tran.sine.va;
            params[off + 3] = sp.tran.sine.freq;
            params[off + 4] = sp.tran.sine.td;
            params[off + 5] = sp.tran.sine.phi;
            if (sp.tran.type == WaveformType::SIN) flags[i] |= CSIM_FLAG_SRC_SIN;
        };
        if (auto r = std::dynamic_pointer_cast<Resistor>(e)) {
            type[i] = CSIM_ELEM_R;
            params[off] = r->R;
        } else if (auto cp = std::dynamic_pointer_cast<CapacitorElement>(e)) {
            type[i] = CSIM_ELEM_C;
            params[off] = cp->getC();
        } else if (auto l = std::dynamic_pointer_cast<Inductor>(e)) {
            type[i] = CSIM_ELEM_L;
            params[off] = l->getL();
            branch[i] = l->getBranchEqIndex();
        } else if (auto v = std::dynamic_pointer_cast<VoltageSource>(e)) {
            type[i] = CSIM_ELEM_V;
            branch[i] = v->getBranchEqIndex();
            put_src(v->getSpec());
        } else if (auto is = std::dynamic_pointer_cast<CurrentSource>(e)) {
            type[i] = CSIM_ELEM_I;
            put_src(is->setSpec());
        } else if (auto m = std::dynamic_pointer_cast<MosfetBase>(e)) {
            type[i] = CSIM_ELEM_MOS;
            if (m->isP) flags[i] |= CSIM_FLAG_PMOS;
            params[off + 0] = m->Vth;
            params[off + 1] = m->K;
            params[off + 2] = m->lambda;
            params[off + 3] = m->Cj0;
        } else {
            return -1;
        }
        off += param_count(e);
    }
    return 0;
}

// node name owning equation eq (for netlist regeneration / CSV headers)
int ref_node_name(void* hv, int eq, char* out, int cap) {
    auto* h = static_cast<Handle*>(hv);
    for (auto& n : h->ckt.nodes)
        if (n.eqIndex == eq) {
            std::snprintf(out, cap, "%s", n.name.c_str());
            return 0;
        }
    return -1;
}
int ref_elem_name(void* hv, int idx, char* out, int cap) {
    auto* h = static_cast<Handle*>(hv);
    if (idx < 0 || idx >= (int)h->ckt.elements.size()) return -1;
    std::snprintf(out, cap, "%s", h->ckt.elements[idx]->getName().c_str());
    return 0;
}

int ref_tran_cfg(void* hv, double* tstep, double* tstop, double* tstart, int32_t* enabled) {
    auto* h = static_cast<Handle*>(hv);
    *tstep = h->sim.tran.tstep;
    *tstop = h->sim.tran.tstop;
    *tstart = h->sim.tran.tstart;
    *enabled = h->sim.tran.enabled ? 1 : 0;
    return 0;
}
int ref_hb_cfg(void* hv, double* f0, int32_t* nharm, int32_t* enabled) {
    auto* h = static_cast<Handle*>(hv);
    *f0 = h->sim.hb.f0;
    *nharm = h->sim.hb.nHarm;
    *enabled = h->sim.hb.enabled ? 1 : 0;
    return 0;
}

// dcSolveLU (src/dcanalysis.cpp:242-248), raw doubles, optional trace
int ref_dc(void* hv, double* x, double* xraw, int32_t* perm, double* err, int64_t cap, int64_t* n_iters,
           int64_t* n_err) {
    auto* h = static_cast<Handle*>(hv);
    Silence s;
    Recorder r;
    r.n = h->ckt.numUnknowns();
    r.reset_perm();
    r.xraw = xraw;
    r.perm = perm;
    r.err = err;
    r.cap = cap;
    Eigen::VectorXd sol;
    {
        HookGuard g(&r);
        sol = dcSolveLU(h->ckt);
    }
    for (int i = 0; i < r.n; ++i) x[i] = sol(i);
    if (n_iters) *n_iters = r.n_iters;
    if (n_err) *n_err = r.n_err;
    return 0;
}

// runTransientAnalysisBackwardEuler (src/tanalisis.cpp:83-424) including its own
// DC solve; the trace therefore starts with the DC iterations.
int ref_tran(void* hv, const char* csv_path, double* xraw, int32_t* perm, double* err, int64_t cap,
             int64_t* n_iters, int64_t* n_err) {
    auto* h = static_cast<Handle*>(hv);
    Silence s;
    Recorder r;
    r.n = h->ckt.numUnknowns();
    r.reset_perm();
    r.xraw = xraw;
    r.perm = perm;
    r.err = err;
    r.cap = cap;
    int rc = 0;
    {
        HookGuard g(&r);
        try {
            runTransientAnalysisBackwardEuler(h->ckt, h->sim, csv_path);
        } catch (const std::exception&) {
            rc = 1;
        }
    }
    if (n_iters) *n_iters = r.n_iters;
    if (n_err) *n_err = r.n_err;
    return rc;
}

// One assembly with the reference's own Element::stamp in netlist order
// (the DC assembly of src/dcanalysis.cpp:126-128 when tran == 0; for tran != 0 only
// the elements the transient driver stamps through Element::stamp, i.e. steps 1-2
// of src/tanalisis.cpp:269-279).  G row-major N*N.
int ref_stamp(void* hv, int tran, const double* x, double scale, double time, double* G, double* I) {
    auto* h = static_cast<Handle*>(hv);
    Silence s;
    const int n = h->ckt.numUnknowns();
    Eigen::MatrixXd Gm = Eigen::MatrixXd::Zero(n, n);
    Eigen::VectorXd Iv = Eigen::VectorXd::Zero(n);
    Eigen::VectorXd xv(n);
    for (int i = 0; i < n; ++i) xv(i) = x[i];
    AnalysisContext ctx;
    ctx.type = tran ? AnalysisType::TRAN : AnalysisType::DC;
    ctx.sourceScale = scale;
    ctx.time = time;
    if (!tran) {
        for (const auto& e : h->ckt.elements) e->stamp(Gm, Iv, h->ckt, xv, ctx);
    } else {
        for (const auto& e : h->ckt.elements) {
            if (std::dynamic_pointer_cast<CapacitorElement>(e)) continue;
            if (std::dynamic_pointer_cast<Inductor>(e)) continue;
            if (std::dynamic_pointer_cast<MosfetBase>(e)) continue;
            e->stamp(Gm, Iv, h->ckt, xv, ctx);
        }
        for (const auto& e : h->ckt.elements)
            if (std::dynamic_pointer_cast<MosfetBase>(e)) e->stamp(Gm, Iv, h->ckt, xv, ctx);
    }
    for (int i = 0; i < n; ++i) {
        I[i] = Iv(i);
        for (int j = 0; j < n; ++j) G[i * n + j] = Gm(i, j);
    }
    return 0;
}

// Solver::solveLinearSystemLU (include/solver.hpp:83-131)
int ref_lu_solve(int n, const double* A, const double* b, double* x, int32_t* perm) {
    Silence s;
    Eigen::MatrixXd Am(n, n);
    Eigen::VectorXd bv(n);
    for (int i = 0; i < n; ++i) {
        bv(i) = b[i];
        for (int j = 0; j < n; ++j) Am(i, j) = A[i * n + j];
    }
    if (perm) {
        Eigen::MatrixXd LU;
        std::vector<int> p;
        Solver::luDecompose(Am, LU, p);
        for (int i = 0; i < n && i < (int)p.size(); ++i) perm[i] = p[i];
    }
    Eigen::VectorXd xs = Solver::solveLinearSystemLU(Am, bv);
    for (int i = 0; i < n; ++i) x[i] = xs(i);
    return 0;
}

// Solver::luDecompose (include/solver.hpp:30-80): packed L\U as the reference leaves it, its permutation and its
// return value (1 = true).  LU is written only when the reference resized it (it does so before any early return
// that matters here).
int ref_lu_decompose(int n, const double* A, double* LU_out, int32_t* perm) {
    Silence s;
    Eigen::MatrixXd Am(n, n), LU;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) Am(i, j) = A[i * n + j];
    std::vector<int> p;
    const bool ok = Solver::luDecompose(Am, LU, p);
    for (int i = 0; i < n && i < (int)p.size(); ++i) perm[i] = p[i];
    if (LU.rows() == n && LU.cols() == n)
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) LU_out[i * n + j] = LU(i, j);
    return ok ? 1 : 0;
}

// dcSolveGaussSeidel (src/dcanalysis.cpp:251-257 -> 71-92 / 166-237): the reference's own Gauss-Seidel DC driver,
// warnings silenced; *n_warnings = lines it wrote to std::cerr.
int ref_dc_gauss_seidel(void* hv, double* x, int64_t* n_warnings) {
    auto* h = static_cast<Handle*>(hv);
    std::ostringstream cap;
    std::streambuf* old_err = std::cerr.rdbuf(cap.rdbuf());
    Eigen::VectorXd sol;
    {
        Silence s2;  // (stdout)
        std::cerr.rdbuf(cap.rdbuf());
        sol = dcSolveGaussSeidel(h->ckt);
    }
    std::cerr.rdbuf(old_err);
    const int n = h->ckt.numUnknowns();
    for (int i = 0; i < n; ++i) x[i] = sol(i);
    if (n_warnings) {
        const std::string t = cap.str();
        *n_warnings = (int64_t)std::count(t.begin(), t.end(), '\n');
    }
    return 0;
}

// Solver::solveLinearSystemGaussSeidel (include/solver.hpp:139-192), x in: x0, out: the last iterate
int ref_gauss_seidel(int n, const double* A, const double* b, double* x, int max_iters, double tol) {
    Silence s;
    Eigen::MatrixXd Am(n, n);
    Eigen::VectorXd bv(n), x0(n);
    for (int i = 0; i < n; ++i) {
        bv(i) = b[i];
        x0(i) = x[i];
        for (int j = 0; j < n; ++j) Am(i, j) = A[i * n + j];
    }
    Eigen::VectorXd xs = Solver::solveLinearSystemGaussSeidel(Am, bv, x0, max_iters, tol);
    for (int i = 0; i < n; ++i) x[i] = xs(i);
    return 0;
}

// SourceSpec::evalTran of the idx-th independent source (V and I in netlist order)
double ref_source_tran(void* hv, int idx, double t) {
    auto* h = static_cast<Handle*>(hv);
    int s = 0;
    for (const auto& e : h->ckt.elements) {
        const SourceSpec* sp = nullptr;
        if (auto v = std::dynamic_pointer_cast<VoltageSource>(e)) sp = &v->getSpec();
        else if (auto c = std::dynamic_pointer_cast<CurrentSource>(e)) sp = &c->setSpec();
        if (!sp) continue;
        if (s == idx) return sp->evalTran(t);
        ++s;
    }
    return NAN;
}

// Gives the idx-th independent source (V and I in netlist order) a PULSE (type 1: spec = v1 v2 td tr tf ton per)
// or PWL (type 3: spec = n times then n values) waveform.  The reference's parser only ever produces SIN
// (src/parser.cpp:346-351), but TranWaveform::eval implements all three (include/sim.hpp:80-138): this sets the
// reference's OWN SourceSpec object, so its drivers then run its own PULSE / PWL code.
int ref_set_source_waveform(void* hv, int idx, int type, const double* spec, int n) {
    auto* h = static_cast<Handle*>(hv);
    int s = 0;
    for (const auto& e : h->ckt.elements) {
        SourceSpec* sp = nullptr;
        if (auto v = std::dynamic_pointer_cast<VoltageSource>(e)) sp = const_cast<SourceSpec*>(&v->getSpec());
        else if (auto c = std::dynamic_pointer_cast<CurrentSource>(e)) sp = const_cast<SourceSpec*>(&c->setSpec());
        if (!sp) continue;
        if (s++ != idx) continue;
        if (type == 1) {
            sp->tran.type = WaveformType::PULSE;
            sp->tran.pulse = PulseSpec{spec[0], spec[1], spec[2], spec[3], spec[4], spec[5], spec[6]};
        } else if (type == 3) {
            sp->tran.type = WaveformType::PWL;
            sp->tran.pwl.t.assign(spec, spec + n);
            sp->tran.pwl.v.assign(spec + n, spec + 2 * n);
        } else {
            return -2;
        }
        return 0;
    }
    return -1;
}

// Timed CPU baseline.  netlists[i] = full netlist text of instance i.  n_threads
// host threads (pinned round-robin onto the CPUs this process may use) pull
// instance indices; each builds its own Circuit and runs the reference transient
// driver writing its CSV to /dev/null.  Returns wall seconds; *steps_done gets the
// total number of timepoints simulated.
double ref_run_batch(const char* const* netlists, int n_instances, int n_threads, int64_t* steps_done) {
    Silence s;
    std::atomic<int> next(0);
    std::atomic<int64_t> steps(0);
    cpu_set_t allowed;
    CPU_ZERO(&allowed);
    sched_getaffinity(0, sizeof(allowed), &allowed);
    std::vector<int> cpus;
    for (int c = 0; c < CPU_SETSIZE; ++c)
        if (CPU_ISSET(c, &allowed)) cpus.push_back(c);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int w = 0; w < n_threads; ++w) {
        th.emplace_back([&, w]() {
            if (!cpus.empty()) {
                cpu_set_t one;
                CPU_ZERO(&one);
                CPU_SET(cpus[w % cpus.size()], &one);
                pthread_setaffinity_np(pthread_self(), sizeof(one), &one);
            }
            for (;;) {
                int i = next.fetch_add(1);
                if (i >= n_instances) break;
                Circuit ckt;
                SimulationConfig sim;
                std::istringstream in(netlists[i]);
                NetlistParser parser(ckt, sim);
                parser.parseStream(in, "<memory>");
                ckt.assignEquationIndices();
                try {
                    runTransientAnalysisBackwardEuler(ckt, sim, "/dev/null");
                } catch (const std::exception&) {
                }
                if (sim.tran.enabled && sim.tran.tstep > 0)
                    steps += (int64_t)std::floor(sim.tran.tstop / sim.tran.tstep + 1e-12);
            }
        });
    }
    for (auto& t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    if (steps_done) *steps_done = steps.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

int ref_hardware_threads(void) {
    cpu_set_t allowed;
    CPU_ZERO(&allowed);
    if (sched_getaffinity(0, sizeof(allowed), &allowed) == 0) return CPU_COUNT(&allowed);
    return (int)std::thread::hardware_concurrency();
}

}  // extern "C"
