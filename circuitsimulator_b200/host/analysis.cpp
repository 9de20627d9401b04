// analysis.cpp — reference-named single-instance entry points (Solver::*, dcSolve*,
// computeDcOperatingPoint, runTransientAnalysisBackwardEuler) on top of the batched CUDA
// path with B = 1.  Error behaviour follows the reference: messages on std::cerr and early
// return; std::runtime_error only for the transient NaN/Inf case (src/tanalisis.cpp:360-362)
// and — new — when no CUDA device is available (this library has no CPU solver).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <iostream>
#include <limits>
#include <stdexcept>

#include "csim/dcanalysis.hpp"
#include "csim/lower.hpp"
#include "csim/solver.hpp"
#include "csim/tanalisis.hpp"

namespace Solver {

static bool deviceSolve(const MatrixXd& A, const VectorXd* b, VectorXd* x, std::vector<int>& perm) {
    const int n = static_cast<int>(A.rows());
    std::vector<double> a(static_cast<std::size_t>(n) * n), rhs(n, 0.0), sol(n, 0.0);
    for (int i = 0; i < n; ++i) {
        if (b) rhs[i] = (*b)(i);
        for (int j = 0; j < n; ++j) a[static_cast<std::size_t>(i) * n + j] = A(i, j);
    }
    std::vector<int32_t> p(n);
    int32_t singular = 0;
    if (csim_host_lu_solve(n, 1, a.data(), rhs.data(), sol.data(), p.data(), &singular, 1e-15))
        throw std::runtime_error(std::string("Solver: ") + csim_last_error());
    perm.assign(p.begin(), p.end());
    if (x)
        for (int i = 0; i < n; ++i) (*x)(i) = sol[i];
    return singular == 0;
}

// include/solver.hpp:30-80 — the packed factors come back exactly as the reference leaves them: U on and above
// the diagonal, the multipliers in the strict lower triangle, rows exchanged physically (csim_host_lu_factor).
bool luDecompose(const MatrixXd& A, MatrixXd& LU, std::vector<int>& perm) {
    const int n = static_cast<int>(A.rows());
    if (n == 0) return false;
    if (A.cols() != n) {
        std::cerr << "LU: matrix is not square.\n";
        return false;
    }
    std::vector<double> a(static_cast<std::size_t>(n) * n), lu(static_cast<std::size_t>(n) * n);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) a[static_cast<std::size_t>(i) * n + j] = A(i, j);
    std::vector<int32_t> p(n);
    int32_t singular = 0;
    if (csim_host_lu_factor(n, 1, a.data(), lu.data(), p.data(), &singular, 1e-15))
        throw std::runtime_error(std::string("Solver: ") + csim_last_error());
    perm.assign(p.begin(), p.end());
    LU = MatrixXd(n, n);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) LU(i, j) = lu[static_cast<std::size_t>(i) * n + j];
    if (singular) {
        std::cerr << "LU: zero (or tiny) pivot.\n";
        return false;
    }
    return true;
}

VectorXd solveLinearSystemLU(const MatrixXd& A, const VectorXd& b) {
    const int n = static_cast<int>(A.rows());
    VectorXd x = VectorXd::Zero(n);
    if (n == 0) return x;
    if (A.cols() != n || b.size() != n) {
        std::cerr << "LU solve: dimension mismatch.\n";
        return x;
    }
    std::vector<int> perm;
    if (!deviceSolve(A, &b, &x, perm)) std::cerr << "LU solve: decomposition failed.\n";  // x is the zero vector
    return x;
}

// include/solver.hpp:139-192 of the reference, one system on the device (csim_host_gauss_seidel); same dimension
// checks and messages; systems with more than 200 unknowns are refused (CSIM_E_UNSUPPORTED -> exception).
VectorXd solveLinearSystemGaussSeidel(const MatrixXd& A, const VectorXd& b, const VectorXd& x0, int maxIters, double tol) {
    const int n = static_cast<int>(A.rows());
    VectorXd x = x0;
    if (n == 0) return x;
    if (A.cols() != n || b.size() != n) {
        std::cerr << "Gauss-Seidel: dimension mismatch.\n";
        return VectorXd::Zero(std::max(n, 0));
    }
    if (x.size() != n) x = VectorXd::Zero(n);
    std::vector<double> a(static_cast<std::size_t>(n) * n), rhs(n), sol(n);
    for (int i = 0; i < n; ++i) {
        rhs[i] = b(i);
        sol[i] = x(i);
        for (int j = 0; j < n; ++j) a[static_cast<std::size_t>(i) * n + j] = A(i, j);
    }
    if (csim_host_gauss_seidel(n, 1, a.data(), rhs.data(), sol.data(), maxIters, tol, nullptr))
        throw std::runtime_error(std::string("Solver: ") + csim_last_error());
    for (int i = 0; i < n; ++i) x(i) = sol[i];
    return x;
}
VectorXd solveLinearSystemGaussSeidel(const MatrixXd& A, const VectorXd& b, int maxIters, double tol) {
    return solveLinearSystemGaussSeidel(A, b, VectorXd::Zero(b.size()), maxIters, tol);
}

}  // namespace Solver

Eigen::VectorXd dcSolveLU(const Circuit& ckt) {
    const int N = ckt.numUnknowns();
    Eigen::VectorXd x = Eigen::VectorXd::Zero(N > 0 ? N : 1);
    if (N == 0) {
        std::cerr << "DC solve (LU): no unknowns.\n";
        return x;
    }
    csim::Lowered L = csim::lower(ckt);
    std::vector<double> sol = csim::dcSolveBatch(ckt, L.nominal, 1);
    x.setZero(N);
    for (int i = 0; i < N; ++i) x(i) = sol[i];
    return x;
}

Eigen::VectorXd dcSolve(const Circuit& ckt) { return dcSolveLU(ckt); }  // reference: src/dcanalysis.cpp:260-262

namespace {

// ||v||_2 in the summation order the kernels use (32 lane-strided partial sums, halving tree)
double norm2_defined(const std::vector<double>& v) {
    double acc[32];
    for (double& a : acc) a = 0.0;
    for (std::size_t i = 0; i < v.size(); ++i) {
        const double sq = v[i] * v[i];
        acc[i & 31] = acc[i & 31] + sq;
    }
    for (int s = 16; s >= 1; s >>= 1)
        for (int i = 0; i < s; ++i) acc[i] = acc[i] + acc[i + s];
    return std::sqrt(acc[0]);
}

bool hasNonlinearDevices(const Circuit& ckt) {  // src/dcanalysis.cpp:26-33
    for (const auto& e : ckt.elements)
        if (dynamic_cast<const MosfetBase*>(e.get())) return true;
    return false;
}

// One DC assembly on the device (csim_host_stamp_csr: src/dcanalysis.cpp:120-130 with the given gmin; gmin < 0: the
// linear direct path, no gmin stamp) scattered into a dense row-major G.
struct DcAssembler {
    csim::Lowered L;
    csim_plan* plan = nullptr;
    std::vector<int32_t> rowptr, colidx;
    int N = 0;
    explicit DcAssembler(const Circuit& ckt) : L(csim::lower(ckt)) {
        csim_topology t = L.view();
        if (csim_plan_create(&t, nullptr, 0, &plan)) throw std::runtime_error(std::string("dcSolveGaussSeidel: ") + csim_last_error());
        N = L.n_unknowns;
        int32_t nnz = 0;
        csim_plan_pattern(plan, CSIM_MODE_DC, &nnz, nullptr, nullptr);
        rowptr.resize(N + 1);
        colidx.resize(nnz);
        csim_plan_pattern(plan, CSIM_MODE_DC, &nnz, rowptr.data(), colidx.data());
    }
    ~DcAssembler() { csim_plan_destroy(plan); }
    void assemble(const std::vector<double>& x, double scale, double gmin, std::vector<double>& G, std::vector<double>& I) {
        std::vector<double> vals(colidx.size());
        I.assign(N, 0.0);
        if (csim_host_stamp_csr(plan, CSIM_MODE_DC, 1, L.nominal.data(), x.data(), nullptr, nullptr, 0.0, scale, gmin,
                                vals.data(), I.data()))
            throw std::runtime_error(std::string("dcSolveGaussSeidel: ") + csim_last_error());
        G.assign(static_cast<std::size_t>(N) * N, 0.0);
        for (int i = 0; i < N; ++i)
            for (int k = rowptr[i]; k < rowptr[i + 1]; ++k) G[static_cast<std::size_t>(i) * N + colidx[k]] = vals[k];
    }
};

void gsSolve(int N, const std::vector<double>& G, const std::vector<double>& I, std::vector<double>& x, int maxIters, double tol) {
    if (csim_host_gauss_seidel(N, 1, G.data(), I.data(), x.data(), maxIters, tol, nullptr))
        throw std::runtime_error(std::string("dcSolveGaussSeidel: ") + csim_last_error());
}

// src/dcanalysis.cpp:71-92: one assembly at scale 1 (no gmin), Gauss-Seidel from the zero vector, 2000 sweeps
Eigen::VectorXd dcSolveDirectGS(const Circuit& ckt) {
    const int N = ckt.numUnknowns();
    Eigen::VectorXd x = Eigen::VectorXd::Zero(N);
    if (N == 0) {
        std::cerr << "DC solve (GS): no unknowns.\n";
        return x;
    }
    DcAssembler as(ckt);
    std::vector<double> G, I, xs(N, 0.0);
    as.assemble(xs, 1.0, 0.0, G, I);
    // (gmin 0.0 adds +0.0 to the node diagonals: exact, and the direct path's matrix never holds a -0 diagonal
    //  that the addition could turn into +0 — every diagonal is a sum starting from +0)
    gsSolve(N, G, I, xs, 2000, 1e-10);
    for (int i = 0; i < N; ++i) x(i) = xs[i];
    return x;
}

// src/dcanalysis.cpp:166-237: the source ramp with Gauss-Seidel as the inner solver (warm start from x), 60 Newton
// iterations per ramp step and 120 on the last, the same ConvController (264-307), the same warnings.  The
// reference's own result on circuits with voltage sources is a vector near zero after ~660 "non-finite"
// warnings (zero diagonals on the branch rows, SURVEY.md §8 row a-); this does the same thing, bit for bit.
Eigen::VectorXd dcSolveNewtonGS(const Circuit& ckt) {
    const int N = ckt.numUnknowns();
    Eigen::VectorXd xo = Eigen::VectorXd::Zero(std::max(N, 1));
    if (N == 0) {
        std::cerr << "DC solve (Newton + GS): no unknowns.\n";
        return xo;
    }
    csim_constants c;
    csim_default_constants(&c);
    const int rampSteps = 10, maxNewtonIters = 60;
    const double tol = 1e-9;
    DcAssembler as(ckt);
    std::vector<double> x(N, 0.0), G, I, xRaw, xNew(N), d(N);
    auto baseGmin = [&](double s) {  // include/dcanalysis.hpp:45-48
        s = std::min(std::max(s, 0.0), 1.0);
        return c.gmin_high_base * (1.0 - s) + c.gmin_low_base * s;
    };
    for (int step = 1; step <= rampSteps; ++step) {
        const double scale = static_cast<double>(step) / rampSteps;
        double alpha = 0.7;  // initialAlphaGS(); dead like alphaNext: update() always restarts from 0.35
        double gmin = baseGmin(scale);
        double prevErr = std::numeric_limits<double>::infinity();
        const int maxIterThisStep = step == rampSteps ? maxNewtonIters * 2 : maxNewtonIters;
        for (int iter = 0; iter < maxIterThisStep; ++iter) {
            as.assemble(x, scale, gmin, G, I);
            xRaw = x;  // warm start
            gsSolve(N, G, I, xRaw, 2000, 1e-10);
            bool finite = true;
            for (double v : xRaw) finite = finite && std::isfinite(v);
            if (!finite) {
                gmin = std::min(gmin * 10.0, 1e-2);
                std::cerr << "WARNING: GS produced non-finite x, increasing gmin to " << gmin << " at ramp step " << step
                          << ", iter " << iter << "\n";
                continue;
            }
            // ConvController::update (src/dcanalysis.cpp:268-307)
            double a = std::min(std::max(0.35, 0.1), 0.5);
            for (int i = 0; i < N; ++i) {
                xNew[i] = x[i] + a * (xRaw[i] - x[i]);
                d[i] = xNew[i] - x[i];
            }
            const double err = norm2_defined(d);
            const double gminBase = baseGmin(scale);
            double gminNext = gminBase;
            if (!(iter == 0 || !std::isfinite(prevErr))) {
                if (err > prevErr * c.slow_conv_ratio) {
                    a = std::max(a * 0.7, 0.1);
                    gminNext = std::min(gmin * 2.0, c.gmin_abs_max);
                } else if (err < prevErr * c.fast_conv_ratio) {
                    a = std::min(a * 1.1, 0.5);
                    gminNext = 0.5 * gmin + 0.5 * gminBase;
                } else {
                    gminNext = 0.7 * gmin + 0.3 * gminBase;
                }
            }
            x = xNew;
            alpha = a;
            gmin = gminNext;
            prevErr = err;
            if (err < tol) break;
            if (iter == maxNewtonIters - 1)
                std::cerr << "WARNING: Newton (GS) did not converge at ramp step " << step << " (err=" << err
                          << ", alpha=" << alpha << ", gmin=" << gmin << ")\n";
        }
    }
    for (int i = 0; i < N; ++i) xo(i) = x[i];
    return xo;
}

}  // namespace

// src/dcanalysis.cpp:251-257
Eigen::VectorXd dcSolveGaussSeidel(const Circuit& ckt) {
    return hasNonlinearDevices(ckt) ? dcSolveNewtonGS(ckt) : dcSolveDirectGS(ckt);
}

Eigen::VectorXd computeDcOperatingPoint(const Circuit& ckt) { return dcSolve(ckt); }

void runTransientAnalysisBackwardEuler(const Circuit& ckt, const SimulationConfig& sim, const std::string& outFile) {
    const TranConfig& cfg = sim.tran;
    if (!cfg.enabled) {
        std::cerr << "Transient analysis is not enabled (.TRAN missing).\n";
        return;
    }
    if (cfg.tstep <= 0.0 || cfg.tstop <= 0.0) {
        std::cerr << "Invalid .TRAN card: tstep and tstop must be > 0.\n";
        return;
    }
    const int N = ckt.numUnknowns();
    if (N <= 0) {
        std::cerr << "Transient: circuit has no unknowns.\n";
        return;
    }
    csim::Lowered L = csim::lower(ckt);
    std::FILE* f = std::fopen(outFile.c_str(), "w");
    if (!f) {
        std::cerr << "Cannot open transient output file '" << outFile << "'.\n";
        return;
    }
    std::vector<int32_t> probes(N);
    for (int i = 0; i < N; ++i) probes[i] = i;
    csim::BatchResult R;
    try {
        R = csim::runTransientAnalysisBackwardEulerBatch(ckt, sim, L.nominal, 1, probes, 1);
    } catch (...) {
        std::fclose(f);
        throw;
    }
    std::printf("[TRAN] tstep=%e, tstop=%e, tstart=%e\n[TRAN] total steps = %d\n", cfg.tstep, cfg.tstop, cfg.tstart,
                R.n_steps);
    const std::vector<int32_t> cols = L.csv_columns();
    std::fprintf(f, "%s\n", L.csv_header().c_str());
    auto row = [&](double t, auto value_of) {
        std::fprintf(f, "%.9e", t);
        for (int32_t c : cols) std::fprintf(f, ",%.9e", value_of(c));
        std::fputc('\n', f);
    };
    if (!(0.0 < cfg.tstart)) row(0.0, [&](int32_t c) { return R.x_dc[c]; });
    const int first = R.n_steps - R.n_out;  // out_stride == 1
    const bool dead = (R.status[0] & CSIM_ST_NONFINITE) != 0;
    for (int r = 0; r < R.n_out; ++r) {
        const double v0 = R.wave[static_cast<std::size_t>(cols[0]) * R.n_out + r];
        if (dead && std::isnan(v0)) break;  // rows after the failure were never produced
        row((first + r + 1) * cfg.tstep, [&](int32_t c) { return R.wave[static_cast<std::size_t>(c) * R.n_out + r]; });
    }
    std::fclose(f);
    if (dead) throw std::runtime_error("Transient: LU produced NaN/Inf.");
    std::printf("Transient analysis (Backward Euler) finished. Results written to '%s'.\n", outFile.c_str());
}
