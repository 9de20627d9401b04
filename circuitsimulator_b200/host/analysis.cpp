// analysis.cpp — reference-named single-instance entry points (Solver::*, dcSolve*,
// computeDcOperatingPoint, runTransientAnalysisBackwardEuler) on top of the batched CUDA
// path with B = 1.  Error behaviour follows the reference: messages on std::cerr and early
// return; std::runtime_error only for the transient NaN/Inf case (src/tanalisis.cpp:360-362)
// and — new — when no CUDA device is available (this library has no CPU solver).
#include <cmath>
#include <cstdio>
#include <iostream>
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

bool luDecompose(const MatrixXd& A, MatrixXd& LU, std::vector<int>& perm) {
    const int n = static_cast<int>(A.rows());
    if (n == 0) return false;
    if (A.cols() != n) {
        std::cerr << "LU: matrix is not square.\n";
        return false;
    }
    const bool ok = deviceSolve(A, nullptr, nullptr, perm);
    if (!ok) {
        std::cerr << "LU: zero (or tiny) pivot.\n";
        return false;
    }
    LU = MatrixXd(n, n);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) LU(i, j) = A(perm[i], j);
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

Eigen::VectorXd dcSolveGaussSeidel(const Circuit&) {
    throw std::runtime_error("dcSolveGaussSeidel: not implemented on the device (SURVEY.md §8(f) rank 4)");
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
