// solver.hpp — namespace Solver of the reference (include/solver.hpp), device-backed.
// luDecompose / solveLinearSystemLU keep their signatures and error behaviour (false / zero
// vector + message on std::cerr, no exceptions for singular systems) but run the batched
// CUDA LU (csim_dev_lu_factor / csim_dev_lu_solve) with B = 1: on chip for n <= 32, the dense
// large-matrix kernels up to n = 1280 (luDecompose) / 1536 (solve).  They throw
// std::runtime_error only when there is no CUDA device (there is no CPU solver in this library).
// solveLinearSystemGaussSeidel (on the reference's interface, off its default path:
// src/dcanalysis.cpp:260-262) runs csim_dev_gauss_seidel with B = 1, n <= 200, bit-identical sweeps.
#pragma once
#include <vector>

#include "dense.hpp"

namespace Solver {
using Eigen::MatrixXd;
using Eigen::VectorXd;

enum class LinearSolver { DirectLU, GaussSeidel };

// perm[i] = original row at position i; LU = the reference's packed factors (include/solver.hpp:64-76): U on and
// above the diagonal, multipliers in the strict lower triangle, rows exchanged physically.
bool luDecompose(const MatrixXd& A, MatrixXd& LU, std::vector<int>& perm);
VectorXd solveLinearSystemLU(const MatrixXd& A, const VectorXd& b);
VectorXd solveLinearSystemGaussSeidel(const MatrixXd& A, const VectorXd& b, const VectorXd& x0, int maxIters = 1000,
                                      double tol = 1e-10);
VectorXd solveLinearSystemGaussSeidel(const MatrixXd& A, const VectorXd& b, int maxIters = 1000, double tol = 1e-10);
}  // namespace Solver
