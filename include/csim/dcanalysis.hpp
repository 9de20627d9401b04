// dcanalysis.hpp — DC operating point entry points of the reference (include/dcanalysis.hpp).
// dcSolve / dcSolveLU lower the circuit and run the CUDA DC kernel with B = 1.
#pragma once
#include "dense.hpp"

class Circuit;

Eigen::VectorXd dcSolve(const Circuit& ckt);
Eigen::VectorXd dcSolveLU(const Circuit& ckt);
// Gauss-Seidel variant: not on the reference's default path; throws std::runtime_error
// ("not implemented on the device") — SURVEY.md §8(f) rank 4.
Eigen::VectorXd dcSolveGaussSeidel(const Circuit& ckt);
