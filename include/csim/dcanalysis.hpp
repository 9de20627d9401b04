// dcanalysis.hpp — DC operating point entry points of the reference (include/dcanalysis.hpp).
// dcSolve / dcSolveLU lower the circuit and run the CUDA DC kernel with B = 1.
#pragma once
#include "dense.hpp"

class Circuit;

Eigen::VectorXd dcSolve(const Circuit& ckt);
Eigen::VectorXd dcSolveLU(const Circuit& ckt);
// Gauss-Seidel variant (src/dcanalysis.cpp:71-92,166-237,251-257; off the reference's default path): the direct
// solve for linear circuits, the source ramp with 60 (120 on the last step) Newton iterations otherwise; assembly
// and the Gauss-Seidel sweeps run on the device, the warnings and the (possibly unconverged) result are the
// reference's.
Eigen::VectorXd dcSolveGaussSeidel(const Circuit& ckt);
