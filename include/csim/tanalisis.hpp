// tanalisis.hpp — transient driver of the reference (include/tanalisis.hpp; the file name
// keeps the reference's spelling).  B = 1 through the CUDA path; CSV identical in layout
// to src/tanalisis.cpp:183-231 ("time,V(node)...,I(Vsrc|L)...", %.9e, rows with t >= tstart).
#pragma once
#include <string>

#include "circuit.hpp"
#include "dense.hpp"
#include "sim.hpp"

Eigen::VectorXd computeDcOperatingPoint(const Circuit& ckt);
void runTransientAnalysisBackwardEuler(const Circuit& ckt, const SimulationConfig& sim, const std::string& outFile);
