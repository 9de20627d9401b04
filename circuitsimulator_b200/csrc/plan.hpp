// plan.hpp — topology -> stamp programs (host side of csim_plan).
//
// The reference assembles G and I by walking polymorphic elements and adding
// scalars into a dense matrix (DC: src/dcanalysis.cpp:120-130; TRAN:
// src/tanalisis.cpp:259-356).  Which scalar lands in which G(i,j)/I(i), and in
// which ORDER, is a pure function of the topology.  The plan records it once:
//
//   value table  : every distinct scalar an assembly adds (1/R, C/dt, -L/dt, 1,
//                  gmin, cap/inductor history terms, source values, and per MOSFET
//                  gd gg gs cst), with a recipe ("value op") per entry;
//   contributions: (row, col|RHS, value, sign) in the reference's order.
//
// Kernels re-evaluate value ops and then replay contributions per row (fused
// Newton kernels) or per CSR slot (standalone stamp kernel); accumulation order
// per slot is the reference's, so sums are bit-identical.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/csim_b200.h"
#include "plan_pod.h"

namespace csim {

struct Contribution {
    int32_t row, col;  // col == -1 : RHS
    int32_t value;
    int32_t sub;  // 1: G -= value
};

struct ModeProgram {
    int32_t n_values = 0;
    std::vector<StaticOp> static_ops;
    std::vector<StepOp> step_ops;
    std::vector<SrcOp> src_ops;
    std::vector<MosOp> mos_ops;
    std::vector<Contribution> contribs;  // reference order
    // structural pattern of G (CSR, rows/cols ascending)
    std::vector<int32_t> rowptr, colidx;
    // per-slot ordered lists for the standalone stamp kernel: slots 0..nnz-1 = G entries in CSR
    // order, slots nnz..nnz+N-1 = RHS rows
    std::vector<int32_t> slot_ptr;
    std::vector<uint32_t> slot_entries;  // contrib_word(col ignored)
    // per-row ordered lists for the fused kernels, [row_len_max][32]
    int32_t row_len_max = 0;
    std::vector<uint32_t> row_prog;  // 0xFFFFFFFF = no-op padding
};

struct HostPlan {
    int32_t N = 0, n_node_eq = 0, n_params = 0, n_sources = 0, n_mos = 0;
    bool has_mos = false;
    csim_constants consts{};
    ModeProgram mode[2];  // CSIM_MODE_DC, CSIM_MODE_TRAN
    std::vector<int32_t> elem_type, elem_flags, elem_nodes, elem_branch, elem_param;
};

// Throws std::runtime_error with a message on inconsistent topologies.
HostPlan build_host_plan(const csim_topology& topo, const csim_constants* consts);

}  // namespace csim
