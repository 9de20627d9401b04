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

namespace csim {

// ---- value ops ------------------------------------------------------------
enum StaticKind : int32_t {
    SK_ONE = 0,       // 1.0
    SK_GMIN = 1,      // gmin (rewritten per Newton iteration in DC)
    SK_RES = 2,       // R==0 ? 0 : 1.0/R                 (src/element.cpp:20-24)
    SK_CAP = 3,       // C<=0||dt<=0 ? 0 : C/dt           (src/tanalisis.cpp:65-67)
    SK_CAPHALF = 4,   // c=0.5*Cj0; same rule             (src/tanalisis.cpp:338-339)
    SK_IND_ONE = 5,   // L<=0 ? 0 : 1.0                   (src/tanalisis.cpp:296, 311-315)
    SK_IND_NREQ = 6,  // L<=0 ? 0 : -(L/dt)               (src/tanalisis.cpp:304, 317)
};
struct StaticOp {  // out = f(kind, params[param])
    int32_t kind, param, out, _pad;
};
enum StepKind : int32_t {
    TK_CAP_HIST = 0,  // out = (-vt[g]) * (xprev[a] - xprev[b])   I_hist, src/tanalisis.cpp:77
    TK_IND_HIST = 1,  // out = vt[g] * xprev[a]                   V_hist = -R_eq*iPrev, src/tanalisis.cpp:305-306
};
struct StepOp {
    int32_t kind, g, a, b, out, _p0, _p1, _p2;  // a, b: equation index or -1 (ground -> 0.0)
};
struct SrcOp {  // one per independent source, element order
    int32_t out, param, is_sin, index;
};
struct MosOp {  // out..out+3 = gd, gg, gs, cst   (src/element.cpp:181-274)
    int32_t out, param, d, g, s, is_p, _p0, _p1;
};

// contribution word: bits 0-5 column (32 = RHS), bit 6 subtract, bits 8.. value index
inline uint32_t contrib_word(int col, bool sub, int v) { return (uint32_t)col | (sub ? 64u : 0u) | ((uint32_t)v << 8); }
constexpr int kRhsCol = 32;

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
