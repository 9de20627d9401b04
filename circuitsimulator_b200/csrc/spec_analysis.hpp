// spec_analysis.hpp — symbolic analysis shared by the kernel generators (internal).
#pragma once
#include <cstdint>
#include <vector>

#include "plan.hpp"

namespace csim {

struct Term {
    int value;
    bool sub;
};

// Assembly (TRAN by default, or DC) + LU of one topology under one (predicted) pivot permutation.  "Position" i holds
// original row perm[i]; columns keep their original numbering.
struct Analysis {
    int N = 0;
    std::vector<int> pos_of_row;
    // per (position, column): ordered contributions; column N = RHS
    std::vector<std::vector<std::vector<Term>>> slot;       // [pos][col]
    std::vector<std::vector<std::vector<Term>>> rest;       // terms added every Newton iteration
    std::vector<std::vector<std::vector<Term>>> prefix;     // terms folded into the static base
    std::vector<int> K;                                     // [N+1] generated adds per column (N = RHS)
    std::vector<std::vector<bool>> pat;                     // structural pattern incl. fill, permuted
    std::vector<std::vector<int>> rows_k, cols_k, ucols;    // elimination rows / pivot-row columns / U columns
    std::vector<bool> col_used_above;
    // before_pivot[k][i]: at elimination step k the reference's pivot search (include/solver.hpp:46-56: positions k..N-1
    // in its CURRENT row order, first maximum wins, strict '>') meets the row that ends at position i BEFORE the row
    // the permutation predicts as pivot.  Such a row takes the pivot on a TIE, so the prediction check must reject
    // |a(i,k)| >= |pivot| for it and |a(i,k)| > |pivot| for the rows scanned after the pivot.
    std::vector<std::vector<bool>> before_pivot;
};

// Throws std::runtime_error when perm is not a permutation of 0..N-1.
// mode CSIM_MODE_DC: the gmin contribution is treated as iterate-dependent (the convergence controller changes it
// between Newton iterations, src/dcanalysis.cpp:268-307).
Analysis analyse(const HostPlan& hp, const std::vector<int32_t>& perm, int mode = CSIM_MODE_TRAN);

}  // namespace csim
