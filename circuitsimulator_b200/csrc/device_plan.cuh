// device_plan.cuh — device-side view of a csim_plan (POD, passed to kernels by value).
#pragma once
#include <stdint.h>

#include "../../include/csim_b200.h"
#include "plan_pod.h"

namespace csim {

struct DevMode {
    int32_t n_values, n_static, n_step, n_src, n_mos, row_len, nnz, n_slots;
    const StaticOp* static_ops;
    const StepOp* step_ops;
    const SrcOp* src_ops;
    const MosOp* mos_ops;
    const uint32_t* row_prog;      // [row_len][32]
    const int32_t* slot_ptr;       // [n_slots+1]
    const uint32_t* slot_entries;  // contribution words per slot, reference order
    // every contribution in the reference's order: word = contrib_word(col | RHS, sub, value), row separately
    int32_t n_contribs, _pad;
    const uint32_t* contrib_words;  // (N <= 32 only: the column field has 6 bits)
    const int32_t* contrib_rows;
    // row / column of every slot (column -1: right-hand side), for the large-matrix path
    const int32_t* slot_row;
    const int32_t* slot_col;
};

struct DevPlan {
    int32_t N, n_node_eq, n_params, n_sources, has_mos, _pad;
    csim_constants c;
    DevMode m[2];
};

// transient launch description (device pointers)
struct DevTran {
    double dt;
    int32_t step_begin, step_end;
    int32_t table_steps, n_probes;
    int32_t out_stride, out_begin_step;
    int64_t wave_ld;
    const double* src_table;     // [n_groups][table_steps]
    const int32_t* src_mode;     // [n_sources]
    const int32_t* src_group;    // [n_sources]
    const int32_t* src_group_b;  // [n_sources][B] or nullptr
    const int32_t* probes;       // [n_probes]
    const uint8_t* skip;         // [B] or nullptr: instances with skip[b] != 0 are left untouched (steady-state driver)
    unsigned long long* miss;    // thread-per-instance kernel: log of the permutations its redo path met, or nullptr
    double* margin;              // [B] or nullptr (generic kernels only): min over every comparison of a step norm with a
                                 // threshold of |err - threshold| / threshold (csim_dev_margin_probe)
};

// Device-side tables of a specialised kernel (built by specialize.cpp).
struct SpecTables {
    const uint16_t* gidx;      // [n_adds][32]  index into vv (value/negated-value pairs) per lane
    const uint32_t* prefix;    // [prefix_len][32] contribution words (col | sub<<6 | value<<8), 0xFFFFFFFF = none
    const int32_t* mos_lane;   // [32][8]: param offset, d, g, s, is_p, out, -, -   (lane m = m-th MOSFET)
    int32_t n_adds, prefix_len, n_mos, n_values;
    uint32_t hash_lo, hash_hi; // sum_i mix(i, perm[i]) of the predicted permutation
};

}  // namespace csim
