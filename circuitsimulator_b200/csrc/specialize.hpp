// specialize.hpp — run-time code generation of the transient kernel for ONE netlist topology
// and ONE predicted pivot permutation (host side).
//
// Why: ncu on the generic kernel (profiles/r1_v1_*) shows 9,557 warp-instructions per Newton
// iteration of which ~13 % are FP64 — the rest is index arithmetic, loop control and pivot
// bookkeeping that depend only on the topology.  Given the topology and the permutation the
// reference's partial pivoting is expected to pick, every address, loop trip count and
// structural zero is known, so the generator emits straight-line CUDA with the matrix rows in
// registers and only the structurally non-zero operations.  The permutation is a PREDICTION:
// the generated code verifies the reference's first-max-wins rule at every elimination step
// and re-does the solve with the generic pivoting path when it does not hold, so results are
// bit-identical to the generic kernel in every case.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "plan.hpp"

namespace csim {

struct SpecHost {
    std::vector<int32_t> perm;       // predicted: perm[i] = original row at position i
    std::string source;              // CUDA C++ translation unit (entry: csim_tran_spec)
    std::vector<uint16_t> gidx;      // [n_adds][32]
    std::vector<uint32_t> prefix;    // [prefix_len][32]
    std::vector<int32_t> mos_lane;   // [32][8]
    int32_t n_adds = 0, prefix_len = 0, n_mos = 0;
    uint32_t hash_lo = 0, hash_hi = 0;
    int32_t warps_per_block = 4, blocks_per_sm = 4;
    // statistics of the generated code (per Newton iteration)
    int32_t n_elim_updates = 0, n_lu_steps = 0, n_back_terms = 0, n_g_adds = 0, n_rhs_adds = 0;
};

// Why a topology cannot be specialised (empty string = it can).
std::string spec_unsupported_reason(const HostPlan& hp);

// perm must be a permutation of 0..N-1.  Throws std::runtime_error on inconsistent input.
SpecHost generate_tran_spec(const HostPlan& hp, const std::vector<int32_t>& perm);

// Per-block dynamic shared memory of the generated kernel.
size_t spec_smem_bytes(const HostPlan& hp, const SpecHost& sh);

// ---- thread-per-instance transient kernel (specialize_tpi.cpp, tpi_runtime.cuh) -----------------
struct TpiHost {
    std::vector<int32_t> perm;                 // variant 0
    std::vector<std::vector<int32_t>> perms;   // all variants
    std::vector<uint64_t> hashes;              // pivot-hash contribution of each variant's permutation
    std::string source;  // entry: csim_tran_tpi
    int32_t threads = 32, replicas = 1, n_slots = 0, n_scratch_rows = 0;
    int32_t inst_per_cta = 32;  // instances one CTA advances: threads x replicas, or 32 with a warp PAIR per 32 instances
    bool pair = false;          // two warps per 32 instances (TpiOptions::pair)
    int32_t pair_events = 0;    // cross-warp hand-overs per Newton iteration (named barriers)
    int32_t pair_cost[3] = {0, 0, 0};  // cost model: LU phase of one warp alone, and of the master / helper warp of the pair
    size_t smem_bytes = 0;
    uint32_t hash_lo = 0, hash_hi = 0;
    // statistics of the generated code (per Newton iteration, per instance)
    int32_t n_elim_updates = 0, n_lu_divs = 0, n_back_terms = 0, n_g_adds = 0, n_rhs_adds = 0, n_rcp = 0;
    // longest dependency chain of one Newton iteration in cycles (operation latencies of
    // profiles/r1_microbench_fp64.txt): what one thread cannot go below however its code is scheduled
    int32_t crit_cycles = 0, crit_lu_cycles = 0;
};
struct TpiOptions {
    bool const_prop = true;  // track +-1 / gmin entries symbolically (exact rewrites only)
    bool x_in_smem = true;   // Newton iterate in shared memory instead of registers
    bool mos_in_smem = false;    // MOSFET parameters and linearisation values parked in shared memory
    bool q_in_smem = false;      // raw solution parked in the (dead) Y slots
    bool mos_branchless = false; // evaluate both MOSFET regions and select (measured slower on dbmixer)
    bool fold_guards = true;     // divisions whose numerator cannot be an exact zero: guards folded into one unsigned accumulator, pivot rule judged on the quotient (tdiv_nz / tdiv_mul)
    bool int_pivot_guard = false; // (with fold_guards) pivot range test on the high word in an unsigned accumulator instead of two FP64 compares per row
    bool cand_retry_tran = false; // ... the same in the transient kernel (B200, buffer.sp, 65,536 instances: 210 ms against 117 ms with the successor hint: resolved misses stop the learning of further variants and every retry costs the warp extra bodies)
    bool cand_retry = true;      // DC kernel, several variants (buffer.sp DC ramp, 65,536 instances: 61 ms against 85 ms): a lane whose variant fails at elimination step k retries variants with the same pivot prefix and another pivot at k (else: one global successor hint, folded pivot guards)
    bool opaque_park = false;    // parked LU rows (pivot, reciprocal, U entries, y) stored by asm blocks and re-read by asm loads: the compiler cannot forward the stored registers to later rows (long live ranges, spills)
    int rows_in_flight = 2;      // matrix rows whose elimination chains are interleaved in the generated source (B200, dbmixer: 1 -> 111.4, 2 -> 113.9, 3 -> 113.4, 4 -> slower; M instance-timepoints/s)
    int warps_per_cta = 0;       // 0 = whatever packs most warps into an SM's shared memory (1, 2 or 4)
    int replicas = 1;            // instances advanced together by one thread (interleaved instruction streams)
    // Two warps per 32 instances (transient kernel, one predicted permutation): lane l of the MASTER warp runs the
    // Newton loop of instance l as before but leaves part of the matrix rows (and the MOSFET linearisations only they
    // need) to lane l of the HELPER warp; parked rows cross through the shared-memory slots they already live in,
    // ordered by named barriers (bar.arrive / bar.sync).  Which warp computes a row changes no operand and no
    // operation order: bit-identical.  Why: one warp per SM sub-partition is bound by its own dependency stalls
    // (1.77 stall cycles per issue), and shared memory, not registers, caps residency at 3-4 warps per SM.
    bool pair = false;
    // One-warp-per-32-instances kernel in CTAs of several warps: the warps vote TOGETHER on the end of every Newton
    // loop (__syncthreads_and), so they run the straight-line body in step and share the instruction lines one of
    // them fetched (the body is larger than the SM's instruction cache).  Costs the iterations a warp idles while
    // another warp of its CTA still iterates, and a generic redo stalls the whole CTA.
    // Measured on B200, dbmixer, 18,944 instances (4 warps per SM): 228.2 M (2-warp CTAs) / 227.3 M (4-warp CTAs) in step
    // against 231.6 / 231.1 M free-running — sharing fetches does not pay where every sub-partition is busy: off.
    bool lockstep = false;
};
TpiHost generate_tran_tpi(const HostPlan& hp, const std::vector<int32_t>& perm, const TpiOptions& opt = TpiOptions());
TpiHost generate_tran_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms,
                          const TpiOptions& opt = TpiOptions());
// The same generator for the DC operating point (dcSolveNewtonLU, src/dcanalysis.cpp:95-163): entry csim_dc_tpi.
// Source ramp and ConvController on the device per lane, gmin a per-lane variable, x starts at zero.
TpiHost generate_dc_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms,
                        const TpiOptions& opt = TpiOptions());

// Device-side log of the permutations the generic redo path met (tpi_generic_solve): kTpiMissEntries entries of
// kTpiMissStride uint64 = { pivot-hash contribution (0 = free), count, perm[0..31] }.
// Behind them kTpiSuccEntries uint64: succ[v] = 1 + the variant a lane moved to when it last left variant v (0 = none).
constexpr int kTpiMissEntries = 16, kTpiMissStride = 34, kTpiSuccEntries = 16;
constexpr int kTpiMissWords = kTpiMissEntries * kTpiMissStride + kTpiSuccEntries;

// ---- standalone assembly kernel (specialize_stamp.cpp) ---------------------------------------
struct StampSpecHost {
    int32_t mode = 0;
    std::string source;  // CUDA C++ translation unit (entry: csim_stamp_spec)
    int32_t threads = 128, blocks_per_sm = 3;
};
// mode: CSIM_MODE_DC / CSIM_MODE_TRAN.  Any topology (no N <= 32 restriction).
StampSpecHost generate_stamp_spec(const HostPlan& hp, int mode);

// Distinct doubles one assembly reads / writes per instance (what the kernels — generic and
// specialised alike — must move; repeated reads of one input hit L1).
struct StampTraffic {
    int64_t doubles_read = 0, doubles_written = 0;
};
StampTraffic stamp_traffic(const HostPlan& hp, int mode, bool has_src_val);

}  // namespace csim
