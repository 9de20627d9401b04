/* csim_oracle.h — TEST INFRASTRUCTURE.  CPU restatement of the reference's
 * Newton/MNA path over the flat topology of include/csim_b200.h.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use
 * anything under oracle/.  It is the CHECKER, never the product: the product
 * path (circuitsimulator_b200/) fails loudly without its CUDA library.
 *
 * Parity status: PINNED against the unmodified reference sources compiled into
 * oracle/_ref (see oracle/Makefile, oracle/ref_harness.cpp): every raw LU
 * solution of every Newton iteration, every pivot swap and every Newton step
 * norm of the unmodified drivers is compared bit-for-bit with this restatement
 * (tests/golden/make_golden.py, run in the build container).  One call is
 * "parity unpinned" against UPSTREAM Eigen: VectorXd::norm() — its summation
 * order is defined in oracle/csim_norm2.h (see there).
 */
#ifndef CSIM_ORACLE_H
#define CSIM_ORACLE_H

#include <stdint.h>
#include "../include/csim_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Optional recorder filled by oracle_dc / oracle_tran. All pointers may be NULL. */
typedef struct oracle_trace {
    double* xraw;        /* [cap_iters][N] raw LU solution of every Newton iteration (what allFinite() sees) */
    double* err;         /* [cap_iters]    Newton step norm of every iteration that reached the update      */
    int32_t* perm;       /* [cap_iters][N] pivot permutation of every factorisation                          */
    int64_t cap_iters;
    int64_t n_iters;     /* out: Newton iterations (LU solves) performed                                     */
    int64_t n_err;       /* out: entries written to err                                                      */
    double min_margin;   /* out: min over all threshold comparisons of |err-thr|/thr  (init to +inf)         */
} oracle_trace;

/* One assembly pass into DENSE row-major G[N*N], I[N].
 * DC  : src/dcanalysis.cpp:120-130 ; TRAN: src/tanalisis.cpp:259-356.
 * src_val[n_sources] = value of each independent source in element order
 * (V and I sources together, in netlist order). */
void oracle_stamp(const csim_topology* t, const csim_constants* c, int mode, const double* params,
                  const double* x, const double* x_prev, const double* src_val, double dt, double gmin,
                  double* G, double* I);

/* evalDC(scale) of every source (include/sim.hpp:152-158). */
void oracle_src_dc(const csim_topology* t, const double* params, double scale, double* src_val);
/* evalTran of every source at step s via the table convention of csim_src_mode. */
void oracle_src_tran(const csim_topology* t, const double* params, const int32_t* src_mode,
                     const int32_t* src_group, const double* table, int32_t table_steps, int32_t step,
                     double dt, double* src_val);
/* S = sin(2*pi*freq*(t-td)+phi) at t=(s+1)*dt, s in [0,n_steps)  (include/sim.hpp:117-122). */
void oracle_sin_table(double freq, double td, double phi, double dt, int32_t n_steps, double* out);
/* CSIM_SRC_VAL_TABLE row: out[s] = TranWaveform::eval((s+1)*dt) for a PULSE (type 1, spec = v1 v2 td tr tf ton per)
 * or PWL (type 3, spec = n times then n values) waveform  (include/sim.hpp:80-115,124-138). */
void oracle_waveform_table(int32_t type, const double* spec, int32_t n, double dt, int32_t n_steps, double* out);

/* Solver::luDecompose (include/solver.hpp:30-80): packed L\U + perm; returns 1 when the reference returns false. */
int oracle_lu_factor(int32_t n, const double* A, double* LU, int32_t* perm, double lu_eps);
/* Solver::solveLinearSystemLU (include/solver.hpp:30-131). A is row-major N*N
 * (not modified), returns 1 when the factorisation failed (x = 0 then). */
int oracle_lu_solve(int32_t n, const double* A, const double* b, double* x, int32_t* perm, double lu_eps);
/* Solver::solveLinearSystemGaussSeidel (include/solver.hpp:139-192): x in = x0, out = last iterate; returns sweeps done */
int32_t oracle_gauss_seidel(int32_t n, const double* A, const double* b, double* x, int32_t max_iters, double tol);

/* dcSolveLU (src/dcanalysis.cpp:242-248). iters[dc_ramp_steps] (NULL ok). */
void oracle_dc(const csim_topology* t, const csim_constants* c, const double* params, double* x,
               int32_t* iters, uint64_t* pivot_hash, uint32_t* status, oracle_trace* tr);

/* Steps [step_begin, step_end) of runTransientAnalysisBackwardEuler
 * (src/tanalisis.cpp:255-420). x in/out [N]. wave_all [step_end-step_begin][N]
 * (NULL ok) receives x after every step. iters_step [..] uint8 (NULL ok).
 * Returns the total number of Newton iterations. */
int64_t oracle_tran(const csim_topology* t, const csim_constants* c, const double* params, double dt,
                    int32_t step_begin, int32_t step_end, const int32_t* src_mode,
                    const int32_t* src_group, const double* table, int32_t table_steps, double* x,
                    double* wave_all, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                    oracle_trace* tr);

uint64_t oracle_pivot_hash_update(uint64_t h, const int32_t* perm, int32_t n);
int32_t oracle_count_sources(const csim_topology* t);
double oracle_norm2(const double* v, int64_t n);

#ifdef __cplusplus
}
#endif
#endif
