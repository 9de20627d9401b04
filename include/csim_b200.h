/* csim_b200.h — C ABI of the B200-native batched Newton/MNA hot path.
 *
 * Drop-in boundary for ONE path of ZyuRao/CircuitSimulator: the damped
 * Newton-Raphson / modified-nodal-analysis inner loop (device stamps, dense-
 * semantics LU with partial pivoting, DC source ramp, fixed-step backward-Euler
 * transient), run batched over B independent instances of one netlist
 * topology.  The reference has no FFI: its boundary is the in-process C++
 * surface listed per entry point below (file:line into the reference tree).
 * The C++ host layer in include/csim/ (same class and function names as the
 * reference) lowers a parsed Circuit into `csim_topology` and calls this ABI.
 *
 * Conventions
 *   - every call returns 0 on success, a negative CSIM_E_* code on failure and
 *     never throws; csim_last_error() returns a thread-local message.
 *   - all per-instance arrays are INSTANCE-MINOR ("SoA"): element (k, b) of a
 *     [K][B] array lives at k*B + b.  A shard of the batch is therefore a
 *     contiguous slice of every row.
 *   - arithmetic is IEEE-754 binary64, round-to-nearest, NO fused multiply-add
 *     on the reference's expressions, operations in the reference's order.
 *   - `csim_dev_*` take DEVICE pointers and a CUDA stream (cudaStream_t passed
 *     as void*) and only enqueue work; `csim_batch_*` take HOST pointers and
 *     include the host<->device copies.
 *   - there is no CPU fallback: without a CUDA device every compute entry point
 *     fails with CSIM_E_CUDA.
 *   - prototypes are hidden from NVRTC (__CUDACC_RTC__): the run-time-compiled
 *     specialised kernels include this header for the shared types only.
 */
#ifndef CSIM_B200_H
#define CSIM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSIM_ABI_VERSION 1

/* ---- status codes ------------------------------------------------------- */
#define CSIM_OK 0
#define CSIM_E_INVALID (-1)     /* bad argument / inconsistent topology        */
#define CSIM_E_UNSUPPORTED (-2) /* topology outside what the kernels cover     */
#define CSIM_E_CUDA (-3)        /* CUDA runtime error or no device             */
#define CSIM_E_NOMEM (-4)

/* ---- per-instance status word (bit mask) -------------------------------- *
 * Reference behaviour mirrored per instance:                 */
#define CSIM_ST_LU_SINGULAR 0x1u     /* pivot < lu_eps seen: that solve returned the ZERO vector (include/solver.hpp:58-61,94-97) */
#define CSIM_ST_NONFINITE 0x2u       /* raw LU solution had NaN/Inf. DC: gmin*=10 and retry (src/dcanalysis.cpp:135-138); TRAN: the reference throws (src/tanalisis.cpp:360-362) -> instance is frozen at its last accepted x */
#define CSIM_ST_DC_UNCONVERGED 0x4u  /* some ramp step hit dc_max_iters (src/dcanalysis.cpp:153-158) */
#define CSIM_ST_TRAN_UNCONVERGED 0x8u/* some timestep hit tran_max_iters; step accepted anyway (src/tanalisis.cpp:372-376) */

/* ---- element kinds (reference classes, include/element.hpp) ------------- */
enum csim_elem_type {
    CSIM_ELEM_R = 0,   /* Resistor          element.hpp:41-52  params: R                        */
    CSIM_ELEM_C = 1,   /* CapacitorElement  element.hpp:93-109 params: C                        */
    CSIM_ELEM_L = 2,   /* Inductor          element.hpp:112-131 params: L ; owns a branch row   */
    CSIM_ELEM_V = 3,   /* VoltageSource     element.hpp:71-90  params: dc v0 va freq td phi ; owns a branch row */
    CSIM_ELEM_I = 4,   /* CurrentSource     element.hpp:55-69  params: dc v0 va freq td phi     */
    CSIM_ELEM_MOS = 5  /* MosfetBase        element.hpp:134-171 params: Vth K lambda Cj0 ; nodes D G S B */
};
/* elem_flags */
#define CSIM_FLAG_PMOS 0x1       /* MOS: isP (sign of .MODEL VT, src/parser.cpp:755-761) */
#define CSIM_FLAG_SRC_SIN 0x1    /* V/I: tran.type == SIN -> evalDC adds v0 (include/sim.hpp:152-158) */

#define CSIM_PARAMS_R 1
#define CSIM_PARAMS_C 1
#define CSIM_PARAMS_L 1
#define CSIM_PARAMS_SRC 6 /* dc, v0, va, freq, td, phi  (SourceSpec / SinSpec, include/sim.hpp:56-62,145-149) */
#define CSIM_PARAMS_MOS 4 /* Vth, K, lambda, Cj0        (MosfetBase, include/element.hpp:134-140)             */

/* How the transient value of an independent source is obtained at step s
 * (t = (s+1)*dt, src/tanalisis.cpp:256).  glibc sin() and CUDA sin() are not
 * bit-identical, so the parity path tabulates on the HOST with the host libm:
 *   CSIM_SRC_CONST      value = dc + 0.0                      (WaveformType::NONE, sim.hpp:77-78,160-162)
 *   CSIM_SRC_SIN_TABLE  value = dc + (t < td ? v0 : v0 + va * S[group][s])
 *                       with S = sin(2*pi*freq*(t-td)+phi) tabulated per group (sim.hpp:117-122)
 *   CSIM_SRC_VAL_TABLE  value = dc + S[group][s]              (S = TranWaveform::eval(t): PULSE / PWL, sim.hpp:80-138)
 */
enum csim_src_mode { CSIM_SRC_CONST = 0, CSIM_SRC_SIN_TABLE = 1, CSIM_SRC_VAL_TABLE = 2 };

/* ---- constants block: every literal of the reference loops -------------- */
typedef struct csim_constants {
    int32_t dc_ramp_steps;   /* 10     src/dcanalysis.cpp:104 */
    int32_t dc_max_iters;    /* 50     src/dcanalysis.cpp:105 */
    double dc_tol;           /* 1e-9   src/dcanalysis.cpp:106 */
    double dc_alpha;         /* 0.35 = clamp(0.35, 0.1, 0.5)  src/dcanalysis.cpp:274 (alphaNext is dead) */
    double gmin_high_base;   /* 1e-6   src/dcanalysis.cpp:264 */
    double gmin_low_base;    /* 3.35e-7 */
    double gmin_abs_max;     /* 1e-4   */
    double fast_conv_ratio;  /* 0.7    */
    double slow_conv_ratio;  /* 1.05   */
    double dc_nonfinite_mul; /* 10.0   src/dcanalysis.cpp:136 */
    double dc_nonfinite_cap; /* 1e-2   */
    int32_t tran_max_iters;  /* 50     src/tanalisis.cpp:241 */
    int32_t _pad0;
    double tran_tol;         /* 1e-6   src/tanalisis.cpp:242 */
    double tran_gmin;        /* 1e-6   src/tanalisis.cpp:243 */
    double tran_alpha;       /* 0.45   src/tanalisis.cpp:244 */
    double lu_eps;           /* 1e-15  include/solver.hpp:31  */
    double mos_off_gds;      /* 1e-12  src/element.cpp:246    */
} csim_constants;

/* Fills *c with the reference's literals. */
#ifndef __CUDACC_RTC__
void csim_default_constants(csim_constants* c);
#endif

/* ---- flat topology: what Circuit + assignEquationIndices() define ------- *
 * (include/circuit.hpp:33-66, src/circuit.cpp:42-61).  Elements are listed in
 * NETLIST ORDER: with the equation numbering it fixes the MNA sparsity pattern
 * AND the order in which contributions are accumulated into each G(i,j)/I(i)
 * (src/dcanalysis.cpp:126-130, src/tanalisis.cpp:268-356).                    */
typedef struct csim_topology {
    int32_t n_unknowns;         /* N = node equations + branch equations (Circuit::numUnknowns)       */
    int32_t n_node_eq;          /* rows 0..n_node_eq-1 are node rows (receive gmin, dcanalysis.cpp:36-43) */
    int32_t n_elems;
    int32_t n_params;           /* P = parameter doubles per instance                                  */
    const int32_t* elem_type;   /* [n_elems] csim_elem_type                                            */
    const int32_t* elem_flags;  /* [n_elems]                                                           */
    const int32_t* elem_nodes;  /* [n_elems][4] equation index of each terminal, -1 = ground/unused    */
    const int32_t* elem_branch; /* [n_elems] branch equation index (V, L) or -1                        */
    const int32_t* elem_param;  /* [n_elems] offset of the element's first parameter in [0,P)          */
} csim_topology;

/* csim_plan: a topology lowered to device tables (immutable) plus the kernels specialised for it at run time and
 * their selection statistics (learned state, guarded by a lock inside the plan from the choice of a kernel to its
 * launch; a kernel is retired only after the device is idle).  Host threads may share one plan, each with its own
 * csim_batch / stream.                                                                                        */
typedef struct csim_plan csim_plan;
typedef struct csim_batch csim_batch; /* B instances on one device: owns device buffers + a stream */

#ifndef __CUDACC_RTC__
const char* csim_last_error(void);
int csim_abi_version(void);
#endif
/* Number of visible CUDA devices, or CSIM_E_CUDA. */
#ifndef __CUDACC_RTC__
int csim_device_count(void);
#endif

/* Builds the stamp programs (ordered per-slot contribution lists for the DC
 * and the TRAN assembly), the CSR patterns and uploads them to `device`.
 * consts == NULL -> reference defaults.                                        */
#ifndef __CUDACC_RTC__
int csim_plan_create(const csim_topology* topo, const csim_constants* consts, int device, csim_plan** out);
void csim_plan_destroy(csim_plan* plan);
#endif

/* Host-side introspection (no device work).  mode: 0 = DC assembly, 1 = TRAN
 * assembly.  Structural pattern of G in CSR, rows ascending, columns ascending
 * within a row.  Pass NULL arrays to query *nnz only.                          */
#define CSIM_MODE_DC 0
#define CSIM_MODE_TRAN 1
#ifndef __CUDACC_RTC__
int csim_plan_pattern(const csim_plan* plan, int mode, int32_t* nnz, int32_t* rowptr /*[N+1]*/, int32_t* colidx /*[nnz]*/);
int csim_plan_dims(const csim_plan* plan, int32_t* n_unknowns, int32_t* n_params, int32_t* n_sources);
#endif

/* ---- run-time specialisation of the transient kernel ------------------------- *
 * For N <= 32 the library generates (NVRTC, sm_100a) a transient kernel for this
 * topology and a PREDICTED pivot permutation: matrix rows in registers, only the
 * structurally non-zero operations of stamping, LU and the triangular solves, in
 * the reference's order.  Every factorisation verifies the reference's
 * first-max-wins pivot rule (include/solver.hpp:46-56); where the prediction does
 * not hold, that solve is redone with the generic pivoting code, so results do
 * not depend on the prediction.  By default csim_dev_tran specialises on first
 * use, predicting the permutation from instance 0 at its start state.           */
#ifndef __CUDACC_RTC__
/* mode 0: generic kernels only; 1 (default): automatic — the thread-per-instance kernel, replaced by the
 * warp-per-instance one when more than 2 % of the solves miss the predicted pivot order; 2: warp-per-instance
 * kernel (csim_tran_spec: one warp per instance, rows in registers); 3: thread-per-instance kernel
 * (csim_tran_tpi: one thread per instance, state in shared memory).  CSIM_JIT=0|2|3 in the environment
 * overrides.  All four paths give bit-identical results.
 * A transient batch that fits one wave of CTAs runs the thread-per-instance kernel in its warp-PAIR form (two warps
 * per 32 instances: the second warp computes part of the matrix rows and the MOSFET linearisations of the next
 * iteration; same operations in the same order, so the same bits).  CSIM_TPI_PAIR=0 turns that off.            */
int csim_plan_jit_mode(csim_plan* plan, int mode);
/* Specialise now for perm[N] (perm[i] = original row expected at pivot position i). */
int csim_plan_specialize(csim_plan* plan, const int32_t* perm);
/* Thread-per-instance kernel with several predicted permutations, perms[n_variants][N] (1 <= n_variants <= 16):
 * every lane runs the variant whose permutation its last solve took; a solve no variant predicts is redone
 * generically, as always.  In automatic mode the library adds variants itself (one per re-evaluation, the
 * permutation the redo path met most often) while more than 0.2 % of the solves miss.                 */
int csim_plan_specialize_variants(csim_plan* plan, int32_t n_variants, const int32_t* perms);
/* The DC operating point of circuits with MOSFETs (dcSolveNewtonLU, src/dcanalysis.cpp:95-163) runs on a generated
 * thread-per-instance kernel too (csim_dc_tpi: source ramp and ConvController per lane, gmin a per-lane variable),
 * unless the JIT mode is 0 or 2, CSIM_JIT=0|2 or CSIM_DC_TPI=0 is set, or N > 32.  The ramp walks through the circuit's
 * operating regions, so the kernel collects predicted permutations (up to 16) over the first launches exactly like
 * the transient kernel.  csim_plan_specialize_dc_variants installs a set directly; csim_plan_dc_spec_info reports
 * whether the kernel is loaded, its variants, and solves redone generically / all solves since it was built.       */
/* CSIM_PLAN_CACHE=<directory> (opt-in): the permutations a plan's thread-per-instance kernels were last built with are
 * stored there per topology (hash of the csim_topology arrays and the constants block) and kernel; a new plan of the
 * same topology starts from them instead of the single pilot prediction.  Predictions only: results cannot depend on
 * the files, a stale or corrupt one is ignored or costs redos.                                                  */
int csim_plan_specialize_dc_variants(csim_plan* plan, int32_t n_variants, const int32_t* perms);
int csim_plan_dc_spec_info(const csim_plan* plan, int32_t* active, int32_t* n_variants, uint64_t* fallback_solves,
                           uint64_t* solves, char* log, int32_t log_cap);
/* active: 0 none / 2 / 3 = kind of the specialised kernel in use; perm [N] (NULL ok); counters since load:
 * Newton solves done by the generic fallback / all Newton solves; log: generator + NVRTC messages. */
int csim_plan_spec_info(const csim_plan* plan, int32_t* active, int32_t* perm, uint64_t* fallback_solves,
                        uint64_t* solves, char* log, int32_t log_cap);
/* Generated CUDA source of the loaded kernel; returns its length (out may be NULL). */
int csim_plan_spec_source(const csim_plan* plan, char* out, int32_t cap);
/* Host-only: source the generator emits for perm (no device needed); returns its length. */
int csim_plan_generate_source(const csim_plan* plan, const int32_t* perm, char* out, int32_t cap);
/* Same for either specialised transient kernel: kind 2 = warp per instance (csim_tran_spec),
 * kind 3 = thread per instance (csim_tran_tpi); kind 4 = the thread-per-instance DC kernel (csim_dc_tpi). */
int csim_plan_generate_source_kind(const csim_plan* plan, int kind, const int32_t* perm, char* out, int32_t cap);
/* Host-only: thread-per-instance source for perms[n_variants][N]. */
int csim_plan_generate_source_variants(const csim_plan* plan, int32_t n_variants, const int32_t* perms, char* out,
                                       int32_t cap);
#endif

/* Systems with more than 32 unknowns (the 200-stage ladder has 404) run on the large-matrix path, the
 * matrix in HBM/L2 and the reference's partial pivoting with logical row exchanges.  Two kernel families:
 *   dense (default, N <= 1280): one CTA per instance, blocked right-looking LU on [A | b], 16-column panel
 *     and a strip of U staged in shared memory, 4 x 4 register tiles for the trailing update, no zero
 *     skipping (the ladder's factorisation is fully dense under the reference's unknown order);
 *   warp (CSIM_BIG_KIND=warp, or N > 1280): one warp per instance, row- and column-major copies,
 *     exact-zero skipping and a fill-in log (pays off only when the factorisation stays sparse).
 * on = 1 forces the large path (dense kernels) for any N, on = 2 forces it with the warp kernels,
 * on = 0 returns to the choice by size; CSIM_BIG=1 forces it process-wide (tests run the small
 * netlists through it).                                                                             */
#ifndef __CUDACC_RTC__
int csim_plan_force_large_path(csim_plan* plan, int on);
#endif

/* Batched Solver::solveLinearSystemGaussSeidel (include/solver.hpp:139-192; on the Solver interface, off the
 * reference's default path: dcSolve -> LU, src/dcanalysis.cpp:260-262).  A [n*n][B] row-major per system, b [n][B],
 * x [n][B] in: the initial iterate x0, out: the last iterate (the reference returns it whether or not it converged);
 * sweeps [B] (NULL ok): sweeps executed.  Per system exactly the reference's sweep: row i subtracts a(i,j)*x(j) in
 * ascending j, newest values for j < i; |diag| < 1e-12 is replaced by +-1e-12; stop when ||x - xOld||_2 < tol
 * (norm order: oracle/csim_norm2.h).  n <= 200.  The reference's defaults: max_iters 1000, tol 1e-10. */
#ifndef __CUDACC_RTC__
int csim_dev_gauss_seidel(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t max_iters,
                          double tol, int32_t* sweeps, void* stream);
/* the same with host buffers (copies inside); used by the host layer's Solver::solveLinearSystemGaussSeidel */
int csim_host_gauss_seidel(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t max_iters,
                           double tol, int32_t* sweeps);
#endif

/* Harmonic extraction over one period (the `.hb f0 nharm` card the reference parses at
 * src/parser.cpp:551-572 and never uses; row (f)2 of SURVEY 8).  For every signal (a row of n_samples
 * doubles, row stride ld; e.g. the last period of one probe of one instance in csim_dev_tran's wave)
 *     X_h = (1/M) * sum_{n=0}^{M-1} x[n] * (cos_tab[(h*n) mod M] - j * sin_tab[(h*n) mod M]),  h = 0..n_harm,
 * M = n_samples, cos_tab/sin_tab [M] = cos/sin(2*pi*n/M) tabulated on the HOST (csimh_twiddle_table: host
 * libm, like the source tables).  The amplitude of harmonic h >= 1 is 2*|X_h|, X_0 is the mean.
 * Summation order (part of the definition): partial sum l (l = 0..31) adds the samples n = l, l+32, ... in
 * ascending order, every product and sum separately rounded (no FMA); the 32 partial sums are combined
 * pairwise by the xor butterfly with strides 16, 8, 4, 2, 1; then one division by M.
 * re, im: [n_harm+1][n_signals] device.  n_harm <= 64 (the reference's own card, tests/dbmixer.sp:45, asks for 50).
 * Harmonics are taken four per pass over the samples.  Cost: 4 separately rounded FP64 operations per sample and
 * harmonic against 8 bytes per sample and pass, so only n_harm <= 3 (one pass) is HBM-bound; from there on the
 * bound is the FP64 pipe (no FMA: the definition rounds every product and sum), e.g. 204 operations per sample
 * at n_harm = 50.  The reference has no counterpart: parity unpinned. */
#ifndef __CUDACC_RTC__
int csim_dev_harmonics(int64_t n_signals, int32_t n_samples, int64_t ld, const double* samples, int32_t n_harm,
                       const double* cos_tab, const double* sin_tab, double* re, double* im, void* stream);
#endif

/* The standalone assembly kernel (csim_dev_stamp_csr) is specialised the same way on first use
 * (straight-line code, every distinct input loaded once, values in registers); any N.
 * bytes_read / bytes_written: distinct input / output bytes ONE assembly of `mode` moves per
 * instance (has_src_val: sources given per instance instead of evalDC(scale) from the parameters) —
 * the algorithmic traffic the roofline figure is computed from.                                  */
#ifndef __CUDACC_RTC__
int csim_plan_stamp_info(const csim_plan* plan, int mode, int has_src_val, int32_t* spec_active, int64_t* bytes_read,
                         int64_t* bytes_written, char* log, int32_t log_cap);
/* Host-only: source the generator emits for the assembly kernel of `mode`; returns its length. */
int csim_plan_generate_stamp_source(const csim_plan* plan, int mode, char* out, int32_t cap);
#endif

#ifndef __CUDACC_RTC__
/* Self-test: q_split[i] = the specialised kernels' "reciprocal once + 3-operation quotient"
 * division of a[i]/b[i], q_ref[i] = the compiler's IEEE a[i]/b[i]; they must be bit-identical. */
int csim_dev_selftest_div(int64_t n, const double* a, const double* b, double* q_split, double* q_ref,
                          void* stream);
/* Same for the thread-per-instance kernel's division (2^-1021 <= |b| <= 2^1021, the kernel's pivot test): q[i], the IEEE quotient and
 * in_range[i] = 1 where the division's own range guards accept it — there q must equal q_ref bit for bit
 * (elsewhere the kernel redoes the solve with true divisions). */
int csim_dev_selftest_tdiv(int64_t n, const double* a, const double* b, double* q, double* q_ref, int32_t* in_range,
                           void* stream);
/* The multiplier division of the thread-per-instance kernel with its guards folded into unsigned accumulators
 * (2^-1021 <= |b| <= 2^1021): flags[i] bit 0 = the range guard accepts the division — q must then equal q_ref bit
 * for bit; bit 1 = the pivot guard fires (|q| >= 1 judged on the high word): it must fire whenever |a| > |b| and
 * may only fire when |a| >= |b|.                                                                        */
int csim_dev_selftest_tdiv_mul(int64_t n, const double* a, const double* b, double* q, double* q_ref, int32_t* flags,
                               void* stream);
#endif

/* ===================== device-pointer entry points ======================== */

/* Standalone assembly (the HBM-roofline kernel; also the parity probe for the
 * stamp programs).  Replaces one pass of the assembly blocks
 *   DC  : src/dcanalysis.cpp:120-130   (elements in netlist order, then gmin)
 *   TRAN: src/tanalisis.cpp:259-356    (R/V/I, MOS channel, C, L, MOS caps, gmin)
 * for every instance.  params [P][B]; x [N][B] = current Newton iterate;
 * x_prev [N][B] = last accepted timepoint (TRAN only, may be NULL for DC);
 * src_val [n_sources][B] = value of each independent source (already scaled
 * for DC, evaluated at t for TRAN), or NULL to use evalDC(scale) from params;
 * vals [nnz(mode)][B], rhs [N][B].                                             */
#ifndef __CUDACC_RTC__
int csim_dev_stamp_csr(const csim_plan* plan, int mode, int64_t B, const double* params, const double* x,
                       const double* x_prev, const double* src_val, double dt, double source_scale,
                       double gmin, double* vals, double* rhs, void* stream);
#endif

/* Batched Solver::solveLinearSystemLU (include/solver.hpp:83-131) on dense
 * N x N systems (N <= 32: matrix in shared memory; larger: the large-matrix path): A [N*N][B] row-major over (i,j), b [N][B] ->
 * x [N][B], perm [N][B] (perm[i] = original row now at position i,
 * solver.hpp:41-44,66), singular [B] (1 -> x is the zero vector).              */
#ifndef __CUDACC_RTC__
int csim_dev_lu_solve(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                      int32_t* singular, double lu_eps, void* stream);
#endif

/* Allocation-free variant of csim_dev_lu_solve.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and results as csim_dev_lu_solve (always the dense large-matrix kernel,
 * whatever n), but the kernel's per-CTA [A | b] workspaces live in a static 32 MiB device array inside the library
 * (every GPU that loads the library holds it until exit) and the call performs no cudaMalloc / cudaFree /
 * stream-ordered allocation.  Limit: 1 <= n <= 1280 (one n x (n + 1) workspace must fit the array; it does up to
 * kDenseMaxN); larger n is rejected with CSIM_E_UNSUPPORTED.  Fewer workspaces than resident CTAs only means fewer
 * CTAs loop over the batch: results are bit-identical.  All calls share the array, which keeps the last call's
 * contents (nothing needs to start at zero: every workspace is fully rewritten before it is read); calls that
 * OVERLAP on the GPU (different streams, or different host threads on one device) race on it — serialise them. */
#ifndef __CUDACC_RTC__
int csim_dev_lu_solve_noalloc(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                              int32_t* singular, double lu_eps, void* stream);
#endif

/* Batched Solver::luDecompose (include/solver.hpp:30-80): LU [N*N][B] = the packed factors exactly as the
 * reference leaves them — U on and above the diagonal, the multipliers L(i,k) in the strict lower triangle
 * (solver.hpp:70-72), rows exchanged physically, multipliers included (solver.hpp:64-67); perm [N][B];
 * singular [B] = 1 where the reference returns false (LU then holds the reference's partial state for
 * N <= 32 and is unspecified for larger N).  N <= 1280.                                              */
#ifndef __CUDACC_RTC__
int csim_dev_lu_factor(int32_t n, int64_t B, const double* A, double* LU, int32_t* perm, int32_t* singular,
                       double lu_eps, void* stream);
int csim_host_lu_factor(int32_t n, int64_t B, const double* A, double* LU, int32_t* perm, int32_t* singular,
                        double lu_eps);
#endif

/* DC operating point for every instance (dcSolveLU, src/dcanalysis.cpp:242-248:
 * Newton ramp 95-163 when the netlist has MOSFETs, direct solve 46-68 otherwise).
 * x [N][B] out; iters [dc_ramp_steps][B] Newton iterations spent per ramp step
 * (may be NULL); pivot_hash [B] in/out running hash of every pivot permutation
 * (may be NULL; see csim_pivot_hash_* below); status [B] OR-ed in.              */
#ifndef __CUDACC_RTC__
int csim_dev_dc(const csim_plan* plan, int64_t B, const double* params, double* x, int32_t* iters,
                uint64_t* pivot_hash, uint32_t* status, void* stream);
#endif

typedef struct csim_tran_args {
    double dt;              /* .TRAN tstep                                                   */
    int32_t step_begin;     /* first step of this launch; t = (step+1)*dt                    */
    int32_t step_end;       /* one past the last step: the reference runs [0, nSteps)        */
    /* source tables (see csim_src_mode) */
    int32_t n_groups;       /* rows of src_table                                             */
    int32_t table_steps;    /* columns of src_table (>= step_end)                            */
    const double* src_table;    /* [n_groups][table_steps] device                            */
    const int32_t* src_mode;    /* [n_sources] host array, csim_src_mode                     */
    const int32_t* src_group;   /* [n_sources] host array: group used by every instance, or  */
    const int32_t* src_group_b; /* [n_sources][B] DEVICE array: per-instance group (overrides src_group when non-NULL);
                                 * entries must lie in [0, n_groups) — csim_batch_tran checks its host copy, the
                                 * device-pointer entry points take them as given.  The tabulated rows carry the
                                 * sine's freq and phi: the per-instance freq / phi in `params` are not read on the
                                 * device (td is, for the t < td test), so a swept frequency needs its own row.  */
    /* outputs */
    int32_t n_probes;       /* probed unknowns                                               */
    const int32_t* probes;  /* [n_probes] host array of equation indices                     */
    int32_t out_stride;     /* record step s when (s+1) % out_stride == 0                    */
    int32_t out_begin_step; /* and s >= out_begin_step (the reference's t >= tstart filter, src/tanalisis.cpp:209) */
    int64_t wave_ld;        /* timepoints allocated per (instance, probe) row of wave        */
} csim_tran_args;

/* Fixed-step backward-Euler transient for every instance, steps
 * [step_begin, step_end) of runTransientAnalysisBackwardEuler
 * (src/tanalisis.cpp:255-420).  x [N][B] in: state at step_begin (the DC
 * operating point for step_begin == 0), out: state after step_end-1.
 * wave [B][n_probes][wave_ld] TIME-MINOR: the value of probe p of instance b at
 * recorded point r (r counts recorded steps from step 0) is
 * wave[(b*n_probes + p)*wave_ld + r]; row 0 of the reference CSV (t = 0, the DC
 * point) is NOT written here.  iters_total [B] += Newton iterations;
 * iters_step [nSteps][B] uint8 per-step counts (may be NULL).                   */
#ifndef __CUDACC_RTC__
int csim_dev_tran(const csim_plan* plan, int64_t B, const double* params, const csim_tran_args* args,
                  double* x, double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash,
                  uint32_t* status, void* stream);
#endif

/* ---- periodic steady state ----------------------------------------------------------------- *
 * The reference names a steady-state driver (src/sstateanalysis.cpp) but the file is EMPTY and the
 * `.hb f0 nHarm` card (src/parser.cpp:551-572, include/sim.hpp:187-191) is parsed and never read,
 * so there is no reference behaviour to match ("parity unpinned").  DEFINITION used here, built
 * only from the reference's own transient step: with T = 1/f0 and steps_per_period = round(T/dt),
 * integrate period after period with runTransientAnalysisBackwardEuler's step loop
 * (src/tanalisis.cpp:255-420) from the given start state; after period k (k = 1, 2, ...) form
 * delta_k = || x(kT) - x((k-1)T) ||_2 over all unknowns (summation order: oracle/csim_norm2.h);
 * the instance has reached its periodic steady state at the first k with delta_k < tol and is
 * frozen there (state, iteration counters, pivot hash, status).  Instances that never get there
 * stop after max_periods with periods[b] = -1.  Step s of the run is t = (s+1)*dt counted from
 * the start state, so the source table must cover max_periods * steps_per_period steps.        */
typedef struct csim_sstate_args {
    int32_t steps_per_period;
    int32_t max_periods;
    double tol;
} csim_sstate_args;
/* tran: dt and source tables as for csim_dev_tran (step range, probes and output fields ignored).
 * x [N][B] in: start state (normally the DC operating point), out: state at convergence (or after
 * max_periods); periods [B] out; delta [B] out: last delta_k of each instance; iters_total,
 * pivot_hash, status as for csim_dev_tran.  Synchronises the stream once per period.           */
#ifndef __CUDACC_RTC__
int csim_dev_sstate(const csim_plan* plan, int64_t B, const double* params, const csim_tran_args* tran,
                    const csim_sstate_args* ss, double* x, int32_t* periods, double* delta, int32_t* iters_total,
                    uint64_t* pivot_hash, uint32_t* status, void* stream);
#endif

/* ======================= host-pointer entry points ======================== */
#ifndef __CUDACC_RTC__
int csim_batch_create(const csim_plan* plan, int64_t B, csim_batch** out);
void csim_batch_destroy(csim_batch* batch);
int csim_batch_set_params(csim_batch* batch, const double* params /*[P][B] host*/);
#endif
/* DC operating point; x [N][B], iters [ramp][B] (NULL ok), status [B] (NULL ok). */
#ifndef __CUDACC_RTC__
int csim_batch_dc(csim_batch* batch, double* x, int32_t* iters, uint64_t* pivot_hash, uint32_t* status);
#endif
/* DC operating point followed by the whole transient, as
 * runTransientAnalysisBackwardEuler does (src/tanalisis.cpp:112 then 255-420).
 * args->src_table / src_group_b are HOST pointers here.  wave as above but
 * host; x_dc [N][B] and x_final [N][B] may be NULL.                             */
#ifndef __CUDACC_RTC__
int csim_batch_tran(csim_batch* batch, const csim_tran_args* args, double* x_dc, double* x_final, double* wave,
                    int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status);
#endif
/* The transient alone, from given states: x [N][B] in = the state at step_begin (t = step_begin * dt), out = the
 * state after step_end.  What the shooting driver (include/csim/sstateanalysis.hpp) integrates its candidates and
 * perturbation columns with.  Other arguments as csim_batch_tran.                                    */
#ifndef __CUDACC_RTC__
int csim_batch_tran_from(csim_batch* batch, const csim_tran_args* args, double* x, double* wave,
                         int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status);
#endif
/* DC operating point followed by csim_dev_sstate from it; tran->src_table / src_group_b are HOST
 * pointers; x_dc, x [N][B], periods, delta, iters_total, pivot_hash, status: host (NULL ok except x). */
#ifndef __CUDACC_RTC__
int csim_batch_sstate(csim_batch* batch, const csim_tran_args* tran, const csim_sstate_args* ss, double* x_dc,
                      double* x, int32_t* periods, double* delta, int32_t* iters_total, uint64_t* pivot_hash,
                      uint32_t* status);
#endif
/* Host-pointer forms of the two standalone kernels (copies inside; for the C++ drop-in
 * classes Element::stamp and Solver::solveLinearSystemLU and for small tests).            */
#ifndef __CUDACC_RTC__
int csim_host_stamp_csr(const csim_plan* plan, int mode, int64_t B, const double* params, const double* x,
                        const double* x_prev, const double* src_val, double dt, double source_scale,
                        double gmin, double* vals, double* rhs);
int csim_host_lu_solve(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                       int32_t* singular, double lu_eps);
#endif
/* ---- one batch over several GPUs of one box ------------------------------------------------------------ *
 * SURVEY.md 8(b) "Threading" / 8(e): the reference is single-threaded and has no counterpart; its CPU baseline runs
 * one instance per host core, this runs one contiguous shard of the batch per GPU.  B instances are split into
 * n_devices contiguous shards (remainder to the low ranks: csim_multi_shard tells which); every device gets its own
 * plan, batch and stream, every call below runs one host thread per device, there is no data-path collective and
 * the "final gather" is each device copying its slice into the caller's host buffers.  devices == NULL: 0..n-1; a
 * device may be listed more than once (two shards share it).  All arrays are the GLOBAL host arrays of
 * csim_batch_* ([K][B] instance-minor, wave [B][n_probes][wave_ld]); per-step Newton counts are not returned.
 * Results are bit-identical to one csim_batch over all B instances (instances never interact).           */
typedef struct csim_multi csim_multi;
#ifndef __CUDACC_RTC__
int csim_multi_create(const csim_topology* topo, const csim_constants* consts, int32_t n_devices, const int32_t* devices,
                      int64_t B, csim_multi** out);
void csim_multi_destroy(csim_multi* m);
int csim_multi_shard(const csim_multi* m, int32_t rank, int32_t* device, int64_t* first, int64_t* count);
int csim_multi_set_params(csim_multi* m, const double* params /*[P][B] host*/);
int csim_multi_dc(csim_multi* m, double* x, int32_t* iters, uint64_t* pivot_hash, uint32_t* status);
int csim_multi_tran(csim_multi* m, const csim_tran_args* args, double* x_dc, double* x_final, double* wave,
                    int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status);
float csim_multi_last_kernel_ms(const csim_multi* m); /* slowest device of the last call */
void csim_set_last_error(const char* msg);            /* internal: lets the multi-GPU layer report through csim_last_error() */
#endif

/* Milliseconds of the kernels of the last csim_batch_* call, measured with CUDA
 * events on the batch's stream (copies excluded).                               */
#ifndef __CUDACC_RTC__
float csim_batch_last_kernel_ms(const csim_batch* batch);
#endif

/* FP64 pipe probe: sustained Tflop/s of SEPARATE multiplies and adds over the whole chip (register-resident, eight
 * independent chains per thread; FMA is not allowed on the reference's expressions, so its peak does not apply).
 * The denominator of the fused kernels' roofline fraction, measured in the run that reports it.   */
#ifndef __CUDACC_RTC__
int csim_dev_fp64_peak(int device, double* tflops_mul_add);
#endif

/* Diagnostic of the one unpinned call on the path, Eigen's norm() (summation order defined in oracle/csim_norm2.h;
 * SURVEY.md 8(c)): while `margin` ([B] doubles on the device, initialised to +inf by the caller) is set for the calling
 * thread, csim_dev_dc and csim_dev_tran (N <= 32) run on the generic kernels and record per instance the smallest
 *     |err - threshold| / threshold
 * over every comparison of a Newton step norm with a threshold: err < tol (DC 1e-9, TRAN 1e-6) and the controller's
 * err > 1.05 prevErr / err < 0.7 prevErr (src/dcanalysis.cpp:286-297, src/tanalisis.cpp:372).  A last-ulp difference in
 * the norm can only flip a comparison whose margin is ~1e-16; results are bit-identical across kernel families, so the
 * margins hold for all of them.  margin == NULL switches the probe off again.                                       */
#ifndef __CUDACC_RTC__
int csim_dev_margin_probe(double* margin);
#endif

/* ---- pivot-sequence hash (shared definition, host-callable) ------------- *
 * Every factorisation contributes f = sum_i mix(i, perm[i]) (mod 2^64) and the
 * running hash becomes h' = (h XOR f) * 0x100000001B3.  h starts at
 * CSIM_PIVOT_HASH_INIT.  Two instances have equal hashes iff (with
 * overwhelming probability) they chose the same pivot rows in the same order
 * in every factorisation.                                                       */
#define CSIM_PIVOT_HASH_INIT 0xCBF29CE484222325ull
#ifndef __CUDACC_RTC__
uint64_t csim_pivot_hash_update(uint64_t h, const int32_t* perm, int32_t n);
#endif

#ifdef __cplusplus
}
#endif
#endif /* CSIM_B200_H */
