/* csim_host.h — C window onto the C++ host layer (include/csim/*.hpp) for bindings that cannot
 * include C++ (Python ctypes): the product's own netlist parser (reference grammar,
 * src/parser.cpp), lowering to csim_topology and host-libm source tabulation.
 * No numerics of the Newton path live here.                                                    */
#ifndef CSIM_HOST_H
#define CSIM_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct csimh_netlist csimh_netlist;

const char* csimh_last_error(void);
/* parseStream + ensureDefaultOp + assignEquationIndices + lower (src/main.cpp:29-34). NULL on error. */
csimh_netlist* csimh_parse(const char* netlist_text);
void csimh_free(csimh_netlist* h);
/* PULSE (type 1: spec = v1 v2 td tr tf ton per) or PWL (type 3: spec = n times then n values) waveform on the idx-th
 * independent source (V and I in netlist order): the two TranWaveform kinds of include/sim.hpp:80-138 that the
 * reference's parser never produces.  csimh_tabulate_sources then emits a CSIM_SRC_VAL_TABLE row for it. */
int csimh_set_source_waveform(csimh_netlist* h, int32_t idx, int32_t type, const double* spec, int32_t n);
int csimh_dims(const csimh_netlist* h, int32_t* n_unknowns, int32_t* n_node_eq, int32_t* n_elems,
               int32_t* n_params, int32_t* n_sources);
/* arrays sized per csimh_dims; nodes is [n_elems][4]; nominal is [n_params] */
int csimh_topology(const csimh_netlist* h, int32_t* type, int32_t* flags, int32_t* nodes, int32_t* branch,
                   int32_t* param_off, double* nominal);
int csimh_node_name(const csimh_netlist* h, int32_t eq, char* out, int32_t cap);
int csimh_elem_name(const csimh_netlist* h, int32_t idx, char* out, int32_t cap);
int csimh_csv_header(const csimh_netlist* h, char* out, int32_t cap);
int csimh_tran_cfg(const csimh_netlist* h, double* tstep, double* tstop, double* tstart, int32_t* enabled);
int csimh_hb_cfg(const csimh_netlist* h, double* f0, int32_t* n_harm, int32_t* enabled);
/* `.MC instances [seed [rel]]` card (superset of the reference grammar; include/csim/montecarlo.hpp): 1 = present */
int csimh_mc_cfg(const csimh_netlist* h, int64_t* instances, uint64_t* seed, double* rel);
/* Monte-Carlo parameters, out [P][B] instance-minor: instances [first, first + B) of the batch that
 * numpy.random.Generator(PCG64(seed)).uniform(-1, 1, (n, P)) defines (SURVEY.md 8(d) recipe), restated in C++ */
int csimh_monte_carlo(const csimh_netlist* h, int64_t B, uint64_t seed, double rel, int64_t first, double* out);
/* The `.MC` card as one batched DC + transient (csim::runMonteCarloTransientBatch); returns the recorded points per
 * probe (n_out) or -1.  x_final [N][B], wave [B][n_probes][n_out], iters_total / pivot_hash / status [B]; NULL ok. */
int csimh_run_mc_tran(const csimh_netlist* h, int32_t n_probes, const int32_t* probes, int32_t out_stride,
                      double* x_final, double* wave, int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status);
int csimh_run_mc_tran_devices(const csimh_netlist* h, int32_t n_devices, const int32_t* devices, int32_t n_probes,
                              const int32_t* probes, int32_t out_stride, double* x_final, double* wave,
                              int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status);
int csimh_plot_probes(const csimh_netlist* h, int32_t* out, int32_t cap);
int csimh_tabulate_sources(const csimh_netlist* h, double dt, int32_t n_steps, int32_t* mode, int32_t* group,
                           double* table, int32_t* n_groups);
int csimh_sin_table(double freq, double td, double phi, double dt, int32_t n_steps, double* out);
int csimh_run_tran_csv(const csimh_netlist* h, const char* path);
/* dcSolveGaussSeidel (src/dcanalysis.cpp:251-257) of the nominal circuit through the host layer; x [N]. */
int csimh_dc_gauss_seidel(const csimh_netlist* h, double* x);
/* Solver::luDecompose through the host layer: A, LU row-major [n*n], perm [n]; 1 = true, 0 = false, -1 = error. */
int csimh_lu_decompose(int32_t n, const double* A, double* LU, int32_t* perm);
/* runSteadyStateAnalysisBatch (include/csim/sstateanalysis.hpp); f0/tstep > 0 override .hb/.TRAN.
 * Returns steps per period, or -1. */
int csimh_run_sstate(const csimh_netlist* h, int64_t B, const double* params, double f0, double tstep,
                     int32_t max_periods, double tol, double* x_dc, double* x, int32_t* periods, double* delta,
                     int32_t* iters_total);
/* runSteadyStateShootingBatch (include/csim/sstateanalysis.hpp): Newton shooting on Phi_T(x) = x with the
 * perturbation columns as batch instances.  Returns steps per period, or -1. */
int csimh_run_shooting(const csimh_netlist* h, int64_t C, const double* params, double f0, double tstep,
                       int32_t max_iters, double tol, double rel_step, double* x_dc, double* x, int32_t* iterations,
                       double* residual, int64_t* transient_instances);
/* Row (f)1: one instance of a batched transient in the reference's CSV layout (csim::writeInstanceCsv,
 * include/csim/batchio.hpp).  x_dc: [N][B] or NULL (no t = 0 row); wave: [B][n_probes][n_out]; status: [B] or NULL.
 * Recorded rows are the last n_out multiples of out_stride among n_steps steps of the netlist's .TRAN tstep. */
int csimh_write_instance_csv(const csimh_netlist* h, const char* path, int64_t B, int64_t instance, int32_t n_probes,
                             const int32_t* probes, int32_t n_steps, int32_t n_out, int32_t out_stride,
                             const double* x_dc, const double* wave, const uint32_t* status);
/* Row (f)3: the netlist's index-th `.DC src start stop step` card as one batch of DC operating points
 * (csim::runDcSweepBatch).  Returns the number of points K (also with NULL outputs, to size them), -1 on error.
 * base: [P] parameters of every point or NULL for the nominal ones; values [K], x [N][K], iters [ramp][K], status [K]. */
int csimh_n_dc_sweeps(const csimh_netlist* h);
/* cos/sin(2*pi*n/M), n = 0..M-1, with the host libm: the tables csim_dev_harmonics takes */
int csimh_twiddle_table(int32_t M, double* cos_out, double* sin_out);
int csimh_run_dc_sweep(const csimh_netlist* h, int32_t index, const double* base, int32_t cap_points, double* values,
                       double* x, int32_t* iters, uint32_t* status);
#ifdef __cplusplus
}
#endif
#endif
