// kernels.hpp — launch wrappers implemented in kernels.cu (callable from plain C++).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "device_plan.cuh"

namespace csim {

constexpr int kWarpsPerBlock = 4;   // fused Newton kernels: 4 instances per CTA
constexpr int kStampThreads = 128;  // standalone assembly: instances per CTA

cudaError_t launch_stamp_csr(const DevPlan& P, int mode, int64_t B, const double* params, const double* x,
                             const double* x_prev, const double* src_val, double dt, double scale, double gmin,
                             double* vals, double* rhs, int sm_count, cudaStream_t stream);
// lu_out (may be null): packed L\U of Solver::luDecompose, [N*N][B]
cudaError_t launch_lu_solve(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                            int32_t* singular, double eps, cudaStream_t stream, double* lu_out = nullptr);
cudaError_t launch_dc(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                      uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream, double* margin = nullptr);
cudaError_t launch_tran(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x, double* wave,
                        int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                        cudaStream_t stream);

// large-matrix path (big_kernels.cu): any N whose per-warp vectors fit in shared memory (N <= ~1700)
constexpr int kBigMaxN = 1536;
cudaError_t launch_tran_big(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x, double* wave,
                            int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                            cudaStream_t stream);
cudaError_t launch_dc_big(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                          uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream);
cudaError_t launch_lu_big(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                          int32_t* singular, double eps, cudaStream_t stream);
size_t big_bytes_per_instance(int N, int n_values);

// large-matrix path, dense (dense_kernels.cu): one CTA per instance, blocked LU on [A | b] in HBM/L2; the default
// for 32 < N <= kDenseMaxN (the 17-column panel of all N rows has to fit in shared memory)
constexpr int kDenseMaxN = 1280;
cudaError_t launch_tran_dense(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x,
                              double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash,
                              uint32_t* status, cudaStream_t stream);
cudaError_t launch_dc_dense(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                            uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream);
cudaError_t launch_lu_dense(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                            int32_t* singular, double eps, cudaStream_t stream, double* lu_out = nullptr);

// allocation-free dense LU solve: scratch in a file-scope device array of kDnNoAllocDoubles doubles (32 MiB per GPU)
constexpr size_t kDnNoAllocDoubles = (size_t)4 << 20;
int lu_dense_noalloc_max_n();
cudaError_t launch_lu_dense_noalloc(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                                    int32_t* singular, double eps, cudaStream_t stream);

cudaError_t launch_sstate_check(int N, int64_t B, const double* x, const double* x_prev, double tol, int period,
                                uint8_t* skip, int32_t* periods, double* delta, unsigned long long* n_open,
                                cudaStream_t stream);

constexpr int kGaussSeidelMaxN = 200;  // two iterates of n doubles for 64 systems per CTA in shared memory
cudaError_t launch_gauss_seidel(int n, int64_t B, const double* A, const double* rhs, double* x, int max_iters,
                                double tol, int32_t* sweeps, cudaStream_t stream);

constexpr int kHarmonicsMax = 64;  // the reference's own card asks for 50 (tests/dbmixer.sp:45 ".hb 100e6 50")
cudaError_t launch_harmonics(int64_t n_signals, int M, int64_t ld, const double* x, int H, const double* ct,
                             const double* st, double* re, double* im, cudaStream_t stream);

cudaError_t launch_tdiv_selftest(int64_t n, const double* a, const double* b, double* q, double* q_ref,
                                 int32_t* in_range, cudaStream_t stream);
cudaError_t launch_tdiv_mul_selftest(int64_t n, const double* a, const double* b, double* q, double* q_ref,
                                     int32_t* flags, cudaStream_t stream);
cudaError_t launch_div_selftest(int64_t n, const double* a, const double* b, double* q_split, double* q_ref,
                                cudaStream_t stream);

// FP64 pipe probe (separate DMUL + DADD, register-resident): Tflop/s of the whole chip, best of three timed launches
cudaError_t measure_fp64_peak(int sm_count, double* tflops, cudaStream_t stream);

}  // namespace csim
