// kernels.cu — sm_100a kernels of the batched Newton/MNA path.
//
//   stamp_csr_kernel : standalone assembly, thread-per-instance, HBM-bound
//   lu_solve_kernel  : batched Solver::solveLinearSystemLU, warp-per-system
//   dc_kernel        : DC operating point (source ramp + damped Newton), warp-per-instance
//   tran_kernel      : backward-Euler step loop + damped Newton, warp-per-instance
#include <cuda_runtime.h>
#include <math_constants.h>

#include <cstdint>

#include "kernels.hpp"
#include "newton_warp.cuh"
#include "spec_runtime.cuh"

namespace csim {

// =============================================================================
// standalone assembly
// =============================================================================
// One thread per instance; the value table of a thread lives in shared memory as
// vt[v][threadIdx.x] (conflict-free).  Inputs are read once, coalesced across the
// warp (instance-minor layout); every output slot is written once, coalesced.
template <int MODE>
__global__ void __launch_bounds__(kStampThreads)
stamp_csr_kernel(DevPlan P, int64_t B, const double* __restrict__ params, const double* __restrict__ x,
                 const double* __restrict__ x_prev, const double* __restrict__ src_val, double dt, double scale,
                 double gmin, double* __restrict__ vals, double* __restrict__ rhs) {
    extern __shared__ double smem[];
    const DevMode& m = P.m[MODE];
    const int tid = threadIdx.x;
    double* vt = smem + tid;  // stride blockDim.x
    const int TS = blockDim.x;
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x + tid; b < B; b += (int64_t)gridDim.x * blockDim.x) {
        for (int i = 0; i < m.n_static; ++i) {
            StaticOp op = m.static_ops[i];
            double p = op.param >= 0 ? __ldg(params + (int64_t)op.param * B + b) : 0.0;
            double v;
            switch (op.kind) {
                case SK_ONE: v = 1.0; break;
                case SK_GMIN: v = gmin; break;
                case SK_RES: v = (p == 0.0) ? 0.0 : 1.0 / p; break;
                case SK_CAP: v = (p <= 0.0 || dt <= 0.0) ? 0.0 : p / dt; break;
                case SK_CAPHALF: {
                    double c = d_mul(0.5, p);
                    v = (c <= 0.0 || dt <= 0.0) ? 0.0 : c / dt;
                } break;
                case SK_IND_ONE: v = (p <= 0.0) ? 0.0 : 1.0; break;
                case SK_IND_NREQ: v = (p <= 0.0) ? 0.0 : -(p / dt); break;
                default: v = 0.0;
            }
            vt[op.out * TS] = v;
        }
        if (MODE == CSIM_MODE_TRAN) {
            for (int i = 0; i < m.n_step; ++i) {
                StepOp op = m.step_ops[i];
                double va = op.a >= 0 ? __ldg(x_prev + (int64_t)op.a * B + b) : 0.0;
                double v;
                if (op.kind == TK_CAP_HIST) {
                    double vb = op.b >= 0 ? __ldg(x_prev + (int64_t)op.b * B + b) : 0.0;
                    v = d_mul(-vt[op.g * TS], d_sub(va, vb));
                } else {
                    v = d_mul(vt[op.g * TS], va);
                }
                vt[op.out * TS] = v;
            }
        }
        for (int i = 0; i < m.n_src; ++i) {
            SrcOp op = m.src_ops[i];
            double v;
            if (src_val) {
                v = __ldg(src_val + (int64_t)op.index * B + b);
            } else {
                double base = __ldg(params + (int64_t)op.param * B + b);
                if (op.is_sin) base = d_add(base, __ldg(params + (int64_t)(op.param + 1) * B + b));
                v = d_mul(base, scale);
            }
            vt[op.out * TS] = v;
        }
        for (int i = 0; i < m.n_mos; ++i) {
            MosOp op = m.mos_ops[i];
            double Vth = __ldg(params + (int64_t)op.param * B + b);
            double K = __ldg(params + (int64_t)(op.param + 1) * B + b);
            double lambda = __ldg(params + (int64_t)(op.param + 2) * B + b);
            double Vd = op.d >= 0 ? __ldg(x + (int64_t)op.d * B + b) : 0.0;
            double Vg = op.g >= 0 ? __ldg(x + (int64_t)op.g * B + b) : 0.0;
            double Vs = op.s >= 0 ? __ldg(x + (int64_t)op.s * B + b) : 0.0;
            double gd, gg, gs, cst;
            mos_eval(Vth, K, lambda, op.is_p != 0, P.c.mos_off_gds, Vd, Vg, Vs, gd, gg, gs, cst);
            vt[op.out * TS] = gd;
            vt[(op.out + 1) * TS] = gg;
            vt[(op.out + 2) * TS] = gs;
            vt[(op.out + 3) * TS] = cst;
        }
        // ordered gather per slot
        int e = m.slot_ptr[0];
        for (int s = 0; s < m.n_slots; ++s) {
            const int e1 = m.slot_ptr[s + 1];
            double acc = 0.0;
            for (; e < e1; ++e) {
                uint32_t w = m.slot_entries[e];
                double v = vt[(w >> 8) * TS];
                acc = (w & 64u) ? d_sub(acc, v) : d_add(acc, v);
            }
            if (s < m.nnz) vals[(int64_t)s * B + b] = acc;
            else rhs[(int64_t)(s - m.nnz) * B + b] = acc;
        }
    }
}

// =============================================================================
// batched dense LU solve (parity probe for the solver)
// =============================================================================
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
lu_solve_kernel(int N, int64_t B, const double* __restrict__ A, const double* __restrict__ bvec,
                double* __restrict__ xout, int32_t* __restrict__ perm_out, int32_t* __restrict__ singular,
                double eps, double* __restrict__ lu_out) {
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    double* S = smem + (size_t)wib * (32 * kLD + 32);
    double* bs = S + 32 * kLD;
    const int64_t b = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
    if (b >= B) return;
    for (int idx = lane; idx < N * N; idx += 32) {
        int i = idx / N, j = idx - i * N;
        S[i * kLD + j] = A[(int64_t)idx * B + b];
    }
    bs[lane] = (lane < N && bvec) ? bvec[(int64_t)lane * B + b] : 0.0;
    __syncwarp();
    int perm;
    bool sing = lu_solve_warp(S, bs, N, lane, eps, perm);
    if (lu_out) {  // Solver::luDecompose: S holds the packed L\U, rows exchanged physically as the reference does
        __syncwarp();
        for (int idx = lane; idx < N * N; idx += 32) {
            int i = idx / N, j = idx - i * N;
            lu_out[(int64_t)idx * B + b] = S[i * kLD + j];
        }
    }
    if (lane < N) {
        if (xout) xout[(int64_t)lane * B + b] = bs[lane];
        if (perm_out) perm_out[(int64_t)lane * B + b] = perm;
    }
    if (lane == 0 && singular) singular[b] = sing ? 1 : 0;
}

// =============================================================================
// fused Newton kernels
// =============================================================================
struct BlockProg {  // block-shared copy of the row program
    uint32_t* words;
};

__device__ __forceinline__ const uint32_t* load_prog(const DevMode& m, uint32_t* dst) {
    for (int i = threadIdx.x; i < m.row_len * 32; i += blockDim.x) dst[i] = m.row_prog[i];
    __syncthreads();
    return dst;
}

// ConvController::baseGmin (include/dcanalysis.hpp:45-48)
__device__ __forceinline__ double base_gmin(const csim_constants& c, double s) {
    if (s < 0.0) s = 0.0;
    if (s > 1.0) s = 1.0;
    return d_add(d_mul(c.gmin_high_base, d_sub(1.0, s)), d_mul(c.gmin_low_base, s));
}

// relative distance of a step norm from a threshold it is compared with (diagnostic of the one unpinned call, ||.||_2)
__device__ __forceinline__ double rel_margin(double err, double thr) { return fabs(err - thr) / thr; }

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
dc_kernel(DevPlan P, int64_t B, const double* __restrict__ params, double* __restrict__ xout,
          int32_t* __restrict__ iters, uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status,
          double* __restrict__ margin) {
    extern __shared__ double smem[];
    const DevMode& m = P.m[CSIM_MODE_DC];
    const csim_constants& c = P.c;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int N = P.N;
    const size_t wd = warp_mem_doubles(m.n_values);
    uint32_t* progbuf = reinterpret_cast<uint32_t*>(smem + wd * kWarpsPerBlock);
    const uint32_t* prog = load_prog(m, progbuf);
    const int64_t b = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
    if (b >= B) return;
    WarpMem w = carve(smem + wd * wib, m.n_values);

    uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    double mg = CUDART_INF;
    double x = 0.0;  // lane i holds x(i)
    w.xs[lane] = 0.0;
    eval_static(m, params, B, b, 0.0, 0.0, w.vt, lane);
    __syncwarp();
    int gmin_slot = -1;
    for (int i = 0; i < m.n_static; ++i)
        if (m.static_ops[i].kind == SK_GMIN) gmin_slot = m.static_ops[i].out;

    if (!P.has_mos) {
        // dcSolveDirectLU (src/dcanalysis.cpp:46-68): one solve at scale 1, no gmin, no damping
        eval_src_dc(m, params, B, b, 1.0, w.vt, lane);
        __syncwarp();
        assemble(m, prog, w, N, lane);
        int perm;
        if (lu_solve_warp(w.S, w.bs, N, lane, c.lu_eps, perm)) st |= CSIM_ST_LU_SINGULAR;
        h = hash_update(h, perm, N, lane);
        x = w.bs[lane];
        if (iters && lane < c.dc_ramp_steps) iters[(int64_t)lane * B + b] = lane == 0 ? 1 : 0;
    } else {
        // dcSolveNewtonLU (src/dcanalysis.cpp:95-163)
        for (int step = 1; step <= c.dc_ramp_steps; ++step) {
            const double scale = (double)step / (double)c.dc_ramp_steps;
            double gmin = base_gmin(c, scale);
            double prevErr = CUDART_INF;
            eval_src_dc(m, params, B, b, scale, w.vt, lane);
            int count = 0;
            for (int iter = 0; iter < c.dc_max_iters; ++iter) {
                ++count;
                if (lane == 0 && gmin_slot >= 0) w.vt[gmin_slot] = gmin;
                eval_mos(m, params, B, b, c.mos_off_gds, w.xs, w.vt, lane);
                __syncwarp();
                assemble(m, prog, w, N, lane);
                int perm;
                if (lu_solve_warp(w.S, w.bs, N, lane, c.lu_eps, perm)) st |= CSIM_ST_LU_SINGULAR;
                h = hash_update(h, perm, N, lane);
                const double xr = lane < N ? w.bs[lane] : 0.0;
                if (__any_sync(kFull, !lane_finite(xr))) {  // dcanalysis.cpp:135-138
                    st |= CSIM_ST_NONFINITE;
                    double g10 = d_mul(gmin, c.dc_nonfinite_mul);
                    gmin = g10 < c.dc_nonfinite_cap ? g10 : c.dc_nonfinite_cap;
                    continue;
                }
                // ConvController::update (src/dcanalysis.cpp:268-307)
                const double xn = d_add(x, d_mul(c.dc_alpha, d_sub(xr, x)));
                const double err = warp_norm2(d_sub(xn, x));
                const double gbase = base_gmin(c, scale);
                double gnext = gbase;
                if (margin) {  // every threshold this iteration compares err with
                    mg = fmin(mg, rel_margin(err, c.dc_tol));
                    if (!(iter == 0 || !lane_finite(prevErr))) {
                        mg = fmin(mg, rel_margin(err, d_mul(prevErr, c.slow_conv_ratio)));
                        if (!(err > d_mul(prevErr, c.slow_conv_ratio))) mg = fmin(mg, rel_margin(err, d_mul(prevErr, c.fast_conv_ratio)));
                    }
                }
                if (!(iter == 0 || !lane_finite(prevErr))) {
                    if (err > d_mul(prevErr, c.slow_conv_ratio)) {
                        double g2 = d_mul(gmin, 2.0);
                        gnext = g2 < c.gmin_abs_max ? g2 : c.gmin_abs_max;
                    } else if (err < d_mul(prevErr, c.fast_conv_ratio)) {
                        gnext = d_add(d_mul(0.5, gmin), d_mul(0.5, gbase));
                    } else {
                        gnext = d_add(d_mul(0.7, gmin), d_mul(0.3, gbase));
                    }
                }
                x = xn;
                __syncwarp();
                w.xs[lane] = x;
                __syncwarp();
                gmin = gnext;
                prevErr = err;
                if (err < c.dc_tol) break;
                if (iter == c.dc_max_iters - 1) st |= CSIM_ST_DC_UNCONVERGED;
            }
            if (iters && lane == 0) iters[(int64_t)(step - 1) * B + b] = count;
        }
    }
    if (lane < N) xout[(int64_t)lane * B + b] = x;
    if (lane == 0) {
        if (pivot_hash) pivot_hash[b] = h;
        if (status) status[b] |= st;
        if (margin) margin[b] = fmin(margin[b], mg);
    }
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
tran_kernel(DevPlan P, DevTran T, int64_t B, const double* __restrict__ params, double* __restrict__ xio,
            double* __restrict__ wave, int32_t* __restrict__ iters_total, uint8_t* __restrict__ iters_step,
            uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status) {
    extern __shared__ double smem[];
    const DevMode& m = P.m[CSIM_MODE_TRAN];
    const csim_constants& c = P.c;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int N = P.N;
    const size_t wd = warp_mem_doubles(m.n_values);
    uint32_t* progbuf = reinterpret_cast<uint32_t*>(smem + wd * kWarpsPerBlock);
    const uint32_t* prog = load_prog(m, progbuf);
    const int64_t b = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
    if (b >= B || (T.skip && T.skip[b])) return;
    WarpMem w = carve(smem + wd * wib, m.n_values);

    uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    int total = 0;
    double mg = CUDART_INF;
    double x = lane < N ? xio[(int64_t)lane * B + b] : 0.0;
    w.xs[lane] = x;
    eval_static(m, params, B, b, T.dt, c.tran_gmin, w.vt, lane);
    __syncwarp();
    const int r0 = T.out_begin_step / T.out_stride;
    const int probe = lane < T.n_probes ? T.probes[lane] : 0;

    for (int step = T.step_begin; step < T.step_end; ++step) {
        // companion-model state = functions of the last accepted x (tanalisis.cpp:139-180, 381-417)
        w.xp[lane] = x;
        __syncwarp();
        eval_step(m, w.xp, w.vt, lane);
        eval_src_tran(m, T, params, B, b, step, w.vt, lane);
        int count = 0;
        bool dead = false;
        for (int iter = 0; iter < c.tran_max_iters; ++iter) {  // tanalisis.cpp:258
            ++count;
            eval_mos(m, params, B, b, c.mos_off_gds, w.xs, w.vt, lane);
            __syncwarp();
            assemble(m, prog, w, N, lane);
            int perm;
            if (lu_solve_warp(w.S, w.bs, N, lane, c.lu_eps, perm)) st |= CSIM_ST_LU_SINGULAR;
            h = hash_update(h, perm, N, lane);
            const double xr = lane < N ? w.bs[lane] : 0.0;
            if (__any_sync(kFull, !lane_finite(xr))) {  // tanalisis.cpp:360-362 (the reference throws)
                st |= CSIM_ST_NONFINITE;
                dead = true;
                break;
            }
            const double xn = d_add(x, d_mul(c.tran_alpha, d_sub(xr, x)));  // tanalisis.cpp:365
            const double err = warp_norm2(d_sub(xn, x));                   // tanalisis.cpp:366
            x = xn;
            __syncwarp();
            w.xs[lane] = x;
            __syncwarp();
            if (T.margin) mg = fmin(mg, rel_margin(err, c.tran_tol));
            if (err < c.tran_tol) break;
            if (iter == c.tran_max_iters - 1) st |= CSIM_ST_TRAN_UNCONVERGED;
        }
        total += count;
        if (iters_step && lane == 0) iters_step[(int64_t)step * B + b] = (uint8_t)count;
        if (dead) break;
        if ((step + 1) % T.out_stride == 0 && step >= T.out_begin_step && lane < T.n_probes) {
            const int64_t r = (step + 1) / T.out_stride - 1 - r0;
            wave[((int64_t)b * T.n_probes + lane) * T.wave_ld + r] = w.xs[probe];
        }
    }
    if (lane < N) xio[(int64_t)lane * B + b] = x;
    if (lane == 0) {
        if (iters_total) iters_total[b] += total;
        if (pivot_hash) pivot_hash[b] = h;
        if (status) status[b] |= st;
        if (T.margin) T.margin[b] = fmin(T.margin[b], mg);
    }
}

// =============================================================================
// steady-state driver: periodicity check after every period (definition: include/csim_b200.h)
// =============================================================================
// Thread per instance.  delta = || x - x_prev ||_2 in the summation order of oracle/csim_norm2.h.
__global__ void sstate_check_kernel(int N, int64_t B, const double* __restrict__ x, const double* __restrict__ x_prev,
                                    double tol, int period, uint8_t* __restrict__ skip, int32_t* __restrict__ periods,
                                    double* __restrict__ delta, unsigned long long* __restrict__ n_open) {
    const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B || skip[b]) return;
    double acc[32];
#pragma unroll
    for (int l = 0; l < 32; ++l) acc[l] = 0.0;
    for (int i = 0; i < N; ++i) {
        const double d = d_sub(x[(int64_t)i * B + b], x_prev[(int64_t)i * B + b]);
        acc[i & 31] = d_add(acc[i & 31], d_mul(d, d));
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1)
#pragma unroll
        for (int i = 0; i < s; ++i) acc[i] = d_add(acc[i], acc[i + s]);
    const double nrm = sqrt(acc[0]);
    delta[b] = nrm;
    if (nrm < tol) {
        periods[b] = period;
        skip[b] = 1;
    } else {
        atomicAdd(n_open, 1ull);
    }
}

cudaError_t launch_sstate_check(int N, int64_t B, const double* x, const double* x_prev, double tol, int period,
                                uint8_t* skip, int32_t* periods, double* delta, unsigned long long* n_open,
                                cudaStream_t stream) {
    sstate_check_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(N, B, x, x_prev, tol, period, skip, periods,
                                                                         delta, n_open);
    return cudaGetLastError();
}

// =============================================================================
// Solver::solveLinearSystemGaussSeidel (include/solver.hpp:139-192), batched: row (f)4 of SURVEY 8
// =============================================================================
// A sweep is one sequential chain under the reference's order (row i subtracts a(i,j) x(j) for ascending j, the
// newest x for j < i), so the parallelism is across systems only: one THREAD per system, the two iterates in
// shared memory at [j][tid] (conflict-free), the matrix streamed from HBM/L2 every sweep (instance-minor:
// consecutive threads read consecutive addresses).  ||x - xOld||_2 in the order of oracle/csim_norm2.h.
constexpr int kGsThreads = 64;

__global__ void __launch_bounds__(kGsThreads)
gauss_seidel_kernel(int n, int64_t B, const double* __restrict__ A, const double* __restrict__ rhs,
                    double* __restrict__ xio, int max_iters, double tol, int32_t* __restrict__ sweeps) {
    extern __shared__ double gs_sm[];
    const int tid = threadIdx.x;
    const int64_t b = (int64_t)blockIdx.x * kGsThreads + tid;
    if (b >= B) return;
    double* x = gs_sm + tid;                          // x[j] at x[j * kGsThreads]
    double* xo = gs_sm + (size_t)n * kGsThreads + tid;
    for (int j = 0; j < n; ++j) x[j * kGsThreads] = xio[(int64_t)j * B + b];
    const double diagEps = 1e-12;  // solver.hpp:160
    int done = 0;
    for (int iter = 0; iter < max_iters; ++iter) {  // solver.hpp:162
        for (int j = 0; j < n; ++j) xo[j * kGsThreads] = x[j * kGsThreads];
        for (int i = 0; i < n; ++i) {
            const double* row = A + (int64_t)i * n * B + b;
            double diag = row[(int64_t)i * B];
            if (fabs(diag) < diagEps) diag = (diag >= 0.0 ? 1.0 : -1.0) * diagEps;  // solver.hpp:169-173
            double sum = rhs[(int64_t)i * B + b];
#pragma unroll 4
            for (int j = 0; j < i; ++j) sum = d_sub(sum, d_mul(row[(int64_t)j * B], x[j * kGsThreads]));
#pragma unroll 4
            for (int j = i + 1; j < n; ++j) sum = d_sub(sum, d_mul(row[(int64_t)j * B], xo[j * kGsThreads]));
            x[i * kGsThreads] = sum / diag;
        }
        // || x - xOld ||_2: 32 leaves (element k goes to leaf k mod 32, ascending k), then the pairwise tree
        // 16, 8, 4, 2, 1 — leaf 0's path, evaluated by this one thread
        double leaf[32];
#pragma unroll
        for (int l = 0; l < 32; ++l) {
            double acc = 0.0;
            for (int k = l; k < n; k += 32) {
                const double d = d_sub(x[k * kGsThreads], xo[k * kGsThreads]);
                acc = d_add(acc, d_mul(d, d));
            }
            leaf[l] = acc;
        }
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1)
#pragma unroll
            for (int l = 0; l < s; ++l) leaf[l] = d_add(leaf[l], leaf[l + s]);
        done = iter + 1;
        if (sqrt(leaf[0]) < tol) break;  // solver.hpp:188-192
    }
    for (int j = 0; j < n; ++j) xio[(int64_t)j * B + b] = x[j * kGsThreads];
    if (sweeps) sweeps[b] = done;
}

cudaError_t launch_gauss_seidel(int n, int64_t B, const double* A, const double* rhs, double* x, int max_iters,
                                double tol, int32_t* sweeps, cudaStream_t stream) {
    const size_t smem = (size_t)2 * n * kGsThreads * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(gauss_seidel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gauss_seidel_kernel<<<(unsigned)((B + kGsThreads - 1) / kGsThreads), kGsThreads, smem, stream>>>(
        n, B, A, rhs, x, max_iters, tol, sweeps);
    return cudaGetLastError();
}

// =============================================================================
// harmonic extraction over one period (row (f)2 of SURVEY 8: the `.hb f0 nharm` card's nHarm)
// =============================================================================
// X_h = (1/M) sum_{n<M} x[n] (c[(h n) mod M] - j s[(h n) mod M]),  h = 0..H, c/s tabulated on the host.
// Summation order (part of the definition, include/csim_b200.h): lane l adds n = l, l+32, ... in ascending
// order, each product and each sum separately rounded; the 32 partial sums are combined by the xor
// butterfly 16, 8, 4, 2, 1; the result is divided by M.
// Mapping: one warp per 8 signals, 4 harmonics per pass.  The accumulators (8 x 4 x re/im doubles per lane) fill
// the register file, so loads held in registers could keep only ~20 KB per SM in flight (2.3 TB/s measured);
// instead every lane copies ITS OWN samples with 8-byte cp.async into a private 3-stage ring in shared memory
// (8 signals x 2 rows per stage; no barrier needed, a lane reads back only what it copied itself) and waits
// on the oldest group.  The 4 + 4 table values of each of its sample indices travel the same way (they come out
// of L1/L2, but a load-then-use in the compute loop stalled every row on the L2 latency: ncu long-scoreboard
// 4.3 per issue), so the compute loop touches shared memory only.  Per signal the order of the definition is
// untouched.
static_assert(kHarmonicsMax == 64, "csim_b200.h documents n_harm <= 64");
constexpr int kHarmSig = 8;     // signals per warp
constexpr int kHarmGrp = 4;     // harmonics per pass over the samples (H <= 3: one pass = 8 bytes of HBM per sample)
constexpr int kHarmWarps = 4;   // warps per CTA
constexpr int kHarmRows = 2;    // 32-sample rows per stage
constexpr int kHarmStages = 3;
constexpr int kHarmSlots = kHarmSig + 2 * kHarmGrp;  // per row and lane: 8 samples + 4 cos + 4 sin table values
constexpr int kHarmStageDoubles = kHarmSlots * kHarmRows * 32;  // per warp

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__global__ void __launch_bounds__(kHarmWarps * 32)
harmonics_kernel(int64_t n_signals, int M, int64_t ld, const double* __restrict__ x, int H,
                 const double* __restrict__ ct, const double* __restrict__ st, double* __restrict__ re,
                 double* __restrict__ im) {
    extern __shared__ double harm_sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t sig0 = ((int64_t)blockIdx.x * kHarmWarps + warp) * kHarmSig;
    if (sig0 >= n_signals) return;
    double* ring = harm_sm + (size_t)warp * kHarmStages * kHarmStageDoubles + lane;  // [stage][s][k][lane]
    const double* row[kHarmSig];
#pragma unroll
    for (int s = 0; s < kHarmSig; ++s) row[s] = x + (sig0 + s < n_signals ? sig0 + s : sig0) * ld;
    const int n_rows = (M + 31) / 32, n_chunks = (n_rows + kHarmRows - 1) / kHarmRows;
    int idx[kHarmGrp], inc[kHarmGrp];
    auto issue = [&](int chunk) {  // this lane's samples + table values of chunk -> stage chunk % kHarmStages
        if (chunk < n_chunks) {
            double* dst = ring + (size_t)(chunk % kHarmStages) * kHarmStageDoubles;
#pragma unroll
            for (int k = 0; k < kHarmRows; ++k) {
                const int n = (chunk * kHarmRows + k) * 32 + lane;
                if (n < M) {
#pragma unroll
                    for (int s = 0; s < kHarmSig; ++s) cp_async8(dst + (s * kHarmRows + k) * 32, row[s] + n);
#pragma unroll
                    for (int g = 0; g < kHarmGrp; ++g) {
                        cp_async8(dst + ((kHarmSig + g) * kHarmRows + k) * 32, ct + idx[g]);
                        cp_async8(dst + ((kHarmSig + kHarmGrp + g) * kHarmRows + k) * 32, st + idx[g]);
                        idx[g] += inc[g];  // index of this lane's next sample: (h * (n + 32)) mod M
                        if (idx[g] >= M) idx[g] -= M;
                    }
                }
            }
        }
        cp_async_commit();  // (an empty group keeps the wait count uniform)
    };
    for (int h0 = 0; h0 <= H; h0 += kHarmGrp) {
        double ar[kHarmSig][kHarmGrp], ai[kHarmSig][kHarmGrp];
#pragma unroll
        for (int g = 0; g < kHarmGrp; ++g) {
            idx[g] = (int)(((long long)(h0 + g) * lane) % M);
            inc[g] = (int)(((long long)(h0 + g) * 32) % M);
#pragma unroll
            for (int s = 0; s < kHarmSig; ++s) ar[s][g] = ai[s][g] = 0.0;
        }
#pragma unroll
        for (int c = 0; c < kHarmStages - 1; ++c) issue(c);
        for (int chunk = 0; chunk < n_chunks; ++chunk) {
            issue(chunk + kHarmStages - 1);         // refills the stage consumed in the previous iteration
            cp_async_wait<kHarmStages - 1>();       // this chunk's group has landed
            const double* src = ring + (size_t)(chunk % kHarmStages) * kHarmStageDoubles;
#pragma unroll
            for (int k = 0; k < kHarmRows; ++k) {
                const int n = (chunk * kHarmRows + k) * 32 + lane;
                if (n >= M) break;
                double c[kHarmGrp], sn[kHarmGrp];
#pragma unroll
                for (int g = 0; g < kHarmGrp; ++g) {
                    c[g] = src[((kHarmSig + g) * kHarmRows + k) * 32];
                    sn[g] = src[((kHarmSig + kHarmGrp + g) * kHarmRows + k) * 32];
                }
#pragma unroll
                for (int s = 0; s < kHarmSig; ++s) {
                    const double v = src[(s * kHarmRows + k) * 32];
#pragma unroll
                    for (int g = 0; g < kHarmGrp; ++g) {
                        ar[s][g] = d_add(ar[s][g], d_mul(v, c[g]));
                        ai[s][g] = d_add(ai[s][g], d_mul(v, sn[g]));
                    }
                }
            }
        }
        cp_async_wait<0>();
        __syncwarp();
#pragma unroll
        for (int s = 0; s < kHarmSig; ++s)
#pragma unroll
            for (int g = 0; g < kHarmGrp; ++g) {
                double a = ar[s][g], b = ai[s][g];
#pragma unroll
                for (int t = 16; t >= 1; t >>= 1) {
                    a = d_add(a, __shfl_xor_sync(kFull, a, t));
                    b = d_add(b, __shfl_xor_sync(kFull, b, t));
                }
                if (lane == 0 && h0 + g <= H && sig0 + s < n_signals) {
                    re[(int64_t)(h0 + g) * n_signals + sig0 + s] = a / (double)M;
                    im[(int64_t)(h0 + g) * n_signals + sig0 + s] = -(b / (double)M);
                }
            }
    }
}

cudaError_t launch_harmonics(int64_t n_signals, int M, int64_t ld, const double* x, int H, const double* ct,
                             const double* st, double* re, double* im, cudaStream_t stream) {
    const int64_t per_cta = (int64_t)kHarmWarps * kHarmSig;
    const size_t smem = (size_t)kHarmWarps * kHarmStages * kHarmStageDoubles * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(harmonics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    harmonics_kernel<<<(unsigned)((n_signals + per_cta - 1) / per_cta), kHarmWarps * 32, smem, stream>>>(
        n_signals, M, ld, x, H, ct, st, re, im);
    return cudaGetLastError();
}

// =============================================================================
// self-test of the split IEEE division used by the specialised kernels
// =============================================================================
__global__ void div_selftest_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b,
                                    double* __restrict__ q_split, double* __restrict__ q_ref) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double y = rcp_seq(b[i]);
        q_split[i] = fdiv(a[i], b[i], y, true);
        q_ref[i] = a[i] / b[i];
    }
}

// tdiv (thread-per-instance kernel): quotient, IEEE quotient, and whether tdiv's own range guards accept the
// division (in_range == 1 promises q == q_ref bit for bit)
__global__ void tdiv_selftest_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b,
                                     double* __restrict__ q, double* __restrict__ q_ref, int32_t* __restrict__ in_range) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned int glo = 0xFFFFFFFFu, ghi = 0u;
        q[i] = tdiv(a[i], b[i], rcp_seq(b[i]), glo, ghi);
        q_ref[i] = a[i] / b[i];
        in_range[i] = !(glo < 0x03600000u || ghi > 0x7FEFFFFFu);
    }
}

// tdiv_mul (the multiplier division with folded guards): flags bit 0 = the range guard accepts the division (then q
// must equal q_ref bit for bit), bit 1 = the pivot guard fires (|q| >= 1 on the high word)
__global__ void tdiv_mul_selftest_kernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b,
                                         double* __restrict__ q, double* __restrict__ q_ref, int32_t* __restrict__ flags) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned int gacc = 0u, pmx = 0u;
        q[i] = tdiv_mul(a[i], b[i], rcp_seq(b[i]), gacc, pmx);
        q_ref[i] = a[i] / b[i];
        flags[i] = (gacc <= kTdivTMax ? 1 : 0) | (pmx >= kTdivTOne ? 2 : 0);
    }
}

cudaError_t launch_tdiv_mul_selftest(int64_t n, const double* a, const double* b, double* q, double* q_ref,
                                     int32_t* flags, cudaStream_t stream) {
    tdiv_mul_selftest_kernel<<<1184, 256, 0, stream>>>(n, a, b, q, q_ref, flags);
    return cudaGetLastError();
}

cudaError_t launch_tdiv_selftest(int64_t n, const double* a, const double* b, double* q, double* q_ref,
                                 int32_t* in_range, cudaStream_t stream) {
    tdiv_selftest_kernel<<<1184, 256, 0, stream>>>(n, a, b, q, q_ref, in_range);
    return cudaGetLastError();
}

cudaError_t launch_div_selftest(int64_t n, const double* a, const double* b, double* q_split, double* q_ref,
                                cudaStream_t stream) {
    div_selftest_kernel<<<1184, 256, 0, stream>>>(n, a, b, q_split, q_ref);
    return cudaGetLastError();
}

// =============================================================================
// FP64 pipe probe: the roofline denominator of the fused kernels, measured in the run.  Register-resident, eight
// independent chains per thread of SEPARATE multiply and add (FMA is not allowed on the reference's expressions,
// so the FMA peak does not apply).  Returns Tflop/s over the whole chip.
// =============================================================================
__global__ void __launch_bounds__(256) fp64_peak_kernel(double* __restrict__ out, double a, double b, int iters) {
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = a + k + threadIdx.x * 1e-6;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = __dadd_rn(__dmul_rn(v[k], b), a);
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

cudaError_t measure_fp64_peak(int sm_count, double* tflops, cudaStream_t stream) {
    const int blocks = sm_count * 8, iters = 20000;
    double* out = nullptr;
    cudaError_t e = cudaMalloc((void**)&out, sizeof(double) * (size_t)blocks * 256);
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {  // first pass warms the clocks up
        cudaEventRecord(e0, stream);
        fp64_peak_kernel<<<blocks, 256, 0, stream>>>(out, 1.0, 0.9999999, iters);
        cudaEventRecord(e1, stream);
        e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flop = 2.0 * (double)blocks * 256 * 8 * iters;
        if (rep > 0 && ms > 0.f) best = flop / ms / 1e9 > best ? flop / ms / 1e9 : best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return e == cudaSuccess ? cudaGetLastError() : e;
}

// =============================================================================
// launch wrappers
// =============================================================================
static size_t newton_smem_bytes(const DevMode& m) {
    return (warp_mem_doubles(m.n_values) * kWarpsPerBlock) * sizeof(double) + (size_t)m.row_len * 32 * sizeof(uint32_t);
}

cudaError_t launch_stamp_csr(const DevPlan& P, int mode, int64_t B, const double* params, const double* x,
                             const double* x_prev, const double* src_val, double dt, double scale, double gmin,
                             double* vals, double* rhs, int sm_count, cudaStream_t stream) {
    const DevMode& m = P.m[mode];
    size_t smem = (size_t)m.n_values * kStampThreads * sizeof(double);
    int64_t blocks = (B + kStampThreads - 1) / kStampThreads;
    (void)sm_count;
    if (mode == CSIM_MODE_DC) {
        cudaFuncSetAttribute(stamp_csr_kernel<CSIM_MODE_DC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        stamp_csr_kernel<CSIM_MODE_DC><<<(unsigned)blocks, kStampThreads, smem, stream>>>(
            P, B, params, x, x_prev, src_val, dt, scale, gmin, vals, rhs);
    } else {
        cudaFuncSetAttribute(stamp_csr_kernel<CSIM_MODE_TRAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        stamp_csr_kernel<CSIM_MODE_TRAN><<<(unsigned)blocks, kStampThreads, smem, stream>>>(
            P, B, params, x, x_prev, src_val, dt, scale, gmin, vals, rhs);
    }
    return cudaGetLastError();
}

cudaError_t launch_lu_solve(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                            int32_t* singular, double eps, cudaStream_t stream, double* lu_out) {
    size_t smem = (size_t)kWarpsPerBlock * (32 * kLD + 32) * sizeof(double);
    cudaFuncSetAttribute(lu_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int64_t blocks = (B + kWarpsPerBlock - 1) / kWarpsPerBlock;
    lu_solve_kernel<<<(unsigned)blocks, kWarpsPerBlock * 32, smem, stream>>>(N, B, A, b, x, perm, singular, eps, lu_out);
    return cudaGetLastError();
}

cudaError_t launch_dc(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                      uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream, double* margin) {
    size_t smem = newton_smem_bytes(P.m[CSIM_MODE_DC]);
    cudaFuncSetAttribute(dc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int64_t blocks = (B + kWarpsPerBlock - 1) / kWarpsPerBlock;
    dc_kernel<<<(unsigned)blocks, kWarpsPerBlock * 32, smem, stream>>>(P, B, params, x, iters, pivot_hash, status, margin);
    return cudaGetLastError();
}

cudaError_t launch_tran(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x, double* wave,
                        int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                        cudaStream_t stream) {
    size_t smem = newton_smem_bytes(P.m[CSIM_MODE_TRAN]);
    cudaFuncSetAttribute(tran_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int64_t blocks = (B + kWarpsPerBlock - 1) / kWarpsPerBlock;
    tran_kernel<<<(unsigned)blocks, kWarpsPerBlock * 32, smem, stream>>>(P, T, B, params, x, wave, iters_total,
                                                                         iters_step, pivot_hash, status);
    return cudaGetLastError();
}

}  // namespace csim
