// big_kernels.cu — the large-matrix path (N > 32, e.g. the 200-stage ladder, N = 404).
//
// One warp owns one instance.  The dense N x N matrix does not fit on chip (404^2 * 8 B = 1.3 MB), so
// it lives in HBM/L2, twice: a row-major copy Ar (pivot-row scans and updates are coalesced) and a
// column-major copy Ac (pivot searches down a column are coalesced).  The reference's dense LU with
// value-dependent partial pivoting (include/solver.hpp:30-131) is run with LOGICAL row exchanges
// (position <-> row tables in shared memory) and with exact-zero skipping, which is bit-identical
// to the dense loops for finite data: per elimination step the warp scans one column
// (pivot = first maximum in POSITION order, strict '>'), one pivot row, and touches only rows /
// columns whose entries are not exactly zero.  Per solve that is ~N^2 coalesced 8-byte loads
// (1.3 MB at N = 404) — HBM/L2-bandwidth bound, not FP64 bound.  Entries written during a solve are
// logged and re-zeroed afterwards, so the matrix is never cleared as a whole between Newton steps.
//
// Small per-instance vectors (iterate, right-hand side, solution, permutation tables) sit in shared
// memory; the value table sits in global memory, instance-major.
#include <cuda_runtime.h>
#include <math_constants.h>

#include <cstdint>

#include "kernels.hpp"
#include "newton_warp.cuh"

namespace csim {

constexpr int kBigWarps = 4;      // instances per CTA
constexpr int kBigFList = 64;     // rows eliminated per pass of one elimination step
constexpr int kBigBatch = 8;      // independent 8-byte loads per lane issued before the first use (256 entries per batch)
constexpr int kBigUList = 32 * kBigBatch;  // non-zero pivot-row entries compacted per pass (one batch of columns)
constexpr int kBigPairs = 4;      // (row, column) updates per lane in flight at once

struct BigWarp {  // per-warp shared memory
    int32_t* rowat;  // [N] position -> row
    int32_t* pos;    // [N] row -> position
    double* bs;      // [N] right-hand side by ROW (forward substitution is fused)
    double* xr;      // [N] raw solution
    double* xs;      // [N] Newton iterate
    int32_t* frow;   // [kBigFList] rows being eliminated in this pass
    double* fval;    // [kBigFList] their multipliers
    int32_t* ucol;   // [kBigUList] columns of the pivot row's non-zeros in the current batch of columns
    double* uval;    // [kBigUList] their values
};

__host__ __device__ inline size_t big_warp_bytes(int N) {
    return (size_t)N * (4 + 4 + 8 + 8 + 8) + (size_t)(kBigFList + kBigUList) * 12 + 64;
}

__device__ __forceinline__ BigWarp big_carve(unsigned char* base, int N) {
    BigWarp w;
    double* d = reinterpret_cast<double*>(base);
    w.bs = d;
    w.xr = d + N;
    w.xs = d + 2 * N;
    w.fval = d + 3 * N;
    w.uval = d + 3 * N + kBigFList;
    int32_t* q = reinterpret_cast<int32_t*>(d + 3 * N + kBigFList + kBigUList);
    w.rowat = q;
    w.pos = q + N;
    w.frow = q + 2 * N;
    w.ucol = q + 2 * N + kBigFList;
    return w;
}

// per-instance global memory of the large-matrix path
struct BigMem {
    double* Ar;       // [B][N*N] row-major
    double* Ac;       // [B][N*N] column-major
    double* vt;       // [B][n_values]
    int32_t* fill;    // [B][fill_cap] linear indices (row*N+col) of entries created by elimination
    int32_t fill_cap;
};

// ordered gather per slot (reference accumulation order) into Ar / Ac / bs; G must be all-zero on entry
__device__ __forceinline__ void big_assemble(const DevMode& m, const double* vt, double* Ar, double* Ac, double* bs,
                                             int N, int lane) {
    for (int s = lane; s < m.n_slots; s += 32) {
        double acc = 0.0;
        for (int e = m.slot_ptr[s]; e < m.slot_ptr[s + 1]; ++e) {
            const uint32_t w = m.slot_entries[e];
            const double v = vt[w >> 8];
            acc = (w & 64u) ? d_sub(acc, v) : d_add(acc, v);
        }
        const int row = m.slot_row[s], col = m.slot_col[s];
        if (col < 0) {
            bs[row] = acc;
        } else {
            Ar[(size_t)row * N + col] = acc;
            Ac[(size_t)col * N + row] = acc;
        }
    }
    __syncwarp();
}

// back to all-zero: the structural slots plus the logged fill-in (or everything when the log overflowed)
__device__ __forceinline__ void big_clear(const DevMode& m, double* Ar, double* Ac, const int32_t* fill, int n_fill,
                                          bool overflow, int N, int lane) {
    if (overflow) {
        for (size_t i = lane; i < (size_t)N * N; i += 32) {
            Ar[i] = 0.0;
            Ac[i] = 0.0;
        }
    } else {
        for (int s = lane; s < m.nnz; s += 32) {
            const int row = m.slot_row[s], col = m.slot_col[s];
            Ar[(size_t)row * N + col] = 0.0;
            Ac[(size_t)col * N + row] = 0.0;
        }
        for (int i = lane; i < n_fill; i += 32) {
            const int idx = fill[i], row = idx / N, col = idx - row * N;
            Ar[idx] = 0.0;
            Ac[(size_t)col * N + row] = 0.0;
        }
    }
    __syncwarp();
}

// Solver::solveLinearSystemLU (include/solver.hpp:83-131).  In: Ar/Ac assembled, w.bs = b (by row).
// Out: w.xr = solution (zero vector when singular); hash updated with the pivot permutation.
// Returns true when singular.  n_fill / overflow: log of entries that became non-zero.
__device__ bool big_lu_solve(const BigWarp& w, double* Ar, double* Ac, int32_t* fill, int fill_cap, int& n_fill,
                             bool& overflow, int N, double eps, int lane, uint64_t& h) {
    for (int i = lane; i < N; i += 32) {
        w.rowat[i] = i;
        w.pos[i] = i;
    }
    __syncwarp();
    bool singular = false;
    for (int k = 0; k < N; ++k) {
        // ---- pivot search down column k: first maximum in position order (solver.hpp:46-56)
        const double* colk = Ac + (size_t)k * N;
        double bestv = -1.0;  // larger |v| wins, then the lower position
        int bestpos = 0x7FFFFFFF;
        // kBigBatch independent loads per lane are in flight before the first compare: one memory latency per
        // 32 * kBigBatch rows instead of one per 32 rows
        for (int base = 0; base < N; base += 32 * kBigBatch) {
            double v[kBigBatch];
            int pp[kBigBatch];
#pragma unroll
            for (int c = 0; c < kBigBatch; ++c) {
                const int row = base + c * 32 + lane;
                pp[c] = row < N ? w.pos[row] : -1;
                v[c] = pp[c] >= k ? colk[row] : 0.0;
            }
#pragma unroll
            for (int c = 0; c < kBigBatch; ++c) {
                if (pp[c] < k) continue;
                const double a = fabs(v[c]);
                // NaN never wins a '>' against the running maximum; a NaN already AT position k is never beaten
                const double key = (a != a) ? (pp[c] == k ? CUDART_INF : -1.0) : a;
                if (key > bestv || (key == bestv && pp[c] < bestpos)) {
                    bestv = key;
                    bestpos = pp[c];
                }
            }
        }
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            const double ov = __shfl_xor_sync(kFull, bestv, s);
            const int op = __shfl_xor_sync(kFull, bestpos, s);
            if (ov > bestv || (ov == bestv && op < bestpos)) {
                bestv = ov;
                bestpos = op;
            }
        }
        // the reference compares with strict '>' starting from position k: position k keeps the pivot on ties,
        // which "lower position wins" reproduces; a NaN at k is kept (key = +inf), other NaNs never win
        const int prow_old = w.rowat[k];
        const double akk = fabs(colk[prow_old]);
        const double maxAbs = (akk != akk) ? akk : bestv;
        if (maxAbs < eps) {  // solver.hpp:58-61 (false for NaN)
            singular = true;
            break;
        }
        const int p = bestpos;
        const int prow = w.rowat[p];
        __syncwarp();
        if (lane == 0 && p != k) {  // logical row exchange (solver.hpp:64-67)
            w.rowat[p] = prow_old;
            w.rowat[k] = prow;
            w.pos[prow_old] = p;
            w.pos[prow] = k;
        }
        __syncwarp();
        const double pv = colk[prow];
        const double* rowk = Ar + (size_t)prow * N;
        const double bk = w.bs[prow];
        // ---- rows below the pivot with a non-zero in column k: (row, multiplier) list, kBigFList at a time;
        //      then the pivot row, only columns whose entry is not exactly zero
        int base = 0, c_next = 0;
        double v[kBigBatch];
        int pp[kBigBatch];
        bool loaded = false;
        while (base < N) {
            int n_f = 0;
            while (base < N && n_f <= kBigFList - 32) {
                if (!loaded) {
#pragma unroll
                    for (int c = 0; c < kBigBatch; ++c) {
                        const int row = base + c * 32 + lane;
                        pp[c] = row < N ? w.pos[row] : -1;
                        v[c] = pp[c] > k ? colk[row] : 0.0;
                    }
                    loaded = true;
                    c_next = 0;
                }
#pragma unroll
                for (int c = 0; c < kBigBatch; ++c) {
                    if (c < c_next || n_f > kBigFList - 32) continue;  // (uniform across the warp)
                    const bool take = pp[c] > k && v[c] != 0.0;  // f = 0 leaves the row unchanged (exact for finite data)
                    const unsigned mk = __ballot_sync(kFull, take);
                    if (take) {
                        const int at = n_f + __popc(mk & ((1u << lane) - 1u));
                        w.frow[at] = base + c * 32 + lane;
                        w.fval[at] = v[c] / pv;  // solver.hpp:71
                    }
                    n_f += __popc(mk);
                    c_next = c + 1;
                }
                if (c_next >= kBigBatch) {
                    base += 32 * kBigBatch;
                    loaded = false;
                }
            }
            __syncwarp();
            if (n_f == 0) continue;
            for (int j0 = k + 1; j0 < N; j0 += 32 * kBigBatch) {
                // the pivot row's non-zeros in this batch of columns, compacted: (column, value) list
                double u[kBigBatch];
#pragma unroll
                for (int c = 0; c < kBigBatch; ++c) {
                    const int j = j0 + c * 32 + lane;
                    u[c] = j < N ? rowk[j] : 0.0;
                }
                int n_u = 0;
#pragma unroll
                for (int c = 0; c < kBigBatch; ++c) {
                    const bool act = u[c] != 0.0;
                    const unsigned mk = __ballot_sync(kFull, act);
                    if (act) {
                        const int at = n_u + __popc(mk & ((1u << lane) - 1u));
                        w.ucol[at] = j0 + c * 32 + lane;
                        w.uval[at] = u[c];
                    }
                    n_u += __popc(mk);
                }
                __syncwarp();
                if (n_u == 0) continue;
                // every (row, column) pair is a different matrix entry: a lane takes a pair, kBigPairs pairs per lane are
                // loaded before the first is used (solver.hpp:74 per entry; the order between entries is free)
                const int total = n_f * n_u;
                for (int p0 = 0; p0 < total; p0 += 32 * kBigPairs) {
                    double old[kBigPairs], prod[kBigPairs];
                    int row[kBigPairs], col[kBigPairs];
#pragma unroll
                    for (int q = 0; q < kBigPairs; ++q) {
                        const int p = p0 + q * 32 + lane;
                        row[q] = -1;
                        old[q] = 1.0;
                        if (p < total) {
                            const int t = p / n_u, c = p - t * n_u;
                            row[q] = w.frow[t];
                            col[q] = w.ucol[c];
                            prod[q] = d_mul(w.fval[t], w.uval[c]);
                            old[q] = Ar[(size_t)row[q] * N + col[q]];
                        }
                    }
#pragma unroll
                    for (int q = 0; q < kBigPairs; ++q) {
                        if (p0 + q * 32 >= total) continue;  // (uniform)
                        if (row[q] >= 0) {
                            const double nv = d_sub(old[q], prod[q]);
                            Ar[(size_t)row[q] * N + col[q]] = nv;
                            Ac[(size_t)col[q] * N + row[q]] = nv;
                        }
                        // an entry that was exactly zero may be new fill-in: log it for the clean-up
                        const bool isnew = old[q] == 0.0;
                        const unsigned mk = __ballot_sync(kFull, isnew);
                        if (mk) {
                            const int at = n_fill + __popc(mk & ((1u << lane) - 1u));
                            if (isnew && at < fill_cap) fill[at] = row[q] * N + col[q];
                            n_fill += __popc(mk);
                            if (n_fill > fill_cap) overflow = true;
                        }
                    }
                }
                __syncwarp();
            }
            // forward substitution fused (solver.hpp:106-113): b_row -= f * b_pivot
            for (int t = lane; t < n_f; t += 32) w.bs[w.frow[t]] = d_sub(w.bs[w.frow[t]], d_mul(w.fval[t], bk));
            __syncwarp();
        }
    }
    // pivot-sequence hash (definition: include/csim_b200.h)
    {
        uint32_t lo = 0, hi = 0;
        for (int i = lane; i < N; i += 32) {
            const uint32_t key = (uint32_t)i * 1024u + (uint32_t)w.rowat[i] + 1u;
            uint32_t l = key * 0x9E3779B1u;
            l ^= l >> 15;
            uint32_t g = key * 0x85EBCA77u;
            g ^= g >> 13;
            lo += l;
            hi += g;
        }
        lo = __reduce_add_sync(kFull, lo);
        hi = __reduce_add_sync(kFull, hi);
        h = (h ^ (((uint64_t)hi << 32) | lo)) * 0x100000001B3ull;
    }
    if (singular) {  // solver.hpp:94-97
        for (int i = lane; i < N; i += 32) w.xr[i] = 0.0;
        __syncwarp();
        return true;
    }
    // ---- back substitution (solver.hpp:116-128): positions bottom-up, columns ascending, zeros skipped
    for (int i = N - 1; i >= 0; --i) {
        const int prow = w.rowat[i];
        const double* rowi = Ar + (size_t)prow * N;
        double sum = w.bs[prow];
        for (int j0 = i + 1; j0 < N; j0 += 32 * kBigBatch) {
            double u[kBigBatch];
#pragma unroll
            for (int c = 0; c < kBigBatch; ++c) {
                const int j = j0 + c * 32 + lane;
                u[c] = j < N ? rowi[j] : 0.0;
            }
#pragma unroll
            for (int c = 0; c < kBigBatch; ++c) {
                unsigned mk = __ballot_sync(kFull, u[c] != 0.0);
                while (mk) {
                    const int l = __ffs(mk) - 1;
                    mk &= mk - 1;
                    const double ul = __shfl_sync(kFull, u[c], l);
                    sum = d_sub(sum, d_mul(ul, w.xr[j0 + c * 32 + l]));
                }
            }
        }
        const double diag = rowi[i];
        const double xi = (fabs(diag) < 1e-15) ? 0.0 : sum / diag;
        __syncwarp();
        if (lane == 0) w.xr[i] = xi;
        __syncwarp();
    }
    return false;
}


// one Newton linear solve: assemble (value table -> matrix, reference order), factor + solve, clear
__device__ __forceinline__ bool big_newton_solve(const DevMode& m, const BigWarp& w, const BigMem& M, int64_t b, int N,
                                                 double eps, int lane, uint64_t& h) {
    double* Ar = M.Ar + (size_t)b * N * N;
    double* Ac = M.Ac + (size_t)b * N * N;
    int32_t* fill = M.fill + (size_t)b * M.fill_cap;
    const double* vt = M.vt + (size_t)b * m.n_values;
    big_assemble(m, vt, Ar, Ac, w.bs, N, lane);
    int n_fill = 0;
    bool overflow = false;
    const bool sing = big_lu_solve(w, Ar, Ac, fill, M.fill_cap, n_fill, overflow, N, eps, lane, h);
    __syncwarp();
    big_clear(m, Ar, Ac, fill, n_fill < M.fill_cap ? n_fill : M.fill_cap, overflow, N, lane);
    return sing;
}

// || xn - x ||_2 in the order of oracle/csim_norm2.h while committing xn = x + alpha (xr - x)
__device__ __forceinline__ double big_damp(const BigWarp& w, int N, double alpha, int lane) {
    double acc = 0.0;
    for (int i = lane; i < N; i += 32) {
        const double x = w.xs[i];
        const double xn = d_add(x, d_mul(alpha, d_sub(w.xr[i], x)));
        const double d = d_sub(xn, x);
        acc = d_add(acc, d_mul(d, d));
        w.xs[i] = xn;
    }
    __syncwarp();
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) acc = d_add(acc, __shfl_xor_sync(kFull, acc, s));
    return sqrt(acc);
}

__device__ __forceinline__ bool big_all_finite(const BigWarp& w, int N, int lane) {
    bool ok = true;
    for (int i = lane; i < N; i += 32) ok = ok && lane_finite(w.xr[i]);
    return __all_sync(kFull, ok);
}

__device__ __forceinline__ double big_base_gmin(const csim_constants& c, double s) {  // include/dcanalysis.hpp:45-48
    if (s < 0.0) s = 0.0;
    if (s > 1.0) s = 1.0;
    return d_add(d_mul(c.gmin_high_base, d_sub(1.0, s)), d_mul(c.gmin_low_base, s));
}

__global__ void __launch_bounds__(kBigWarps * 32)
tran_big_kernel(DevPlan P, DevTran T, int64_t B, const double* __restrict__ params, double* __restrict__ xio,
                double* __restrict__ wave, int32_t* __restrict__ iters_total, uint8_t* __restrict__ iters_step,
                uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status, BigMem M) {
    extern __shared__ unsigned char big_smem[];
    const DevMode& m = P.m[CSIM_MODE_TRAN];
    const csim_constants& c = P.c;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int N = P.N;
    const int64_t b = (int64_t)blockIdx.x * kBigWarps + wib;
    if (b >= B || (T.skip && T.skip[b])) return;
    const BigWarp w = big_carve(big_smem + big_warp_bytes(N) * wib, N);
    double* vt = M.vt + (size_t)b * m.n_values;
    uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    int total = 0;
    for (int i = lane; i < N; i += 32) w.xs[i] = xio[(int64_t)i * B + b];
    eval_static(m, params, B, b, T.dt, c.tran_gmin, vt, lane);
    __syncwarp();
    const int r0 = T.out_begin_step / T.out_stride;
    for (int step = T.step_begin; step < T.step_end; ++step) {
        eval_step(m, w.xs, vt, lane);  // xs still holds the last accepted timepoint here
        eval_src_tran(m, T, params, B, b, step, vt, lane);
        int count = 0;
        bool dead = false;
        for (int iter = 0; iter < c.tran_max_iters; ++iter) {  // src/tanalisis.cpp:258
            ++count;
            eval_mos(m, params, B, b, c.mos_off_gds, w.xs, vt, lane);
            __syncwarp();
            if (big_newton_solve(m, w, M, b, N, c.lu_eps, lane, h)) st |= CSIM_ST_LU_SINGULAR;
            if (!big_all_finite(w, N, lane)) {  // src/tanalisis.cpp:360-362 (the reference throws)
                st |= CSIM_ST_NONFINITE;
                dead = true;
                break;
            }
            const double err = big_damp(w, N, c.tran_alpha, lane);  // src/tanalisis.cpp:365-366
            if (err < c.tran_tol) break;
            if (iter == c.tran_max_iters - 1) st |= CSIM_ST_TRAN_UNCONVERGED;
        }
        total += count;
        if (iters_step && lane == 0) iters_step[(int64_t)step * B + b] = (uint8_t)count;
        if (dead) break;
        if ((step + 1) % T.out_stride == 0 && step >= T.out_begin_step) {
            const int64_t r = (step + 1) / T.out_stride - 1 - r0;
            for (int p = lane; p < T.n_probes; p += 32)
                wave[((int64_t)b * T.n_probes + p) * T.wave_ld + r] = w.xs[T.probes[p]];
        }
    }
    for (int i = lane; i < N; i += 32) xio[(int64_t)i * B + b] = w.xs[i];
    if (lane == 0) {
        if (iters_total) iters_total[b] += total;
        if (pivot_hash) pivot_hash[b] = h;
        if (status) status[b] |= st;
    }
}

__global__ void __launch_bounds__(kBigWarps * 32)
dc_big_kernel(DevPlan P, int64_t B, const double* __restrict__ params, double* __restrict__ xout,
              int32_t* __restrict__ iters, uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status, BigMem M) {
    extern __shared__ unsigned char big_smem[];
    const DevMode& m = P.m[CSIM_MODE_DC];
    const csim_constants& c = P.c;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int N = P.N;
    const int64_t b = (int64_t)blockIdx.x * kBigWarps + wib;
    if (b >= B) return;
    const BigWarp w = big_carve(big_smem + big_warp_bytes(N) * wib, N);
    double* vt = M.vt + (size_t)b * m.n_values;
    uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    for (int i = lane; i < N; i += 32) w.xs[i] = 0.0;
    eval_static(m, params, B, b, 0.0, 0.0, vt, lane);
    __syncwarp();
    int gmin_slot = -1;
    for (int i = 0; i < m.n_static; ++i)
        if (m.static_ops[i].kind == SK_GMIN) gmin_slot = m.static_ops[i].out;
    if (!P.has_mos) {  // dcSolveDirectLU (src/dcanalysis.cpp:46-68)
        eval_src_dc(m, params, B, b, 1.0, vt, lane);
        __syncwarp();
        if (big_newton_solve(m, w, M, b, N, c.lu_eps, lane, h)) st |= CSIM_ST_LU_SINGULAR;
        for (int i = lane; i < N; i += 32) w.xs[i] = w.xr[i];
        if (iters && lane < c.dc_ramp_steps) iters[(int64_t)lane * B + b] = lane == 0 ? 1 : 0;
    } else {  // dcSolveNewtonLU (src/dcanalysis.cpp:95-163)
        for (int step = 1; step <= c.dc_ramp_steps; ++step) {
            const double scale = (double)step / (double)c.dc_ramp_steps;
            double gmin = big_base_gmin(c, scale);
            double prevErr = CUDART_INF;
            eval_src_dc(m, params, B, b, scale, vt, lane);
            int count = 0;
            for (int iter = 0; iter < c.dc_max_iters; ++iter) {
                ++count;
                if (lane == 0 && gmin_slot >= 0) vt[gmin_slot] = gmin;
                eval_mos(m, params, B, b, c.mos_off_gds, w.xs, vt, lane);
                __syncwarp();
                if (big_newton_solve(m, w, M, b, N, c.lu_eps, lane, h)) st |= CSIM_ST_LU_SINGULAR;
                if (!big_all_finite(w, N, lane)) {  // dcanalysis.cpp:135-138
                    st |= CSIM_ST_NONFINITE;
                    const double g10 = d_mul(gmin, c.dc_nonfinite_mul);
                    gmin = g10 < c.dc_nonfinite_cap ? g10 : c.dc_nonfinite_cap;
                    continue;
                }
                const double err = big_damp(w, N, c.dc_alpha, lane);  // ConvController::update (dcanalysis.cpp:268-307)
                const double gbase = big_base_gmin(c, scale);
                double gnext = gbase;
                if (!(iter == 0 || !lane_finite(prevErr))) {
                    if (err > d_mul(prevErr, c.slow_conv_ratio)) {
                        const double g2 = d_mul(gmin, 2.0);
                        gnext = g2 < c.gmin_abs_max ? g2 : c.gmin_abs_max;
                    } else if (err < d_mul(prevErr, c.fast_conv_ratio)) {
                        gnext = d_add(d_mul(0.5, gmin), d_mul(0.5, gbase));
                    } else {
                        gnext = d_add(d_mul(0.7, gmin), d_mul(0.3, gbase));
                    }
                }
                gmin = gnext;
                prevErr = err;
                if (err < c.dc_tol) break;
                if (iter == c.dc_max_iters - 1) st |= CSIM_ST_DC_UNCONVERGED;
            }
            if (iters && lane == 0) iters[(int64_t)(step - 1) * B + b] = count;
        }
    }
    __syncwarp();
    for (int i = lane; i < N; i += 32) xout[(int64_t)i * B + b] = w.xs[i];
    if (lane == 0) {
        if (pivot_hash) pivot_hash[b] = h;
        if (status) status[b] |= st;
    }
}

// batched Solver::solveLinearSystemLU for n > 32 (parity probe): A [N*N][B], b [N][B] instance-minor
__global__ void __launch_bounds__(kBigWarps * 32)
lu_big_kernel(int N, int64_t B, const double* __restrict__ A, const double* __restrict__ bvec, double* __restrict__ xout,
              int32_t* __restrict__ perm_out, int32_t* __restrict__ singular, double eps, BigMem M) {
    extern __shared__ unsigned char big_smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t b = (int64_t)blockIdx.x * kBigWarps + wib;
    if (b >= B) return;
    const BigWarp w = big_carve(big_smem + big_warp_bytes(N) * wib, N);
    double* Ar = M.Ar + (size_t)b * N * N;
    double* Ac = M.Ac + (size_t)b * N * N;
    for (int idx = lane; idx < N * N; idx += 32) {
        const int i = idx / N, j = idx - i * N;
        const double v = A[(int64_t)idx * B + b];
        Ar[idx] = v;
        Ac[(size_t)j * N + i] = v;
    }
    for (int i = lane; i < N; i += 32) w.bs[i] = bvec[(int64_t)i * B + b];
    __syncwarp();
    int n_fill = 0;
    bool overflow = false;
    uint64_t h = 0;
    const bool sing = big_lu_solve(w, Ar, Ac, M.fill + (size_t)b * M.fill_cap, M.fill_cap, n_fill, overflow, N, eps, lane, h);
    __syncwarp();
    for (int i = lane; i < N; i += 32) {
        xout[(int64_t)i * B + b] = w.xr[i];
        if (perm_out) perm_out[(int64_t)i * B + b] = w.rowat[i];
    }
    if (lane == 0 && singular) singular[b] = sing ? 1 : 0;
}

// ---- launch wrappers (scratch is stream-ordered) ---------------------------------------------------
static cudaError_t big_alloc(int N, int n_values, int64_t B, bool zero, cudaStream_t stream, BigMem& M) {
    M = BigMem{};
    M.fill_cap = 16 * N;
    cudaError_t e;
    const size_t nn = (size_t)N * N * B * sizeof(double);
    if ((e = cudaMallocAsync((void**)&M.Ar, nn, stream)) != cudaSuccess) return e;
    if ((e = cudaMallocAsync((void**)&M.Ac, nn, stream)) != cudaSuccess) return e;
    if ((e = cudaMallocAsync((void**)&M.vt, (size_t)(n_values > 0 ? n_values : 1) * B * sizeof(double), stream)) != cudaSuccess) return e;
    if ((e = cudaMallocAsync((void**)&M.fill, (size_t)M.fill_cap * B * sizeof(int32_t), stream)) != cudaSuccess) return e;
    if (zero) {
        if ((e = cudaMemsetAsync(M.Ar, 0, nn, stream)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(M.Ac, 0, nn, stream)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}
static void big_free(BigMem& M, cudaStream_t stream) {
    if (M.Ar) cudaFreeAsync(M.Ar, stream);
    if (M.Ac) cudaFreeAsync(M.Ac, stream);
    if (M.vt) cudaFreeAsync(M.vt, stream);
    if (M.fill) cudaFreeAsync(M.fill, stream);
}
static size_t big_smem_bytes(int N) { return big_warp_bytes(N) * kBigWarps; }

cudaError_t launch_tran_big(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x, double* wave,
                            int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                            cudaStream_t stream) {
    BigMem M;
    cudaError_t e = big_alloc(P.N, P.m[CSIM_MODE_TRAN].n_values, B, true, stream, M);
    if (e == cudaSuccess) {
        const size_t smem = big_smem_bytes(P.N);
        e = cudaFuncSetAttribute(tran_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) {
            tran_big_kernel<<<(unsigned)((B + kBigWarps - 1) / kBigWarps), kBigWarps * 32, smem, stream>>>(
                P, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, M);
            e = cudaGetLastError();
        }
    }
    big_free(M, stream);
    return e;
}

cudaError_t launch_dc_big(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                          uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream) {
    BigMem M;
    cudaError_t e = big_alloc(P.N, P.m[CSIM_MODE_DC].n_values, B, true, stream, M);
    if (e == cudaSuccess) {
        const size_t smem = big_smem_bytes(P.N);
        e = cudaFuncSetAttribute(dc_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) {
            dc_big_kernel<<<(unsigned)((B + kBigWarps - 1) / kBigWarps), kBigWarps * 32, smem, stream>>>(
                P, B, params, x, iters, pivot_hash, status, M);
            e = cudaGetLastError();
        }
    }
    big_free(M, stream);
    return e;
}

cudaError_t launch_lu_big(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                          int32_t* singular, double eps, cudaStream_t stream) {
    BigMem M;
    cudaError_t e = big_alloc(N, 0, B, false, stream, M);
    if (e == cudaSuccess) {
        const size_t smem = big_smem_bytes(N);
        e = cudaFuncSetAttribute(lu_big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) {
            lu_big_kernel<<<(unsigned)((B + kBigWarps - 1) / kBigWarps), kBigWarps * 32, smem, stream>>>(
                N, B, A, b, x, perm, singular, eps, M);
            e = cudaGetLastError();
        }
    }
    big_free(M, stream);
    return e;
}

size_t big_bytes_per_instance(int N, int n_values) {
    return 2 * (size_t)N * N * sizeof(double) + (size_t)n_values * sizeof(double) + (size_t)16 * N * sizeof(int32_t);
}

}  // namespace csim
