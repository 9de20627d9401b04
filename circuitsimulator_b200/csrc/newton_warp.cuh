// newton_warp.cuh — generic warp-per-instance Newton/MNA machinery (N <= 32).
//
// One warp owns one circuit instance.  Its dense N x N matrix lives in shared
// memory (row stride 33 doubles -> conflict-free row AND column access), lane i
// owns row i during assembly and elimination, lane j owns column j during row
// swaps.  All reference arithmetic goes through __dadd_rn/__dsub_rn/__dmul_rn and
// IEEE division, so nothing is ever contracted into an FMA, and
// every sum is formed in the reference's order (F5).  The LU performs the
// reference's value-dependent partial pivoting on every factorisation (F2) and
// skips exact zeros of the pivot row, which is bit-identical to the dense loops
// for finite data (F3).
#pragma once
#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#endif

#include "device_plan.cuh"

namespace csim {

constexpr int kLD = 33;  // row stride of the per-warp matrix, in doubles
constexpr unsigned kFull = 0xFFFFFFFFu;

__device__ __forceinline__ double d_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double d_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double d_mul(double a, double b) { return __dmul_rn(a, b); }

struct WarpMem {
    double* S;   // [32*kLD] matrix
    double* xs;  // [32] current Newton iterate
    double* xp;  // [32] last accepted timepoint
    double* bs;  // [32] right-hand side -> y -> raw solution
    double* vt;  // [n_values] value table
};

__host__ __device__ inline size_t warp_mem_doubles(int n_values) { return 32 * kLD + 3 * 32 + ((n_values + 3) & ~3); }

__device__ __forceinline__ WarpMem carve(double* base, int n_values) {
    WarpMem w;
    w.S = base;
    w.xs = base + 32 * kLD;
    w.xp = w.xs + 32;
    w.bs = w.xp + 32;
    w.vt = w.bs + 32;
    (void)n_values;
    return w;
}

// ---- value ops ---------------------------------------------------------------
__device__ __forceinline__ void eval_static(const DevMode& m, const double* __restrict__ params, int64_t B,
                                            int64_t b, double dt, double gmin, double* vt, int lane,
                                            int stride = 32) {
    for (int i = lane; i < m.n_static; i += stride) {
        StaticOp op = m.static_ops[i];
        double p = op.param >= 0 ? params[(int64_t)op.param * B + b] : 0.0;
        double v;
        switch (op.kind) {
            case SK_ONE: v = 1.0; break;
            case SK_GMIN: v = gmin; break;
            case SK_RES: v = (p == 0.0) ? 0.0 : 1.0 / p; break;                           // element.cpp:20-24
            case SK_CAP: v = (p <= 0.0 || dt <= 0.0) ? 0.0 : p / dt; break;               // tanalisis.cpp:65-67
            case SK_CAPHALF: {
                double c = d_mul(0.5, p);                                                  // tanalisis.cpp:338
                v = (c <= 0.0 || dt <= 0.0) ? 0.0 : c / dt;
            } break;
            case SK_IND_ONE: v = (p <= 0.0) ? 0.0 : 1.0; break;                           // tanalisis.cpp:296
            case SK_IND_NREQ: v = (p <= 0.0) ? 0.0 : -(p / dt); break;                    // tanalisis.cpp:304,317
            default: v = 0.0;
        }
        vt[op.out] = v;
    }
}

__device__ __forceinline__ void eval_step(const DevMode& m, const double* xp, double* vt, int lane,
                                          int stride = 32) {
    for (int i = lane; i < m.n_step; i += stride) {
        StepOp op = m.step_ops[i];
        double va = op.a >= 0 ? xp[op.a] : 0.0;
        double v;
        if (op.kind == TK_CAP_HIST) {
            double vb = op.b >= 0 ? xp[op.b] : 0.0;
            v = d_mul(-vt[op.g], d_sub(va, vb));  // I_hist = -Gc * vPrev
        } else {
            v = d_mul(vt[op.g], va);             // V_hist = -R_eq * iPrev  (vt[g] = -R_eq)
        }
        vt[op.out] = v;
    }
}

// SourceSpec::evalDC (include/sim.hpp:152-158)
__device__ __forceinline__ void eval_src_dc(const DevMode& m, const double* __restrict__ params, int64_t B, int64_t b,
                                            double scale, double* vt, int lane, int stride = 32) {
    for (int i = lane; i < m.n_src; i += stride) {
        SrcOp op = m.src_ops[i];
        double base = params[(int64_t)op.param * B + b];
        if (op.is_sin) base = d_add(base, params[(int64_t)(op.param + 1) * B + b]);
        vt[op.out] = d_mul(base, scale);
    }
}

// SourceSpec::evalTran (include/sim.hpp:160-162) through the host-tabulated waveform
__device__ __forceinline__ void eval_src_tran(const DevMode& m, const DevTran& T, const double* __restrict__ params,
                                              int64_t B, int64_t b, int step, double* vt, int lane,
                                              int stride = 32) {
    const double tnow = d_mul((double)(step + 1), T.dt);  // tanalisis.cpp:256
    for (int i = lane; i < m.n_src; i += stride) {
        SrcOp op = m.src_ops[i];
        const int s = op.index;
        const int mode = T.src_mode[s];
        double w = 0.0;
        if (mode != CSIM_SRC_CONST) {
            int grp = T.src_group_b ? T.src_group_b[(int64_t)s * B + b] : T.src_group[s];
            double tab = T.src_table[(int64_t)grp * T.table_steps + step];
            if (mode == CSIM_SRC_SIN_TABLE) {
                double v0 = params[(int64_t)(op.param + 1) * B + b];
                double td = params[(int64_t)(op.param + 4) * B + b];
                if (tnow < td) w = v0;
                else w = d_add(v0, d_mul(params[(int64_t)(op.param + 2) * B + b], tab));
            } else {
                w = tab;
            }
        }
        vt[op.out] = d_add(params[(int64_t)op.param * B + b], w);
    }
}

// MosfetBase::stamp linearisation (src/element.cpp:181-274)
__device__ __forceinline__ void mos_eval(double Vth, double K, double lambda, bool is_p, double off_gds, double Vd,
                                         double Vg, double Vs, double& gd, double& gg, double& gs, double& cst) {
    const double pp = is_p ? -1.0 : 1.0;
    const double Vgs = d_mul(pp, d_sub(Vg, Vs));
    const double Vds = d_mul(pp, d_sub(Vd, Vs));
    double Ids0, gds0, gm0;
    if (Vgs > Vth && Vds >= 0.0) {
        const double Vov = d_sub(Vgs, Vth);
        if (Vds < Vov) {  // triode
            Ids0 = d_mul(K, d_sub(d_mul(Vov, Vds), d_mul(d_mul(0.5, Vds), Vds)));
            gds0 = d_mul(K, d_sub(Vov, Vds));
            gm0 = d_mul(K, Vds);
        } else {  // saturation
            Ids0 = d_mul(d_mul(d_mul(0.5, K), Vov), Vov);
            gds0 = 0.0;
            gm0 = d_mul(K, Vov);
        }
    } else {
        Ids0 = 0.0;
        gm0 = 0.0;
        gds0 = off_gds;
    }
    double factor = d_add(1.0, d_mul(lambda, Vds));
    if (factor < 0.0) factor = 0.0;
    const double Ids_eff = d_mul(Ids0, factor);
    const double dVds = d_add(d_mul(gds0, factor), d_mul(Ids0, lambda));
    const double dVgs = d_mul(gm0, factor);
    const double Ids = d_mul(pp, Ids_eff);
    gd = dVds;
    gg = dVgs;
    gs = -d_add(dVds, dVgs);
    cst = d_sub(d_sub(d_sub(Ids, d_mul(gd, Vd)), d_mul(gg, Vg)), d_mul(gs, Vs));
}

__device__ __forceinline__ void eval_mos(const DevMode& m, const double* __restrict__ params, int64_t B, int64_t b,
                                         double off_gds, const double* xs, double* vt, int lane,
                                         int stride = 32) {
    for (int i = lane; i < m.n_mos; i += stride) {
        MosOp op = m.mos_ops[i];
        double Vth = params[(int64_t)op.param * B + b];
        double K = params[(int64_t)(op.param + 1) * B + b];
        double lambda = params[(int64_t)(op.param + 2) * B + b];
        double Vd = op.d >= 0 ? xs[op.d] : 0.0;
        double Vg = op.g >= 0 ? xs[op.g] : 0.0;
        double Vs = op.s >= 0 ? xs[op.s] : 0.0;
        double gd, gg, gs, cst;
        mos_eval(Vth, K, lambda, op.is_p != 0, off_gds, Vd, Vg, Vs, gd, gg, gs, cst);
        vt[op.out] = gd;
        vt[op.out + 1] = gg;
        vt[op.out + 2] = gs;
        vt[op.out + 3] = cst;
    }
}

// ---- assembly: replay the row program into S / bs -----------------------------
__device__ __forceinline__ void assemble(const DevMode& m, const uint32_t* prog, const WarpMem& w, int N, int lane) {
    for (int idx = lane; idx < N * kLD; idx += 32) w.S[idx] = 0.0;
    __syncwarp();
    double rb = 0.0;
    double* row = w.S + lane * kLD;
    for (int c = 0; c < m.row_len; ++c) {
        uint32_t word = prog[c * 32 + lane];
        if (word != 0xFFFFFFFFu) {
            double v = w.vt[word >> 8];
            if (word & 64u) v = -v;
            int col = word & 63u;
            if (col == kRhsCol) rb = d_add(rb, v);
            else row[col] = d_add(row[col], v);
        }
    }
    w.bs[lane] = rb;
    __syncwarp();
}

// ---- Solver::solveLinearSystemLU (include/solver.hpp:30-131) -------------------
// In: S (N x N), bs (rhs).  Out: bs = solution (zero vector when singular),
// perm (lane i: original row now at position i).  Returns true when singular.
__device__ __forceinline__ bool lu_solve_warp(double* S, double* bs, int N, int lane, double eps, int& perm) {
    perm = lane;
    bool singular = false;
    for (int k = 0; k < N; ++k) {
        // pivot search: first maximum of |S[i][k]|, i >= k (solver.hpp:46-56)
        const bool cand = (lane >= k) && (lane < N);
        double a = cand ? S[lane * kLD + k] : 0.0;
        unsigned long long key = (unsigned long long)__double_as_longlong(fabs(a));
        if (a != a) key = (lane == k) ? ~0ull : 0ull;  // NaN never wins a '>' ; a NaN at k is never beaten
        if (!cand) key = 0ull;
        const unsigned hi = (unsigned)(key >> 32);
        const unsigned mh = __reduce_max_sync(kFull, hi);
        const unsigned lo = (hi == mh) ? (unsigned)key : 0u;
        const unsigned ml = __reduce_max_sync(kFull, lo);
        const unsigned win = __ballot_sync(kFull, cand && hi == mh && (unsigned)key == ml);
        const int p = __ffs(win) - 1;
        const double maxAbs = __longlong_as_double((long long)(((unsigned long long)mh << 32) | ml));
        if (maxAbs < eps) {  // solver.hpp:58-61
            singular = true;
            break;
        }
        if (p != k) {  // row swap (solver.hpp:64-67): lane j swaps column j
            if (lane < N) {
                double t1 = S[k * kLD + lane], t2 = S[p * kLD + lane];
                S[k * kLD + lane] = t2;
                S[p * kLD + lane] = t1;
            }
            if (lane == 0) {
                double t1 = bs[k], t2 = bs[p];
                bs[k] = t2;
                bs[p] = t1;
            }
            int other = __shfl_sync(kFull, perm, lane == k ? p : k);
            if (lane == k || lane == p) perm = other;
            __syncwarp();
        }
        const double pv = S[k * kLD + k];
        const bool act = (lane > k) && (lane < N);
        double f = 0.0;
        if (act) {
            f = S[lane * kLD + k] / pv;  // solver.hpp:71
            S[lane * kLD + k] = f;
        }
        // columns of the pivot row that are not exactly zero
        const double urow = act ? S[k * kLD + lane] : 0.0;
        unsigned cm = __ballot_sync(kFull, act && urow != 0.0);
        const double bk = bs[k];
        if (act) {
            double* row = S + lane * kLD;
            unsigned mm = cm;
            while (mm) {
                int j = __ffs(mm) - 1;
                mm &= mm - 1;
                row[j] = d_sub(row[j], d_mul(f, S[k * kLD + j]));  // solver.hpp:74
            }
            bs[lane] = d_sub(bs[lane], d_mul(f, bk));  // forward substitution, fused (solver.hpp:106-113)
        }
        __syncwarp();
    }
    if (singular) {  // solver.hpp:94-97: zero vector
        __syncwarp();
        bs[lane] = 0.0;
        __syncwarp();
        return true;
    }
    // back substitution (solver.hpp:116-128): rows bottom-up, columns ascending
    for (int i = N - 1; i >= 0; --i) {
        const bool in = (lane > i) && (lane < N);
        const double u = in ? S[i * kLD + lane] : 0.0;
        unsigned cm = __ballot_sync(kFull, in && u != 0.0);
        double sum = bs[i];
        while (cm) {
            int j = __ffs(cm) - 1;
            cm &= cm - 1;
            sum = d_sub(sum, d_mul(S[i * kLD + j], bs[j]));
        }
        const double diag = S[i * kLD + i];
        const double xi = (fabs(diag) < 1e-15) ? 0.0 : sum / diag;
        __syncwarp();
        if (lane == 0) bs[i] = xi;
        __syncwarp();
    }
    return false;
}

// pivot-sequence hash (definition: include/csim_b200.h)
__device__ __forceinline__ uint64_t hash_update(uint64_t h, int perm, int N, int lane) {
    uint32_t lo = 0, hi = 0;
    if (lane < N) {
        uint32_t key = (uint32_t)lane * 1024u + (uint32_t)perm + 1u;
        lo = key * 0x9E3779B1u;
        lo ^= lo >> 15;
        hi = key * 0x85EBCA77u;
        hi ^= hi >> 13;
    }
    lo = __reduce_add_sync(kFull, lo);
    hi = __reduce_add_sync(kFull, hi);
    uint64_t f = ((uint64_t)hi << 32) | (uint64_t)lo;
    return (h ^ f) * 0x100000001B3ull;
}

// sqrt of the 32-leaf pairwise tree sum of squares (norm definition: include/csim_b200.h,
// oracle/csim_norm2.h); every lane returns the same value.
__device__ __forceinline__ double warp_norm2(double d) {
    double v = d_mul(d, d);
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) v = d_add(v, __shfl_xor_sync(kFull, v, s));
    return sqrt(v);
}

__device__ __forceinline__ bool lane_finite(double v) { return fabs(v) <= 1.7976931348623157e308; }

}  // namespace csim
