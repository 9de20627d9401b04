// dense_kernels.cu — the large-matrix path for systems whose elimination fills in (N > 32): one CTA per
// instance, blocked right-looking LU with partial pivoting on the augmented matrix [A | b] held in HBM/L2.
//
// Why a second large path: with the reference's natural unknown order (node "vdd" first) the first pivot
// row and column of the 200-stage ladder are dense, so the factorisation of its 404 x 404 transient matrix
// is FULLY dense (161,612 of 163,216 entries non-zero, 21.5 M multiply-subtract pairs per solve, measured
// with numpy on the oracle's matrices).  The warp-per-instance kernel of big_kernels.cu walks entries one
// dependent global read-modify-write at a time and is latency-bound there; this kernel stages a 16-column
// panel and a 16 x 128 strip of U in shared memory and updates the trailing matrix in 4 x 4 register tiles,
// so every matrix entry is read and written once per PANEL instead of once per elimination step.
//
// Bit-exactness (Solver::solveLinearSystemLU, include/solver.hpp:30-131 of the reference): an entry a(i,j)
// receives the updates  a -= f(i,k) * a(k,j)  for k = 0, 1, 2 ... in ascending k, each one a separately
// rounded multiply and subtract.  Blocking only regroups WHICH entries are updated when; per entry the
// sequence of k is unchanged (panel: k inside the panel ascending; strip: triangular solve ascending;
// trailing tile: the 16 k of the panel ascending), so the bits are the reference's.  The pivot search sees
// fully updated columns (right-looking) and keeps the reference's rule: first maximum from position k down,
// strict '>'.  The right-hand side is column N of the augmented matrix (the reference's forward substitution
// applies the same  b(i) -= f * b(k)  in the same order).  Back substitution runs in ascending column order
// per row, sequentially, as the reference does.  No zero skipping: 0 * inf gives NaN as in the reference.
//
// Row exchanges are logical (rowat[position] = physical row); multipliers live only in the shared-memory
// panel (L is never written back: nothing reads it again).
#include <cuda_runtime.h>
#include <math_constants.h>

#include "kernels.hpp"
#include "newton_warp.cuh"

namespace csim {

constexpr int kDnThreads = 256;
constexpr int kDnNB = 16;            // panel width
constexpr int kDnLDP = kDnNB + 1;    // panel row stride in shared memory (odd: column reads are conflict-free)
constexpr int kDnCW = 128;           // columns of U staged per pass (strip = kDnNB x kDnCW)
constexpr int kDnRM = 4, kDnCN = 4;  // register tile of the trailing update: rows x columns per thread
constexpr int kDnRB = 16;            // rows per block of the back substitution (kDnRB * (N + 1) <= N * kDnLDP)

#ifdef CSIM_DN_PROF  // tools/dn_prof.cu: cycles per phase of dn_lu_solve, summed over CTAs (thread 0's clock)
__device__ unsigned long long dn_prof[8];
#define DN_T(i)                                  \
    do {                                         \
        if (tid == 0) {                          \
            const long long t_ = clock64();      \
            prof_acc[i] += t_ - prof_t;          \
            prof_t = t_;                         \
        }                                        \
    } while (0)
#else
#define DN_T(i) \
    do {        \
    } while (0)
#endif

// ---- TMA (bulk asynchronous copy) staging of the strip of U ------------------------------------------------------------
// The strip rows (positions k0..k0+15 of one 128-column block) are 16 contiguous 1 KB segments of [A | b]: one elected
// thread issues 16 cp.async.bulk copies global -> shared that complete on an mbarrier (SASS: UBLKCP.S.G +
// SYNCS.ARRIVE.TRANS64), into the buffer the CTA is NOT computing from, so the next block's rows arrive while the
// trailing update of the current block runs.
__device__ __forceinline__ unsigned dn_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void dn_mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dn_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void dn_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dn_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dn_bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dn_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(dn_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void dn_mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n.reg .pred p;\nDN_WAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DN_DONE_%=;\nbra DN_WAIT_%=;\nDN_DONE_%=:\n}" ::"r"(
            dn_smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// orders the CTA's earlier generic-proxy accesses (st.global of the trailing update, reads of the buffer about to be
// overwritten) before the bulk copies the calling thread issues next
__device__ __forceinline__ void dn_fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct DnSmem {
    double* xs;      // [N] Newton iterate
    double* xr;      // [N] raw solution
    double* Pn;      // [N][kDnLDP] panel: positions k0.. (relative), columns k0..k0+NB-1; also scratch [N]
    double* Us;      // [2][kDnNB][kDnCW] strip of U, double-buffered (16-byte aligned: bulk-copy destination)
    uint64_t* bar;   // [2] mbarriers of the two strip buffers
    uint32_t* ph;    // [2] their current phase parity (uniform over the CTA; thread 0 flips it behind a barrier)
    double* redv;    // [8] per-warp pivot candidates / broadcast slot
    int32_t* redp;   // [8]
    int32_t* rowat;  // [N] position -> physical row
};

__host__ __device__ inline size_t dn_smem_bytes(int N) {
    return (size_t)N * (8 + 8 + 8 * kDnLDP + 4) + (size_t)2 * kDnNB * kDnCW * 8 + 16 + 16 + 8 * 8 + 8 * 4 + 64;
}
__host__ __device__ inline int dn_ld(int N) { return (N + 1 + 3) & ~3; }  // row stride of [A | b], 32-byte multiple

__device__ __forceinline__ DnSmem dn_carve(unsigned char* base, int N) {
    DnSmem s;
    double* d = reinterpret_cast<double*>(base);
    s.xs = d;
    s.xr = d + N;
    s.Pn = d + 2 * N;
    s.Us = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(s.Pn + (size_t)N * kDnLDP) + 15) & ~(uintptr_t)15);
    s.bar = reinterpret_cast<uint64_t*>(s.Us + 2 * kDnNB * kDnCW);
    s.ph = reinterpret_cast<uint32_t*>(s.bar + 2);
    s.redv = reinterpret_cast<double*>(s.bar + 3);
    s.redp = reinterpret_cast<int32_t*>(s.redv + 8);
    s.rowat = s.redp + 8;
    return s;
}

// once per kernel, after dn_carve: the strip buffers' mbarriers (one arrival each: the elected thread's expect_tx)
__device__ __forceinline__ void dn_init_sync(const DnSmem& S, int tid) {
    if (tid == 0) {
        dn_mbar_init(&S.bar[0], 1);
        dn_mbar_init(&S.bar[1], 1);
        S.ph[0] = S.ph[1] = 0;
        dn_fence_proxy_async();
    }
    __syncthreads();
}

// [A | b] := 0, then the ordered gather per slot (reference accumulation order)
__device__ __forceinline__ void dn_assemble(const DevMode& m, const double* vt, double* A, int N, int LD, int tid) {
    double2* A2 = reinterpret_cast<double2*>(A);
    for (int i = tid; i < N * LD / 2; i += kDnThreads) A2[i] = make_double2(0.0, 0.0);
    __syncthreads();
    for (int s = tid; s < m.n_slots; s += kDnThreads) {
        double acc = 0.0;
        for (int e = m.slot_ptr[s]; e < m.slot_ptr[s + 1]; ++e) {
            const uint32_t w = m.slot_entries[e];
            const double v = vt[w >> 8];
            acc = (w & 64u) ? d_sub(acc, v) : d_add(acc, v);
        }
        const int row = m.slot_row[s], col = m.slot_col[s];
        A[(size_t)row * LD + (col < 0 ? N : col)] = acc;
    }
    __syncthreads();
}

// Factor [A | b] in place and solve.  Out: S.xr (zero vector when singular), S.rowat = pivot permutation,
// h updated.  Returns true when singular (uniform over the CTA).
// keepL: also write the multipliers back into the strict lower triangle of A (Solver::luDecompose's packed L\U,
// solver.hpp:70-72); the Newton kernels never read L again and skip the traffic.
__device__ bool dn_lu_solve(double* __restrict__ A, int N, int LD, const DnSmem& S, double eps, int tid, uint64_t& h,
                            bool keepL = false) {
    const int lane = tid & 31, warp = tid >> 5;
#ifdef CSIM_DN_PROF
    long long prof_acc[6] = {0, 0, 0, 0, 0, 0}, prof_t = clock64();
#endif
    for (int i = tid; i < N; i += kDnThreads) S.rowat[i] = i;
    __syncthreads();
    bool singular = false;
    for (int k0 = 0; k0 < N; k0 += kDnNB) {
        const int nb = (N - k0 < kDnNB) ? N - k0 : kDnNB, mrows = N - k0;
        // ---- panel: rows at positions k0.., columns k0..k0+nb-1, gathered through rowat (128 B per row)
        for (int e = tid; e < mrows * kDnNB; e += kDnThreads) {
            const int r = e / kDnNB, c = e - r * kDnNB;
            if (c < nb) S.Pn[r * kDnLDP + c] = A[(size_t)S.rowat[k0 + r] * LD + k0 + c];
        }
        __syncthreads();
        DN_T(0);
        for (int kk = 0; kk < nb; ++kk) {
            // pivot search down column kk from the diagonal: first maximum in position order (solver.hpp:46-56)
            const double akk = fabs(S.Pn[kk * kDnLDP + kk]);
            double bestv = -1.0;
            int bestr = 0x7FFFFFFF;
            for (int r = kk + tid; r < mrows; r += kDnThreads) {
                const double a = fabs(S.Pn[r * kDnLDP + kk]);
                // NaN never wins a '>' against the running maximum; a NaN already on the diagonal is never beaten
                const double key = (a != a) ? (r == kk ? CUDART_INF : -1.0) : a;
                if (key > bestv) {  // r ascending per thread: strict '>' keeps the first
                    bestv = key;
                    bestr = r;
                }
            }
#pragma unroll
            for (int s = 16; s >= 1; s >>= 1) {
                const double ov = __shfl_xor_sync(kFull, bestv, s);
                const int orr = __shfl_xor_sync(kFull, bestr, s);
                if (ov > bestv || (ov == bestv && orr < bestr)) {
                    bestv = ov;
                    bestr = orr;
                }
            }
            if (lane == 0) {
                S.redv[warp] = bestv;
                S.redp[warp] = bestr;
            }
            __syncthreads();
            bestv = S.redv[0];
            bestr = S.redp[0];
#pragma unroll
            for (int wv = 1; wv < kDnThreads / 32; ++wv) {
                const double ov = S.redv[wv];
                const int orr = S.redp[wv];
                if (ov > bestv || (ov == bestv && orr < bestr)) {
                    bestv = ov;
                    bestr = orr;
                }
            }
            const double maxAbs = (akk != akk) ? akk : bestv;
            if (maxAbs < eps) {  // solver.hpp:58-61 (false for NaN)
                singular = true;
                break;
            }
            // row exchange (solver.hpp:64-67): the whole panel row, multipliers included, and the permutation
            if (bestr != kk) {
                if (tid < kDnNB) {
                    const double t0 = S.Pn[kk * kDnLDP + tid];
                    S.Pn[kk * kDnLDP + tid] = S.Pn[bestr * kDnLDP + tid];
                    S.Pn[bestr * kDnLDP + tid] = t0;
                } else if (tid == 32) {
                    const int t0 = S.rowat[k0 + kk];
                    S.rowat[k0 + kk] = S.rowat[k0 + bestr];
                    S.rowat[k0 + bestr] = t0;
                }
            }
            __syncthreads();
            // multipliers and the rest of the panel (solver.hpp:69-76)
            const double pv = S.Pn[kk * kDnLDP + kk];
            for (int r = kk + 1 + tid; r < mrows; r += kDnThreads) {
                double* row = S.Pn + r * kDnLDP;
                const double f = row[kk] / pv;
                row[kk] = f;
                for (int c = kk + 1; c < nb; ++c) row[c] = d_sub(row[c], d_mul(f, S.Pn[kk * kDnLDP + c]));
            }
            __syncthreads();
        }
        if (singular) break;
        DN_T(1);
        // the U part of the panel's own rows goes back to the matrix (back substitution reads it)
        for (int e = tid; e < nb * kDnNB; e += kDnThreads) {
            const int t = e / kDnNB, c = e - t * kDnNB;
            if (c >= t && c < nb) A[(size_t)S.rowat[k0 + t] * LD + k0 + c] = S.Pn[t * kDnLDP + c];
        }
        if (keepL)  // multipliers of this panel: every row below the diagonal, through rowat (later exchanges move them along)
            for (int e = tid; e < mrows * kDnNB; e += kDnThreads) {
                const int r = e / kDnNB, c = e - r * kDnNB;
                if (c < nb && c < r) A[(size_t)S.rowat[k0 + r] * LD + k0 + c] = S.Pn[r * kDnLDP + c];
            }
        // ---- columns right of the panel (the right-hand side, column N, included), kDnCW at a time
        // Full panels: the strip rows travel by TMA (see above), block j0 + kDnCW while block j0 is being used; the
        // last, partial panel (nb < 16: only column N is left, its address is not 16-byte aligned) loads them directly.
        const bool tma = nb == kDnNB;
        auto issue_strip = [&](int j0, int b) {  // thread 0 only; every earlier access to buffer b is behind a barrier
            const int w = (N + 1 - j0 < kDnCW) ? N + 1 - j0 : kDnCW;
            const unsigned bytes = (unsigned)((w + 1) & ~1) * 8u;  // (even count: LD pads every row to a multiple of 4)
            dn_fence_proxy_async();
            dn_mbar_expect_tx(&S.bar[b], bytes * kDnNB);
            for (int t = 0; t < kDnNB; ++t)
                dn_bulk_g2s(S.Us + ((size_t)b * kDnNB + t) * kDnCW, A + (size_t)S.rowat[k0 + t] * LD + j0, bytes, &S.bar[b]);
        };
        int buf = 0;
        if (tma && tid == 0 && k0 + nb <= N) issue_strip(k0 + nb, 0);
        for (int j0 = k0 + nb; j0 <= N; j0 += kDnCW) {
            double* Ub = S.Us + (size_t)buf * kDnNB * kDnCW;
            if (tma) dn_mbar_wait(&S.bar[buf], S.ph[buf]);
            // strip of U: rows k0..k0+nb-1 of these columns, triangular solve against the panel's L (ascending k)
            if (tid < kDnCW) {
                const int j = j0 + tid;
                if (j <= N) {
                    double* us = Ub + tid;
                    if (!tma) {
#pragma unroll
                        for (int t = 0; t < kDnNB; ++t)  // independent loads first
                            if (t < nb) us[t * kDnCW] = A[(size_t)S.rowat[k0 + t] * LD + j];
                    }
#pragma unroll 1
                    for (int t = 1; t < nb; ++t) {
                        double a = us[t * kDnCW];
                        const double* lt = S.Pn + t * kDnLDP;
#pragma unroll 4
                        for (int q = 0; q < t; ++q) a = d_sub(a, d_mul(lt[q], us[q * kDnCW]));
                        us[t * kDnCW] = a;
                    }
#pragma unroll
                    for (int t = 1; t < kDnNB; ++t)  // (row k0 itself is unchanged)
                        if (t < nb) A[(size_t)S.rowat[k0 + t] * LD + j] = us[t * kDnCW];
                }
            }
            __syncthreads();
            if (tma && tid == 0) {  // everyone is past the wait on this buffer and past the last use of the other one
                S.ph[buf] ^= 1u;
                if (j0 + kDnCW <= N) issue_strip(j0 + kDnCW, buf ^ 1);
            }
            DN_T(2);
            // trailing update: rows at positions >= k0+nb (only reached with a full panel, nb == kDnNB)
            const int tx = tid & 31, ty = tid >> 5;
            for (int p0 = k0 + nb; p0 < N; p0 += (kDnThreads / 32) * kDnRM) {
                double acc[kDnRM][kDnCN];
                int prow[kDnRM], prel[kDnRM];
#pragma unroll
                for (int r = 0; r < kDnRM; ++r) {
                    const int p = p0 + ty * kDnRM + r;
                    prow[r] = p < N ? S.rowat[p] : -1;
                    prel[r] = (p < N ? p : N - 1) - k0;
                }
#pragma unroll
                for (int r = 0; r < kDnRM; ++r)
#pragma unroll
                    for (int c = 0; c < kDnCN; ++c) {
                        const int j = j0 + tx + 32 * c;
                        acc[r][c] = (prow[r] >= 0 && j <= N) ? A[(size_t)prow[r] * LD + j] : 0.0;
                    }
#pragma unroll 4
                for (int t = 0; t < kDnNB; ++t) {
                    double l[kDnRM], u[kDnCN];
#pragma unroll
                    for (int r = 0; r < kDnRM; ++r) l[r] = S.Pn[prel[r] * kDnLDP + t];
#pragma unroll
                    for (int c = 0; c < kDnCN; ++c) u[c] = Ub[t * kDnCW + tx + 32 * c];
#pragma unroll
                    for (int r = 0; r < kDnRM; ++r)
#pragma unroll
                        for (int c = 0; c < kDnCN; ++c) acc[r][c] = d_sub(acc[r][c], d_mul(l[r], u[c]));  // solver.hpp:74
                }
#pragma unroll
                for (int r = 0; r < kDnRM; ++r)
#pragma unroll
                    for (int c = 0; c < kDnCN; ++c) {
                        const int j = j0 + tx + 32 * c;
                        if (prow[r] >= 0 && j <= N) A[(size_t)prow[r] * LD + j] = acc[r][c];
                    }
            }
            __syncthreads();
            DN_T(3);
            buf ^= 1;
        }
    }
    __syncthreads();  // (a singular break leaves the pivot-search slots still being read)
    // pivot-sequence hash (definition: include/csim_b200.h), warp 0
    if (warp == 0) {
        uint32_t lo = 0, hi = 0;
        for (int i = lane; i < N; i += 32) {
            const uint32_t key = (uint32_t)i * 1024u + (uint32_t)S.rowat[i] + 1u;
            uint32_t l = key * 0x9E3779B1u;
            l ^= l >> 15;
            uint32_t g = key * 0x85EBCA77u;
            g ^= g >> 13;
            lo += l;
            hi += g;
        }
        lo = __reduce_add_sync(kFull, lo);
        hi = __reduce_add_sync(kFull, hi);
        if (lane == 0) {
            S.redp[0] = (int32_t)lo;
            S.redp[1] = (int32_t)hi;
        }
    }
    __syncthreads();
    h = (h ^ (((uint64_t)(uint32_t)S.redp[1] << 32) | (uint32_t)S.redp[0])) * 0x100000001B3ull;
    if (singular) {  // solver.hpp:94-97
        for (int i = tid; i < N; i += kDnThreads) S.xr[i] = 0.0;
        __syncthreads();
        return true;
    }
    // ---- back substitution (solver.hpp:116-128): positions bottom-up, per row the subtraction chain runs in
    //      ascending column order on one thread (the reference's order; the chain cannot start before x(i+1) is
    //      known, so it is the critical path whatever the thread count).  Rows are taken kDnRB at a time: the
    //      whole CTA loads them (one memory latency per block, coalesced) and pre-multiplies the columns whose
    //      x is already known; the chain then runs out of shared memory.
    double* Ub = S.Pn;  // [kDnRB][N + 1]
    const int ldu = N + 1;
    for (int hi = N - 1; hi >= 0; hi -= kDnRB) {
        const int lo = hi - kDnRB + 1 > 0 ? hi - kDnRB + 1 : 0, nr = hi - lo + 1, ncol = N + 1 - lo;
#pragma unroll 4
        for (int e = tid; e < nr * ncol; e += kDnThreads) {
            const int r = e / ncol, j = lo + (e - r * ncol);
            double v = A[(size_t)S.rowat[lo + r] * LD + j];
            if (j > hi && j < N) v = d_mul(v, S.xr[j]);
            Ub[r * ldu + j] = v;
        }
        __syncthreads();
        if (tid == 0) {
            for (int i = hi; i >= lo; --i) {
                const double* u = Ub + (i - lo) * ldu;
                double sum = u[N];
                for (int j = i + 1; j <= hi; ++j) sum = d_sub(sum, d_mul(u[j], S.xr[j]));
                // the long part of the chain: 8 values are loaded while the previous 8 are being subtracted, so the
                // loop runs at the latency of the dependent subtraction, not of shared memory
                int j = hi + 1;
                double nx[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) nx[q] = (j + q < N) ? u[j + q] : 0.0;
                for (; j + 8 <= N; j += 8) {
                    double cu[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) cu[q] = nx[q];
#pragma unroll
                    for (int q = 0; q < 8; ++q) nx[q] = (j + 8 + q < N) ? u[j + 8 + q] : 0.0;
#pragma unroll
                    for (int q = 0; q < 8; ++q) sum = d_sub(sum, cu[q]);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (j + q < N) sum = d_sub(sum, nx[q]);
                const double diag = u[i];
                S.xr[i] = (fabs(diag) < 1e-15) ? 0.0 : sum / diag;
            }
        }
        __syncthreads();
    }
    DN_T(4);
#ifdef CSIM_DN_PROF
    if (tid == 0)
        for (int i = 0; i < 5; ++i) atomicAdd(&dn_prof[i], (unsigned long long)prof_acc[i]);
#endif
    return false;
}

// one Newton linear solve: assemble (value table -> [A | b], reference order), factor, solve
__device__ __forceinline__ bool dn_newton_solve(const DevMode& m, const double* vt, double* A, int N, int LD,
                                                const DnSmem& S, double eps, int tid, uint64_t& h) {
    dn_assemble(m, vt, A, N, LD, tid);
    return dn_lu_solve(A, N, LD, S, eps, tid, h);
}

// || xn - x ||_2 in the order of oracle/csim_norm2.h while committing xn = x + alpha (xr - x); warp 0 computes
// (lane-strided partial sums, xor butterfly), every thread gets the value
__device__ __forceinline__ double dn_damp(const DnSmem& S, int N, double alpha, int tid) {
    if (tid < 32) {
        double acc = 0.0;
        for (int i = tid; i < N; i += 32) {
            const double x = S.xs[i];
            const double xn = d_add(x, d_mul(alpha, d_sub(S.xr[i], x)));
            const double d = d_sub(xn, x);
            acc = d_add(acc, d_mul(d, d));
            S.xs[i] = xn;
        }
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) acc = d_add(acc, __shfl_xor_sync(kFull, acc, s));
        if (tid == 0) S.redv[0] = sqrt(acc);
    }
    __syncthreads();
    const double err = S.redv[0];
    __syncthreads();
    return err;
}

__device__ __forceinline__ bool dn_all_finite(const DnSmem& S, int N, int tid) {
    bool ok = true;
    for (int i = tid; i < N; i += kDnThreads) ok = ok && lane_finite(S.xr[i]);
    return __syncthreads_and(ok) != 0;
}

__device__ __forceinline__ double dn_base_gmin(const csim_constants& c, double s) {  // include/dcanalysis.hpp:45-48
    if (s < 0.0) s = 0.0;
    if (s > 1.0) s = 1.0;
    return d_add(d_mul(c.gmin_high_base, d_sub(1.0, s)), d_mul(c.gmin_low_base, s));
}

// per-CTA global workspace (CTAs are persistent: instance b = blockIdx.x, blockIdx.x + gridDim.x, ...)
struct DnMem {
    double* A;   // [grid][N * LD]
    double* vt;  // [grid][n_values]
};

__global__ void __launch_bounds__(kDnThreads, 2)
tran_dense_kernel(DevPlan P, DevTran T, int64_t B, const double* __restrict__ params, double* __restrict__ xio,
                  double* __restrict__ wave, int32_t* __restrict__ iters_total, uint8_t* __restrict__ iters_step,
                  uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status, DnMem M) {
    extern __shared__ unsigned char dn_smem[];
    const DevMode& m = P.m[CSIM_MODE_TRAN];
    const csim_constants& c = P.c;
    const int tid = threadIdx.x, N = P.N, LD = dn_ld(N);
    const DnSmem S = dn_carve(dn_smem, N);
    dn_init_sync(S, tid);
    double* A = M.A + (size_t)blockIdx.x * N * LD;
    double* vt = M.vt + (size_t)blockIdx.x * m.n_values;
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        if (T.skip && T.skip[b]) continue;
        uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
        uint32_t st = 0;
        int total = 0;
        __syncthreads();
        for (int i = tid; i < N; i += kDnThreads) S.xs[i] = xio[(int64_t)i * B + b];
        eval_static(m, params, B, b, T.dt, c.tran_gmin, vt, tid, kDnThreads);
        __syncthreads();
        const int r0 = T.out_begin_step / T.out_stride;
        for (int step = T.step_begin; step < T.step_end; ++step) {
            eval_step(m, S.xs, vt, tid, kDnThreads);  // xs still holds the last accepted timepoint here
            eval_src_tran(m, T, params, B, b, step, vt, tid, kDnThreads);
            int count = 0;
            bool dead = false;
            for (int iter = 0; iter < c.tran_max_iters; ++iter) {  // src/tanalisis.cpp:258
                ++count;
                eval_mos(m, params, B, b, c.mos_off_gds, S.xs, vt, tid, kDnThreads);
                __syncthreads();
                if (dn_newton_solve(m, vt, A, N, LD, S, c.lu_eps, tid, h)) st |= CSIM_ST_LU_SINGULAR;
                if (!dn_all_finite(S, N, tid)) {  // src/tanalisis.cpp:360-362 (the reference throws)
                    st |= CSIM_ST_NONFINITE;
                    dead = true;
                    break;
                }
                const double err = dn_damp(S, N, c.tran_alpha, tid);  // src/tanalisis.cpp:365-366
                if (err < c.tran_tol) break;
                if (iter == c.tran_max_iters - 1) st |= CSIM_ST_TRAN_UNCONVERGED;
            }
            total += count;
            if (iters_step && tid == 0) iters_step[(int64_t)step * B + b] = (uint8_t)count;
            if (dead) break;
            if ((step + 1) % T.out_stride == 0 && step >= T.out_begin_step) {
                const int64_t r = (step + 1) / T.out_stride - 1 - r0;
                for (int p = tid; p < T.n_probes; p += kDnThreads)
                    wave[((int64_t)b * T.n_probes + p) * T.wave_ld + r] = S.xs[T.probes[p]];
            }
        }
        for (int i = tid; i < N; i += kDnThreads) xio[(int64_t)i * B + b] = S.xs[i];
        if (tid == 0) {
            if (iters_total) iters_total[b] += total;
            if (pivot_hash) pivot_hash[b] = h;
            if (status) status[b] |= st;
        }
    }
}

__global__ void __launch_bounds__(kDnThreads, 2)
dc_dense_kernel(DevPlan P, int64_t B, const double* __restrict__ params, double* __restrict__ xout,
                int32_t* __restrict__ iters, uint64_t* __restrict__ pivot_hash, uint32_t* __restrict__ status, DnMem M) {
    extern __shared__ unsigned char dn_smem[];
    const DevMode& m = P.m[CSIM_MODE_DC];
    const csim_constants& c = P.c;
    const int tid = threadIdx.x, N = P.N, LD = dn_ld(N);
    const DnSmem S = dn_carve(dn_smem, N);
    dn_init_sync(S, tid);
    double* A = M.A + (size_t)blockIdx.x * N * LD;
    double* vt = M.vt + (size_t)blockIdx.x * m.n_values;
    int gmin_slot = -1;
    for (int i = 0; i < m.n_static; ++i)
        if (m.static_ops[i].kind == SK_GMIN) gmin_slot = m.static_ops[i].out;
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        uint64_t h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;
        uint32_t st = 0;
        __syncthreads();
        for (int i = tid; i < N; i += kDnThreads) S.xs[i] = 0.0;
        eval_static(m, params, B, b, 0.0, 0.0, vt, tid, kDnThreads);
        __syncthreads();
        if (!P.has_mos) {  // dcSolveDirectLU (src/dcanalysis.cpp:46-68)
            eval_src_dc(m, params, B, b, 1.0, vt, tid, kDnThreads);
            __syncthreads();
            if (dn_newton_solve(m, vt, A, N, LD, S, c.lu_eps, tid, h)) st |= CSIM_ST_LU_SINGULAR;
            for (int i = tid; i < N; i += kDnThreads) S.xs[i] = S.xr[i];
            if (iters && tid < c.dc_ramp_steps) iters[(int64_t)tid * B + b] = tid == 0 ? 1 : 0;
        } else {  // dcSolveNewtonLU (src/dcanalysis.cpp:95-163)
            for (int step = 1; step <= c.dc_ramp_steps; ++step) {
                const double scale = (double)step / (double)c.dc_ramp_steps;
                double gmin = dn_base_gmin(c, scale);
                double prevErr = CUDART_INF;
                eval_src_dc(m, params, B, b, scale, vt, tid, kDnThreads);
                int count = 0;
                for (int iter = 0; iter < c.dc_max_iters; ++iter) {
                    ++count;
                    if (tid == 0 && gmin_slot >= 0) vt[gmin_slot] = gmin;
                    eval_mos(m, params, B, b, c.mos_off_gds, S.xs, vt, tid, kDnThreads);
                    __syncthreads();
                    if (dn_newton_solve(m, vt, A, N, LD, S, c.lu_eps, tid, h)) st |= CSIM_ST_LU_SINGULAR;
                    if (!dn_all_finite(S, N, tid)) {  // dcanalysis.cpp:135-138
                        st |= CSIM_ST_NONFINITE;
                        const double g10 = d_mul(gmin, c.dc_nonfinite_mul);
                        gmin = g10 < c.dc_nonfinite_cap ? g10 : c.dc_nonfinite_cap;
                        continue;
                    }
                    const double err = dn_damp(S, N, c.dc_alpha, tid);  // ConvController::update (dcanalysis.cpp:268-307)
                    const double gbase = dn_base_gmin(c, scale);
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
                if (iters && tid == 0) iters[(int64_t)(step - 1) * B + b] = count;
            }
        }
        __syncthreads();
        for (int i = tid; i < N; i += kDnThreads) xout[(int64_t)i * B + b] = S.xs[i];
        if (tid == 0) {
            if (pivot_hash) pivot_hash[b] = h;
            if (status) status[b] |= st;
        }
    }
}

// csim_dev_lu_solve for n > 32: A [n*n][B], rhs [n][B] instance-minor
__global__ void __launch_bounds__(kDnThreads, 2)
lu_dense_kernel(int N, int64_t B, const double* __restrict__ Ain, const double* __restrict__ bin,
                double* __restrict__ xout, int32_t* __restrict__ perm_out, int32_t* __restrict__ singular, double eps,
                DnMem M, double* __restrict__ lu_out) {
    extern __shared__ unsigned char dn_smem[];
    const int tid = threadIdx.x, LD = dn_ld(N);
    const DnSmem S = dn_carve(dn_smem, N);
    dn_init_sync(S, tid);
    double* A = M.A + (size_t)blockIdx.x * N * LD;
    for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
        __syncthreads();
        for (int e = tid; e < N * (N + 1); e += kDnThreads) {
            const int i = e / (N + 1), j = e - i * (N + 1);
            A[(size_t)i * LD + j] = j < N ? Ain[((int64_t)i * N + j) * B + b] : (bin ? bin[(int64_t)i * B + b] : 0.0);
        }
        __syncthreads();
        uint64_t h = 0;
        const bool sing = dn_lu_solve(A, N, LD, S, eps, tid, h, lu_out != nullptr);
        __syncthreads();
        for (int i = tid; i < N; i += kDnThreads) {
            if (xout) xout[(int64_t)i * B + b] = S.xr[i];
            if (perm_out) perm_out[(int64_t)i * B + b] = S.rowat[i];
        }
        if (lu_out)  // packed L\U in position order (the reference exchanges rows physically, solver.hpp:64-67)
            for (int e = tid; e < N * N; e += kDnThreads) {
                const int i = e / N, j = e - i * N;
                lu_out[(int64_t)e * B + b] = A[(size_t)S.rowat[i] * LD + j];
            }
        if (tid == 0 && singular) singular[b] = sing ? 1 : 0;
    }
}

// ---- launch wrappers (scratch is stream-ordered) ---------------------------------------------------
template <typename K>
static cudaError_t dn_grid(K kernel, int N, int64_t B, size_t& smem, int& grid) {
    smem = dn_smem_bytes(N);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 0, occ = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kDnThreads, smem)) != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    const int64_t slots = (int64_t)sms * occ;
    grid = (int)(B < slots ? B : slots);
    return cudaSuccess;
}
static cudaError_t dn_alloc(int N, int n_values, int grid, cudaStream_t stream, DnMem& M) {
    M = DnMem{};
    cudaError_t e;
    if ((e = cudaMallocAsync((void**)&M.A, (size_t)N * dn_ld(N) * grid * sizeof(double), stream)) != cudaSuccess) return e;
    return cudaMallocAsync((void**)&M.vt, (size_t)(n_values > 0 ? n_values : 1) * grid * sizeof(double), stream);
}
static void dn_free(DnMem& M, cudaStream_t stream) {
    if (M.A) cudaFreeAsync(M.A, stream);
    if (M.vt) cudaFreeAsync(M.vt, stream);
}

cudaError_t launch_tran_dense(const DevPlan& P, const DevTran& T, int64_t B, const double* params, double* x,
                              double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash,
                              uint32_t* status, cudaStream_t stream) {
    size_t smem;
    int grid;
    cudaError_t e = dn_grid(tran_dense_kernel, P.N, B, smem, grid);
    if (e != cudaSuccess) return e;
    DnMem M;
    e = dn_alloc(P.N, P.m[CSIM_MODE_TRAN].n_values, grid, stream, M);
    if (e == cudaSuccess) {
        tran_dense_kernel<<<grid, kDnThreads, smem, stream>>>(P, T, B, params, x, wave, iters_total, iters_step,
                                                              pivot_hash, status, M);
        e = cudaGetLastError();
    }
    dn_free(M, stream);
    return e;
}

cudaError_t launch_dc_dense(const DevPlan& P, int64_t B, const double* params, double* x, int32_t* iters,
                            uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream) {
    size_t smem;
    int grid;
    cudaError_t e = dn_grid(dc_dense_kernel, P.N, B, smem, grid);
    if (e != cudaSuccess) return e;
    DnMem M;
    e = dn_alloc(P.N, P.m[CSIM_MODE_DC].n_values, grid, stream, M);
    if (e == cudaSuccess) {
        dc_dense_kernel<<<grid, kDnThreads, smem, stream>>>(P, B, params, x, iters, pivot_hash, status, M);
        e = cudaGetLastError();
    }
    dn_free(M, stream);
    return e;
}

cudaError_t launch_lu_dense(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                            int32_t* singular, double eps, cudaStream_t stream, double* lu_out) {
    size_t smem;
    int grid;
    cudaError_t e = dn_grid(lu_dense_kernel, N, B, smem, grid);
    if (e != cudaSuccess) return e;
    DnMem M;
    e = dn_alloc(N, 0, grid, stream, M);
    if (e == cudaSuccess) {
        lu_dense_kernel<<<grid, kDnThreads, smem, stream>>>(N, B, A, b, x, perm, singular, eps, M, lu_out);
        e = cudaGetLastError();
    }
    dn_free(M, stream);
    return e;
}

// ---- allocation-free variant of the dense LU solve (csim_dev_lu_solve_noalloc) -----------------------------------
// Callers under CUDA graph capture, tight serving loops or harnesses that forbid cudaMalloc cannot take
// launch_lu_dense's stream-ordered scratch.  Here the per-CTA [A | b] workspaces live in a file-scope device array of
// a fixed, modest size; every GPU that loads the library holds it until exit.  The array's name is a host-side
// stand-in in host code, NOT a device address: the address is asked for once per device (with that device current)
// through cudaGetSymbolAddress and cached.  Every call shares the array and it keeps the last call's contents;
// nothing has to start at zero (each CTA rewrites all N x (N + 1) entries it reads before factorising, the padding
// columns of a row are never read), but two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) double dn_noalloc_scratch[kDnNoAllocDoubles];

static cudaError_t dn_noalloc_address(double** out) {
    constexpr int kMaxDev = 64;
    static double* addr[kMaxDev] = {};
    static bool have[kMaxDev] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
    if (!have[dev]) {  // (a race between host threads writes the same value twice)
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, dn_noalloc_scratch);
        if (e != cudaSuccess) return e;
        addr[dev] = p;
        have[dev] = true;
    }
    *out = addr[dev];
    return cudaSuccess;
}

// largest n whose single [A | b] workspace fits the static array
int lu_dense_noalloc_max_n() {
    int n = kDenseMaxN;
    while (n > 1 && (size_t)n * dn_ld(n) > kDnNoAllocDoubles) --n;
    return n;
}

cudaError_t launch_lu_dense_noalloc(int N, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                                    int32_t* singular, double eps, cudaStream_t stream) {
    size_t smem;
    int grid;
    cudaError_t e = dn_grid(lu_dense_kernel, N, B, smem, grid);
    if (e != cudaSuccess) return e;
    const size_t per_cta = (size_t)N * dn_ld(N);
    const size_t fit = kDnNoAllocDoubles / per_cta;  // persistent CTAs loop over the batch: fewer CTAs, same results
    if (fit < 1) return cudaErrorInvalidValue;
    if ((size_t)grid > fit) grid = (int)fit;
    DnMem M{};
    if ((e = dn_noalloc_address(&M.A)) != cudaSuccess) return e;
    lu_dense_kernel<<<grid, kDnThreads, smem, stream>>>(N, B, A, b, x, perm, singular, eps, M, nullptr);
    return cudaGetLastError();
}

#ifdef CSIM_DN_PROF
cudaError_t dn_prof_read(unsigned long long* out) { return cudaMemcpyFromSymbol(out, dn_prof, sizeof(dn_prof)); }
#endif

}  // namespace csim
