/* csim_oracle.c — TEST INFRASTRUCTURE (see csim_oracle.h).
 *
 * Plain-C restatement of the reference's Newton/MNA hot path.  Every function
 * names the reference lines it follows (paths relative to /root/reference).
 * Build with -ffp-contract=off: the reference results are FMA-sensitive at the
 * parity tolerance.  Dense, scalar, single-threaded — exactly
 * the reference's operation order; no attempt at speed.
 */
#include "csim_oracle.h"
#include "csim_norm2.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define REF_M_PI 3.14159265358979323846 /* include/sim.hpp:8 */

double oracle_norm2(const double* v, int64_t n) { return csim_norm2(v, (long)n); }

/* same definition as csim_pivot_hash_update (include/csim_b200.h) */
uint64_t oracle_pivot_hash_update(uint64_t h, const int32_t* perm, int32_t n) {
    uint32_t flo = 0, fhi = 0;
    for (int32_t i = 0; i < n; ++i) {
        uint32_t key = (uint32_t)i * 1024u + (uint32_t)perm[i] + 1u;
        uint32_t lo = key * 0x9E3779B1u;
        lo ^= lo >> 15;
        uint32_t hi = key * 0x85EBCA77u;
        hi ^= hi >> 13;
        flo += lo;
        fhi += hi;
    }
    uint64_t f = ((uint64_t)fhi << 32) | (uint64_t)flo;
    return (h ^ f) * 0x100000001B3ull;
}

int32_t oracle_count_sources(const csim_topology* t) {
    int32_t n = 0;
    for (int32_t e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_V || t->elem_type[e] == CSIM_ELEM_I) ++n;
    return n;
}

/* getV lambda of src/element.cpp:197-200 / getNodeVoltage of src/tanalisis.cpp:19-26 */
static double node_v(const double* x, int eq) { return eq >= 0 ? x[eq] : 0.0; }

/* SourceSpec::evalDC, include/sim.hpp:152-158 */
void oracle_src_dc(const csim_topology* t, const double* params, double scale, double* src_val) {
    int32_t s = 0;
    for (int32_t e = 0; e < t->n_elems; ++e) {
        int ty = t->elem_type[e];
        if (ty != CSIM_ELEM_V && ty != CSIM_ELEM_I) continue;
        const double* p = params + t->elem_param[e];
        double base = p[0];
        if (t->elem_flags[e] & CSIM_FLAG_SRC_SIN) base += p[1];
        src_val[s++] = base * scale;
    }
}

/* TranWaveform::eval SIN branch, include/sim.hpp:117-122, tabulated part only */
void oracle_sin_table(double freq, double td, double phi, double dt, int32_t n_steps, double* out) {
    for (int32_t s = 0; s < n_steps; ++s) {
        double t = (s + 1) * dt; /* src/tanalisis.cpp:256 */
        double tau = t - td;
        double w = 2.0 * REF_M_PI * freq;
        out[s] = sin(w * tau + phi);
    }
}

/* clamp01, include/sim.hpp (helper of the PULSE branch) */
static double clamp01(double k) { return k < 0.0 ? 0.0 : (k > 1.0 ? 1.0 : k); }

/* TranWaveform::eval PULSE branch, include/sim.hpp:80-115.  spec = v1 v2 td tr tf ton per */
static double pulse_eval(const double* p, double t) {
    const double v1 = p[0], v2 = p[1], td = p[2], tr = p[3], tf = p[4], ton = p[5], per = p[6];
    if (per <= 0.0) { /* single shot, sim.hpp:81-96 */
        double tau = t - td;
        if (tau <= 0.0) return v1;
        if (tau < tr) {
            double k = clamp01(tau / tr);
            return v1 + k * (v2 - v1);
        } else if (tau < tr + ton) {
            return v2;
        } else {
            double tfall = tau - (tr + ton);
            double k = clamp01(tfall / tf);
            return v2 + k * (v1 - v2);
        }
    }
    /* periodic, sim.hpp:97-114 */
    if (t < td) return v1;
    double tau = fmod(t - td, per);
    if (tau < 0.0) tau += per;
    if (tau < tr) {
        double k = clamp01(tau / tr);
        return v1 + (v2 - v1) * k;
    } else if (tau < tr + ton) {
        return v2;
    } else if (tau < tr + ton + tf) {
        double tfall = tau - (tr + ton);
        double k = clamp01(tfall / tf);
        return v2 + (v1 - v2) * k;
    }
    return v1;
}

/* TranWaveform::eval PWL branch, include/sim.hpp:124-138.  spec = t[0..n) then v[0..n) */
static double pwl_eval(const double* tt, const double* vv, int32_t n, double t) {
    if (n <= 0) return 0.0;
    if (t <= tt[0]) return vv[0];
    if (t >= tt[n - 1]) return vv[n - 1];
    for (int32_t i = 0; i + 1 < n; ++i) {
        if (t > tt[i] && t <= tt[i + 1]) {
            double k = (t - tt[i]) / (tt[i + 1] - tt[i]);
            return vv[i] + (vv[i + 1] - vv[i]) * k;
        }
    }
    return vv[n - 1];
}

/* CSIM_SRC_VAL_TABLE row: out[s] = TranWaveform::eval((s+1)*dt).  type 1 = PULSE (spec: 7 doubles, n ignored),
 * type 3 = PWL (spec: n times then n values) — the reference's WaveformType numbering (sim.hpp:25-30). */
void oracle_waveform_table(int32_t type, const double* spec, int32_t n, double dt, int32_t n_steps, double* out) {
    for (int32_t s = 0; s < n_steps; ++s) {
        double t = (s + 1) * dt; /* src/tanalisis.cpp:256 */
        out[s] = type == 1 ? pulse_eval(spec, t) : (type == 3 ? pwl_eval(spec, spec + n, n, t) : 0.0);
    }
}

/* SourceSpec::evalTran, include/sim.hpp:160-162 (+ eval 75-143 through the table) */
void oracle_src_tran(const csim_topology* t, const double* params, const int32_t* src_mode,
                     const int32_t* src_group, const double* table, int32_t table_steps, int32_t step,
                     double dt, double* src_val) {
    int32_t s = 0;
    double tnow = (step + 1) * dt;
    for (int32_t e = 0; e < t->n_elems; ++e) {
        int ty = t->elem_type[e];
        if (ty != CSIM_ELEM_V && ty != CSIM_ELEM_I) continue;
        const double* p = params + t->elem_param[e];
        double w;
        switch (src_mode[s]) {
            case CSIM_SRC_SIN_TABLE:
                if (tnow < p[4]) w = p[1];
                else w = p[1] + p[2] * table[(int64_t)src_group[s] * table_steps + step];
                break;
            case CSIM_SRC_VAL_TABLE:
                w = table[(int64_t)src_group[s] * table_steps + step];
                break;
            default:
                w = 0.0;
        }
        src_val[s++] = p[0] + w;
    }
}

/* stampCapBE, src/tanalisis.cpp:59-80 */
static void stamp_cap_be(int n, int eq1, int eq2, double C, double dt, double vPrev, double* G, double* I) {
    if (C <= 0.0 || dt <= 0.0) return;
    double Gc = C / dt;
    if (eq1 >= 0) G[eq1 * n + eq1] += Gc;
    if (eq2 >= 0) G[eq2 * n + eq2] += Gc;
    if (eq1 >= 0 && eq2 >= 0) {
        G[eq1 * n + eq2] -= Gc;
        G[eq2 * n + eq1] -= Gc;
    }
    double I_hist = -Gc * vPrev;
    if (eq1 >= 0) I[eq1] -= I_hist;
    if (eq2 >= 0) I[eq2] += I_hist;
}

/* MosfetBase::stamp, src/element.cpp:181-307 */
static void stamp_mos(const csim_constants* c, int n, const int32_t* nd, int isP, const double* p,
                      const double* x, double* G, double* I) {
    int eqD = nd[0], eqG = nd[1], eqS = nd[2];
    double Vth = p[0], K = p[1], lambda = p[2];
    double Vd = node_v(x, eqD), Vg = node_v(x, eqG), Vs = node_v(x, eqS);
    double pp = isP ? -1.0 : 1.0;
    double Vgs_eff = pp * (Vg - Vs);
    double Vds_eff = pp * (Vd - Vs);
    double Ids0 = 0.0, gds0 = 0.0, gm0 = 0.0;
    int on = 0;
    if (Vgs_eff > Vth && Vds_eff >= 0) {
        on = 1;
        double Vov = Vgs_eff - Vth;
        if (Vds_eff < Vov) {
            Ids0 = K * (Vov * Vds_eff - 0.5 * Vds_eff * Vds_eff);
            gds0 = K * (Vov - Vds_eff);
            gm0 = K * Vds_eff;
        } else {
            Ids0 = 0.5 * K * Vov * Vov;
            gds0 = 0.0;
            gm0 = K * Vov;
        }
    }
    if (!on) {
        Ids0 = 0.0;
        gm0 = 0.0;
        gds0 = c->mos_off_gds;
    }
    double factor = 1.0 + lambda * Vds_eff;
    if (factor < 0.0) factor = 0.0;
    double Ids_eff = Ids0 * factor;
    double dId_dVds_eff = gds0 * factor + Ids0 * lambda;
    double dId_dVgs_eff = gm0 * factor;
    double Ids = pp * Ids_eff;
    double gd = dId_dVds_eff;
    double gg = dId_dVgs_eff;
    double gs = -(dId_dVds_eff + dId_dVgs_eff);
    double cst = Ids - gd * Vd - gg * Vg - gs * Vs;
    if (eqD >= 0) {
        if (eqD >= 0) G[eqD * n + eqD] += gd;
        if (eqG >= 0) G[eqD * n + eqG] += gg;
        if (eqS >= 0) G[eqD * n + eqS] += gs;
        I[eqD] -= cst;
    }
    if (eqS >= 0) {
        if (eqD >= 0) G[eqS * n + eqD] += -gd;
        if (eqG >= 0) G[eqS * n + eqG] += -gg;
        if (eqS >= 0) G[eqS * n + eqS] += -gs;
        I[eqS] += cst;
    }
}

/* the "0 V source" part shared by VoltageSource::stamp (src/element.cpp:113-120),
 * Inductor::stamp (172-176) and the inductor companion (src/tanalisis.cpp:311-316) */
static void stamp_branch(int n, int eqP, int eqM, int k, double* G) {
    if (eqP >= 0) G[eqP * n + k] += 1.0;
    if (eqM >= 0) G[eqM * n + k] -= 1.0;
    if (eqP >= 0) G[k * n + eqP] += 1.0;
    if (eqM >= 0) G[k * n + eqM] -= 1.0;
}

/* Element::stamp for R / I / V, src/element.cpp:9-32, 34-66, 83-123 */
static void stamp_rvi(int n, int ty, const int32_t* nd, int k, const double* p, double srcv, double* G, double* I) {
    if (ty == CSIM_ELEM_R) {
        int eq1 = nd[0], eq2 = nd[1];
        if (p[0] == 0.0) return;
        double g = 1.0 / p[0];
        if (eq1 >= 0) G[eq1 * n + eq1] += g;
        if (eq2 >= 0) G[eq2 * n + eq2] += g;
        if (eq1 >= 0 && eq2 >= 0) {
            G[eq1 * n + eq2] -= g;
            G[eq2 * n + eq1] -= g;
        }
    } else if (ty == CSIM_ELEM_I) {
        if (nd[0] >= 0) I[nd[0]] -= srcv;
        if (nd[1] >= 0) I[nd[1]] += srcv;
    } else if (ty == CSIM_ELEM_V) {
        if (k < 0 || k >= n) return;
        stamp_branch(n, nd[0], nd[1], k, G);
        I[k] += srcv;
    }
}

/* stampGlobalGmin, src/dcanalysis.cpp:36-43 and src/tanalisis.cpp:30-41 */
static void stamp_gmin(const csim_topology* t, double gmin, double* G) {
    int n = t->n_unknowns;
    for (int eq = 0; eq < t->n_node_eq; ++eq) G[eq * n + eq] += gmin;
}

void oracle_stamp(const csim_topology* t, const csim_constants* c, int mode, const double* params,
                  const double* x, const double* x_prev, const double* src_val, double dt, double gmin,
                  double* G, double* I) {
    const int n = t->n_unknowns;
    memset(G, 0, sizeof(double) * (size_t)n * (size_t)n);
    memset(I, 0, sizeof(double) * (size_t)n);
    if (mode == CSIM_MODE_DC) {
        /* src/dcanalysis.cpp:126-130: every element in netlist order, then gmin */
        int s = 0;
        for (int e = 0; e < t->n_elems; ++e) {
            int ty = t->elem_type[e];
            const int32_t* nd = t->elem_nodes + 4 * e;
            const double* p = params + t->elem_param[e];
            double sv = 0.0;
            if (ty == CSIM_ELEM_V || ty == CSIM_ELEM_I) sv = src_val[s++];
            if (ty == CSIM_ELEM_R || ty == CSIM_ELEM_V || ty == CSIM_ELEM_I) {
                stamp_rvi(n, ty, nd, t->elem_branch[e], p, sv, G, I);
            } else if (ty == CSIM_ELEM_L) { /* Inductor::stamp, src/element.cpp:156-178 */
                int k = t->elem_branch[e];
                if (k >= 0 && k < n) stamp_branch(n, nd[0], nd[1], k, G);
            } else if (ty == CSIM_ELEM_MOS) {
                stamp_mos(c, n, nd, t->elem_flags[e] & CSIM_FLAG_PMOS, p, x, G, I);
            } /* CapacitorElement::stamp is a no-op, include/element.hpp:103-108 */
        }
        stamp_gmin(t, gmin, G);
        return;
    }
    /* TRAN, src/tanalisis.cpp:268-356 */
    /* 1) everything that is not C / L / MOS (269-274) */
    int s = 0;
    for (int e = 0; e < t->n_elems; ++e) {
        int ty = t->elem_type[e];
        double sv = 0.0;
        if (ty == CSIM_ELEM_V || ty == CSIM_ELEM_I) sv = src_val[s++];
        if (ty == CSIM_ELEM_C || ty == CSIM_ELEM_L || ty == CSIM_ELEM_MOS) continue;
        stamp_rvi(n, ty, t->elem_nodes + 4 * e, t->elem_branch[e], params + t->elem_param[e], sv, G, I);
    }
    /* 2) MOS channel (277-279) */
    for (int e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_MOS)
            stamp_mos(c, n, t->elem_nodes + 4 * e, t->elem_flags[e] & CSIM_FLAG_PMOS,
                      params + t->elem_param[e], x, G, I);
    /* 3) explicit capacitors (282-291), vPrev = v1 - v2 at the last accepted point (139-147, 382-389) */
    for (int e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_C) {
            const int32_t* nd = t->elem_nodes + 4 * e;
            double vPrev = node_v(x_prev, nd[0]) - node_v(x_prev, nd[1]);
            stamp_cap_be(n, nd[0], nd[1], params[t->elem_param[e]], dt, vPrev, G, I);
        }
    /* 4) inductors, BE Thevenin companion (294-319) */
    for (int e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_L) {
            double Lval = params[t->elem_param[e]];
            if (Lval <= 0.0) continue;
            const int32_t* nd = t->elem_nodes + 4 * e;
            int k = t->elem_branch[e];
            if (k < 0 || k >= n) continue;
            double R_eq = Lval / dt;
            double iPrev = x_prev[k];
            double V_hist = -R_eq * iPrev;
            stamp_branch(n, nd[0], nd[1], k, G);
            G[k * n + k] += -R_eq;
            I[k] += V_hist;
        }
    /* 5) MOS parasitic capacitors (322-353), state from (161-180, 399-417) */
    for (int e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_MOS) {
            const int32_t* nd = t->elem_nodes + 4 * e;
            int eqD = nd[0], eqG = nd[1], eqS = nd[2], eqB = nd[3];
            double vD = node_v(x_prev, eqD), vG = node_v(x_prev, eqG);
            double vS = node_v(x_prev, eqS), vB = node_v(x_prev, eqB);
            double Cj0 = params[t->elem_param[e] + 3];
            double Cgs = 0.5 * Cj0, Cgd = 0.5 * Cj0, CsJ = Cj0, CdJ = Cj0;
            stamp_cap_be(n, eqG, eqS, Cgs, dt, vG - vS, G, I);
            stamp_cap_be(n, eqG, eqD, Cgd, dt, vG - vD, G, I);
            stamp_cap_be(n, eqS, eqB, CsJ, dt, vS - vB, G, I);
            stamp_cap_be(n, eqD, eqB, CdJ, dt, vD - vB, G, I);
        }
    /* 6) gmin (356) */
    stamp_gmin(t, gmin, G);
}

/* Solver::luDecompose + solveLinearSystemLU, include/solver.hpp:30-131 */
/* Solver::luDecompose, include/solver.hpp:30-80: LU = packed factors (multipliers in the strict lower triangle,
 * rows exchanged physically), perm; returns 1 when the reference returns false (LU holds its partial state). */
int oracle_lu_factor(int32_t n, const double* A, double* LU, int32_t* perm, double lu_eps) {
    memcpy(LU, A, sizeof(double) * (size_t)n * (size_t)n);
    for (int i = 0; i < n; ++i) perm[i] = i;
    for (int k = 0; k < n; ++k) {
        int pivot = k;
        double maxAbs = fabs(LU[k * n + k]);
        for (int i = k + 1; i < n; ++i) {
            double val = fabs(LU[i * n + k]);
            if (val > maxAbs) {
                maxAbs = val;
                pivot = i;
            }
        }
        if (maxAbs < lu_eps) return 1;
        if (pivot != k) {
            for (int j = 0; j < n; ++j) {
                double tmp = LU[k * n + j];
                LU[k * n + j] = LU[pivot * n + j];
                LU[pivot * n + j] = tmp;
            }
            int32_t tp = perm[k];
            perm[k] = perm[pivot];
            perm[pivot] = tp;
        }
        for (int i = k + 1; i < n; ++i) {
            double factor = LU[i * n + k] / LU[k * n + k];
            LU[i * n + k] = factor;
            for (int j = k + 1; j < n; ++j) LU[i * n + j] -= factor * LU[k * n + j];
        }
    }
    return 0;
}

int oracle_lu_solve(int32_t n, const double* A, const double* b, double* x, int32_t* perm_out, double lu_eps) {
    double* LU = (double*)malloc(sizeof(double) * (size_t)n * (size_t)n);
    double* y = (double*)malloc(sizeof(double) * (size_t)n);
    int32_t* perm = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    memcpy(LU, A, sizeof(double) * (size_t)n * (size_t)n);
    for (int i = 0; i < n; ++i) {
        perm[i] = i;
        x[i] = 0.0;
    }
    int failed = 0;
    for (int k = 0; k < n; ++k) {
        int pivot = k;
        double maxAbs = fabs(LU[k * n + k]);
        for (int i = k + 1; i < n; ++i) {
            double val = fabs(LU[i * n + k]);
            if (val > maxAbs) {
                maxAbs = val;
                pivot = i;
            }
        }
        if (maxAbs < lu_eps) {
            failed = 1;
            break;
        }
        if (pivot != k) {
            for (int j = 0; j < n; ++j) {
                double tmp = LU[k * n + j];
                LU[k * n + j] = LU[pivot * n + j];
                LU[pivot * n + j] = tmp;
            }
            int32_t tp = perm[k];
            perm[k] = perm[pivot];
            perm[pivot] = tp;
        }
        for (int i = k + 1; i < n; ++i) {
            double factor = LU[i * n + k] / LU[k * n + k];
            LU[i * n + k] = factor;
            for (int j = k + 1; j < n; ++j) LU[i * n + j] -= factor * LU[k * n + j];
        }
    }
    if (perm_out) memcpy(perm_out, perm, sizeof(int32_t) * (size_t)n);
    if (!failed) {
        for (int i = 0; i < n; ++i) {
            double sum = b[perm[i]];
            for (int j = 0; j < i; ++j) sum -= LU[i * n + j] * y[j];
            y[i] = sum;
        }
        for (int i = n - 1; i >= 0; --i) {
            double sum = y[i];
            for (int j = i + 1; j < n; ++j) sum -= LU[i * n + j] * x[j];
            double diag = LU[i * n + i];
            if (fabs(diag) < 1e-15) x[i] = 0.0; /* solver.hpp:122-125, literal */
            else x[i] = sum / diag;
        }
    }
    free(LU);
    free(y);
    free(perm);
    return failed;
}

/* Solver::solveLinearSystemGaussSeidel, include/solver.hpp:139-192 */
int32_t oracle_gauss_seidel(int32_t n, const double* A, const double* b, double* x, int32_t max_iters, double tol) {
    if (n <= 0) return 0;
    double* xOld = (double*)malloc(sizeof(double) * (size_t)n);
    double* d = (double*)malloc(sizeof(double) * (size_t)n);
    const double diagEps = 1e-12; /* 160 */
    int32_t done = 0;
    for (int iter = 0; iter < max_iters; ++iter) { /* 162 */
        memcpy(xOld, x, sizeof(double) * (size_t)n);
        for (int i = 0; i < n; ++i) {
            double diag = A[i * n + i];
            if (fabs(diag) < diagEps) { /* 169-173 */
                double sign = (diag >= 0.0 ? 1.0 : -1.0);
                diag = sign * diagEps;
            }
            double sum = b[i];
            for (int j = 0; j < i; ++j) sum -= A[i * n + j] * x[j];        /* 178-180: newest values */
            for (int j = i + 1; j < n; ++j) sum -= A[i * n + j] * xOld[j]; /* 181-183: previous sweep */
            x[i] = sum / diag;
        }
        for (int i = 0; i < n; ++i) d[i] = x[i] - xOld[i];
        double err = oracle_norm2(d, n); /* 188 */
        done = iter + 1;
        if (err < tol) break;
    }
    free(xOld);
    free(d);
    return done;
}

static int all_finite(const double* v, int n) {
    for (int i = 0; i < n; ++i)
        if (!isfinite(v[i])) return 0;
    return 1;
}

static void margin(oracle_trace* tr, double err, double thr) {
    if (!tr || !(thr > 0.0) || !isfinite(thr) || !isfinite(err)) return;
    double m = fabs(err - thr) / thr;
    if (m < tr->min_margin) tr->min_margin = m;
}

static void trace_iter(oracle_trace* tr, int n, const double* xraw, const int32_t* perm) {
    if (!tr) return;
    if (tr->n_iters < tr->cap_iters) {
        if (tr->xraw) memcpy(tr->xraw + tr->n_iters * n, xraw, sizeof(double) * (size_t)n);
        if (tr->perm) memcpy(tr->perm + tr->n_iters * n, perm, sizeof(int32_t) * (size_t)n);
    }
    tr->n_iters++;
}
static void trace_err(oracle_trace* tr, double err) {
    if (!tr) return;
    if (tr->err && tr->n_err < tr->cap_iters) tr->err[tr->n_err] = err;
    tr->n_err++;
}

/* ConvController::baseGmin, include/dcanalysis.hpp:45-48 */
static double base_gmin(const csim_constants* c, double s) {
    if (s < 0.0) s = 0.0;
    if (s > 1.0) s = 1.0;
    return c->gmin_high_base * (1.0 - s) + c->gmin_low_base * s;
}

void oracle_dc(const csim_topology* t, const csim_constants* c, const double* params, double* x,
               int32_t* iters, uint64_t* pivot_hash, uint32_t* status, oracle_trace* tr) {
    const int n = t->n_unknowns;
    const int nsrc = oracle_count_sources(t);
    double* G = (double*)malloc(sizeof(double) * (size_t)n * (size_t)n);
    double* I = (double*)malloc(sizeof(double) * (size_t)n);
    double* xRaw = (double*)malloc(sizeof(double) * (size_t)n);
    double* xNew = (double*)malloc(sizeof(double) * (size_t)n);
    double* d = (double*)malloc(sizeof(double) * (size_t)n);
    double* sv = (double*)malloc(sizeof(double) * (size_t)(nsrc > 0 ? nsrc : 1));
    int32_t* perm = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    uint64_t h = pivot_hash ? *pivot_hash : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    for (int i = 0; i < n; ++i) x[i] = 0.0;
    if (iters)
        for (int s = 0; s < c->dc_ramp_steps; ++s) iters[s] = 0;

    int has_mos = 0; /* hasNonlinearDevices, src/dcanalysis.cpp:26-33 */
    for (int e = 0; e < t->n_elems; ++e)
        if (t->elem_type[e] == CSIM_ELEM_MOS) has_mos = 1;

    if (!has_mos) {
        /* dcSolveDirectLU, src/dcanalysis.cpp:46-68: one solve at scale 1, NO gmin */
        oracle_src_dc(t, params, 1.0, sv);
        /* no stampGlobalGmin call at all on this path */
        oracle_stamp(t, c, CSIM_MODE_DC, params, x, NULL, sv, 0.0, 0.0, G, I);
        if (oracle_lu_solve(n, G, I, xRaw, perm, c->lu_eps)) st |= CSIM_ST_LU_SINGULAR;
        h = oracle_pivot_hash_update(h, perm, n);
        trace_iter(tr, n, xRaw, perm);
        memcpy(x, xRaw, sizeof(double) * (size_t)n);
        if (iters) iters[0] = 1;
    } else {
        /* dcSolveNewtonLU, src/dcanalysis.cpp:95-163 */
        for (int step = 1; step <= c->dc_ramp_steps; ++step) {
            double scale = (double)step / c->dc_ramp_steps;
            double gmin = base_gmin(c, scale);
            double prevErr = INFINITY;
            int it_count = 0;
            for (int iter = 0; iter < c->dc_max_iters; ++iter) {
                ++it_count;
                oracle_src_dc(t, params, scale, sv);
                oracle_stamp(t, c, CSIM_MODE_DC, params, x, NULL, sv, 0.0, gmin, G, I);
                if (oracle_lu_solve(n, G, I, xRaw, perm, c->lu_eps)) st |= CSIM_ST_LU_SINGULAR;
                h = oracle_pivot_hash_update(h, perm, n);
                trace_iter(tr, n, xRaw, perm);
                if (!all_finite(xRaw, n)) {
                    st |= CSIM_ST_NONFINITE;
                    double g10 = gmin * c->dc_nonfinite_mul;
                    gmin = g10 < c->dc_nonfinite_cap ? g10 : c->dc_nonfinite_cap;
                    continue;
                }
                /* ConvController::update, src/dcanalysis.cpp:268-307 */
                double alpha = c->dc_alpha;
                for (int i = 0; i < n; ++i) {
                    double dd = xRaw[i] - x[i];
                    double ad = alpha * dd;
                    xNew[i] = x[i] + ad;
                }
                for (int i = 0; i < n; ++i) d[i] = xNew[i] - x[i];
                double err = oracle_norm2(d, n);
                trace_err(tr, err);
                double gminBase = base_gmin(c, scale);
                double gminNext = gminBase;
                if (iter == 0 || !isfinite(prevErr)) {
                    gminNext = gminBase;
                } else {
                    margin(tr, err, prevErr * c->slow_conv_ratio);
                    margin(tr, err, prevErr * c->fast_conv_ratio);
                    if (err > prevErr * c->slow_conv_ratio) {
                        double g2 = gmin * 2.0;
                        gminNext = g2 < c->gmin_abs_max ? g2 : c->gmin_abs_max;
                    } else if (err < prevErr * c->fast_conv_ratio) {
                        gminNext = 0.5 * gmin + 0.5 * gminBase;
                    } else {
                        gminNext = 0.7 * gmin + 0.3 * gminBase;
                    }
                }
                memcpy(x, xNew, sizeof(double) * (size_t)n);
                gmin = gminNext;
                prevErr = err;
                margin(tr, err, c->dc_tol);
                if (err < c->dc_tol) break;
                if (iter == c->dc_max_iters - 1) st |= CSIM_ST_DC_UNCONVERGED;
            }
            if (iters) iters[step - 1] = it_count;
        }
    }
    if (pivot_hash) *pivot_hash = h;
    if (status) *status |= st;
    free(G); free(I); free(xRaw); free(xNew); free(d); free(sv); free(perm);
}

int64_t oracle_tran(const csim_topology* t, const csim_constants* c, const double* params, double dt,
                    int32_t step_begin, int32_t step_end, const int32_t* src_mode,
                    const int32_t* src_group, const double* table, int32_t table_steps, double* x,
                    double* wave_all, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                    oracle_trace* tr) {
    const int n = t->n_unknowns;
    const int nsrc = oracle_count_sources(t);
    double* G = (double*)malloc(sizeof(double) * (size_t)n * (size_t)n);
    double* I = (double*)malloc(sizeof(double) * (size_t)n);
    double* xRaw = (double*)malloc(sizeof(double) * (size_t)n);
    double* xNew = (double*)malloc(sizeof(double) * (size_t)n);
    double* d = (double*)malloc(sizeof(double) * (size_t)n);
    double* xprev = (double*)malloc(sizeof(double) * (size_t)n);
    double* sv = (double*)malloc(sizeof(double) * (size_t)(nsrc > 0 ? nsrc : 1));
    int32_t* perm = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    uint64_t h = pivot_hash ? *pivot_hash : CSIM_PIVOT_HASH_INIT;
    uint32_t st = 0;
    int64_t total = 0;
    int dead = 0;

    for (int32_t step = step_begin; step < step_end && !dead; ++step) {
        /* companion state = functions of the last accepted x (src/tanalisis.cpp:139-180, 381-417) */
        memcpy(xprev, x, sizeof(double) * (size_t)n);
        oracle_src_tran(t, params, src_mode, src_group, table, table_steps, step, dt, sv);
        int it_count = 0;
        for (int iter = 0; iter < c->tran_max_iters; ++iter) { /* 258 */
            ++it_count;
            oracle_stamp(t, c, CSIM_MODE_TRAN, params, x, xprev, sv, dt, c->tran_gmin, G, I);
            if (oracle_lu_solve(n, G, I, xRaw, perm, c->lu_eps)) st |= CSIM_ST_LU_SINGULAR;
            h = oracle_pivot_hash_update(h, perm, n);
            trace_iter(tr, n, xRaw, perm);
            if (!all_finite(xRaw, n)) { /* 360-362: the reference throws; the instance stops here */
                st |= CSIM_ST_NONFINITE;
                dead = 1;
                break;
            }
            for (int i = 0; i < n; ++i) { /* 365 */
                double dd = xRaw[i] - x[i];
                double ad = c->tran_alpha * dd;
                xNew[i] = x[i] + ad;
            }
            for (int i = 0; i < n; ++i) d[i] = xNew[i] - x[i]; /* 366 */
            double err = oracle_norm2(d, n);
            trace_err(tr, err);
            memcpy(x, xNew, sizeof(double) * (size_t)n);
            margin(tr, err, c->tran_tol);
            if (err < c->tran_tol) break;
            if (iter == c->tran_max_iters - 1) st |= CSIM_ST_TRAN_UNCONVERGED;
        }
        total += it_count;
        if (iters_step) iters_step[step - step_begin] = (uint8_t)it_count;
        if (dead) break; /* the reference never reaches its state update / dumpRow for this step */
        if (wave_all) memcpy(wave_all + (int64_t)(step - step_begin) * n, x, sizeof(double) * (size_t)n);
    }
    if (pivot_hash) *pivot_hash = h;
    if (status) *status |= st;
    free(G); free(I); free(xRaw); free(xNew); free(d); free(xprev); free(sv); free(perm);
    return total;
}
