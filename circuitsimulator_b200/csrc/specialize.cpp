// specialize.cpp — generator of the specialised transient kernel (see specialize.hpp).
#include "specialize.hpp"
#include "spec_analysis.hpp"

#include <algorithm>
#include <map>
#include <sstream>
#include <cstdio>
#include <stdexcept>


namespace csim {

std::string spec_unsupported_reason(const HostPlan& hp) {
    const ModeProgram& mp = hp.mode[CSIM_MODE_TRAN];
    if (hp.N > 32) return "N > 32";
    if (hp.n_mos > 32) return "more than 32 MOSFETs";
    if (2 * (mp.n_values + 1) > 65535) return "value table too large for 16-bit indices";
    return "";
}

Analysis analyse(const HostPlan& hp, const std::vector<int32_t>& perm, int mode) {
    const ModeProgram& mp = hp.mode[mode];
    Analysis A;
    const int N = A.N = hp.N;
    if ((int)perm.size() != N) throw std::runtime_error("specialize: permutation has the wrong length");
    A.pos_of_row.assign(N, -1);
    for (int i = 0; i < N; ++i) {
        if (perm[i] < 0 || perm[i] >= N || A.pos_of_row[perm[i]] != -1)
            throw std::runtime_error("specialize: not a permutation");
        A.pos_of_row[perm[i]] = i;
    }
    std::vector<bool> dynamic(mp.n_values, false);  // values that change with the Newton iterate
    for (const MosOp& m : mp.mos_ops)
        for (int t = 0; t < 4; ++t) dynamic[m.out + t] = true;
    if (mode == CSIM_MODE_DC)
        for (const StaticOp& so : mp.static_ops)
            if (so.kind == SK_GMIN) dynamic[so.out] = true;

    A.slot.assign(N, std::vector<std::vector<Term>>(N + 1));
    for (const Contribution& c : mp.contribs) {
        const int i = A.pos_of_row[c.row];
        const int j = c.col >= 0 ? c.col : N;
        A.slot[i][j].push_back({c.value, c.sub != 0});
    }
    A.rest = A.prefix = A.slot;
    A.K.assign(N + 1, 0);
    A.pat.assign(N, std::vector<bool>(N, false));
    for (int i = 0; i < N; ++i)
        for (int j = 0; j <= N; ++j) {
            const auto& all = A.slot[i][j];
            size_t first_dyn = all.size();
            if (j == N) first_dyn = 0;  // the right-hand side is rebuilt from zero every iteration
            else
                for (size_t t = 0; t < all.size(); ++t)
                    if (dynamic[all[t].value]) {
                        first_dyn = t;
                        break;
                    }
            A.prefix[i][j].assign(all.begin(), all.begin() + first_dyn);
            A.rest[i][j].assign(all.begin() + first_dyn, all.end());
            A.K[j] = std::max<int>(A.K[j], (int)A.rest[i][j].size());
            if (j < N && !all.empty()) A.pat[i][j] = true;
        }
    // symbolic elimination in the predicted order
    A.rows_k.assign(N, {});
    A.cols_k.assign(N, {});
    for (int k = 0; k < N; ++k) {
        for (int i = k + 1; i < N; ++i)
            if (A.pat[i][k]) A.rows_k[k].push_back(i);
        for (int j = k + 1; j < N; ++j)
            if (A.pat[k][j]) A.cols_k[k].push_back(j);
        for (int i : A.rows_k[k])
            for (int j : A.cols_k[k]) A.pat[i][j] = true;
    }
    // the reference's row order at every step: rows k and p are exchanged physically when position p wins (solver.hpp:64-67)
    A.before_pivot.assign(N, std::vector<bool>(N, false));
    {
        std::vector<int> cur(N);
        for (int i = 0; i < N; ++i) cur[i] = i;
        for (int k = 0; k < N; ++k) {
            int p = k;
            while (cur[p] != perm[k]) ++p;
            for (int q = k; q < p; ++q) A.before_pivot[k][A.pos_of_row[cur[q]]] = true;
            std::swap(cur[k], cur[p]);
        }
    }
    A.ucols.assign(N, {});
    A.col_used_above.assign(N, false);
    for (int i = 0; i < N; ++i)
        for (int j = i + 1; j < N; ++j)
            if (A.pat[i][j]) {
                A.ucols[i].push_back(j);
                A.col_used_above[j] = true;
            }
    return A;
}

namespace {
// rows the reference scans before the predicted pivot of step k win a tie (Analysis::before_pivot): lane mask
std::string tie_clause(const Analysis& A, int k) {
    unsigned mask = 0;
    for (int i = k + 1; i < A.N; ++i)
        if (A.before_pivot[k][i]) mask |= 1u << i;
    if (!mask) return "";
    char buf[96];
    std::snprintf(buf, sizeof buf, " || (((0x%xu >> lane) & 1u) && fabs(a%d) == fabs(pk))", mask, k);
    return buf;
}
// must match warp_mem_doubles() in newton_warp.cuh and the carve-up in the generated kernel
size_t warp_doubles_generic(int n_values) { return 32 * 33 + 3 * 32 + ((n_values + 3) & ~3); }
size_t warp_doubles_spec(int n_values) { return warp_doubles_generic(n_values) + ((2 * (n_values + 1) + 1) & ~1); }
}  // namespace

size_t spec_smem_bytes(const HostPlan& hp, const SpecHost& sh) {
    const ModeProgram& mp = hp.mode[CSIM_MODE_TRAN];
    size_t bytes = warp_doubles_spec(mp.n_values) * sizeof(double) * sh.warps_per_block;
    bytes += (size_t)mp.row_len_max * 32 * sizeof(uint32_t);  // generic row program (fallback path)
    bytes += (size_t)sh.n_adds * 32 * sizeof(uint16_t);       // per-lane value indices of the generated adds
    return (bytes + 15) & ~(size_t)15;
}

SpecHost generate_tran_spec(const HostPlan& hp, const std::vector<int32_t>& perm) {
    const std::string why = spec_unsupported_reason(hp);
    if (!why.empty()) throw std::runtime_error("specialize: " + why);
    const ModeProgram& mp = hp.mode[CSIM_MODE_TRAN];
    const Analysis A = analyse(hp, perm);
    const int N = A.N;
    SpecHost S;
    S.perm = perm;

    // ---- pivot-hash contribution of the predicted permutation (definition: include/csim_b200.h)
    for (int i = 0; i < N; ++i) {
        uint32_t key = (uint32_t)i * 1024u + (uint32_t)perm[i] + 1u;
        uint32_t lo = key * 0x9E3779B1u;
        lo ^= lo >> 15;
        uint32_t hi = key * 0x85EBCA77u;
        hi ^= hi >> 13;
        S.hash_lo += lo;
        S.hash_hi += hi;
    }

    // ---- per-lane MOSFET table
    S.n_mos = (int)mp.mos_ops.size();
    S.mos_lane.assign(32 * 8, 0);
    for (int m = 0; m < S.n_mos; ++m) {
        const MosOp& op = mp.mos_ops[m];
        int32_t* r = &S.mos_lane[m * 8];
        r[0] = op.param;
        r[1] = op.d;
        r[2] = op.g;
        r[3] = op.s;
        r[4] = op.is_p;
        r[5] = op.out;
    }

    // ---- static prefix program, per lane (position), reference order within a slot
    std::vector<std::vector<uint32_t>> pre(32);
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j)
            for (const Term& t : A.prefix[i][j]) pre[i].push_back(contrib_word(j, t.sub, t.value));
    for (auto& v : pre) S.prefix_len = std::max<int>(S.prefix_len, (int)v.size());
    S.prefix.assign((size_t)S.prefix_len * 32, 0xFFFFFFFFu);
    for (int i = 0; i < 32; ++i)
        for (size_t c = 0; c < pre[i].size(); ++c) S.prefix[c * 32 + i] = pre[i][c];

    // ---- generated adds: for column j, K[j] rounds; lane i adds its t-th remaining term or +0.0
    const uint16_t zero_idx = (uint16_t)(2 * mp.n_values);
    std::vector<int> add_col;  // column of every generated add (N = RHS)
    for (int j = 0; j <= N; ++j)
        for (int t = 0; t < A.K[j]; ++t) {
            add_col.push_back(j);
            for (int i = 0; i < 32; ++i) {
                uint16_t idx = zero_idx;
                if (i < N && t < (int)A.rest[i][j].size())
                    idx = (uint16_t)(2 * A.rest[i][j][t].value + (A.rest[i][j][t].sub ? 1 : 0));
                S.gidx.push_back(idx);
            }
        }
    S.n_adds = (int)add_col.size();
    for (int c : add_col) (c == N ? S.n_rhs_adds : S.n_g_adds)++;

    // ---- source text ---------------------------------------------------------------------------
    std::ostringstream o;
    o << "// generated by specialize.cpp — transient kernel specialised to one topology and one predicted\n"
         "// pivot permutation; rows in registers, structural non-zeros only, prediction verified per solve.\n"
         "#include \"spec_runtime.cuh\"\n"
         "using namespace csim;\n"
         "#define SPEC_N " << N << "\n"
         "#define SPEC_WARPS " << S.warps_per_block << "\n"
         "extern \"C\" __global__ void __launch_bounds__(SPEC_WARPS * 32, " << S.blocks_per_sm << ")\n"
         "csim_tran_spec(DevPlan P, DevTran T, SpecTables ST, long long B, const double* __restrict__ params,\n"
         "               double* __restrict__ xio, double* __restrict__ wave, int* __restrict__ iters_total,\n"
         "               unsigned char* __restrict__ iters_step, unsigned long long* __restrict__ pivot_hash,\n"
         "               unsigned int* __restrict__ status, unsigned long long* __restrict__ stats) {\n"
         "    extern __shared__ double smem[];\n"
         "    const DevMode& m = P.m[CSIM_MODE_TRAN];\n"
         "    const csim_constants& c = P.c;\n"
         "    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;\n"
         "    constexpr int N = SPEC_N;\n"
         "    const size_t wdg = warp_mem_doubles(m.n_values);\n"
         "    const size_t wd = wdg + ((2 * (m.n_values + 1) + 1) & ~1);\n"
         "    uint32_t* progbuf = reinterpret_cast<uint32_t*>(smem + wd * SPEC_WARPS);\n"
         "    uint16_t* gidx = reinterpret_cast<uint16_t*>(progbuf + m.row_len * 32);\n"
         "    for (int i = threadIdx.x; i < m.row_len * 32; i += blockDim.x) progbuf[i] = m.row_prog[i];\n"
         "    for (int i = threadIdx.x; i < ST.n_adds * 32; i += blockDim.x) gidx[i] = ST.gidx[i];\n"
         "    __syncthreads();\n"
         "    const uint32_t* prog = progbuf;\n"
         "    const long long b = (long long)blockIdx.x * SPEC_WARPS + wib;\n"
         "    if (b >= B || (T.skip && T.skip[b])) return;\n"
         "    WarpMem w = carve(smem + wd * wib, m.n_values);\n"
         "    double* vv = smem + wd * wib + wdg;\n"
         "    double* base = w.S;  // the generic matrix area doubles as the static-part table\n"
         "    unsigned long long h = pivot_hash ? pivot_hash[b] : CSIM_PIVOT_HASH_INIT;\n"
         "    const unsigned long long F = ((unsigned long long)ST.hash_hi << 32) | ST.hash_lo;\n"
         "    unsigned int st = 0, n_fallback = 0;\n"
         "    int total = 0;\n"
         "    double x = lane < N ? xio[(long long)lane * B + b] : 0.0;\n"
         "    w.xs[lane] = x;\n"
         "    eval_static(m, params, B, b, T.dt, c.tran_gmin, w.vt, lane);\n"
         "    __syncwarp();\n"
         "    build_base(ST, w.vt, base, lane);\n"
         "    const MosLane ml = load_mos_lane(ST, params, B, b, lane);\n"
         "    const int r0 = T.out_begin_step / T.out_stride;\n"
         "    const int probe = lane < T.n_probes ? T.probes[lane] : 0;\n"
         "    for (int step = T.step_begin; step < T.step_end; ++step) {\n"
         "        w.xp[lane] = x;\n"
         "        __syncwarp();\n"
         "        eval_step(m, w.xp, w.vt, lane);\n"
         "        eval_src_tran(m, T, params, B, b, step, w.vt, lane);\n"
         "        __syncwarp();\n"
         "        mirror_values(w.vt, vv, m.n_values, lane);\n"
         "        int count = 0;\n"
         "        bool dead = false;\n"
         "        for (int iter = 0; iter < c.tran_max_iters; ++iter) {\n"
         "            ++count;\n"
         "            double xr = 0.0;\n"
         "            spec_mos_eval(ml, x, c.mos_off_gds, vv);\n"
         "            __syncwarp();\n";
    // registers <- static part
    for (int j = 0; j < N; ++j) o << "            double a" << j << " = base[" << j << " * 32 + lane];\n";
    o << "            double rb = 0.0;\n";
    for (int c = 0; c < S.n_adds; ++c) {
        const int j = add_col[c];
        const std::string reg = j == N ? std::string("rb") : "a" + std::to_string(j);
        o << "            " << reg << " = d_add(" << reg << ", vv[gidx[" << c << " * 32 + lane]]);\n";
    }
    // elimination with fused forward substitution (include/solver.hpp:46-76, 106-113)
    o << "            bool bad = false;\n"
         "            double dg = 1.0, rdg = 1.0;\n";
    for (int k = 0; k < N; ++k) {
        o << "            {  // step " << k << "\n"
          << "                const double pk = bcast(a" << k << ", " << k << ");\n"
          << "                bad = bad || (lane > " << k << " && (fabs(a" << k << ") > fabs(pk)" << tie_clause(A, k)
          << ")) || (fabs(pk) < c.lu_eps);\n"
          << "                const double rk = rcp_seq(pk);\n"
          << "                if (lane == " << k << ") { dg = pk; rdg = rk; }\n";
        if (!A.rows_k[k].empty()) {
            S.n_lu_steps++;
            o << "                const bool div = lane > " << k << " && a" << k << " != 0.0;\n"
              << "                double f = fdiv(a" << k << ", pk, rk, div);\n"
              << "                f = div ? f : 0.0;\n";
            for (int j : A.cols_k[k]) {
                o << "                a" << j << " = d_sub(a" << j << ", d_mul(f, bcast(a" << j << ", " << k << ")));\n";
                S.n_elim_updates++;
            }
            o << "                rb = d_sub(rb, d_mul(f, bcast(rb, " << k << ")));\n";
        }
        o << "            }\n";
    }
    o << "            if (__any_sync(kFull, bad)) {\n"
         "                // the predicted pivot order does not hold for this matrix: generic partial pivoting\n"
         "                ++n_fallback;\n"
         "                w.xs[lane] = x;\n"
         "                __syncwarp();\n"
         "                eval_mos(m, params, B, b, c.mos_off_gds, w.xs, w.vt, lane);\n"
         "                __syncwarp();\n"
         "                assemble(m, prog, w, N, lane);\n"
         "                int perm;\n"
         "                if (lu_solve_warp(w.S, w.bs, N, lane, c.lu_eps, perm)) st |= CSIM_ST_LU_SINGULAR;\n"
         "                h = hash_update(h, perm, N, lane);\n"
         "                xr = lane < N ? w.bs[lane] : 0.0;\n"
         "                __syncwarp();\n"
         "                build_base(ST, w.vt, base, lane);\n"
         "            } else {\n"
         "                h = (h ^ F) * 0x100000001B3ull;\n";
    // back substitution (include/solver.hpp:116-128): rows bottom-up, columns ascending
    for (int i = N - 1; i >= 0; --i) {
        o << "                {  // row " << i << "\n"
          << "                    double s = rb;\n";
        for (int j : A.ucols[i]) {
            o << "                    s = d_sub(s, a" << j << ");\n";
            S.n_back_terms++;
        }
        o << "                    double q = fdiv(s, dg, rdg, lane == " << i << " && s != 0.0);\n"
          << "                    if (s == 0.0) q = (dg < 0.0) ? -s : s;\n"
          << "                    if (fabs(dg) < 1e-15) q = 0.0;\n"
          << "                    if (lane == " << i << ") xr = q;\n";
        if (A.col_used_above[i])
            o << "                    a" << i << " = d_mul(a" << i << ", bcast(q, " << i << "));\n";
        o << "                }\n";
    }
    o << "            }\n"
         "            if (__any_sync(kFull, !lane_finite(xr))) {  // tanalisis.cpp:360-362\n"
         "                st |= CSIM_ST_NONFINITE;\n"
         "                dead = true;\n"
         "                break;\n"
         "            }\n"
         "            const double xn = d_add(x, d_mul(c.tran_alpha, d_sub(xr, x)));  // tanalisis.cpp:365\n"
         "            const double err = warp_norm2(d_sub(xn, x));                   // tanalisis.cpp:366\n"
         "            x = xn;\n"
         "            if (err < c.tran_tol) break;\n"
         "            if (iter == c.tran_max_iters - 1) st |= CSIM_ST_TRAN_UNCONVERGED;\n"
         "        }\n"
         "        total += count;\n"
         "        if (iters_step && lane == 0) iters_step[(long long)step * B + b] = (unsigned char)count;\n"
         "        if (dead) break;\n"
         "        if ((step + 1) % T.out_stride == 0 && step >= T.out_begin_step) {\n"
         "            const double pv = bcast(x, probe);\n"
         "            if (lane < T.n_probes) {\n"
         "                const long long r = (step + 1) / T.out_stride - 1 - r0;\n"
         "                wave[((long long)b * T.n_probes + lane) * T.wave_ld + r] = pv;\n"
         "            }\n"
         "        }\n"
         "    }\n"
         "    if (lane < N) xio[(long long)lane * B + b] = x;\n"
         "    if (lane == 0) {\n"
         "        if (iters_total) iters_total[b] += total;\n"
         "        if (pivot_hash) pivot_hash[b] = h;\n"
         "        if (status) status[b] |= st;\n"
         "        if (stats) {\n"
         "            atomicAdd(stats, (unsigned long long)n_fallback);\n"
         "            atomicAdd(stats + 1, (unsigned long long)total);\n"
         "        }\n"
         "    }\n"
         "}\n";
    S.source = o.str();
    return S;
}

}  // namespace csim
