// specialize_tpi.cpp — generator of the THREAD-per-instance specialised transient kernel
// (design: tpi_runtime.cuh; host interface: specialize.hpp).
//
// Row-wise ("up-looking") LU under the predicted permutation: row i is assembled into registers,
// receives the updates of every earlier pivot row k with a structural non-zero at (i,k) in
// ascending k — per entry exactly the update sequence of the reference's right-looking loop
// (include/solver.hpp:70-76), so every intermediate is bit-identical — and then parks its U part,
// pivot, pivot reciprocal and forward-substituted right-hand side in shared memory.  Multipliers
// are never stored (forward substitution is fused, solver.hpp:106-113).
//
// Constant propagation: entries whose value is the same compile-time constant for every instance
// (the +-1 incidence entries of voltage-source and inductor branch rows, gmin) are tracked
// symbolically through assembly and elimination.  Only IEEE-exact rewrites are applied
// (x*1 = x, x*(-1) = -x, x/1 = x, x/(-1) = -x, x-(-y) = x+y, constant folding in host binary64,
// which rounds like the device); sums with +0 are kept because they normalise -0.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>

#include "spec_analysis.hpp"
#include "specialize.hpp"

namespace csim {

namespace {

constexpr size_t kMaxSmem = 232448;    // 227 KB dynamic shared memory per CTA on sm_100
constexpr size_t kSmemPerSm = 233472;  // 228 KB per SM, of which every resident CTA reserves 1 KB
constexpr size_t kCtaReserve = 1024;

// Warps never talk to each other (no CTA barrier), so the CTA size is free: pick the one that packs the
// most warps into an SM's shared memory.  dbmixer: 226 doubles x 32 lanes = 57,856 B per warp — three
// 1-warp CTAs (1 KB reserve each) but FOUR warps as two 2-warp CTAs (2 x 116,736 B = 228 KB exactly),
// i.e. one warp per SM sub-partition.
int pick_warps_per_cta(size_t warp_bytes) {
    int best_w = 1;
    size_t best = 0;
    for (int w : {1, 2, 4}) {
        if (warp_bytes * w > kMaxSmem) break;
        const size_t warps = std::min<size_t>(32, kSmemPerSm / (warp_bytes * w + kCtaReserve)) * w;
        if (warps > best) {
            best = warps;
            best_w = w;
        }
    }
    return best_w;
}

// Generated text is written for ONE instance with placeholders — identifier suffix '@' and replica
// number '@R' — and then replicated (replicate() below) for the R instances a thread advances together.
std::string S(int slot) { return "SM(" + std::to_string(slot) + ", @R)"; }

// Lines containing '@' are emitted once per replica, back to back (so the instruction streams of the
// replicas interleave at source level); regions between //@BEGIN and //@END lines are replicated as a
// whole; ALL(x@) / ANY(x@) expand to the conjunction / disjunction over replicas.
std::string replicate(const std::string& text, int R) {
    auto subst = [](std::string l, int r) {
        const std::string num = std::to_string(r), suf = "_" + std::to_string(r);
        size_t p;
        while ((p = l.find("@R")) != std::string::npos) l.replace(p, 2, num);
        while ((p = l.find('@')) != std::string::npos) l.replace(p, 1, suf);
        return l;
    };
    auto expand = [&](std::string l, const std::string& key, const std::string& op) {
        size_t p;
        while ((p = l.find(key + "(")) != std::string::npos) {
            const size_t q = l.find(')', p);
            const std::string inner = l.substr(p + key.size() + 1, q - p - key.size() - 1);
            std::string rep = "(";
            for (int r = 0; r < R; ++r) rep += (r ? op : "") + subst(inner, r);
            l.replace(p, q - p + 1, rep + ")");
        }
        return l;
    };
    std::istringstream in(text);
    std::ostringstream out;
    std::string line, region;
    bool in_region = false;
    while (std::getline(in, line)) {
        if (line.find("//@BEGIN") != std::string::npos) {
            in_region = true;
            region.clear();
            continue;
        }
        if (line.find("//@END") != std::string::npos) {
            in_region = false;
            for (int r = 0; r < R; ++r) {
                std::istringstream rin(region);
                std::string l;
                while (std::getline(rin, l)) out << subst(l, r) << "\n";
            }
            continue;
        }
        if (in_region) {
            region += line + "\n";
            continue;
        }
        if (line.find("ALL(") != std::string::npos || line.find("ANY(") != std::string::npos) {
            out << expand(expand(line, "ALL", " && "), "ANY", " || ") << "\n";
            continue;
        }
        if (line.find('@') == std::string::npos) {
            out << line << "\n";
            continue;
        }
        for (int r = 0; r < R; ++r) out << subst(line, r) << "\n";
    }
    return out.str();
}

std::string lit(double v) {  // exact literal
    if (v == 0.0) return std::signbit(v) ? "(-0.0)" : "0.0";
    if (v == 1.0) return "1.0";
    if (v == -1.0) return "(-1.0)";
    char buf[64];
    std::snprintf(buf, sizeof buf, "%a", v);
    return std::string("(") + buf + ")";
}

// a value of the generated program: a compile-time constant or an expression (variable name,
// shared-memory slot or nested call)
struct Val {
    bool c = false;
    double v = 0.0;
    std::string e;
    int t = 0;  // dependency-chain estimate: cycle at which the value is available (kLat* below)
    bool z = false;  // may routinely be an EXACT zero (a MOSFET's gm while it is off, and what is built from it)
    static Val K(double x) {
        Val r;
        r.c = true;
        r.v = x;
        r.e = lit(x);
        r.z = x == 0.0;
        return r;
    }
    static Val E(const std::string& s, int t = 0, bool z = false) {
        Val r;
        r.e = s;
        r.t = t;
        r.z = z;
        return r;
    }
};

// latencies of the dependency-chain estimate (profiles/r1_microbench_fp64.txt): one dependent FP64
// operation, the three-operation quotient of tdiv plus its select, rcp_seq (MUFU + five dependent
// DFMA), one MOSFET linearisation from the iterate, a shared-memory store -> load round trip
constexpr int kLatOp = 8, kLatDiv = 28, kLatRcp = 64, kLatMos = 96, kLatSmem = 35, kLatSqrt = 92;

Val neg(const Val& a) {
    if (a.c) return Val::K(-a.v);
    if (a.e.size() > 3 && a.e.compare(0, 2, "(-") == 0 && a.e.back() == ')' && a.e.find('(', 1) == std::string::npos)
        return Val::E(a.e.substr(2, a.e.size() - 3), a.t, a.z);
    return Val::E("(-" + a.e + ")", a.t, a.z);
}
Val add(const Val& a, const Val& b) {
    if (a.c && b.c) return Val::K(a.v + b.v);
    return Val::E("d_add(" + a.e + ", " + b.e + ")", std::max(a.t, b.t) + kLatOp, a.z && b.z);
}
Val sub(const Val& a, const Val& b) {
    if (a.c && b.c) return Val::K(a.v - b.v);
    return Val::E("d_sub(" + a.e + ", " + b.e + ")", std::max(a.t, b.t) + kLatOp, a.z && b.z);
}
Val mul(const Val& a, const Val& b) {
    if (a.c && b.c) return Val::K(a.v * b.v);
    if (a.c && a.v == 1.0) return b;
    if (b.c && b.v == 1.0) return a;
    if (a.c && a.v == -1.0) return neg(b);
    if (b.c && b.v == -1.0) return neg(a);
    return Val::E("d_mul(" + a.e + ", " + b.e + ")", std::max(a.t, b.t) + kLatOp, a.z || b.z);
}

// ---- warp pair (TpiOptions::pair): which of the two warps computes which matrix row, and in which order ------------
// Cost model in cycles of one warp alone on its sub-partition, fitted to the static stall sums of the generated
// dbmixer kernel per row (tools/sass_phases.py ... rows): assembly 5 + 3 per term, one elimination step 14 + 4 per
// updated entry (right-hand side included), parking 20 + 3 per U entry, one MOSFET linearisation 180; a hand-over
// between the warps (bar.arrive -> released bar.sync) 40.
constexpr int kPairMos = 180;
int kPairSync = 40;       // (development aid: CSIM_PAIR_SYNC / CSIM_PAIR_EVENTS override these two)
// named barriers of a CTA: 0 unused, 1 = iteration start, 2 = helper done, 3 = next iterate at the MOSFET terminals,
// 4..15 = hand-overs
int kPairMaxEvents = 12;
int kPairDefer = 0, kPairMargin = 0;  // (development aids: CSIM_PAIR_DEFER, CSIM_PAIR_MARGIN)

struct PairModel {
    int N = 0, W = 1;  // W: rows in flight per warp
    std::vector<int> asm_cost, park_cost;
    std::vector<std::vector<std::pair<int, int>>> steps;  // per row: (k, cost) ascending k
    std::vector<std::vector<int>> mos;                    // per row: MOSFETs whose linearisation its assembly reads
    int n_mos = 0;
};
struct PairAct {
    enum Kind { ASM, STEP, PARK, WAIT } kind;
    int role, row, k;  // WAIT: row = the producing warp, k = its park position; PARK: k = park position of the row
};
struct PairSchedule {
    std::vector<PairAct> acts;  // in an order in which every needed row is parked before its use
    std::set<std::pair<int, int>> handovers;  // (producing warp, park position) followed by an arrive
    int t[2] = {0, 0};
    int makespan() const { return std::max(t[0], t[1] + kPairSync); }
};
// List scheduling of the two instruction streams: the warp whose modelled clock is behind emits its next segment
// (its open rows round-robin, as the single-warp generator interleaves them).  A step that needs a row of the other
// warp waits for a hand-over — the latest one the other warp has reached by then, so later needs are covered too.
// Every wait points backwards in the emitted order and each stream is a subsequence of it: no circular wait.
PairSchedule pair_schedule(const PairModel& M, const std::vector<int>& owner) {
    PairSchedule S;
    std::vector<char> need[2] = {std::vector<char>(M.n_mos, 0), std::vector<char>(M.n_mos, 0)};
    for (int i = 0; i < M.N; ++i)
        for (int m : M.mos[i]) need[owner[i]][m] = 1;
    for (int r = 0; r < 2; ++r)
        for (char c : need[r]) S.t[r] += c ? kPairMos : 0;
    struct Open {
        int row;
        size_t next = 0;
        bool assembled = false;
    };
    std::vector<Open> open[2];
    std::vector<int> queue[2];
    size_t qnext[2] = {0, 0}, cursor[2] = {0, 0};
    for (int i = 0; i < M.N; ++i) queue[owner[i]].push_back(i);
    std::vector<int> T(M.N, -1), pos(M.N, -1);
    std::vector<int> Tpos[2];
    int waited[2] = {0, 0};
    auto pending = [&](int r) { return qnext[r] < queue[r].size() || !open[r].empty(); };
    while (pending(0) || pending(1)) {
        bool done = false;
        const int first = (pending(0) && (!pending(1) || S.t[0] <= S.t[1])) ? 0 : 1;
        for (int pass = 0; pass < 2 && !done; ++pass) {
            const int r = pass ? 1 - first : first;
            if (!pending(r)) continue;
            std::vector<Open>& op = open[r];
            while (op.size() < (size_t)M.W && qnext[r] < queue[r].size()) op.push_back(Open{queue[r][qnext[r]++]});
            // first a row that can go on without a new hand-over (the wait would also hold up the other rows in flight)
            for (int avoid_wait = kPairDefer ? 1 : 0; avoid_wait >= 0 && !done; --avoid_wait)
            for (size_t n = 0; n < op.size() && !done; ++n) {
                const size_t idx = (cursor[r] + n) % op.size();
                Open& R = op[idx];
                if (avoid_wait && R.assembled && R.next < M.steps[R.row].size()) {
                    const int k = M.steps[R.row][R.next].first;
                    if (T[k] >= 0 && owner[k] != r && pos[k] >= waited[r]) continue;
                }
                if (!R.assembled) {
                    S.acts.push_back({PairAct::ASM, r, R.row, -1});
                    S.t[r] += M.asm_cost[R.row];
                    R.assembled = true;
                    cursor[r] = idx + 1;
                    done = true;
                } else if (R.next < M.steps[R.row].size()) {
                    const int k = M.steps[R.row][R.next].first;
                    if (T[k] < 0) continue;  // not parked yet (a lower row, still open in one of the streams)
                    if (owner[k] != r && pos[k] >= waited[r]) {
                        const int o = 1 - r, ready = std::max(S.t[r], T[k] + kPairSync);
                        int q = pos[k];
                        while (q + 1 < (int)Tpos[o].size() && Tpos[o][q + 1] + kPairSync + kPairMargin <= ready) ++q;
                        waited[r] = q + 1;
                        S.handovers.insert({o, q});
                        S.acts.push_back({PairAct::WAIT, r, o, q});
                        S.t[r] = ready;
                    }
                    S.acts.push_back({PairAct::STEP, r, R.row, k});
                    S.t[r] += M.steps[R.row][R.next].second;
                    ++R.next;
                    cursor[r] = idx + 1;
                    done = true;
                } else {
                    S.t[r] += M.park_cost[R.row];
                    T[R.row] = S.t[r];
                    pos[R.row] = (int)Tpos[r].size();
                    Tpos[r].push_back(S.t[r]);
                    S.acts.push_back({PairAct::PARK, r, R.row, pos[R.row]});
                    op.erase(op.begin() + (long)idx);
                    cursor[r] = idx;
                    done = true;
                }
            }
        }
        if (!done) throw std::runtime_error("specialize: pair schedule made no progress");
    }
    return S;
}
// hill climbing from "everything on the master" over moves of one row or one contiguous run of rows
std::vector<int> pair_partition(const PairModel& M) {
    std::vector<int> owner(M.N, 0);
    auto score = [&](const std::vector<int>& ow) {
        const PairSchedule s = pair_schedule(M, ow);
        return (int)s.handovers.size() > kPairMaxEvents ? 1 << 30 : s.makespan();
    };
    int cur = score(owner);
    for (int round = 0; round < 2 * M.N; ++round) {
        int best = cur, bi = -1, bj = -1;
        for (int i = 0; i < M.N; ++i)
            for (int j = i; j < M.N; ++j) {
                for (int t = i; t <= j; ++t) owner[t] ^= 1;
                const int sc = score(owner);
                if (sc < best) best = sc, bi = i, bj = j;
                for (int t = i; t <= j; ++t) owner[t] ^= 1;
            }
        if (bi < 0) break;
        for (int t = bi; t <= bj; ++t) owner[t] ^= 1;
        cur = best;
    }
    return owner;
}

}  // namespace

TpiHost generate_tran_tpi(const HostPlan& hp, const std::vector<int32_t>& perm, const TpiOptions& opt) {
    return generate_tran_tpi(hp, std::vector<std::vector<int32_t>>{perm}, opt);
}

// Several predicted permutations ("variants"): the Newton body is emitted once per variant and every lane runs
// the variant that held for its last solve (int var).  A lane whose variant misses redoes the solve generically
// — which yields the permutation the reference actually took — and moves to the variant with that
// permutation if there is one.  Circuits whose pivot order depends on the operating region (buffer.sp while
// its stages switch) stay on straight-line code that way; lanes of one warp on different variants execute
// the bodies one after the other.
namespace {
TpiHost generate_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms, const TpiOptions& opt, int mode);
}
TpiHost generate_tran_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms, const TpiOptions& opt) {
    return generate_tpi(hp, perms, opt, CSIM_MODE_TRAN);
}
TpiHost generate_dc_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms, const TpiOptions& opt) {
    if (!hp.has_mos) throw std::runtime_error("specialize: linear circuits take the direct DC solve");
    return generate_tpi(hp, perms, opt, CSIM_MODE_DC);
}

namespace {
TpiHost generate_tpi(const HostPlan& hp, const std::vector<std::vector<int32_t>>& perms, const TpiOptions& opt, int mode) {
    const bool DC = mode == CSIM_MODE_DC;
    const std::string why = spec_unsupported_reason(hp);
    if (!why.empty()) throw std::runtime_error("specialize: " + why);
    if (perms.empty()) throw std::runtime_error("specialize: no permutation");
    const ModeProgram& mp = hp.mode[mode];
    const csim_constants& C = hp.consts;
    const int K = (int)perms.size();
    // candidate retry by first failing step (else: one successor hint); measured per mode, see TpiOptions
    const bool CAND = K > 1 && (mode == CSIM_MODE_DC ? opt.cand_retry : opt.cand_retry_tran);
    std::vector<Analysis> As;
    for (const auto& pm : perms) As.push_back(analyse(hp, pm, mode));
    const int N = As[0].N;
    if (K > 1 && opt.replicas > 1) throw std::runtime_error("specialize: variants and replicas cannot be combined");
    TpiHost H;
    H.perm = perms[0];
    H.perms = perms;
    for (const auto& pm : perms) {  // pivot-hash contribution of each predicted permutation
        uint32_t hlo = 0, hhi = 0;
        for (int i = 0; i < N; ++i) {
            uint32_t key = (uint32_t)i * 1024u + (uint32_t)pm[i] + 1u;
            uint32_t lo = key * 0x9E3779B1u;
            lo ^= lo >> 15;
            uint32_t hi = key * 0x85EBCA77u;
            hi ^= hi >> 13;
            hlo += lo;
            hhi += hi;
        }
        H.hashes.push_back(((uint64_t)hhi << 32) | hlo);
    }
    H.hash_lo = (uint32_t)H.hashes[0];
    H.hash_hi = (uint32_t)(H.hashes[0] >> 32);

    // ---- classify values ------------------------------------------------------------------------
    enum Kind { STATIC, STEP, SRC, MOS, GMINV };
    std::vector<Kind> kind(mp.n_values, STATIC);
    std::vector<bool> is_const(mp.n_values, false);
    std::vector<double> const_val(mp.n_values, 0.0);
    std::vector<int> ind_params;  // inductances: L <= 0 makes SK_IND_ONE 0 -> such instances always solve generically
    for (const StaticOp& s : mp.static_ops) {
        if (s.kind == SK_ONE || (s.kind == SK_IND_ONE && opt.const_prop)) {
            is_const[s.out] = true;
            const_val[s.out] = 1.0;
            if (s.kind == SK_IND_ONE) ind_params.push_back(s.param);
        } else if (s.kind == SK_GMIN && !DC) {
            is_const[s.out] = true;
            const_val[s.out] = C.tran_gmin;
        }
    }
    if (DC)  // the controller's gmin: a per-lane variable, changes between Newton iterations
        for (const StaticOp& s : mp.static_ops)
            if (s.kind == SK_GMIN) kind[s.out] = GMINV;
    if (!opt.const_prop)
        for (const StaticOp& s : mp.static_ops)
            if (s.kind == SK_ONE) is_const[s.out] = false;  // (gmin stays a literal either way)
    for (const StepOp& s : mp.step_ops) kind[s.out] = STEP;
    for (const SrcOp& s : mp.src_ops) kind[s.out] = SRC;
    std::set<int> mos_gm;  // value ids of the gate transconductance terms (gd, gg, gs, cst at m.out + 0..3)
    for (const MosOp& m : mp.mos_ops) {
        for (int t = 0; t < 4; ++t) kind[m.out + t] = MOS;
        mos_gm.insert(m.out + 1);
    }

    // ---- shared-memory slots: Y first and contiguous (probe scratch), the rest on demand -------------
    int n_slots = 0;
    auto alloc = [&]() { return n_slots++; };
    const int Y0 = n_slots;
    n_slots += N;
    int X0 = -1;
    if (opt.x_in_smem) {
        X0 = n_slots;
        n_slots += N;
    }
    auto xrd = [&](int k) -> std::string {
        if (k < 0) return "0.0";
        return opt.x_in_smem ? S(X0 + k) : "x" + std::to_string(k) + "@";
    };
    // static prefix sums: constant-folded, or one slot per distinct ordered term list
    std::map<std::vector<std::pair<int, bool>>, int> base_of;
    std::vector<std::vector<std::vector<Val>>> bases(K, std::vector<std::vector<Val>>(N, std::vector<Val>(N, Val::K(0.0))));
    for (int v = 0; v < K; ++v)
    for (int i = 0; i < N; ++i)
        for (int j = 0; j < N; ++j) {
            const Analysis& A = As[v];
            std::vector<std::vector<Val>>& base = bases[v];
            if (A.prefix[i][j].empty()) continue;
            bool all_c = true;
            for (const Term& t : A.prefix[i][j]) all_c = all_c && is_const[t.value];
            if (all_c) {
                Val acc = Val::K(0.0);
                for (const Term& t : A.prefix[i][j])
                    acc = t.sub ? sub(acc, Val::K(const_val[t.value])) : add(acc, Val::K(const_val[t.value]));
                base[i][j] = acc;
                continue;
            }
            std::vector<std::pair<int, bool>> key;
            for (const Term& t : A.prefix[i][j]) key.push_back({t.value, t.sub});
            auto it = base_of.find(key);
            if (it == base_of.end()) it = base_of.emplace(key, alloc()).first;
            base[i][j] = Val::E(S(it->second));
        }
    // non-MOS, non-constant values referenced by the per-iteration terms
    std::map<int, int> val_slot;
    for (int v = 0; v < K; ++v)
    for (int i = 0; i < N; ++i)
        for (int j = 0; j <= N; ++j)
            for (const Term& t : As[v].rest[i][j])
                if (kind[t.value] != MOS && kind[t.value] != GMINV && !is_const[t.value] && !val_slot.count(t.value))
                    val_slot[t.value] = alloc();
    // static values needed again at every step (companion-model conductances) -> global scratch rows
    std::map<int, int> gs_row;
    for (const StepOp& s : mp.step_ops)
        if (!gs_row.count(s.g)) {
            int r = (int)gs_row.size();
            gs_row[s.g] = r;
        }
    H.n_scratch_rows = DC ? 0 : (int)gs_row.size() + N;  // + x_prev for the generic fallback
    const int XP0 = (int)gs_row.size();

    std::map<int, int> mos_slot;
    if (opt.mos_in_smem)
        for (const MosOp& m : mp.mos_ops)
            for (int t = 0; t < 4; ++t) mos_slot[m.out + t] = alloc();
    std::map<int, int> prm_slot;  // MOSFET parameters (Vth, K, lambda), loaded once
    if (opt.mos_in_smem)
        for (const MosOp& m : mp.mos_ops)
            for (int t = 0; t < 3; ++t)
                if (!prm_slot.count(m.param + t)) prm_slot[m.param + t] = alloc();
    bool pair_mos = false;  // warp pair: the helper linearises every MOSFET ahead of the iteration, values in shared memory
    auto val = [&](int v) -> Val {  // per-iteration operand
        if (is_const[v]) return Val::K(const_val[v]);
        if (kind[v] == GMINV) return Val::E("gmin@");
        if (kind[v] == MOS)  // (gm is exactly zero while the device is off: src/element.cpp:240-247)
            return pair_mos        ? Val::E(S(mos_slot.at(v)), kLatSmem, mos_gm.count(v) > 0)
                   : opt.mos_in_smem ? Val::E(S(mos_slot.at(v)), kLatMos + kLatSmem, mos_gm.count(v) > 0)
                                   : Val::E("m" + std::to_string(v) + "@", kLatMos, mos_gm.count(v) > 0);
        return Val::E(S(val_slot.at(v)));
    };


    const int R = std::max(1, opt.replicas);
    // ---- warp pair: row ownership (0 = master, 1 = helper) and the two hand-shake slots ---------------------------------
    const bool PAIR = opt.pair && K == 1 && !DC && R == 1 && opt.x_in_smem && !opt.opaque_park && !opt.mos_in_smem && !opt.q_in_smem;
    const bool LOCKSTEP = opt.lockstep && !PAIR && !DC;
    std::vector<int> owner(N, 0);
    std::vector<int> mos_of_value(mp.n_values, -1);
    for (size_t m = 0; m < mp.mos_ops.size(); ++m)
        for (int t = 0; t < 4; ++t) mos_of_value[mp.mos_ops[m].out + t] = (int)m;
    std::vector<char> mos_need[2] = {std::vector<char>(mp.mos_ops.size(), 1), std::vector<char>(mp.mos_ops.size(), 0)};
    std::map<int, int> cand_slot;  // unknown -> slot of the NEXT iterate's value at a MOSFET terminal
    int CMD = -1, HB = -1;  // CMD, CMD + 1: the masters' command words of even / odd rounds
    PairSchedule pair_plan;
    if (PAIR) {
        const Analysis& A = As[0];
        PairModel M;
        M.N = N;
        M.n_mos = 0;  // (the helper linearises all MOSFETs between the iterations, below)
        M.W = std::max(1, opt.rows_in_flight);
        M.asm_cost.assign(N, 5);
        M.park_cost.assign(N, 20);
        M.steps.resize(N);
        M.mos.resize(N);
        for (int i = 0; i < N; ++i) {
            std::set<int> ms;
            for (int j = 0; j <= N; ++j)
                for (const Term& t : A.rest[i][j]) {
                    if (is_const[t.value]) continue;
                    M.asm_cost[i] += 3;
                }
            M.mos[i].assign(ms.begin(), ms.end());
            M.park_cost[i] += 3 * (int)A.ucols[i].size();
            for (int k = 0; k < i; ++k)
                if (A.pat[i][k]) M.steps[i].push_back({k, 14 + 4 * ((int)A.cols_k[k].size() + 1)});
        }
        if (const char* e = std::getenv("CSIM_PAIR_SYNC")) kPairSync = std::atoi(e);
        if (const char* e = std::getenv("CSIM_PAIR_DEFER")) kPairDefer = std::atoi(e);
        if (const char* e = std::getenv("CSIM_PAIR_MARGIN")) kPairMargin = std::atoi(e);
        if (const char* e = std::getenv("CSIM_PAIR_EVENTS")) kPairMaxEvents = std::min(12, std::atoi(e));
        owner = pair_partition(M);
        if (const char* e = std::getenv("CSIM_PAIR_ROWS")) {  // development aid: the helper's rows, comma separated
            std::fill(owner.begin(), owner.end(), 0);
            for (const char* c = e; *c;) {
                const int r = std::atoi(c);
                if (r >= 0 && r < N) owner[r] = 1;
                while (*c && *c != ',') ++c;
                if (*c == ',') ++c;
            }
        }
        pair_plan = pair_schedule(M, owner);
        H.pair_cost[0] = pair_schedule(M, std::vector<int>(N, 0)).t[0];
        H.pair_cost[1] = pair_plan.t[0];
        H.pair_cost[2] = pair_plan.t[1];
        if (std::getenv("CSIM_PAIR_DEBUG")) {
            std::fprintf(stderr, "pair: model single %d, master %d, helper %d, hand-overs %d; helper rows:", H.pair_cost[0], H.pair_cost[1], H.pair_cost[2], (int)pair_plan.handovers.size());
            for (int i = 0; i < N; ++i) if (owner[i]) std::fprintf(stderr, " %d", i);
            std::fprintf(stderr, "\n");
        }
        for (int r = 0; r < 2; ++r) std::fill(mos_need[r].begin(), mos_need[r].end(), 0);
        pair_mos = true;
        for (const MosOp& m : mp.mos_ops) {
            for (int t = 0; t < 4; ++t) mos_slot[m.out + t] = alloc();
            for (int node : {m.d, m.g, m.s})
                if (node >= 0 && !cand_slot.count(node)) cand_slot[node] = alloc();
        }
        CMD = alloc();
        alloc();
        HB = alloc();
        H.pair = true;
    }
    std::ostringstream o;  // kernel body after the prologue, written for one replica ('@' placeholders)
    std::ostringstream oh_mos, oh_rows;  // warp pair: the helper's share of a Newton iteration
    std::ostringstream* row_out = &o;    // where the row generators below write
#define O (*row_out)
    // (warp pair) MOSFET linearisations, emitted where the comments at "MOSFET linearisation" below say
    // The master takes the last pair_mos_master of them itself: its tail (pairwise tree, square root, votes) is a chain of
    // dependent operations with issue slots to spare, while the helper's linearisations are what the loop waits for.
    std::ostringstream om_start, om_iter;  // the master's linearisations: before the first step / in the tail of an iteration
    if (PAIR) {
        int pair_mos_master = 0;  // (B200, dbmixer headline run: 142.2 M with 0 against 139.7 M with 1, although the static timeline says 5.9 k against 6.2 k cycles)
        if (const char* e = std::getenv("CSIM_PAIR_MOS_MASTER")) pair_mos_master = std::atoi(e);
        const int n_mos = (int)mp.mos_ops.size();
        for (int mi = 0; mi < n_mos; ++mi) {
            const MosOp& op = mp.mos_ops[mi];
            const bool by_master = mi >= n_mos - pair_mos_master;
            auto mn = [&](int t) { return "m" + std::to_string(op.out + t) + "@"; };
            for (int where = 0; where < (by_master ? 2 : 1); ++where) {
                std::ostringstream& om = by_master ? (where ? om_iter : om_start) : oh_mos;
                auto cand = [&](int node) {
                    if (node < 0) return std::string("0.0");
                    if (!by_master) return S(cand_slot.at(node));
                    return where ? "c" + std::to_string(node) + "@" : xrd(node);
                };
                om << "            {\n                double " << mn(0) << ", " << mn(1) << ", " << mn(2) << ", " << mn(3) << ";\n"
                   << "                " << (opt.mos_branchless ? "mos_eval_sel" : "mos_eval_pp") << "<" << (op.is_p ? -1 : 1) << ">(PRM("
                   << op.param << ", bb@), PRM(" << op.param + 1 << ", bb@), PRM(" << op.param + 2 << ", bb@), " << lit(C.mos_off_gds)
                   << ", " << cand(op.d) << ", " << cand(op.g) << ", " << cand(op.s) << ", " << mn(0) << ", " << mn(1) << ", " << mn(2)
                   << ", " << mn(3) << ");\n";
                for (int t = 0; t < 4; ++t) om << "                " << S(mos_slot.at(op.out + t)) << " = " << mn(t) << ";\n";
                om << "            }\n";
            }
        }
    }
    if (PAIR) o << "    const long long b@ = (long long)blockIdx.x * (32 * PAIRS) + (tid >> 6) * 32 + LTID;\n";
    else o << "    const long long b@ = (long long)blockIdx.x * (LANES * NREP) + @R * LANES + LTID;\n";
    o << ""
         "    const bool live@ = b@ < B && !(T.skip && T.skip[b@ < B ? b@ : B - 1]);\n"
         "    const long long bb@ = b@ < B ? b@ : B - 1;  // idle lanes shadow an instance and never store\n"
         "    unsigned long long h@ = (live@ && pivot_hash) ? pivot_hash[b@] : CSIM_PIVOT_HASH_INIT;\n"
         "    unsigned int st@ = 0, n_fallback@ = 0;\n"
         "    int total@ = 0;\n"
         "    bool dead@ = !live@, degen@ = false;\n";
    if (K > 1) o << "    int var@ = 0;  // variant (predicted permutation) that held for this lane's last solve\n";
    if (PAIR) o << "//%HELPER%\n";
    for (int i = 0; i < N; ++i) {
        const std::string x0 = DC ? std::string("0.0")  // x.setZero(N), src/dcanalysis.cpp:110
                                  : "xio[(long long)" + std::to_string(i) + " * B + bb@]";
        if (opt.x_in_smem) o << "    " << S(X0 + i) << " = " << x0 << ";\n";
        else o << "    double x" << i << "@ = " << x0 << ";\n";
    }
    // ---- one-off: static values, base sums ---------------------------------------------------------
    for (const auto& kv : prm_slot)
        o << "    " << S(kv.second) << " = __ldg(params + (long long)" << kv.first << " * B + bb@);\n";
    o << "    //@BEGIN\n    {\n";
    for (int p : ind_params) o << "        degen@ = degen@ || !(__ldg(params + (long long)" << p << " * B + bb@) > 0.0);\n";
    for (const StaticOp& op : mp.static_ops) {
        o << "        const double v" << op.out << " = ";
        const std::string P = "__ldg(params + (long long)" + std::to_string(op.param) + " * B + bb@)";
        if (is_const[op.out]) o << lit(const_val[op.out]);
        else switch (op.kind) {
            case SK_ONE: o << "1.0"; break;
            case SK_RES: o << "res_g(" << P << ")"; break;
            case SK_CAP: o << "cap_g(" << P << ", dt)"; break;
            case SK_CAPHALF: o << "cap_half(" << P << ", dt)"; break;
            case SK_IND_ONE: o << "ind_one(" << P << ")"; break;
            case SK_IND_NREQ: o << "ind_nreq(" << P << ", dt)"; break;
            default: o << "0.0";
        }
        o << ";\n";
        if (val_slot.count(op.out)) o << "        " << S(val_slot[op.out]) << " = v" << op.out << ";\n";
        if (gs_row.count(op.out))
            o << "        if (live@) scratch[(long long)" << gs_row[op.out] << " * B + b@] = v" << op.out << ";\n";
    }
    for (const auto& kv : base_of) {
        std::string acc = "0.0";
        for (const auto& t : kv.first)
            acc = std::string(t.second ? "d_sub(" : "d_add(") + acc + ", v" + std::to_string(t.first) + ")";
        o << "        " << S(kv.second) << " = " << acc << ";\n";
    }
    o << "    }\n    //@END\n";
    if (DC) {
        // dcSolveNewtonLU's source ramp (src/dcanalysis.cpp:112-117): scale = step / rampSteps, the controller's base
        // gmin for this step, prevErr = infinity; sources at evalDC(scale) (include/sim.hpp:152-158)
        o << "    for (int step = 1; step <= " << C.dc_ramp_steps << "; ++step) {\n"
             "        const double scale = (double)step / (double)" << C.dc_ramp_steps << ";\n"
             "        const double gbase = tpi_base_gmin(" << lit(C.gmin_high_base) << ", " << lit(C.gmin_low_base) << ", scale);\n"
             "        double gmin@ = gbase, prevErr@ = CSIM_TPI_INF;\n";
        for (const SrcOp& op : mp.src_ops) {
            if (!val_slot.count(op.out)) continue;
            o << "        " << S(val_slot[op.out]) << " = src_dc_value(params, B, bb@, " << op.param << ", " << (op.is_sin ? 1 : 0)
              << ", scale);\n";
        }
    } else {
    if (PAIR) {
        for (const auto& kv : cand_slot) o << "    " << S(kv.second) << " = " << xrd(kv.first) << ";\n";
        o << "    PAIR_ARRIVE(3);  // the helpers linearise the MOSFETs at the start iterate\n" << om_start.str();
    }
    o << "    const int r0 = T.out_begin_step / T.out_stride;\n"
         "    for (int step = T.step_begin; step < T.step_end; ++step) {\n"
      << (PAIR ? "" : LOCKSTEP ? "        if (__syncthreads_and(ALL(dead@))) break;\n"
                                : "        if (__all_sync(0xFFFFFFFFu, ALL(dead@))) break;\n");  // (warp pairs of a CTA stay in step)
    // ---- per step: x_prev -> scratch, companion history terms, source values -------------------------
    for (int i = 0; i < N; ++i)
        o << "        if (live@) scratch[(long long)" << (XP0 + i) << " * B + b@] = " << xrd(i) << ";\n";
    for (const StepOp& op : mp.step_ops) {
        if (!val_slot.count(op.out)) continue;
        const std::string g = "scratch[(long long)" + std::to_string(gs_row[op.g]) + " * B + bb@]";
        std::string e;
        if (op.kind == TK_CAP_HIST) e = "d_mul(-" + g + ", d_sub(" + xrd(op.a) + ", " + xrd(op.b) + "))";
        else e = "d_mul(" + g + ", " + xrd(op.a) + ")";
        o << "        " << S(val_slot[op.out]) << " = " << e << ";\n";
    }
    if (!mp.src_ops.empty()) {
        o << "        const double tnow = d_mul((double)(step + 1), dt);  // tanalisis.cpp:256\n";
        for (const SrcOp& op : mp.src_ops) {
            if (!val_slot.count(op.out)) continue;
            o << "        " << S(val_slot[op.out]) << " = src_tran_value(T, params, B, bb@, " << op.index << ", "
              << op.param << ", step, tnow);\n";
        }
    }
    }  // TRAN step head
    const int max_iters = DC ? C.dc_max_iters : C.tran_max_iters;
    o << "        int count@ = 0;\n"
         "        bool done@ = dead@;\n"
         "        for (int iter = 0; iter < " << max_iters << "; ++iter) {\n"
      << (PAIR ? "            // the masters of the CTA decide together (their warps share the named barriers); what they publish —\n"
                 "            // iterate, step values, static sums — is in shared memory before the barrier releases the helpers\n"
                 "            SMG(" + std::to_string(CMD) + " + par, tid >> 6) = __all_sync(0xFFFFFFFFu, done@) ? 2.0 : 1.0;\n"
                 "            PAIR_SYNC(1);\n"
                 "            if (pair_cmd(sm, " + std::to_string(CMD) + " + par, LTID) == 2.0) { par ^= 1; break; }\n"
                 "            par ^= 1;\n"
               : LOCKSTEP ? "            if (__syncthreads_and(ALL(done@))) break;  // the warps of the CTA iterate in step: they share fetched instruction lines\n"
               : "            if (__all_sync(0xFFFFFFFFu, ALL(done@))) break;\n")
      << "            bool bad@ = degen@;\n"
         "            unsigned int glo@ = 0xFFFFFFFFu, ghi@ = 0u, gacc@ = 0u, pmx@ = 0u, dacc@ = 0u;\n";
    // ---- MOSFET linearisation -------------------------------------------------------------------------
    // (warp pair: each warp linearises the MOSFETs the rows it owns read — both of them where both need one)
    // (warp pair: the helper linearises every MOSFET for the NEXT iteration while the master is busy with norm and
    // convergence test — from the damped iterate the master leaves at the MOSFET terminals — so neither warp has the
    // linearisations on its path; see the damping block below)
    for (size_t mi = 0; mi < mp.mos_ops.size(); ++mi)
    for (int role = 0; role < 1; ++role) {
        if (!mos_need[role][mi]) continue;
        const MosOp& op = mp.mos_ops[mi];
        std::ostringstream& om = o;
        auto prm = [&](int k) { return opt.mos_in_smem ? S(prm_slot.at(k)) : "PRM(" + std::to_string(k) + ", bb@)"; };
        auto mn = [&](int t) { return "m" + std::to_string(op.out + t) + "@"; };
        om << "            double " << mn(0) << ", " << mn(1) << ", " << mn(2) << ", " << mn(3) << ";\n"
          << "            " << (opt.mos_branchless ? "mos_eval_sel" : "mos_eval_pp") << "<" << (op.is_p ? -1 : 1) << ">("
          << prm(op.param) << ", " << prm(op.param + 1) << ", " << prm(op.param + 2) << ", " << lit(C.mos_off_gds) << ", "
          << xrd(op.d) << ", " << xrd(op.g) << ", " << xrd(op.s) << ", " << mn(0) << ", " << mn(1) << ", " << mn(2) << ", "
          << mn(3) << ");\n";
        if (opt.mos_in_smem)
            for (int t = 0; t < 4; ++t) om << "            " << S(mos_slot[op.out + t]) << " = " << mn(t) << ";\n";
    }
    // ---- rows: assemble, eliminate, park ----------------------------------------------------------------
    // A row is emitted in SEGMENTS — assembly, one elimination step per earlier pivot row k with a structural
    // non-zero at (i,k), parking — and up to opt.rows_in_flight rows are open at a time, their segments
    // interleaved round-robin.  Step (i,k) only needs row k parked, so any such order is a topological order
    // of the up-looking LU and every entry still receives its updates in ascending k (bit-identical).  The
    // elimination steps of ONE row form a serial chain (multiplier -> update -> next multiplier, ~44 cycles
    // per step); rows that do not feed each other fill those gaps, and the compiler's scheduling window is
    // too short to find them when whole rows follow each other in the source.
    // Several variants: a lane whose variant misses first tries the variant that FOLLOWED it the last time any lane
    // left it (succ[], a hint table behind the miss log, racy on purpose) in a second pass over the same bodies, and
    // only then takes the generic redo.  Instances of one batch walk through the same sequence of operating
    // regions, so after the first of them every change of permutation costs one more straight-line solve instead
    // of a dense generic one.
    if (K > 1) {
        for (int i = 0; i < N; ++i) o << "            double q" << i << "@ = 0.0;\n";  // (done lanes skip all bodies)
        o << "            bool need@ = !done@, missed@ = false;\n"
             "            int var_before@ = var@;\n"
             "#pragma unroll 1\n"
             "            unsigned int tried@ = 0u;\n"
             "            int kf@ = SPEC_N;  // first elimination step at which the running variant's pivot prediction failed\n"
             "            for (int pass = 0; pass < " << (CAND ? 3 : 2) << "; ++pass) {\n"
             "            if (need@) {\n"
             "                kf@ = SPEC_N;\n"
             "                bad@ = degen@;\n"
             "                glo@ = 0xFFFFFFFFu;\n"
             "                ghi@ = 0u;\n"
             "                gacc@ = 0u;\n"
             "                pmx@ = 0u;\n"
             "                dacc@ = 0u;\n"
             "            }\n";
    }
    unsigned dacc_limit = 0xFFFFFFFFu;  // pivot range accumulator's acceptance limit (fold_guards)
    // TPI_CAND[v * N + k]: variants with the same pivots as v before step k and another one at k
    std::string cand_table;
    if (K > 1) {
        std::ostringstream ct;
        ct << "__constant__ unsigned short TPI_CAND[" << K * N << "] = {";
        for (int v = 0; v < K; ++v)
            for (int k = 0; k < N; ++k) {
                unsigned mask = 0;
                for (int u = 0; u < K; ++u) {
                    if (u == v || perms[u][k] == perms[v][k]) continue;
                    bool same = true;
                    for (int t = 0; t < k; ++t) same = same && perms[u][t] == perms[v][t];
                    if (same) mask |= 1u << u;
                }
                ct << (v + k ? ", " : "") << mask;
            }
        ct << "};\n";
        cand_table = ct.str();
    }
    const int n_common = n_slots;  // slots below are shared by all variants; each variant parks its LU above
    int n_slots_max = n_slots;
    TpiHost H0;  // statistics of variant 0
    for (int v = 0; v < K; ++v) {
    const Analysis& A = As[v];
    const std::vector<std::vector<Val>>& base = bases[v];
    const std::vector<int32_t>& perm = perms[v];
    n_slots = n_common;
    if (K > 1)
        o << "            if (__any_sync(0xFFFFFFFFu, var@ == " << v << " && need@)) {\n"
             "            if (var@ == " << v << " && need@) {  // variant " << v << "\n";
    // generator-side description of the factorisation so far
    // [c]: as the code of warp c refers to them (warp pair with register forwarding: the warp that parked a row reads its
    // own registers, the other warp the shared-memory slots; otherwise both entries are the same)
    std::vector<Val> DvA[2] = {std::vector<Val>(N), std::vector<Val>(N)}, RDvA[2] = {std::vector<Val>(N), std::vector<Val>(N)},
                     YvA[2] = {std::vector<Val>(N), std::vector<Val>(N)};
    std::vector<std::map<int, Val>> UvA[2] = {std::vector<std::map<int, Val>>(N), std::vector<std::map<int, Val>>(N)};
    const bool REGFWD = PAIR && std::getenv("CSIM_PAIR_REGFWD") && std::atoi(std::getenv("CSIM_PAIR_REGFWD")) != 0;
#define CR (row_out == &oh_rows ? 1 : 0)
#define Dv DvA[CR]
#define RDv RDvA[CR]
#define Yv YvA[CR]
#define Uv UvA[CR]
    std::vector<bool> parked(N, false);
    char num[64];
    auto fabs_lit = [&](double v) { return lit(std::fabs(v)); };
    // a / pivot(k): exact rewrites for +-1 pivots, constant folding, otherwise the guarded split division
    // true -> divide() emits a division whose guards also judge the pivot rule (tdiv_mul): no explicit comparison needed
    auto folded_mult = [&](const Val& a, int k) {
        const Val& d = Dv[k];
        // (several variants: the failing STEP is needed to pick the next candidate, so the rule is tested explicitly)
        return !CAND && opt.fold_guards && !a.z && !(d.c && (d.v == 1.0 || d.v == -1.0)) && !(a.c && d.c);
    };
    auto divide = [&](const Val& a, int k, const std::string& tmp, const char* ind, bool mult) -> Val {
        const Val& d = Dv[k];
        if (d.c && d.v == 1.0) return a;
        if (d.c && d.v == -1.0) return neg(a);
        if (a.c && d.c) return Val::K(a.v / d.v);
        if (opt.fold_guards && !a.z)
            O << ind << "const double " << tmp << " = " << (mult && !CAND ? "tdiv_mul(" : "tdiv_nz(") << a.e << ", " << d.e
              << ", " << RDv[k].e << ", gacc@" << (mult && !CAND ? ", pmx@" : "") << ");\n";
        else
            O << ind << "const double " << tmp << " = tdiv(" << a.e << ", " << d.e << ", " << RDv[k].e << ", glo@, ghi@);\n";
        H.n_lu_divs++;
        return Val::E(tmp, std::max(a.t, RDv[k].t) + kLatDiv, a.z);
    };
    struct RowGen {
        int i = 0;
        std::map<int, Val> row;
        std::set<int> declared;
        std::vector<int> ks;  // elimination steps still to do, ascending
        size_t next = 0;
        bool assembled = false;
    };
    const char* ind = "            ";
    auto assign = [&](RowGen& R, int j, const Val& v) {  // j == N: right-hand side
        const std::string name =
            (j == N ? "rb" + std::to_string(R.i) : "r" + std::to_string(R.i) + "_" + std::to_string(j)) + "@";
        if (v.c) {
            R.row[j] = v;
            return;
        }
        if (v.e == name) return;
        O << ind << (R.declared.count(j) ? "" : "double ") << name << " = " << v.e << ";\n";
        R.declared.insert(j);
        R.row[j] = Val::E(name, v.t, v.z);
    };
    auto emit_assembly = [&](RowGen& R) {
        const int i = R.i;
        O << ind << "// position " << i << " = row " << perm[i] << "\n";
        for (int j = 0; j <= N; ++j) {
            if (j < N && !A.pat[i][j]) continue;
            Val acc = j < N ? base[i][j] : Val::K(0.0);
            for (const Term& t : A.rest[i][j]) {
                acc = t.sub ? sub(acc, val(t.value)) : add(acc, val(t.value));
                (j < N ? H.n_g_adds : H.n_rhs_adds)++;
            }
            assign(R, j, acc);
        }
        R.assembled = true;
    };
    auto emit_step = [&](RowGen& R, int k) {
        const int i = R.i;
        const Val a = R.row[k];
        // the reference would have picked this row instead of the predicted pivot (solver.hpp:46-56): a larger
        // entry always, an EQUAL one when its scan meets this row before the predicted pivot row (first maximum wins)
        const bool tie_loses = A.before_pivot[k][i];
        if (folded_mult(a, k)) {
            // judged on the quotient by tdiv_mul (|f| >= 1 on the high word: never misses, ties included)
        } else if (a.c && Dv[k].c) {
            if (std::fabs(a.v) > std::fabs(Dv[k].v) || (tie_loses && std::fabs(a.v) == std::fabs(Dv[k].v)))
                O << ind << (CAND ? "kf@ = min(kf@, " + std::to_string(k) + ");\n" : std::string("bad@ = true;\n"));
        } else if (CAND) {
            O << ind << "if (fabs(" << a.e << ") " << (tie_loses ? ">=" : ">") << " "
              << (Dv[k].c ? fabs_lit(Dv[k].v) : "fabs(" + Dv[k].e + ")") << ") kf@ = min(kf@, " << k << ");\n";
        } else {
            O << ind << "bad@ = bad@ || (fabs(" << a.e << ") " << (tie_loses ? ">=" : ">") << " "
              << (Dv[k].c ? fabs_lit(Dv[k].v) : "fabs(" + Dv[k].e + ")") << ");\n";
        }
        std::snprintf(num, sizeof num, "f%d_%d@", i, k);
        const Val f = divide(a, k, num, ind, true);
        for (int j : A.cols_k[k]) {
            assign(R, j, sub(R.row[j], mul(f, Uv[k].at(j))));
            H.n_elim_updates++;
        }
        assign(R, N, sub(R.row[N], mul(f, Yv[k])));
    };
    // opaque parking (opt.opaque_park): (slot, value) pairs of the row, flushed by asm blocks that define tok<i>
    std::vector<std::pair<int, std::string>> park_list;
    int ldp_uid = 0;
    auto PK = [&](int slot, int row) {  // read expression of a parked slot
        return opt.opaque_park ? "ldp<SOFF(" + std::to_string(slot) + ", @R), __LINE__>(sbase, tok" + std::to_string(row) +
                                     "_v" + std::to_string(v) + "@)"
                               : S(slot);
    };
    auto park = [&](int slot, const std::string& value) {
        if (opt.opaque_park) park_list.push_back({slot, value});
        else O << ind << S(slot) << " = " << value << ";\n";
    };
    auto park_flush = [&](int row) {
        if (!opt.opaque_park) return;
        const std::string tok = "tok" + std::to_string(row) + "_v" + std::to_string(v) + "@";
        const size_t kChunk = 12;  // (an asm statement takes at most 30 operands)
        std::string base = "sbase";
        for (size_t c0 = 0, ci = 0;; c0 += kChunk, ++ci) {
            const size_t c1 = std::min(park_list.size(), c0 + kChunk);
            const bool last = c1 >= park_list.size();
            const std::string out = last ? tok : tok + "c" + std::to_string(ci);
            O << ind << "unsigned int " << out << ";\n" << ind << "asm volatile(\"";
            for (size_t e = 0; e < c1 - c0; ++e) O << "st.shared.f64 [%1+%" << (2 * e + 2) << "], %" << (2 * e + 3) << "; ";
            O << "mov.b32 %0, %1;\" : \"=r\"(" << out << ") : \"r\"(" << base << ")";
            for (size_t e = c0; e < c1; ++e)
                O << ", \"n\"(SOFF(" << park_list[e].first << ", @R)), \"d\"(" << park_list[e].second << ")";
            O << ");\n";
            base = out;
            if (last) break;
        }
        park_list.clear();
    };
    (void)ldp_uid;
    auto emit_park = [&](RowGen& R) {
        const int i = R.i;
        // pivot of this position: singular / non-finite pivots go to the generic path
        // (a permutation nobody's matrix can take may put a structural zero on the diagonal: constant 0 -> "bad")
        const Val d = R.row.count(i) ? R.row[i] : Val::K(0.0);
        const int me = CR;  // the warp that parks this row
        auto both = [&](std::vector<Val>* arr, int idx, const Val& shared, const Val& own) {
            for (int c = 0; c < 2; ++c) arr[c][idx] = (REGFWD && c == me) ? own : shared;
        };
        if (d.c && (d.v == 1.0 || d.v == -1.0)) {
            both(DvA, i, d, d);
        } else {
            const int sd = alloc(), srd = alloc();
            const std::string dn = "r" + std::to_string(i) + "_" + std::to_string(i) + "@";
            if (d.c) O << ind << "const double " << dn << " = " << d.e << ";\n";
            // (rcp_seq needs a normal reciprocal: 2^-1021 <= |pivot| <= 2^1021; anything else is redone generically)
            if (opt.fold_guards && opt.int_pivot_guard) {
                // on the high word, conservatively: accept 2^e_lo <= |pivot| < 2^1021 with 2^e_lo the smallest power of
                // two >= max(lu_eps, 2^-1021) — a pivot between lu_eps and 2^e_lo only costs a generic redo.  One
                // unsigned accumulator: t = 2*hi - 2*HI(2^e_lo) wraps above the limit for everything too small
                // (zero and subnormals included) and exceeds it for everything too large, infinities and NaNs included.
                int e_lo = 0;
                std::frexp(std::max(C.lu_eps, 0x1p-1021), &e_lo);  // value = m * 2^e_lo, 0.5 <= m < 1  ->  2^e_lo >= value
                const unsigned hi_lo = (unsigned)(e_lo + 1023) << 20, hi_hi = (unsigned)(1020 + 1023) << 20 | 0xFFFFFu;  // |pivot| < 2^1021
                char buf[160];
                std::snprintf(buf, sizeof buf, "dacc@ = max(dacc@, 2u * (unsigned int)__double2hiint(%s) - 0x%xu);  // <= 0x%xu\n",
                              dn.c_str(), 2u * hi_lo, 2u * (hi_hi - hi_lo) + 1u);
                O << ind << buf;
                dacc_limit = 2u * (hi_hi - hi_lo) + 1u;
            } else {
                O << ind << "bad@ = bad@ || !(fabs(" << dn << ") >= " << lit(std::max(C.lu_eps, 0x1p-1021)) << " && fabs(" << dn
                  << ") <= 0x1p+1021);\n";
            }
            if (opt.opaque_park || REGFWD) O << ind << "const double rd" << i << "@ = rcp_seq(" << dn << ");\n";
            park(sd, dn);
            park(srd, (opt.opaque_park || REGFWD) ? "rd" + std::to_string(i) + "@" : "rcp_seq(" + dn + ")");
            both(DvA, i, Val::E(PK(sd, i), d.t + kLatSmem), Val::E(dn, d.t));
            both(RDvA, i, Val::E(PK(srd, i), d.t + kLatRcp + kLatSmem), Val::E("rd" + std::to_string(i) + "@", d.t + kLatRcp));
            H.n_rcp++;
        }
        for (int j : A.ucols[i]) {
            const Val u = R.row[j];
            if (u.c) {
                for (int c = 0; c < 2; ++c) UvA[c][i][j] = u;
            } else {
                const int su = alloc();
                park(su, u.e);
                for (int c = 0; c < 2; ++c)
                    UvA[c][i][j] = (REGFWD && c == me) ? Val::E(u.e, u.t, u.z) : Val::E(PK(su, i), u.t + kLatSmem, u.z);
            }
        }
        if (R.row[N].c) {
            both(YvA, i, R.row[N], R.row[N]);
        } else {
            park(Y0 + i, "rb" + std::to_string(i) + "@");
            both(YvA, i, Val::E(PK(Y0 + i, i), R.row[N].t + kLatSmem, R.row[N].z),
                 Val::E("rb" + std::to_string(i) + "@", R.row[N].t, R.row[N].z));
        }
        park_flush(i);
        parked[i] = true;
    };
    if (!PAIR) {
        const size_t W = (size_t)std::max(1, opt.rows_in_flight);
        std::vector<RowGen> open;
        int next_row = 0;
        while (next_row < N || !open.empty()) {
            while (open.size() < W && next_row < N) {
                RowGen R;
                R.i = next_row++;
                for (int k = 0; k < R.i; ++k)
                    if (A.pat[R.i][k]) R.ks.push_back(k);
                open.push_back(std::move(R));
            }
            bool progressed = false;
            for (size_t r = 0; r < open.size();) {
                RowGen& R = open[r];
                if (!R.assembled) {
                    emit_assembly(R);
                } else if (R.next < R.ks.size()) {
                    if (!parked[R.ks[R.next]]) {  // waits for a row that is still open (a lower one: no deadlock)
                        ++r;
                        continue;
                    }
                    emit_step(R, R.ks[R.next++]);
                } else {
                    emit_park(R);
                    open.erase(open.begin() + (long)r);
                    progressed = true;
                    continue;
                }
                progressed = true;
                ++r;
            }
            if (!progressed) throw std::runtime_error("specialize: row schedule made no progress");
        }
    } else {
        // Warp pair: replay the list schedule (pair_schedule) into the two streams.  A hand-over is one named barrier,
        // used once per Newton iteration — the iteration-start barrier lies between two uses, so an arrive can never
        // run into the previous use of its barrier.
        std::map<std::pair<int, int>, int> bar_of;
        for (const auto& hp2 : pair_plan.handovers) {
            const int id = 4 + (int)bar_of.size();
            bar_of[hp2] = id;
        }
        H.pair_events = (int)bar_of.size();
        std::map<int, RowGen> rows;
        for (const PairAct& act : pair_plan.acts) {
            row_out = act.role ? &oh_rows : &o;
            if (act.kind == PairAct::WAIT) {
                O << ind << "PAIR_SYNC(" << bar_of.at({act.row, act.k}) << ");\n";
                continue;
            }
            RowGen& R = rows[act.row];
            R.i = act.row;
            if (act.kind == PairAct::ASM) {
                emit_assembly(R);
            } else if (act.kind == PairAct::STEP) {
                if (!parked[act.k]) throw std::runtime_error("specialize: pair schedule uses a row before it is parked");
                emit_step(R, act.k);
            } else {
                emit_park(R);
                auto it = bar_of.find({act.role, act.k});
                if (it != bar_of.end()) O << ind << "PAIR_ARRIVE(" << it->second << ");\n";
                rows.erase(act.row);
            }
        }
        row_out = &o;
    }
    if (PAIR)
        O << ind << "PAIR_SYNC(2);  // the helper's rows are parked\n"
          << ind << "bad@ = bad@ || " << S(HB) << " != 0.0;\n";
    // ---- back substitution (include/solver.hpp:116-128) ---------------------------------------------------
    std::vector<Val> Q(N);
    if (K == 1)
        for (int i = 0; i < N; ++i) o << "            double q" << i << "@;\n";
    for (int i = N - 1; i >= 0; --i) {
        o << "            {\n";
        Val s = Yv[i];
        int t = 0;
        for (int j : A.ucols[i]) {
            s = sub(s, mul(Uv[i].at(j), Q[j]));
            H.n_back_terms++;
            if (!s.c && (++t % 3) == 0) {  // keep expressions shallow
                o << "                const double s" << t << "@ = " << s.e << ";\n";
                s = Val::E("s" + std::to_string(t) + "@", s.t, s.z);
            }
        }
        const Val q = divide(s, i, "qq@", "                ", false);
        o << "                q" << i << "@ = " << q.e << ";\n";
        // solver.hpp:124-127 zeroes the unknown of a pivot below 1e-15; the pivot check above sends every
        // |pivot| < lu_eps to the generic path, so the test is dead code unless lu_eps was lowered
        if (!Dv[i].c && !(C.lu_eps >= 1e-15))
            o << "                if (fabs(" << Dv[i].e << ") < 1e-15) q" << i << "@ = 0.0;\n";
        if (opt.q_in_smem) o << "                " << S(Y0 + i) << " = q" << i << "@;\n";
        o << "            }\n";
        Q[i] = opt.q_in_smem ? Val::E(S(Y0 + i), q.t + kLatSmem) : Val::E("q" + std::to_string(i) + "@", q.t);
        H.crit_cycles = std::max(H.crit_cycles, Q[i].t);
    }
    if (K > 1) o << "            }\n            }\n";
    n_slots_max = std::max(n_slots_max, n_slots);
    if (v == 0) H0 = H;
    }  // variants
    n_slots = n_slots_max;
    {  // the reported statistics are those of variant 0
        H.n_elim_updates = H0.n_elim_updates, H.n_lu_divs = H0.n_lu_divs, H.n_back_terms = H0.n_back_terms;
        H.n_g_adds = H0.n_g_adds, H.n_rhs_adds = H0.n_rhs_adds, H.n_rcp = H0.n_rcp, H.crit_cycles = H0.crit_cycles;
    }
    H.crit_lu_cycles = H.crit_cycles;
    H.crit_cycles += 5 * kLatOp + 5 * kLatOp + kLatSqrt;  // damped update + square, 32-leaf tree, sqrt
    // ---- prediction check / per-lane generic redo ------------------------------------------------------------
    if (K > 1 && !CAND)
        o << "            if (need@) missed@ = bad@ || (glo@ < 0x03600000u) || (ghi@ > 0x7FEFFFFFu) || (gacc@ > kTdivTMax) || (pmx@ >= kTdivTOne) || (dacc@ > DACC_LIMIT);\n"
             "            need@ = false;\n"
             "            if (pass == 0 && missed@ && !degen@ && T.miss) {\n"
             "                const int w = (int)T.miss[TPI_SUCC0 + var@] - 1;\n"
             "                if (w >= 0 && w < " << K << " && w != var@) {\n"
             "                    var@ = w;\n"
             "                    need@ = true;\n"
             "                }\n"
             "            }\n"
             "            if (!__any_sync(0xFFFFFFFFu, need@)) break;\n"
             "            }  // pass\n"
             "            const bool redo@ = missed@;\n";
    else if (K > 1)
        // A lane whose variant failed at elimination step kf retries — at most twice — a variant with the SAME pivots
        // before kf and a different one at kf (TPI_CAND, from the permutations' common prefixes): the matrix up to
        // step kf is the same under all of them, so they are the only known orders that can hold.  Among several
        // candidates the one that last followed this variant (succ[], a racy global hint) goes first, then the
        // earliest-learned (most frequent).  Failures that are not about the order (range guards) go straight to the redo.
        o << "            const bool order_only@ = !bad@ && !((glo@ < 0x03600000u) || (ghi@ > 0x7FEFFFFFu) || (gacc@ > kTdivTMax) || (dacc@ > DACC_LIMIT));\n"
             "            if (need@) missed@ = !order_only@ || kf@ < SPEC_N;\n"
             "            const bool was@ = need@;\n"
             "            need@ = false;\n"
             "            if (was@ && pass < 2 && missed@ && order_only@ && !degen@) {\n"
             "                tried@ |= 1u << var@;\n"
             "                const unsigned int cand = (unsigned int)TPI_CAND[var@ * SPEC_N + kf@] & ~tried@;\n"
             "                if (cand) {\n"
             "                    const int hint = T.miss ? (int)T.miss[TPI_SUCC0 + var@] - 1 : -1;\n"
             "                    var@ = (hint >= 0 && hint < " << K << " && ((cand >> hint) & 1u)) ? hint : (__ffs(cand) - 1);\n"
             "                    need@ = true;\n"
             "                }\n"
             "            }\n"
             "            if (!__any_sync(0xFFFFFFFFu, need@)) break;\n"
             "            }  // pass\n"
             "            const bool redo@ = missed@;\n";
    else
    o << "            const bool redo@ = (bad@ || (glo@ < 0x03600000u) || (ghi@ > 0x7FEFFFFFu) || (gacc@ > kTdivTMax) || (pmx@ >= kTdivTOne) || (dacc@ > DACC_LIMIT)) && !done@;\n";
    o << ""
         "            if (__any_sync(0xFFFFFFFFu, ANY(redo@))) {\n"
         "                //@BEGIN\n"
         "                if (redo@) {\n"
         "                    ++n_fallback@;\n"
         "                    double xl[SPEC_N], xo[SPEC_N];\n";
    if (K > 1) o << "                    unsigned long long pc@ = 0;\n";
    o << "";
    for (int i = 0; i < N; ++i) o << "                    xl[" << i << "] = " << xrd(i) << ";\n";
    if (DC)
        o << "                    tpi_generic_solve_dc<SPEC_N, SPEC_NV>(P, T, params, B, bb@, scale, gmin@, xl, xo, h@, st@"
          << (K > 1 ? ", pc@" : "") << ");\n";
    else
    o << "                    tpi_generic_solve<SPEC_N, SPEC_NV>(P, T, params, B, bb@, step, xl, scratch + (long long)" << XP0
      << " * B, xo, h@, st@" << (K > 1 ? ", pc@" : "") << ");\n";
    for (int i = 0; i < N; ++i) o << "                    q" << i << "@ = xo[" << i << "];\n";
    if (K > 1) {  // the permutation the reference took: run that variant from now on, and leave the hint
        o << "                    int learned@ = -1;\n";
        for (int v = 0; v < K; ++v)
            o << "                    if (pc@ == 0x" << std::hex << H.hashes[v] << std::dec << "ull) learned@ = " << v << ";\n";
        o << "                    var@ = learned@ >= 0 ? learned@ : var_before@;\n"
             "                    if (learned@ >= 0 && learned@ != var_before@ && T.miss)\n"
             "                        T.miss[TPI_SUCC0 + var_before@] = (unsigned long long)(learned@ + 1);\n";
    }
    o << "                }\n"
         "                //@END\n"
         "            }\n";
    {
        std::ostringstream f;
        for (int v = K - 1; v >= 1; --v) f << "var@ == " << v << " ? 0x" << std::hex << H.hashes[v] << std::dec << "ull : ";
        f << "F";
        o << "            if (!redo@ && !done@) h@ = (h@ ^ (" << f.str() << ")) * 0x100000001B3ull;\n";
    }
    // ---- damping, norm, convergence (src/tanalisis.cpp:359-376) -----------------------------------------------
    // The reference throws when a solution entry is not finite.  A non-finite q(i) makes n(i) and the sum of
    // squares non-finite (inf/NaN propagate through every operation below), so "sum of squares finite" proves
    // all 31 entries finite with one comparison; only when it is not (overflow of a square counts too) are the
    // entries themselves inspected, in a branch no lane normally takes.
    // (warp pair: first the unknowns at the MOSFET terminals — the helper linearises for the next iteration from them
    // while this warp finishes damping, norm and convergence test; x stays where the lane is done)
    for (int phase = PAIR ? 0 : 1; phase < 2; ++phase) {
    for (int i = 0; i < N; ++i) {
        if (PAIR && (cand_slot.count(i) > 0) != (phase == 0)) continue;
        const std::string xi = xrd(i), I = std::to_string(i);
        if (opt.x_in_smem) o << "            const double xo" << I << "@ = " << xi << ";\n";
        const std::string xv = opt.x_in_smem ? "xo" + I + "@" : xi;
        o << "            const double n" << I << "@ = d_add(" << xv << ", d_mul(" << lit(DC ? C.dc_alpha : C.tran_alpha) << ", d_sub(q" << I
          << "@, " << xv << ")));\n";
        if (PAIR && phase == 0)
            o << "            const double c" << I << "@ = done@ ? xo" << I << "@ : n" << I << "@;\n"
              << "            " << S(cand_slot.at(i)) << " = c" << I << "@;\n";
        o << "            double e" << I << "@ = d_sub(n" << I << "@, " << xv << ");\n"
          << "            e" << I << "@ = d_mul(e" << I << "@, e" << I << "@);\n";
    }
    if (PAIR && phase == 0) o << "            PAIR_ARRIVE(3);\n" << om_iter.str();
    }
    // 32-leaf pairwise tree (oracle/csim_norm2.h); leaves >= N are +0 and adding +0 to a square is exact
    {
        std::vector<std::string> leaf(32);
        for (int i = 0; i < 32; ++i) leaf[i] = i < N ? "e" + std::to_string(i) + "@" : "";
        for (int s = 16; s >= 1; s >>= 1)
            for (int i = 0; i < s; ++i) {
                if (leaf[i + s].empty()) continue;
                if (leaf[i].empty()) leaf[i] = leaf[i + s];
                else leaf[i] = "d_add(" + leaf[i] + ", " + leaf[i + s] + ")";
            }
        o << "            const double ssq@ = " << leaf[0] << ";\n";
    }
    o << "            bool fin@ = finite_d(ssq@);\n"
         "            if (__any_sync(0xFFFFFFFFu, ANY(!fin@ && !done@))) {\n"
         "                //@BEGIN\n"
         "                if (!fin@) {\n"
         "                    fin@ = true;\n";
    for (int i = 0; i < N; ++i) o << "                    fin@ = fin@ && finite_d(q" << i << "@);\n";
    o << "                }\n"
         "                //@END\n"
         "            }\n"
         "            //@BEGIN\n";
    if (DC)  // src/dcanalysis.cpp:135-138: gmin *= 10 (capped), the iteration is spent, x stays
        o << "            bool nf@ = false;\n"
             "            if (!done@) {\n"
             "                ++count@;\n"
             "                if (!fin@) {\n"
             "                    st@ |= CSIM_ST_NONFINITE;\n"
             "                    const double g10 = d_mul(gmin@, " << lit(C.dc_nonfinite_mul) << ");\n"
             "                    gmin@ = g10 < " << lit(C.dc_nonfinite_cap) << " ? g10 : " << lit(C.dc_nonfinite_cap) << ";\n"
             "                    nf@ = true;\n"
             "                }\n"
             "            }\n"
             "            //@END\n"
             "            const bool upd@ = !done@ && !nf@;\n";
    else
    o << "            if (!done@) {\n"
         "                ++count@;\n"
         "                if (!fin@) {  // the reference throws here\n"
         "                    st@ |= CSIM_ST_NONFINITE;\n"
         "                    dead@ = true;\n"
         "                    done@ = true;\n"
         "                }\n"
         "            }\n"
         "            //@END\n"
         "            const bool upd@ = !done@;\n";
    for (int i = 0; i < N; ++i) {
        const std::string xi = xrd(i), I = std::to_string(i);
        if (opt.x_in_smem) o << "            if (upd@) " << xi << " = n" << I << "@;\n";
        else o << "            x" << I << "@ = upd@ ? n" << I << "@ : x" << I << "@;\n";
    }
    o << "            const double err@ = sqrt(ssq@);\n";
    if (DC) {
        // ConvController::update (src/dcanalysis.cpp:268-307): gmin for the next iteration from the error trend
        o << "            //@BEGIN\n"
             "            if (upd@) {\n"
             "                double gnext = gbase;\n"
             "                if (!(iter == 0 || !finite_d(prevErr@))) {\n"
             "                    if (err@ > d_mul(prevErr@, " << lit(C.slow_conv_ratio) << ")) {\n"
             "                        const double g2 = d_mul(gmin@, 2.0);\n"
             "                        gnext = g2 < " << lit(C.gmin_abs_max) << " ? g2 : " << lit(C.gmin_abs_max) << ";\n"
             "                    } else if (err@ < d_mul(prevErr@, " << lit(C.fast_conv_ratio) << ")) {\n"
             "                        gnext = d_add(d_mul(0.5, gmin@), d_mul(0.5, gbase));\n"
             "                    } else {\n"
             "                        gnext = d_add(d_mul(0.7, gmin@), d_mul(0.3, gbase));\n"
             "                    }\n"
             "                }\n"
             "                gmin@ = gnext;\n"
             "                prevErr@ = err@;\n"
             "                if (err@ < " << lit(C.dc_tol) << ") done@ = true;\n"
             "                else if (iter == " << C.dc_max_iters - 1 << ") st@ |= CSIM_ST_DC_UNCONVERGED;\n"
             "            }\n"
             "            //@END\n"
             "        }\n"
             "        total@ += count@;\n"
             "        if (iters_total && live@) iters_total[(long long)(step - 1) * B + b@] = count@;  // iters [ramp][B]\n"
             "    }\n"
             "    //@BEGIN\n"
             "    if (live@) {\n";
        for (int i = 0; i < N; ++i) o << "        xio[(long long)" << i << " * B + b@] = " << xrd(i) << ";\n";
        o << "        if (pivot_hash) pivot_hash[b@] = h@;\n"
             "        if (status) status[b@] |= st@;\n"
             "        if (stats) {\n"
             "            if (n_fallback@) atomicAdd(stats, (unsigned long long)n_fallback@);\n"
             "            atomicAdd(stats + 1, (unsigned long long)total@);\n"
             "        }\n"
             "    }\n"
             "    //@END\n"
             "}\n";
    } else {
    o << "            //@BEGIN\n"
         "            if (upd@) {\n"
         "                if (err@ < " << lit(C.tran_tol) << ") done@ = true;\n"
         "                else if (iter == " << C.tran_max_iters - 1 << ") st@ |= CSIM_ST_TRAN_UNCONVERGED;\n"
         "            }\n"
         "            //@END\n"
         "        }\n"
         "        total@ += count@;\n"
         "        if (iters_step && live@ && count@ > 0) iters_step[(long long)step * B + b@] = (unsigned char)count@;\n"
         "        if ((step + 1) % T.out_stride == 0 && step >= T.out_begin_step && T.n_probes > 0) {\n"
         "            const long long r = (step + 1) / T.out_stride - 1 - r0;\n";
    const int PB = opt.x_in_smem ? X0 : Y0;  // where the probes are read from
    if (!opt.x_in_smem)
        for (int i = 0; i < N; ++i) o << "            " << S(Y0 + i) << " = x" << i << "@;\n";
    o << "            //@BEGIN\n"
         "            if (live@ && !dead@)\n"
         "                for (int p = 0; p < T.n_probes; ++p)\n"
         "                    wave[((long long)b@ * T.n_probes + p) * T.wave_ld + r] = SM(" << PB << " + T.probes[p], @R);\n"
         "            //@END\n"
         "        }\n"
         "    }\n"
      << (PAIR ? "    SMG(" + std::to_string(CMD) + " + par, tid >> 6) = 0.0;\n    PAIR_SYNC(1);  // dismiss the helpers\n" : std::string())
      << "    //@BEGIN\n"
         "    if (live@) {\n";
    for (int i = 0; i < N; ++i) o << "        xio[(long long)" << i << " * B + b@] = " << xrd(i) << ";\n";
    o << "        if (iters_total) iters_total[b@] += total@;\n"
         "        if (pivot_hash) pivot_hash[b@] = h@;\n"
         "        if (status) status[b@] |= st@;\n"
         "        if (stats) {\n"
         "            if (n_fallback@) atomicAdd(stats, (unsigned long long)n_fallback@);\n"
         "            atomicAdd(stats + 1, (unsigned long long)total@);\n"
         "        }\n"
         "    }\n"
         "    //@END\n"
         "}\n";
    }  // TRAN tail

#undef O
#undef CR
#undef Dv
#undef RDv
#undef Yv
#undef Uv
    // ---- warp pair: hand-overs -> named barriers, helper loop ---------------------------------------------------------------
    std::string body = o.str();
    if (PAIR) {
        std::ostringstream h;
        int n_helper_rows = 0;
        for (int i = 0; i < N; ++i) n_helper_rows += owner[i];
        h << "    if (tid & 32) {  // helper warp: " << n_helper_rows << " of the " << N << " rows of every Newton iteration the master runs\n"
             "        PAIR_SYNC(3);  // linearisation at the start iterate\n"
          << oh_mos.str()
          << "        for (int par = 0;; par ^= 1) {\n"
             "            PAIR_SYNC(1);\n"
             "            const double cmd = pair_cmd(sm, " << CMD << " + par, LTID);\n"
             "            if (cmd == 0.0) break;     // kernel end\n"
             "            if (cmd == 2.0) continue;  // every master's Newton loop has ended\n"
             "            bool bad@ = false;\n"
             "            unsigned int glo@ = 0xFFFFFFFFu, ghi@ = 0u, gacc@ = 0u, pmx@ = 0u, dacc@ = 0u;\n"
          << oh_rows.str()
          << "            " << S(HB) << " = (bad@ || (glo@ < 0x03600000u) || (ghi@ > 0x7FEFFFFFu) || (gacc@ > kTdivTMax) || (pmx@ >= kTdivTOne) || (dacc@ > DACC_LIMIT)) ? 1.0 : 0.0;\n"
             "            PAIR_ARRIVE(2);\n"
             "            PAIR_SYNC(3);  // the master's next iterate is there: linearise for the next iteration\n"
          << oh_mos.str()
          << "        }\n"
             "        return;\n"
             "    }\n";
        const size_t at = body.find("//%HELPER%\n");
        body.replace(at, std::string("//%HELPER%\n").size(), h.str());
    }

    H.n_slots = n_slots;
    H.replicas = R;
    const size_t warp_bytes = (size_t)n_slots * 8 * 32 * R;
    if (warp_bytes > kMaxSmem) throw std::runtime_error("specialize: per-instance state exceeds shared memory");
    // warp pair: the pairs of one CTA share the named barriers (they run the same two instruction streams in step,
    // which is what lets them share instruction-cache lines)
    int n_pairs = opt.warps_per_cta > 0 ? opt.warps_per_cta : 2;
    while (n_pairs > 1 && warp_bytes * n_pairs > kMaxSmem) --n_pairs;
    const int kTpb = PAIR ? 64 * n_pairs : 32 * (opt.warps_per_cta > 0 ? opt.warps_per_cta : pick_warps_per_cta(warp_bytes));
    H.threads = kTpb;
    H.inst_per_cta = PAIR ? 32 * n_pairs : kTpb * R;
    H.smem_bytes = PAIR ? warp_bytes * n_pairs : warp_bytes * (kTpb / 32);
    if (H.smem_bytes > kMaxSmem) throw std::runtime_error("specialize: CTA state exceeds shared memory");
    std::ostringstream pro;
    pro << "// generated by specialize_tpi.cpp — thread-per-instance transient kernel for one topology and one\n"
           "// predicted pivot permutation (" << K << " variant(s)); " << n_slots << " shared-memory doubles per instance, " << R
        << " instance(s) per thread, " << kTpb << " threads per CTA.\n"
           "// longest dependency chain of one Newton iteration: " << H.crit_cycles << " cycles (LU + substitutions: "
        << H.crit_lu_cycles << ")\n"
           "#include \"tpi_runtime.cuh\"\n"
           "using namespace csim;\n"
           "#define TPB " << kTpb << "\n"
           "#define NREP " << R << "\n"
           "#define NSLOTS " << n_slots << "\n"
           "#define SPEC_N " << N << "\n"
           "#define SPEC_NV " << mp.n_values << "\n"
           "#define LANES " << (PAIR ? "32" : "TPB") << "\n"
           "#define LTID " << (PAIR ? "(tid & 31)" : "tid") << "\n"
           "#define PAIRS " << (PAIR ? n_pairs : 0) << "\n"
           "#define SM(k, r) sm[((k) + " << (PAIR ? "(tid >> 6)" : "(r)") << " * NSLOTS) * LANES + LTID]\n"
           "#define SOFF(k, r) ((((k) + (r) * NSLOTS) * LANES) * 8)\n"
           "#define SMG(k, g) sm[((k) + (g) * NSLOTS) * LANES + LTID]\n"
           "#define PAIR_SYNC(id) asm volatile(\"bar.sync %0, %1;\" ::\"n\"(id), \"n\"(TPB) : \"memory\")\n"
           "#define PAIR_ARRIVE(id) asm volatile(\"bar.arrive %0, %1;\" ::\"n\"(id), \"n\"(TPB) : \"memory\")\n"
           "#define PRM(k, bb) ld_nohoist(params + (long long)(k) * B + (bb))\n"
           "#define DACC_LIMIT 0x" << std::hex << dacc_limit << std::dec << "u\n"
        << cand_table <<
           "#define TPI_SUCC0 " << kTpiMissEntries * kTpiMissStride << "\n"
           "// command of the round: 0 = leave as soon as one master says so, 2 = all masters' Newton loops have ended, else 1\n"
           "__device__ __forceinline__ double pair_cmd(const double* sm, int slot, int lane) {\n"
           "    double c = 2.0;\n"
           "    for (int g = 0; g < PAIRS; ++g) {\n"
           "        const double v = sm[(slot + g * NSLOTS) * 32 + lane];\n"
           "        c = (v == 0.0 || c == 0.0) ? 0.0 : (v == 1.0 ? 1.0 : c);\n"
           "    }\n"
           "    return c;\n"
           "}\n"
           "extern \"C\" __global__ void __launch_bounds__(TPB, " << std::max<size_t>(1, std::min<size_t>(16, kSmemPerSm / (H.smem_bytes + kCtaReserve)))
        << ")\n"
           "" << (DC ? "csim_dc_tpi" : "csim_tran_tpi") << "(DevPlan P, DevTran T, long long B, const double* __restrict__ params,\n"
           "              double* __restrict__ xio, double* __restrict__ wave, int* __restrict__ iters_total,\n"
           "              unsigned char* __restrict__ iters_step, unsigned long long* __restrict__ pivot_hash,\n"
           "              unsigned int* __restrict__ status, double* __restrict__ scratch,\n"
           "              unsigned long long* __restrict__ stats) {\n"
           "    extern __shared__ double sm[];\n"
           "    const int tid = threadIdx.x;\n"
           "    const unsigned int sbase = (unsigned int)__cvta_generic_to_shared(sm) + 8u * LTID;\n"

           "    int par = 0;  // (warp pair) parity of the command round\n"
           "    const double dt = T.dt;\n"
           "    const unsigned long long F = 0x" << std::hex << (((unsigned long long)H.hash_hi << 32) | H.hash_lo) << std::dec
        << "ull;\n";
    H.source = pro.str() + replicate(body, R);
    return H;
}
}  // namespace

}  // namespace csim
