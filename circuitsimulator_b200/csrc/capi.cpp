// capi.cpp — the extern "C" boundary declared in include/csim_b200.h.
#include <cuda_runtime.h>

#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/csim_b200.h"
#include <cstdlib>
#include <mutex>
#include <unistd.h>

#include "jit.hpp"
#include "kernels.hpp"
#include "plan.hpp"
#include "specialize.hpp"

using namespace csim;

namespace {
thread_local std::string g_err;
thread_local const uint8_t* g_tran_skip = nullptr;  // per-instance skip mask of the steady-state driver
thread_local double* g_margin = nullptr;  // csim_dev_margin_probe: [B] device buffer the generic kernels record threshold margins into
thread_local bool g_generic_stamp_only = false;  // set while the pivot pilot runs (it holds the plan's JIT lock)
int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    return fail(CSIM_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(call)                                            \
    do {                                                    \
        cudaError_t e__ = (call);                           \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

template <typename T>
cudaError_t upload(const std::vector<T>& v, const T** out, std::vector<void*>& owned) {
    *out = nullptr;
    if (v.empty()) return cudaSuccess;
    void* d = nullptr;
    cudaError_t e = cudaMalloc(&d, v.size() * sizeof(T));
    if (e != cudaSuccess) return e;
    owned.push_back(d);
    e = cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    *out = static_cast<const T*>(d);
    return e;
}
}  // namespace

// Specialised (run-time generated) transient kernel of a plan; see specialize.hpp.
struct SpecState {
    std::recursive_mutex mu;
    // 0 = generic kernels only; 1 = automatic (thread-per-instance kernel; falls back to the warp-per-instance
    // one when too many solves miss the predicted pivot order); 2 = warp-per-instance; 3 = thread-per-instance
    int mode = 1;
    int kind = 0;        // kernel in use: 0 none yet, 2 warp-per-instance, 3 thread-per-instance
    bool tried = false;  // a specialisation attempt has been made
    bool ready = false;  // kernel loaded and usable
    SpecHost host;
    JitKernel k;
    SpecTables dev{};
    size_t smem = 0;
    std::vector<void*> owned;
    unsigned long long* d_stats = nullptr;  // [2]: fallback solves, Newton solves
    std::string log;
};

// Thread-per-instance specialised transient kernel (specialize_tpi.cpp).
struct TpiState {
    bool ready = false;
    TpiHost host;  // one-warp CTAs: the fastest per warp (measured), default
    JitKernel k;
    int warps_per_sm = 0, n_sms = 0;
    // Same code in wider CTAs where that packs more warps into an SM's shared memory (dbmixer: 4 instead of 3).
    // Built on the first launch whose batch needs fewer waves of CTAs that way.
    bool wide_tried = false, wide_ready = false;
    TpiHost host_wide;
    JitKernel k_wide;
    int warps_per_sm_wide = 0;
    // Two warps per 32 instances (TpiOptions::pair): shorter Newton iterations for batches that leave SM sub-partitions
    // idle anyway.  Built on the first launch whose batch fits one wave of its CTAs (transient, one permutation).
    bool pair_tried = false, pair_ready = false;
    TpiHost host_pair;
    JitKernel k_pair;
    int groups_per_sm_pair = 0;
    unsigned long long* d_stats = nullptr;  // [2]: fallback solves, Newton solves
    unsigned long long* d_miss = nullptr;   // permutations met by the generic redo path (specialize.hpp kTpiMiss*)
    std::string history;                    // one line per variant added by the automatic mode
    std::string log;
};

// Specialised standalone assembly kernel, one per mode (specialize_stamp.cpp).
struct StampSpecState {
    bool tried = false, ready = false;
    StampSpecHost host;
    JitKernel k;
    std::string log;
};

struct csim_plan {
    HostPlan host;
    DevPlan dev{};
    int device = -1;
    int sm_count = 0;
    std::vector<void*> owned;
    mutable SpecState spec;
    mutable StampSpecState stamp_spec[2];
    mutable TpiState tpi;
    mutable TpiState tpi_dc;      // thread-per-instance DC kernel (csim_dc_tpi): same machinery, own variants and statistics
    mutable bool dc_tried = false;
    uint64_t topo_hash = 0;  // FNV-1a of the topology arrays + constants: key of the on-disk variant cache
    int force_big = 0;  // 0: by size; 1: large path (dense kernels); 2: large path (sparse-skipping warp kernels)
};

struct csim_batch {
    const csim_plan* plan = nullptr;
    int64_t B = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;  // waveform chunks go home while the next chunk of timepoints is computed
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_chunk = nullptr;
    double* d_params = nullptr;
    double* d_x = nullptr;
    uint64_t* d_hash = nullptr;
    uint32_t* d_status = nullptr;
    int32_t* d_iters = nullptr;  // [max(ramp,1)][B], also reused for iters_total
    double* d_wave = nullptr;    // device waveform of csim_batch_tran, kept between calls (grow-only): allocating and
    size_t wave_bytes = 0;       // freeing gigabytes per call costs more than the copy it feeds
    float last_ms = 0.f;
};

namespace {
struct DevBuf {  // RAII scratch for the host-pointer convenience calls
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t put(const void* src, size_t bytes) {
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 8);
        if (e == cudaSuccess && src && bytes) e = cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice);
        return e;
    }
    template <typename T> T* as() { return static_cast<T*>(p); }
};
}  // namespace

namespace {

// Generates, compiles (NVRTC, sm_100a) and loads the transient kernel specialised to `perm`.
int spec_build(const csim_plan* plan, const std::vector<int32_t>& perm) {
    SpecState& sp = plan->spec;
    sp.ready = false;
    // launches of the old kernel may still be running on other streams (the caller holds spec.mu, so none is being
    // prepared): retire its module and tables only once the device is idle
    cudaDeviceSynchronize();
    for (void* d : sp.owned) cudaFree(d);
    sp.owned.clear();
    jit_unload(sp.k);
    const std::string why = spec_unsupported_reason(plan->host);
    if (!why.empty()) {
        sp.log = "not specialised: " + why;
        return fail(CSIM_E_UNSUPPORTED, sp.log);
    }
    try {
        sp.host = generate_tran_spec(plan->host, perm);
    } catch (const std::exception& e) {
        sp.log = e.what();
        return fail(CSIM_E_INVALID, sp.log);
    }
    std::vector<char> cubin;
    std::string log;
    if (!jit_compile(sp.host.source, "csim_tran_spec.cu", jit_library_dir() + "/csrc", cubin, log)) {
        sp.log = "NVRTC: " + log;
        return fail(CSIM_E_CUDA, sp.log);
    }
    CU(jit_load(cubin, "csim_tran_spec", sp.k));
    sp.smem = spec_smem_bytes(plan->host, sp.host);
    CU(cudaKernelSetAttributeForDevice(sp.k.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp.smem,
                                       plan->device));
    SpecTables& t = sp.dev;
    CU(upload(sp.host.gidx, &t.gidx, sp.owned));
    CU(upload(sp.host.prefix, &t.prefix, sp.owned));
    CU(upload(sp.host.mos_lane, &t.mos_lane, sp.owned));
    t.n_adds = sp.host.n_adds;
    t.prefix_len = sp.host.prefix_len;
    t.n_mos = sp.host.n_mos;
    t.n_values = plan->host.mode[CSIM_MODE_TRAN].n_values;
    t.hash_lo = sp.host.hash_lo;
    t.hash_hi = sp.host.hash_hi;
    void* st = nullptr;
    CU(cudaMalloc(&st, 2 * sizeof(unsigned long long)));
    sp.owned.push_back(st);
    sp.d_stats = static_cast<unsigned long long*>(st);
    CU(cudaMemset(st, 0, 2 * sizeof(unsigned long long)));
    sp.log = "specialised: " + std::to_string(sp.host.n_lu_steps) + " elimination steps, " +
             std::to_string(sp.host.n_elim_updates) + " row updates, " + std::to_string(sp.host.n_back_terms) +
             " back-substitution terms, " + std::to_string(sp.host.n_g_adds) + "+" +
             std::to_string(sp.host.n_rhs_adds) + " assembly adds per Newton iteration; " +
             std::to_string(sp.smem) + " B shared memory per CTA" + (log.size() > 1 ? "\n" + log : "");
    sp.ready = true;
    return CSIM_OK;
}

// ---- on-disk cache of learned pivot-order variants (opt-in: CSIM_PLAN_CACHE=<directory>) -----------------------------
// The permutations a plan's kernels were last built with, per topology (hash of the csim_topology arrays and the
// constants block) and kernel (tran / dc): a new plan of the same topology starts from them instead of the single
// pilot prediction, so a one-shot user does not pay the learning launches again.  They are predictions like any other:
// every factorisation still verifies the reference's pivot rule, a stale or foreign file can only cost redos.
std::string cache_path(const csim_plan* plan, bool dc) {
    const char* dir = std::getenv("CSIM_PLAN_CACHE");
    if (!dir || !dir[0] || !plan->topo_hash) return "";
    char name[64];
    std::snprintf(name, sizeof name, "/csim_%016llx_%s.perms", (unsigned long long)plan->topo_hash, dc ? "dc" : "tran");
    return std::string(dir) + name;
}
void cache_store(const csim_plan* plan, bool dc, const std::vector<std::vector<int32_t>>& perms) {
    const std::string path = cache_path(plan, dc);
    if (path.empty()) return;
    const std::string tmp = path + ".tmp" + std::to_string((long long)getpid());
    if (FILE* f = std::fopen(tmp.c_str(), "w")) {
        std::fprintf(f, "%d %d\n", plan->host.N, (int)perms.size());
        for (const auto& pm : perms) {
            for (int32_t v : pm) std::fprintf(f, "%d ", v);
            std::fprintf(f, "\n");
        }
        std::fclose(f);
        std::rename(tmp.c_str(), path.c_str());
    }
}
bool cache_load(const csim_plan* plan, bool dc, std::vector<std::vector<int32_t>>& perms) {
    const std::string path = cache_path(plan, dc);
    if (path.empty()) return false;
    FILE* f = std::fopen(path.c_str(), "r");
    if (!f) return false;
    int n = 0, k = 0;
    bool ok = std::fscanf(f, "%d %d", &n, &k) == 2 && n == plan->host.N && k >= 1 && k <= 16;
    perms.assign(ok ? k : 0, std::vector<int32_t>(plan->host.N));
    for (auto& pm : perms) {
        std::vector<char> seen(plan->host.N, 0);
        for (int32_t& v : pm) {
            ok = ok && std::fscanf(f, "%d", &v) == 1 && v >= 0 && v < plan->host.N && !seen[v];
            if (ok) seen[v] = 1;
        }
    }
    std::fclose(f);
    if (!ok) perms.clear();
    return ok;
}

// Development aid: CSIM_TPI_OPTS = bit mask of TpiOptions (bit 5: folded division guards, bit 6: opaque row parking; bits 8-9: replicas - 1, bits 12-14: warps per CTA, bits 16-19: rows in flight, bit 20: warp pair per 32 instances, bit 21: warps of a CTA in step); unset = defaults.
TpiOptions tpi_options_from_env() {
    TpiOptions opt;
    if (const char* env = std::getenv("CSIM_TPI_OPTS")) {
        const int m = std::atoi(env);
        opt.const_prop = m & 1;
        opt.x_in_smem = m & 2;
        opt.mos_in_smem = m & 4;
        opt.q_in_smem = m & 8;
        opt.mos_branchless = m & 16;
        opt.fold_guards = m & 32;
        opt.opaque_park = m & 64;
        opt.int_pivot_guard = m & 128;
        opt.cand_retry = !(m & 1024);
        opt.cand_retry_tran = m & 2048;
        opt.replicas = 1 + ((m >> 8) & 3);
        opt.warps_per_cta = (m >> 12) & 7;
        if ((m >> 16) & 15) opt.rows_in_flight = (m >> 16) & 15;
        opt.pair = (m >> 20) & 1;
        opt.lockstep = (m >> 21) & 1;
    }
    return opt;
}

// Generates, compiles and loads one variant of the thread-per-instance kernel specialised to `perm`.
// warps_per_cta: 1, or 0 = the CTA width that packs most warps into an SM's shared memory.
int tpi_build_variant(const csim_plan* plan, const std::vector<std::vector<int32_t>>& perms, int warps_per_cta,
                      TpiHost& host, JitKernel& k, int& warps_per_sm, std::string& log, bool dc = false,
                      bool pair = false) {
    jit_unload(k);
    TpiOptions opt = tpi_options_from_env();
    if (opt.warps_per_cta == 0) opt.warps_per_cta = warps_per_cta;
    if (!std::getenv("CSIM_TPI_OPTS")) opt.pair = pair;  // (with the development mask set, its bit 20 decides)
    try {
        host = dc ? generate_dc_tpi(plan->host, perms, opt) : generate_tran_tpi(plan->host, perms, opt);
    } catch (const std::exception& e) {
        log = e.what();
        return fail(CSIM_E_UNSUPPORTED, log);
    }
    std::vector<char> cubin;
    if (!jit_compile(host.source, dc ? "csim_dc_tpi.cu" : "csim_tran_tpi.cu", jit_library_dir() + "/csrc", cubin, log)) {
        log = "NVRTC: " + log;
        return fail(CSIM_E_CUDA, log);
    }
    CU(jit_load(cubin, dc ? "csim_dc_tpi" : "csim_tran_tpi", k));
    CU(cudaKernelSetAttributeForDevice(k.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)host.smem_bytes,
                                       plan->device));
    int ctas = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas, reinterpret_cast<const void*>(k.kernel), host.threads,
                                                     host.smem_bytes));
    warps_per_sm = ctas * (host.inst_per_cta / (32 * host.replicas));  // groups of 32 x replicas instances resident per SM
    return CSIM_OK;
}

int tpi_build(const csim_plan* plan, const std::vector<std::vector<int32_t>>& perms, bool dc = false);
int tpi_build(const csim_plan* plan, const std::vector<int32_t>& perm) {
    return tpi_build(plan, std::vector<std::vector<int32_t>>{perm});
}
int tpi_build(const csim_plan* plan, const std::vector<std::vector<int32_t>>& perms, bool dc) {
    TpiState& tp = dc ? plan->tpi_dc : plan->tpi;
    tp.ready = false;
    tp.wide_tried = tp.wide_ready = false;
    tp.pair_tried = tp.pair_ready = false;
    cudaDeviceSynchronize();  // see spec_build: running launches of the old kernels finish before they are unloaded
    jit_unload(tp.k_wide);
    jit_unload(tp.k_pair);
    if (tp.d_stats) cudaFree(tp.d_stats);
    if (tp.d_miss) cudaFree(tp.d_miss);
    tp.d_stats = tp.d_miss = nullptr;
    std::string log;
    const int rc = tpi_build_variant(plan, perms, 1, tp.host, tp.k, tp.warps_per_sm, log, dc);
    if (rc) {
        tp.log = log;
        return rc;
    }
    CU(cudaDeviceGetAttribute(&tp.n_sms, cudaDevAttrMultiProcessorCount, plan->device));
    CU(cudaMalloc((void**)&tp.d_stats, 2 * sizeof(unsigned long long)));
    CU(cudaMemset(tp.d_stats, 0, 2 * sizeof(unsigned long long)));
    CU(cudaMalloc((void**)&tp.d_miss, sizeof(unsigned long long) * kTpiMissWords));
    CU(cudaMemset(tp.d_miss, 0, sizeof(unsigned long long) * kTpiMissWords));
    tp.log = std::string(dc ? "thread-per-instance DC: " : "thread-per-instance: ") + std::to_string(tp.host.perms.size()) + " predicted permutation(s), " +
             std::to_string(tp.host.n_slots) + " shared doubles per instance, " +
             std::to_string(tp.host.replicas) + " instance(s) per thread, " + std::to_string(tp.host.threads) +
             " threads per CTA, " + std::to_string(tp.warps_per_sm) + " warps resident per SM; per Newton iteration: " +
             std::to_string(tp.host.n_lu_divs) + " multipliers, " + std::to_string(tp.host.n_elim_updates) +
             " row updates, " + std::to_string(tp.host.n_back_terms) + " back-substitution terms, " +
             std::to_string(tp.host.n_g_adds) + "+" + std::to_string(tp.host.n_rhs_adds) + " assembly adds; longest dependency chain " +
             std::to_string(tp.host.crit_cycles) + " cycles" +
             (tp.host.pair ? "; warp pair per 32 instances, " + std::to_string(tp.host.pair_events) +
                                 " hand-overs per Newton iteration, LU cost model " + std::to_string(tp.host.pair_cost[0]) +
                                 " -> " + std::to_string(tp.host.pair_cost[1]) + " / " + std::to_string(tp.host.pair_cost[2])
                           : std::string()) +
             tp.history + (log.size() > 1 ? "\n" + log : "");
    tp.ready = true;
    cache_store(plan, dc, tp.host.perms);
    return CSIM_OK;
}

// A warp of a 2-warp CTA runs ~14 % slower than a warp of a 1-warp CTA (B200, dbmixer: 4,415 ms against 3,874 ms
// for the same 256 warps), so the wide variant is used only when it saves whole waves of CTAs.
constexpr double kTpiWidePenalty = 1.15;

// true -> launch the wide variant for a batch of B instances (builds it on first need)
bool tpi_use_wide(const csim_plan* plan, int64_t B) {
    TpiState& tp = plan->tpi;
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    if (tp.warps_per_sm <= 0 || tp.n_sms <= 0) return false;
    const int64_t warps = (B + 32 * tp.host.replicas - 1) / (32 * tp.host.replicas);
    const int64_t cap = (int64_t)tp.n_sms * tp.warps_per_sm;
    const int64_t waves = (warps + cap - 1) / cap;
    if (waves <= 1) return false;
    if (!tp.wide_tried) {
        tp.wide_tried = true;
        std::string log;
        if (tpi_build_variant(plan, tp.host.perms, 0, tp.host_wide, tp.k_wide, tp.warps_per_sm_wide, log) == CSIM_OK &&
            tp.host_wide.threads > 32 && tp.warps_per_sm_wide > tp.warps_per_sm) {
            tp.wide_ready = true;
            tp.log += "\nwide variant: " + std::to_string(tp.host_wide.threads) + " threads per CTA, " +
                      std::to_string(tp.warps_per_sm_wide) + " warps resident per SM";
        } else {
            jit_unload(tp.k_wide);
        }
    }
    if (!tp.wide_ready) return false;
    const int64_t cap_w = (int64_t)tp.n_sms * tp.warps_per_sm_wide;
    const int64_t waves_w = (warps + cap_w - 1) / cap_w;
    return (double)waves_w * kTpiWidePenalty < (double)waves;
}

// true -> launch the warp-pair variant for a batch of B instances (builds it on first need).  B200, dbmixer, 8,192
// instances (256 of 592 sub-partitions busy with one warp each): 5.21 us per Newton iteration against 6.16 us.  A batch
// beyond one wave of its CTAs is better off with one warp per 32 instances (shared memory holds 2-3 pairs per SM
// against 3-4 single warps).  CSIM_TPI_PAIR=0 turns it off.
bool tpi_use_pair(const csim_plan* plan, int64_t B) {
    TpiState& tp = plan->tpi;
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    if (tp.host.perms.size() != 1 || tp.n_sms <= 0) return false;
    if (const char* e = std::getenv("CSIM_TPI_PAIR"))
        if (std::atoi(e) == 0) return false;
    if (std::getenv("CSIM_TPI_OPTS")) return false;  // (development mask: tp.host itself is what the mask asks for)
    if (!tp.pair_tried) {
        if (B > (int64_t)tp.n_sms * 3 * 32) return false;  // cannot fit whatever the kernel needs: do not build yet
        tp.pair_tried = true;
        std::string log;
        int pairs_per_cta = 2;  // (development aid: CSIM_PAIR_PER_CTA)
        if (const char* e = std::getenv("CSIM_PAIR_PER_CTA")) pairs_per_cta = std::max(1, std::atoi(e));
        if (tpi_build_variant(plan, tp.host.perms, pairs_per_cta, tp.host_pair, tp.k_pair, tp.groups_per_sm_pair, log, false, true) == CSIM_OK &&
            tp.host_pair.pair) {
            tp.pair_ready = true;
            tp.log += "\nwarp-pair variant: " + std::to_string(tp.host_pair.threads) + " threads per CTA for " +
                      std::to_string(tp.host_pair.inst_per_cta) + " instances, " + std::to_string(tp.host_pair.n_slots) +
                      " shared doubles per instance, " + std::to_string(tp.host_pair.pair_events) +
                      " hand-overs per Newton iteration (LU cost model " + std::to_string(tp.host_pair.pair_cost[0]) + " -> " +
                      std::to_string(tp.host_pair.pair_cost[1]) + " / " + std::to_string(tp.host_pair.pair_cost[2]) + "), " +
                      std::to_string(tp.groups_per_sm_pair) + " groups of 32 instances resident per SM";
        } else {
            jit_unload(tp.k_pair);
        }
    }
    if (!tp.pair_ready) return false;
    return B <= (int64_t)tp.n_sms * tp.groups_per_sm_pair * 32;
}

// Pivot order the reference's LU picks for instance 0 of the batch at its current state: one TRAN
// assembly (device stamp kernel) + one generic LU (device).  Only the ORDER is used, as a prediction.
int spec_pilot(const csim_plan* plan, int64_t B, const double* d_params, const double* d_x, double dt,
               cudaStream_t stream, std::vector<int32_t>& perm) {
    const int N = plan->host.N, P = plan->host.n_params, S = plan->host.n_sources;
    const ModeProgram& mp = plan->host.mode[CSIM_MODE_TRAN];
    std::vector<double> p(P > 0 ? P : 1), x(N), sv(S > 0 ? S : 1, 0.0);
    CU(cudaStreamSynchronize(stream));
    if (P > 0) CU(cudaMemcpy2D(p.data(), sizeof(double), d_params, sizeof(double) * B, sizeof(double), P, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy2D(x.data(), sizeof(double), d_x, sizeof(double) * B, sizeof(double), N, cudaMemcpyDeviceToHost));
    std::vector<double> vals(mp.colidx.size()), rhs(N);
    g_generic_stamp_only = true;
    int rc = csim_host_stamp_csr(plan, CSIM_MODE_TRAN, 1, p.data(), x.data(), x.data(), sv.data(), dt, 1.0,
                                 plan->host.consts.tran_gmin, vals.data(), rhs.data());
    g_generic_stamp_only = false;
    if (rc) return rc;
    std::vector<double> A((size_t)N * N, 0.0), sol(N);
    for (int i = 0; i < N; ++i)
        for (int k = mp.rowptr[i]; k < mp.rowptr[i + 1]; ++k) A[(size_t)i * N + mp.colidx[k]] = vals[k];
    perm.assign(N, 0);
    int32_t singular = 0;
    rc = csim_host_lu_solve(N, 1, A.data(), rhs.data(), sol.data(), perm.data(), &singular, plan->host.consts.lu_eps);
    if (rc) return rc;
    if (singular) return fail(CSIM_E_UNSUPPORTED, "pilot factorisation is singular");
    return CSIM_OK;
}

// In automatic mode the thread-per-instance kernel is kept only while its predicted pivot order holds for
// (almost) every solve: a lane that misses redoes the solve alone while its 31 neighbours wait.
constexpr double kTpiMaxMissRate = 0.02;
constexpr unsigned long long kTpiMinSolves = 20000;
// Before giving the kernel up, further predicted permutations are added (one per re-evaluation, the one the
// redo path met most often) while more than kTpiAdaptMissRate of the solves miss.
constexpr double kTpiAdaptMissRate = 0.002;
constexpr int kTpiMaxVariants = 16;  // (TPI_CAND masks are 16 bits wide)
// How many variants the automatic mode collects by itself: every variant is one more straight-line body the lanes of
// a warp may be spread over.  Transient (lanes stay in one operating region for many solves): 8 (measured optimum on
// buffer.sp); DC ramp (every lane crosses every region once): 16.  CSIM_TPI_MAX_VARIANTS / CSIM_DC_MAX_VARIANTS override.
int auto_variant_cap(bool dc) {
    const char* env = std::getenv(dc ? "CSIM_DC_MAX_VARIANTS" : "CSIM_TPI_MAX_VARIANTS");
    const int v = env ? std::atoi(env) : (dc ? 16 : 8);
    return v < 1 ? 1 : (v > kTpiMaxVariants ? kTpiMaxVariants : v);
}

// Kernel to use for this launch: 0 generic, 2 warp-per-instance, 3 thread-per-instance (compiled on first use).
int spec_prepare(const csim_plan* plan, int64_t B, const double* d_params, const double* d_x, double dt,
                 cudaStream_t stream) {
    SpecState& sp = plan->spec;
    TpiState& tp = plan->tpi;
    std::lock_guard<std::recursive_mutex> lock(sp.mu);
    if (sp.mode == 0) return 0;
    if ((sp.mode == 1 || sp.mode == 3) && sp.kind == 3 && tp.ready && tp.d_stats) {  // re-evaluate the choice
        unsigned long long st[2] = {0, 0};
        cudaStreamSynchronize(stream);
        const bool have = cudaMemcpy(st, tp.d_stats, sizeof(st), cudaMemcpyDeviceToHost) == cudaSuccess &&
                          st[1] >= kTpiMinSolves;
        // first remedy: one more predicted permutation — the one the redo path met most often
        if (have && (double)st[0] > kTpiAdaptMissRate * (double)st[1] && tp.d_miss &&
            (int)tp.host.perms.size() < auto_variant_cap(false)) {
            std::vector<unsigned long long> tab((size_t)kTpiMissEntries * kTpiMissStride, 0);
            int best = -1;
            if (cudaMemcpy(tab.data(), tp.d_miss, tab.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) ==
                cudaSuccess)
                for (int e = 0; e < kTpiMissEntries; ++e) {
                    const unsigned long long hsh = tab[(size_t)e * kTpiMissStride], cnt = tab[(size_t)e * kTpiMissStride + 1];
                    if (!hsh || std::find(tp.host.hashes.begin(), tp.host.hashes.end(), hsh) != tp.host.hashes.end()) continue;
                    if (best < 0 || cnt > tab[(size_t)best * kTpiMissStride + 1]) best = e;
                }
            if (best >= 0 && tab[(size_t)best * kTpiMissStride + 1] * 4 >= st[0] / 10) {  // worth a variant: >= 2.5 % of the misses
                std::vector<std::vector<int32_t>> perms = tp.host.perms;
                std::vector<int32_t> np(plan->host.N);
                for (int i = 0; i < plan->host.N; ++i) np[i] = (int32_t)tab[(size_t)best * kTpiMissStride + 2 + i];
                perms.push_back(np);
                const std::string note = "\n" + std::to_string(st[0]) + " of " + std::to_string(st[1]) +
                                         " solves missed the predicted pivot order -> variant " +
                                         std::to_string(perms.size() - 1) + " added (met " +
                                         std::to_string(tab[(size_t)best * kTpiMissStride + 1]) + " times)";
                const std::vector<std::vector<int32_t>> old = tp.host.perms;
                const std::string hist = tp.history;
                tp.history += note;
                if (tpi_build(plan, perms) != CSIM_OK) {
                    tp.history = hist;
                    if (tpi_build(plan, old) != CSIM_OK) sp.kind = 0;
                }
                st[0] = st[1] = 0;  // statistics restart with the new kernel
            }
        }
        if (sp.mode == 1 && sp.kind == 3 && have && st[1] >= kTpiMinSolves && (double)st[0] > kTpiMaxMissRate * (double)st[1] &&
            (int)tp.host.perms.size() >= auto_variant_cap(false)) {
            tp.log += "\nautomatic mode: " + std::to_string(st[0]) + " of " + std::to_string(st[1]) +
                      " solves missed the predicted pivot order -> warp-per-instance kernel";
            sp.kind = 2;
            if (!sp.ready && spec_build(plan, tp.host.perm) != CSIM_OK) sp.kind = 0;
            sp.tried = true;
        }
    }
    if (sp.kind == 3 && tp.ready) return 3;
    if (sp.kind == 2 && sp.ready) return 2;
    if (sp.tried) return 0;
    sp.tried = true;
    int want = sp.mode == 2 ? 2 : 3;
    const char* env = std::getenv("CSIM_JIT");
    if (env && env[0] == '0') {
        sp.mode = 0;
        sp.log = "disabled by CSIM_JIT=0";
        return 0;
    }
    if (env && (env[0] == '2' || env[0] == '3')) {
        want = env[0] - '0';
        sp.mode = want;
    }
    if (!spec_unsupported_reason(plan->host).empty()) {
        sp.log = "not specialised: " + spec_unsupported_reason(plan->host);
        return 0;
    }
    if (want == 3) {
        std::vector<std::vector<int32_t>> cached;
        if (cache_load(plan, false, cached) && tpi_build(plan, cached) == CSIM_OK) {
            tp.history += "\n" + std::to_string(cached.size()) + " variant(s) from the plan cache";
            return sp.kind = 3;
        }
    }
    std::vector<int32_t> perm;
    if (spec_pilot(plan, B, d_params, d_x, dt, stream, perm) != CSIM_OK) {
        sp.log = std::string("pilot failed: ") + csim_last_error();
        return 0;
    }
    if (want == 3 && tpi_build(plan, perm) == CSIM_OK) return sp.kind = 3;
    if (spec_build(plan, perm) == CSIM_OK) return sp.kind = 2;
    return 0;
}

// ---- thread-per-instance DC kernel (csim_dc_tpi) ------------------------------------------------------------------
// Pivot order of instance 0's FIRST DC solve (x = 0, first ramp step, base gmin): a prediction to start from; the
// ramp then walks through the circuit's operating regions and the redo log teaches the other orders.
int dc_pilot(const csim_plan* plan, int64_t B, const double* d_params, cudaStream_t stream, std::vector<int32_t>& perm) {
    const int N = plan->host.N, P = plan->host.n_params;
    const ModeProgram& mp = plan->host.mode[CSIM_MODE_DC];
    const csim_constants& c = plan->host.consts;
    std::vector<double> p(P > 0 ? P : 1), x(N, 0.0);
    CU(cudaStreamSynchronize(stream));
    if (P > 0) CU(cudaMemcpy2D(p.data(), sizeof(double), d_params, sizeof(double) * B, sizeof(double), P, cudaMemcpyDeviceToHost));
    const double scale = 1.0 / (double)(c.dc_ramp_steps > 0 ? c.dc_ramp_steps : 1);
    const double gmin = c.gmin_high_base * (1.0 - scale) + c.gmin_low_base * scale;
    std::vector<double> vals(mp.colidx.size()), rhs(N);
    g_generic_stamp_only = true;
    int rc = csim_host_stamp_csr(plan, CSIM_MODE_DC, 1, p.data(), x.data(), nullptr, nullptr, 0.0, scale, gmin, vals.data(),
                                 rhs.data());
    g_generic_stamp_only = false;
    if (rc) return rc;
    std::vector<double> A((size_t)N * N, 0.0), sol(N);
    for (int i = 0; i < N; ++i)
        for (int k = mp.rowptr[i]; k < mp.rowptr[i + 1]; ++k) A[(size_t)i * N + mp.colidx[k]] = vals[k];
    perm.assign(N, 0);
    int32_t singular = 0;
    rc = csim_host_lu_solve(N, 1, A.data(), rhs.data(), sol.data(), perm.data(), &singular, c.lu_eps);
    if (rc) return rc;
    if (singular) return fail(CSIM_E_UNSUPPORTED, "DC pilot factorisation is singular");
    return CSIM_OK;
}

// true -> launch csim_dc_tpi (built on first use; between launches one more predicted permutation is added while
// more than kTpiAdaptMissRate of the solves miss, the rule of the transient kernel).  Caller holds spec.mu.
bool dc_prepare(const csim_plan* plan, int64_t B, const double* d_params, cudaStream_t stream) {
    SpecState& sp = plan->spec;
    TpiState& tp = plan->tpi_dc;
    if (sp.mode == 0 || sp.mode == 2 || !plan->host.has_mos) return false;
    if (const char* env = std::getenv("CSIM_JIT"))
        if (env[0] == '0' || env[0] == '2') return false;
    if (const char* env = std::getenv("CSIM_DC_TPI"))
        if (env[0] == '0') return false;
    if (!plan->dc_tried) {
        plan->dc_tried = true;
        if (!spec_unsupported_reason(plan->host).empty()) return false;
        std::vector<std::vector<int32_t>> cached;
        if (cache_load(plan, true, cached) && tpi_build(plan, cached, true) == CSIM_OK) {
            tp.history += "\n" + std::to_string(cached.size()) + " variant(s) from the plan cache";
            return true;
        }
        std::vector<int32_t> perm;
        if (dc_pilot(plan, B, d_params, stream, perm) != CSIM_OK) {
            tp.log = std::string("DC pilot failed: ") + csim_last_error();
            return false;
        }
        if (tpi_build(plan, std::vector<std::vector<int32_t>>{perm}, true) != CSIM_OK) return false;
        return true;
    }
    if (!tp.ready) return false;
    unsigned long long st[2] = {0, 0};
    cudaStreamSynchronize(stream);
    const bool have = tp.d_stats && cudaMemcpy(st, tp.d_stats, sizeof(st), cudaMemcpyDeviceToHost) == cudaSuccess &&
                      st[1] >= kTpiMinSolves;
    if (have && (double)st[0] > kTpiAdaptMissRate * (double)st[1] && tp.d_miss && (int)tp.host.perms.size() < auto_variant_cap(true)) {
        std::vector<unsigned long long> tab((size_t)kTpiMissEntries * kTpiMissStride, 0);
        int best = -1;
        if (cudaMemcpy(tab.data(), tp.d_miss, tab.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess)
            for (int e = 0; e < kTpiMissEntries; ++e) {
                const unsigned long long hsh = tab[(size_t)e * kTpiMissStride], cnt = tab[(size_t)e * kTpiMissStride + 1];
                if (!hsh || std::find(tp.host.hashes.begin(), tp.host.hashes.end(), hsh) != tp.host.hashes.end()) continue;
                if (best < 0 || cnt > tab[(size_t)best * kTpiMissStride + 1]) best = e;
            }
        if (best >= 0 && tab[(size_t)best * kTpiMissStride + 1] * 4 >= st[0] / 10) {
            // every logged order worth a variant (>= 2.5 % of the misses), most frequent first: the ramp crosses many
            // operating regions, and one variant per launch would take a dozen launches (and compilations) to settle
            std::vector<int> order;
            for (int e = 0; e < kTpiMissEntries; ++e) {
                const unsigned long long hsh = tab[(size_t)e * kTpiMissStride], cnt = tab[(size_t)e * kTpiMissStride + 1];
                if (!hsh || cnt * 4 < st[0] / 10) continue;
                if (std::find(tp.host.hashes.begin(), tp.host.hashes.end(), hsh) != tp.host.hashes.end()) continue;
                order.push_back(e);
            }
            std::sort(order.begin(), order.end(), [&](int a, int b) {
                return tab[(size_t)a * kTpiMissStride + 1] > tab[(size_t)b * kTpiMissStride + 1];
            });
            std::vector<std::vector<int32_t>> perms = tp.host.perms;
            for (int e : order) {
                if ((int)perms.size() >= auto_variant_cap(true)) break;
                std::vector<int32_t> np(plan->host.N);
                for (int i = 0; i < plan->host.N; ++i) np[i] = (int32_t)tab[(size_t)e * kTpiMissStride + 2 + i];
                perms.push_back(np);
            }
            const std::vector<std::vector<int32_t>> old = tp.host.perms;
            const std::string hist = tp.history;
            tp.history += "\n" + std::to_string(st[0]) + " of " + std::to_string(st[1]) +
                          " DC solves missed the predicted pivot order -> " + std::to_string(perms.size() - old.size()) +
                          " variant(s) added (" + std::to_string(perms.size()) + " now)";
            if (tpi_build(plan, perms, true) != CSIM_OK) {
                tp.history = hist;
                if (tpi_build(plan, old, true) != CSIM_OK) return false;
            }
        }
    }
    return tp.ready;
}

cudaError_t dc_tpi_launch(const csim_plan* plan, int64_t B, const double* params, double* x, int32_t* iters,
                          uint64_t* pivot_hash, uint32_t* status, cudaStream_t stream) {
    TpiState& tp = plan->tpi_dc;
    DevPlan P = plan->dev;
    DevTran Tc{};
    Tc.miss = tp.d_miss;
    long long Bll = B;
    double* wave = nullptr;
    uint8_t* iters_step = nullptr;
    double* scratch = nullptr;
    unsigned long long* stats = tp.d_stats;
    void* args[] = {&P, &Tc, &Bll, &params, &x, &wave, &iters, &iters_step, &pivot_hash, &status, &scratch, &stats};
    const TpiHost& H = tp.host;
    const int64_t per_cta = H.inst_per_cta;
    const unsigned blocks = (unsigned)((B + per_cta - 1) / per_cta);
    return cudaLaunchKernel(reinterpret_cast<const void*>(tp.k.kernel), dim3(blocks), dim3(H.threads), args, H.smem_bytes,
                            stream);
}

cudaError_t spec_launch(const csim_plan* plan, const DevTran& T, int64_t B, const double* params, double* x,
                        double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash,
                        uint32_t* status, cudaStream_t stream) {
    SpecState& sp = plan->spec;
    DevPlan P = plan->dev;
    DevTran Tc = T;
    SpecTables ST = sp.dev;
    long long Bll = B;
    unsigned long long* stats = sp.d_stats;
    void* args[] = {&P, &Tc, &ST, &Bll, &params, &x, &wave, &iters_total, &iters_step, &pivot_hash, &status, &stats};
    const int wpb = sp.host.warps_per_block;
    const unsigned blocks = (unsigned)((B + wpb - 1) / wpb);
    return cudaLaunchKernel(reinterpret_cast<const void*>(sp.k.kernel), dim3(blocks), dim3(wpb * 32), args, sp.smem,
                            stream);
}

// true -> the specialised assembly kernel of `mode` is loaded (compiled on first use)
bool stamp_spec_prepare(const csim_plan* plan, int mode) {
    SpecState& sp = plan->spec;
    std::lock_guard<std::recursive_mutex> lock(sp.mu);
    StampSpecState& ss = plan->stamp_spec[mode];
    if (sp.mode == 0) return false;
    if (ss.ready) return true;
    if (ss.tried) return false;
    ss.tried = true;
    const char* env = std::getenv("CSIM_JIT");
    if (env && env[0] == '0') {
        ss.log = "disabled by CSIM_JIT=0";
        return false;
    }
    try {
        ss.host = generate_stamp_spec(plan->host, mode);
    } catch (const std::exception& e) {
        ss.log = e.what();
        return false;
    }
    std::vector<char> cubin;
    std::string log;
    if (!jit_compile(ss.host.source, "csim_stamp_spec.cu", jit_library_dir() + "/csrc", cubin, log)) {
        ss.log = "NVRTC: " + log;
        return false;
    }
    if (jit_load(cubin, "csim_stamp_spec", ss.k) != cudaSuccess) {
        ss.log = "cudaLibraryLoadData failed";
        cudaGetLastError();
        return false;
    }
    ss.log = "specialised" + (log.size() > 1 ? "\n" + log : std::string());
    ss.ready = true;
    return true;
}

cudaError_t stamp_spec_launch(const csim_plan* plan, int mode, int64_t B, const double* params, const double* x,
                              const double* x_prev, const double* src_val, double dt, double scale, double gmin,
                              double* vals, double* rhs, cudaStream_t stream) {
    StampSpecState& ss = plan->stamp_spec[mode];
    long long Bll = B;
    double off = plan->host.consts.mos_off_gds;
    void* args[] = {&Bll, &params, &x, &x_prev, &src_val, &dt, &scale, &gmin, &off, &vals, &rhs};
    const int T = ss.host.threads;
    const unsigned blocks = (unsigned)((B + T - 1) / T);
    return cudaLaunchKernel(reinterpret_cast<const void*>(ss.k.kernel), dim3(blocks), dim3(T), args, 0, stream);
}

cudaError_t tpi_launch(const csim_plan* plan, const DevTran& T, int64_t B, const double* params, double* x,
                       double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                       cudaStream_t stream) {
    TpiState& tp = plan->tpi;
    DevPlan P = plan->dev;
    DevTran Tc = T;
    long long Bll = B;
    double* scratch = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&scratch, sizeof(double) * (size_t)tp.host.n_scratch_rows * B, stream);
    if (e != cudaSuccess) return e;
    unsigned long long* stats = tp.d_stats;
    Tc.miss = tp.d_miss;
    void* args[] = {&P, &Tc, &Bll, &params, &x, &wave, &iters_total, &iters_step, &pivot_hash, &status, &scratch, &stats};
    const bool pair = tpi_use_pair(plan, B);
    const bool wide = !pair && tpi_use_wide(plan, B);
    const TpiHost& H = pair ? tp.host_pair : wide ? tp.host_wide : tp.host;
    const int T_ = H.threads;
    const int64_t per_cta = H.inst_per_cta;
    const unsigned blocks = (unsigned)((B + per_cta - 1) / per_cta);
    e = cudaLaunchKernel(reinterpret_cast<const void*>((pair ? tp.k_pair : wide ? tp.k_wide : tp.k).kernel), dim3(blocks),
                         dim3(T_), args, H.smem_bytes, stream);
    cudaFreeAsync(scratch, stream);
    return e;
}

}  // namespace

extern "C" {

const char* csim_last_error(void) { return g_err.c_str(); }
void csim_set_last_error(const char* msg) { g_err = msg ? msg : ""; }  // (multi.cpp reports through the same slot)
int csim_abi_version(void) { return CSIM_ABI_VERSION; }

void csim_default_constants(csim_constants* c) {
    if (!c) return;
    std::memset(c, 0, sizeof(*c));
    c->dc_ramp_steps = 10;
    c->dc_max_iters = 50;
    c->dc_tol = 1e-9;
    c->dc_alpha = 0.35;
    c->gmin_high_base = 1e-6;
    c->gmin_low_base = 3.35e-7;
    c->gmin_abs_max = 1e-4;
    c->fast_conv_ratio = 0.7;
    c->slow_conv_ratio = 1.05;
    c->dc_nonfinite_mul = 10.0;
    c->dc_nonfinite_cap = 1e-2;
    c->tran_max_iters = 50;
    c->tran_tol = 1e-6;
    c->tran_gmin = 1e-6;
    c->tran_alpha = 0.45;
    c->lu_eps = 1e-15;
    c->mos_off_gds = 1e-12;
}

int csim_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    return n;
}

uint64_t csim_pivot_hash_update(uint64_t h, const int32_t* perm, int32_t n) {
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
    return (h ^ (((uint64_t)fhi << 32) | flo)) * 0x100000001B3ull;
}

int csim_plan_create(const csim_topology* topo, const csim_constants* consts, int device, csim_plan** out) {
    if (!topo || !out) return fail(CSIM_E_INVALID, "csim_plan_create: null argument");
    *out = nullptr;
    std::unique_ptr<csim_plan> p(new csim_plan());
    try {
        p->host = build_host_plan(*topo, consts);
    } catch (const std::exception& e) {
        return fail(CSIM_E_INVALID, std::string("csim_plan_create: ") + e.what());
    }
    {
        uint64_t hsh = 0xCBF29CE484222325ull;
        auto mixin = [&](const void* data, size_t bytes) {
            const unsigned char* c = static_cast<const unsigned char*>(data);
            for (size_t i = 0; i < bytes; ++i) hsh = (hsh ^ c[i]) * 0x100000001B3ull;
        };
        const int32_t dims[4] = {topo->n_unknowns, topo->n_node_eq, topo->n_elems, topo->n_params};
        mixin(dims, sizeof dims);
        mixin(topo->elem_type, sizeof(int32_t) * topo->n_elems);
        mixin(topo->elem_flags, sizeof(int32_t) * topo->n_elems);
        mixin(topo->elem_nodes, sizeof(int32_t) * 4 * topo->n_elems);
        mixin(topo->elem_branch, sizeof(int32_t) * topo->n_elems);
        mixin(topo->elem_param, sizeof(int32_t) * topo->n_elems);
        mixin(&p->host.consts, sizeof(csim_constants));
        p->topo_hash = hsh ? hsh : 1;
    }
    p->device = device;
    if (device >= 0) {
        CU(cudaSetDevice(device));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, device));
        p->sm_count = prop.multiProcessorCount;
        DevPlan& d = p->dev;
        d.N = p->host.N;
        d.n_node_eq = p->host.n_node_eq;
        d.n_params = p->host.n_params;
        d.n_sources = p->host.n_sources;
        d.has_mos = p->host.has_mos ? 1 : 0;
        d.c = p->host.consts;
        for (int m = 0; m < 2; ++m) {
            const ModeProgram& mp = p->host.mode[m];
            DevMode& dm = d.m[m];
            dm.n_values = mp.n_values;
            dm.n_static = (int)mp.static_ops.size();
            dm.n_step = (int)mp.step_ops.size();
            dm.n_src = (int)mp.src_ops.size();
            dm.n_mos = (int)mp.mos_ops.size();
            dm.row_len = mp.row_len_max;
            dm.nnz = (int)mp.colidx.size();
            dm.n_slots = (int)mp.slot_ptr.size() - 1;
            CU(upload(mp.static_ops, &dm.static_ops, p->owned));
            CU(upload(mp.step_ops, &dm.step_ops, p->owned));
            CU(upload(mp.src_ops, &dm.src_ops, p->owned));
            CU(upload(mp.mos_ops, &dm.mos_ops, p->owned));
            CU(upload(mp.row_prog, &dm.row_prog, p->owned));
            CU(upload(mp.slot_ptr, &dm.slot_ptr, p->owned));
            CU(upload(mp.slot_entries, &dm.slot_entries, p->owned));
            std::vector<uint32_t> cw;
            std::vector<int32_t> cr;
            if (p->host.N <= 32) for (const Contribution& c : mp.contribs) {
                cw.push_back(contrib_word(c.col >= 0 ? c.col : kRhsCol, c.sub != 0, c.value));
                cr.push_back(c.row);
            }
            dm.n_contribs = (int)cw.size();
            CU(upload(cw, &dm.contrib_words, p->owned));
            CU(upload(cr, &dm.contrib_rows, p->owned));
            std::vector<int32_t> srow, scol;
            for (int i = 0; i < p->host.N; ++i)
                for (int k = mp.rowptr[i]; k < mp.rowptr[i + 1]; ++k) {
                    srow.push_back(i);
                    scol.push_back(mp.colidx[k]);
                }
            for (int i = 0; i < p->host.N; ++i) {
                srow.push_back(i);
                scol.push_back(-1);
            }
            CU(upload(srow, &dm.slot_row, p->owned));
            CU(upload(scol, &dm.slot_col, p->owned));
        }
    }
    *out = p.release();
    return CSIM_OK;
}

void csim_plan_destroy(csim_plan* plan) {
    if (!plan) return;
    if (plan->device >= 0) {
        cudaSetDevice(plan->device);
        for (void* d : plan->owned) cudaFree(d);
        for (void* d : plan->spec.owned) cudaFree(d);
        jit_unload(plan->spec.k);
        jit_unload(plan->tpi.k);
        jit_unload(plan->tpi.k_wide);
        jit_unload(plan->tpi.k_pair);
        if (plan->tpi.d_stats) cudaFree(plan->tpi.d_stats);
        if (plan->tpi.d_miss) cudaFree(plan->tpi.d_miss);
        jit_unload(plan->tpi_dc.k);
        if (plan->tpi_dc.d_stats) cudaFree(plan->tpi_dc.d_stats);
        if (plan->tpi_dc.d_miss) cudaFree(plan->tpi_dc.d_miss);
        jit_unload(plan->stamp_spec[0].k);
        jit_unload(plan->stamp_spec[1].k);
    }
    delete plan;
}

int csim_plan_pattern(const csim_plan* plan, int mode, int32_t* nnz, int32_t* rowptr, int32_t* colidx) {
    if (!plan || mode < 0 || mode > 1) return fail(CSIM_E_INVALID, "csim_plan_pattern: bad argument");
    const ModeProgram& mp = plan->host.mode[mode];
    if (nnz) *nnz = (int32_t)mp.colidx.size();
    if (rowptr) std::memcpy(rowptr, mp.rowptr.data(), mp.rowptr.size() * sizeof(int32_t));
    if (colidx) std::memcpy(colidx, mp.colidx.data(), mp.colidx.size() * sizeof(int32_t));
    return CSIM_OK;
}

int csim_plan_dims(const csim_plan* plan, int32_t* n_unknowns, int32_t* n_params, int32_t* n_sources) {
    if (!plan) return fail(CSIM_E_INVALID, "csim_plan_dims: null plan");
    if (n_unknowns) *n_unknowns = plan->host.N;
    if (n_params) *n_params = plan->host.n_params;
    if (n_sources) *n_sources = plan->host.n_sources;
    return CSIM_OK;
}

// ------------------------------------------------------------------ device API
static int need_device(const csim_plan* plan, const char* who) {
    if (!plan) return fail(CSIM_E_INVALID, std::string(who) + ": null plan");
    if (plan->device < 0) return fail(CSIM_E_CUDA, std::string(who) + ": plan was created without a CUDA device (no CPU fallback)");
    cudaError_t e = cudaSetDevice(plan->device);
    if (e != cudaSuccess) return cuda_fail(e, "cudaSetDevice");
    return CSIM_OK;
}
// N <= 32: warp / thread-per-instance kernels with the matrix on chip; larger systems (or force_big, a test
// switch) take the large-matrix path with the matrix in HBM/L2: the blocked dense kernels of dense_kernels.cu
// (one CTA per instance; the default) or the sparse-skipping kernels of big_kernels.cu (one warp per instance;
// CSIM_BIG_KIND=warp, force_big == 2, or N > kDenseMaxN).
static bool env_big() {
    const char* e = std::getenv("CSIM_BIG");
    return e && e[0] == '1';
}
static bool env_big_warp() {
    const char* e = std::getenv("CSIM_BIG_KIND");
    return e && e[0] == 'w';
}
static bool use_big(const csim_plan* plan) { return plan->host.N > 32 || plan->force_big || env_big(); }
static bool use_dense(const csim_plan* plan) {
    return plan->force_big != 2 && !env_big_warp() && plan->host.N <= kDenseMaxN;
}
static int need_fits(const csim_plan* plan, const char* who) {
    if (plan->host.N > kBigMaxN)
        return fail(CSIM_E_UNSUPPORTED, std::string(who) + ": more than " + std::to_string(kBigMaxN) + " unknowns");
    return CSIM_OK;
}

int csim_dev_stamp_csr(const csim_plan* plan, int mode, int64_t B, const double* params, const double* x,
                       const double* x_prev, const double* src_val, double dt, double source_scale, double gmin,
                       double* vals, double* rhs, void* stream) {
    int rc = need_device(plan, "csim_dev_stamp_csr");
    if (rc) return rc;
    if (mode < 0 || mode > 1 || B <= 0 || !params || !x || !vals || !rhs || (mode == CSIM_MODE_TRAN && !x_prev))
        return fail(CSIM_E_INVALID, "csim_dev_stamp_csr: bad argument");
    if (!g_generic_stamp_only && stamp_spec_prepare(plan, mode))
        CU(stamp_spec_launch(plan, mode, B, params, x, x_prev, src_val, dt, source_scale, gmin, vals, rhs,
                             (cudaStream_t)stream));
    else
        CU(launch_stamp_csr(plan->dev, mode, B, params, x, x_prev, src_val, dt, source_scale, gmin, vals, rhs,
                            plan->sm_count, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_dev_lu_solve(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                      int32_t* singular, double lu_eps, void* stream) {
    if (n <= 0 || n > kBigMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_dev_lu_solve: n out of range");
    if (B <= 0 || !A || !b || !x) return fail(CSIM_E_INVALID, "csim_dev_lu_solve: bad argument");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(CSIM_E_CUDA, "csim_dev_lu_solve: no CUDA device (no CPU fallback)");
    if ((n > 32 || env_big()) && n <= kDenseMaxN && !env_big_warp())
        CU(launch_lu_dense(n, B, A, b, x, perm, singular, lu_eps, (cudaStream_t)stream));
    else if (n > 32 || env_big()) CU(launch_lu_big(n, B, A, b, x, perm, singular, lu_eps, (cudaStream_t)stream));
    else CU(launch_lu_solve(n, B, A, b, x, perm, singular, lu_eps, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_dev_lu_solve_noalloc(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                              int32_t* singular, double lu_eps, void* stream) {
    const int max_n = lu_dense_noalloc_max_n();
    if (n <= 0 || n > max_n)
        return fail(CSIM_E_UNSUPPORTED, "csim_dev_lu_solve_noalloc: n = " + std::to_string(n) + " is outside 1.." +
                                            std::to_string(max_n) + " (one [A | b] workspace must fit the " +
                                            std::to_string(kDnNoAllocDoubles * sizeof(double) >> 20) +
                                            " MiB static scratch array)");
    if (B <= 0 || !A || !b || !x) return fail(CSIM_E_INVALID, "csim_dev_lu_solve_noalloc: bad argument");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(CSIM_E_CUDA, "csim_dev_lu_solve_noalloc: no CUDA device (no CPU fallback)");
    CU(launch_lu_dense_noalloc(n, B, A, b, x, perm, singular, lu_eps, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_dev_lu_factor(int32_t n, int64_t B, const double* A, double* LU, int32_t* perm, int32_t* singular,
                       double lu_eps, void* stream) {
    if (n <= 0 || n > kDenseMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_dev_lu_factor: n out of range");
    if (B <= 0 || !A || !LU) return fail(CSIM_E_INVALID, "csim_dev_lu_factor: bad argument");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(CSIM_E_CUDA, "csim_dev_lu_factor: no CUDA device (no CPU fallback)");
    if (n > 32 || env_big())
        CU(launch_lu_dense(n, B, A, nullptr, nullptr, perm, singular, lu_eps, (cudaStream_t)stream, LU));
    else
        CU(launch_lu_solve(n, B, A, nullptr, nullptr, perm, singular, lu_eps, (cudaStream_t)stream, LU));
    return CSIM_OK;
}

int csim_dev_dc(const csim_plan* plan, int64_t B, const double* params, double* x, int32_t* iters,
                uint64_t* pivot_hash, uint32_t* status, void* stream) {
    int rc = need_device(plan, "csim_dev_dc");
    if (rc) return rc;
    if ((rc = need_fits(plan, "csim_dev_dc"))) return rc;
    if (B <= 0 || !params || !x) return fail(CSIM_E_INVALID, "csim_dev_dc: bad argument");
    if (use_big(plan) && use_dense(plan))
        CU(launch_dc_dense(plan->dev, B, params, x, iters, pivot_hash, status, (cudaStream_t)stream));
    else if (use_big(plan)) CU(launch_dc_big(plan->dev, B, params, x, iters, pivot_hash, status, (cudaStream_t)stream));
    else if (g_margin) {
        CU(launch_dc(plan->dev, B, params, x, iters, pivot_hash, status, (cudaStream_t)stream, g_margin));
    } else {
        // (kernel choice and launch under the plan's lock, as in csim_dev_tran)
        std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
        if (dc_prepare(plan, B, params, (cudaStream_t)stream))
            CU(dc_tpi_launch(plan, B, params, x, iters, pivot_hash, status, (cudaStream_t)stream));
        else
            CU(launch_dc(plan->dev, B, params, x, iters, pivot_hash, status, (cudaStream_t)stream));
    }
    return CSIM_OK;
}

int csim_dev_tran(const csim_plan* plan, int64_t B, const double* params, const csim_tran_args* a, double* x,
                  double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status,
                  void* stream_v) {
    int rc = need_device(plan, "csim_dev_tran");
    if (rc) return rc;
    if ((rc = need_fits(plan, "csim_dev_tran"))) return rc;
    const bool big = use_big(plan);
    if (B <= 0 || !params || !a || !x) return fail(CSIM_E_INVALID, "csim_dev_tran: bad argument");
    const int ns = plan->host.n_sources;
    if (a->dt <= 0.0 || a->step_begin < 0 || a->step_end < a->step_begin)
        return fail(CSIM_E_INVALID, "csim_dev_tran: bad step range or dt");
    if (a->n_probes < 0 || (!big && a->n_probes > 32) || (a->n_probes > 0 && (!a->probes || !wave || a->out_stride <= 0)))
        return fail(CSIM_E_INVALID, "csim_dev_tran: bad probe specification (at most 32 probes for N <= 32)");
    if (ns > 0 && (!a->src_mode || (!a->src_group && !a->src_group_b)))
        return fail(CSIM_E_INVALID, "csim_dev_tran: source modes/groups missing");
    bool uses_table = false;
    for (int s = 0; s < ns; ++s) {
        if (a->src_mode[s] < 0 || a->src_mode[s] > CSIM_SRC_VAL_TABLE)
            return fail(CSIM_E_INVALID, "csim_dev_tran: bad src_mode");
        if (a->src_mode[s] != CSIM_SRC_CONST) {
            uses_table = true;
            if (a->src_group && (a->src_group[s] < 0 || a->src_group[s] >= a->n_groups))
                return fail(CSIM_E_INVALID, "csim_dev_tran: src_group out of range");
        }
    }
    if (uses_table && (!a->src_table || a->table_steps < a->step_end))
        return fail(CSIM_E_INVALID, "csim_dev_tran: source table missing or shorter than step_end");
    for (int p = 0; p < a->n_probes; ++p)
        if (a->probes[p] < 0 || a->probes[p] >= plan->host.N) return fail(CSIM_E_INVALID, "csim_dev_tran: probe out of range");
    if (a->n_probes > 0) {
        int64_t recorded = (int64_t)a->step_end / a->out_stride - (int64_t)a->out_begin_step / a->out_stride;
        if (recorded > a->wave_ld) return fail(CSIM_E_INVALID, "csim_dev_tran: wave_ld too small");
    }
    cudaStream_t stream = (cudaStream_t)stream_v;
    // small host-side descriptors -> stream-ordered device scratch
    std::vector<int32_t> pack((size_t)2 * ns + 32 + (a->n_probes > 0 ? a->n_probes : 0), 0);
    for (int s = 0; s < ns; ++s) {
        pack[s] = a->src_mode[s];
        pack[ns + s] = a->src_group ? a->src_group[s] : 0;
    }
    for (int p = 0; p < a->n_probes; ++p) pack[2 * ns + p] = a->probes[p];
    int32_t* d_pack = nullptr;
    CU(cudaMallocAsync((void**)&d_pack, pack.size() * sizeof(int32_t), stream));
    CU(cudaMemcpyAsync(d_pack, pack.data(), pack.size() * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
    CU(cudaStreamSynchronize(stream));  // `pack` is a pageable temporary
    DevTran T{};
    T.dt = a->dt;
    T.step_begin = a->step_begin;
    T.step_end = a->step_end;
    T.table_steps = a->table_steps;
    T.n_probes = a->n_probes;
    T.out_stride = a->out_stride > 0 ? a->out_stride : 1;
    T.out_begin_step = a->out_begin_step;
    T.wave_ld = a->wave_ld;
    T.src_table = a->src_table;
    T.src_mode = d_pack;
    T.src_group = d_pack + ns;
    T.src_group_b = a->src_group_b;
    T.probes = d_pack + 2 * ns;
    T.skip = g_tran_skip;
    T.margin = g_margin;
    cudaError_t e;
    // The learned kernel-selection state of the plan (which kernel, its module, statistics buffers) is guarded by
    // spec.mu from the choice to the launch: another host thread sharing the plan cannot rebuild or unload a kernel
    // between this thread's spec_prepare and its cudaLaunchKernel (launches themselves are asynchronous).
    std::unique_lock<std::recursive_mutex> plan_lock(plan->spec.mu, std::defer_lock);
    if (!big) plan_lock.lock();
    const int kind = big ? -1 : (g_margin ? 0 : spec_prepare(plan, B, params, x, a->dt, stream));
    if (big && use_dense(plan)) {
        e = launch_tran_dense(plan->dev, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, stream);
    } else if (big) {
        e = launch_tran_big(plan->dev, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, stream);
    } else if (kind == 3) {
        e = tpi_launch(plan, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, stream);
    } else if (kind == 2) {
        e = spec_launch(plan, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, stream);
    } else {
        e = launch_tran(plan->dev, T, B, params, x, wave, iters_total, iters_step, pivot_hash, status, stream);
    }
    cudaFreeAsync(d_pack, stream);
    if (e != cudaSuccess) return cuda_fail(e, "tran_kernel launch");
    return CSIM_OK;
}

int csim_dev_sstate(const csim_plan* plan, int64_t B, const double* params, const csim_tran_args* tran,
                    const csim_sstate_args* ss, double* x, int32_t* periods, double* delta, int32_t* iters_total,
                    uint64_t* pivot_hash, uint32_t* status, void* stream_v) {
    int rc = need_device(plan, "csim_dev_sstate");
    if (rc) return rc;
    if (B <= 0 || !params || !tran || !ss || !x || !periods || !delta)
        return fail(CSIM_E_INVALID, "csim_dev_sstate: null argument");
    if (ss->steps_per_period <= 0 || ss->max_periods <= 0 || !(ss->tol > 0.0))
        return fail(CSIM_E_INVALID, "csim_dev_sstate: steps_per_period, max_periods and tol must be positive");
    if ((int64_t)ss->steps_per_period * ss->max_periods > 2147483647LL)
        return fail(CSIM_E_INVALID, "csim_dev_sstate: step index overflows int32");
    cudaStream_t stream = (cudaStream_t)stream_v;
    const int N = plan->host.N;
    double* x_prev = nullptr;
    uint8_t* skip = nullptr;
    unsigned long long* n_open = nullptr;
    auto cleanup = [&]() {
        cudaFreeAsync(x_prev, stream);
        cudaFreeAsync(skip, stream);
        cudaFreeAsync(n_open, stream);
    };
#define CUS(call)                                  \
    do {                                           \
        cudaError_t e__ = (call);                  \
        if (e__ != cudaSuccess) {                  \
            cleanup();                             \
            return cuda_fail(e__, #call);          \
        }                                          \
    } while (0)
    CUS(cudaMallocAsync((void**)&x_prev, sizeof(double) * (size_t)N * B, stream));
    CUS(cudaMallocAsync((void**)&skip, (size_t)B, stream));
    CUS(cudaMallocAsync((void**)&n_open, sizeof(unsigned long long), stream));
    CUS(cudaMemsetAsync(skip, 0, (size_t)B, stream));
    CUS(cudaMemsetAsync(periods, 0xFF, sizeof(int32_t) * (size_t)B, stream));  // -1
    CUS(cudaMemsetAsync(delta, 0, sizeof(double) * (size_t)B, stream));
    csim_tran_args a = *tran;
    a.n_probes = 0;
    a.probes = nullptr;
    a.out_stride = 1;
    a.out_begin_step = 0;
    a.wave_ld = 0;
    for (int k = 1; k <= ss->max_periods; ++k) {
        CUS(cudaMemcpyAsync(x_prev, x, sizeof(double) * (size_t)N * B, cudaMemcpyDeviceToDevice, stream));
        a.step_begin = (k - 1) * ss->steps_per_period;
        a.step_end = k * ss->steps_per_period;
        g_tran_skip = skip;
        rc = csim_dev_tran(plan, B, params, &a, x, nullptr, iters_total, nullptr, pivot_hash, status, stream);
        g_tran_skip = nullptr;
        if (rc) {
            cleanup();
            return rc;
        }
        CUS(cudaMemsetAsync(n_open, 0, sizeof(unsigned long long), stream));
        CUS(launch_sstate_check(N, B, x, x_prev, ss->tol, k, skip, periods, delta, n_open, stream));
        unsigned long long open = 0;
        CUS(cudaMemcpyAsync(&open, n_open, sizeof(open), cudaMemcpyDeviceToHost, stream));
        CUS(cudaStreamSynchronize(stream));
        if (open == 0) break;
    }
    cleanup();
#undef CUS
    return CSIM_OK;
}

// -------------------------------------------------------------------- host API
int csim_batch_create(const csim_plan* plan, int64_t B, csim_batch** out) {
    if (!out) return fail(CSIM_E_INVALID, "csim_batch_create: null out");
    *out = nullptr;
    int rc = need_device(plan, "csim_batch_create");
    if (rc) return rc;
    if (B <= 0) return fail(CSIM_E_INVALID, "csim_batch_create: B <= 0");
    std::unique_ptr<csim_batch> b(new csim_batch());
    b->plan = plan;
    b->B = B;
    const int N = plan->host.N, P = plan->host.n_params;
    const int ramp = plan->host.consts.dc_ramp_steps > 0 ? plan->host.consts.dc_ramp_steps : 1;
    CU(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&b->ev0));
    CU(cudaEventCreate(&b->ev1));
    CU(cudaMalloc((void**)&b->d_params, sizeof(double) * (size_t)(P > 0 ? P : 1) * B));
    CU(cudaMalloc((void**)&b->d_x, sizeof(double) * (size_t)N * B));
    CU(cudaMalloc((void**)&b->d_hash, sizeof(uint64_t) * B));
    CU(cudaMalloc((void**)&b->d_status, sizeof(uint32_t) * B));
    CU(cudaMalloc((void**)&b->d_iters, sizeof(int32_t) * (size_t)ramp * B));
    *out = b.release();
    return CSIM_OK;
}

void csim_batch_destroy(csim_batch* b) {
    if (!b) return;
    cudaSetDevice(b->plan->device);
    cudaFree(b->d_params);
    cudaFree(b->d_x);
    cudaFree(b->d_hash);
    cudaFree(b->d_status);
    cudaFree(b->d_iters);
    cudaFree(b->d_wave);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->ev_chunk) cudaEventDestroy(b->ev_chunk);
    if (b->copy_stream) cudaStreamDestroy(b->copy_stream);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
}

int csim_batch_set_params(csim_batch* b, const double* params) {
    if (!b || !params) return fail(CSIM_E_INVALID, "csim_batch_set_params: null argument");
    CU(cudaSetDevice(b->plan->device));
    CU(cudaMemcpyAsync(b->d_params, params, sizeof(double) * (size_t)b->plan->host.n_params * b->B,
                       cudaMemcpyHostToDevice, b->stream));
    CU(cudaStreamSynchronize(b->stream));
    return CSIM_OK;
}

static int batch_reset_diag(csim_batch* b) {
    std::vector<uint64_t> h0((size_t)b->B, CSIM_PIVOT_HASH_INIT);
    CU(cudaMemcpyAsync(b->d_hash, h0.data(), sizeof(uint64_t) * b->B, cudaMemcpyHostToDevice, b->stream));
    CU(cudaMemsetAsync(b->d_status, 0, sizeof(uint32_t) * b->B, b->stream));
    CU(cudaStreamSynchronize(b->stream));
    return CSIM_OK;
}

int csim_batch_dc(csim_batch* b, double* x, int32_t* iters, uint64_t* pivot_hash, uint32_t* status) {
    if (!b) return fail(CSIM_E_INVALID, "csim_batch_dc: null batch");
    CU(cudaSetDevice(b->plan->device));
    int rc = batch_reset_diag(b);
    if (rc) return rc;
    const int N = b->plan->host.N, ramp = b->plan->host.consts.dc_ramp_steps;
    CU(cudaEventRecord(b->ev0, b->stream));
    rc = csim_dev_dc(b->plan, b->B, b->d_params, b->d_x, b->d_iters, b->d_hash, b->d_status, b->stream);
    if (rc) return rc;
    CU(cudaEventRecord(b->ev1, b->stream));
    if (x) CU(cudaMemcpyAsync(x, b->d_x, sizeof(double) * (size_t)N * b->B, cudaMemcpyDeviceToHost, b->stream));
    if (iters) CU(cudaMemcpyAsync(iters, b->d_iters, sizeof(int32_t) * (size_t)ramp * b->B, cudaMemcpyDeviceToHost, b->stream));
    if (pivot_hash) CU(cudaMemcpyAsync(pivot_hash, b->d_hash, sizeof(uint64_t) * b->B, cudaMemcpyDeviceToHost, b->stream));
    if (status) CU(cudaMemcpyAsync(status, b->d_status, sizeof(uint32_t) * b->B, cudaMemcpyDeviceToHost, b->stream));
    CU(cudaStreamSynchronize(b->stream));
    CU(cudaEventElapsedTime(&b->last_ms, b->ev0, b->ev1));
    return CSIM_OK;
}

// x_start == nullptr: DC operating point first (csim_batch_tran); else the transient starts from x_start [N][B]
// csim_batch_tran with host buffers: waveforms of at least kChunkMinWaveBytes over at least kChunkMinSteps
// timepoints are produced and copied home in kWaveChunks pieces (see batch_tran_impl)
constexpr size_t kChunkMinWaveBytes = (size_t)256 << 20;
constexpr int64_t kChunkMinSteps = 4096;
constexpr int kWaveChunks = 32;  // (B200, headline run: end-to-end / device-resident throughput 0.963 with 8 chunks, 0.98-0.99 with 32: the
                                 // copy of the LAST chunk is what the caller still waits for)

static int batch_tran_impl(csim_batch* b, const csim_tran_args* a, const double* x_start, double* x_dc, double* x_final,
                           double* wave, int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash,
                           uint32_t* status) {
    if (!b || !a) return fail(CSIM_E_INVALID, "csim_batch_tran: null argument");
    const csim_plan* plan = b->plan;
    CU(cudaSetDevice(plan->device));
    const int N = plan->host.N;
    const int64_t B = b->B;
    int rc = batch_reset_diag(b);
    if (rc) return rc;
    const int nsteps = a->step_end;
    // device copies of the host-side tables / outputs
    double* d_table = nullptr;
    int32_t* d_group_b = nullptr;
    double* d_wave = nullptr;
    uint8_t* d_its = nullptr;
    int32_t* d_total = nullptr;
    size_t wave_elems = (size_t)B * (a->n_probes > 0 ? a->n_probes : 0) * (size_t)(a->wave_ld > 0 ? a->wave_ld : 0);
    auto cleanup = [&]() {
        cudaFree(d_table);
        cudaFree(d_group_b);
        cudaFree(d_its);
        cudaFree(d_total);
    };
#define CUC(call)                                    \
    do {                                             \
        cudaError_t e__ = (call);                    \
        if (e__ != cudaSuccess) {                    \
            cleanup();                               \
            return cuda_fail(e__, #call);            \
        }                                            \
    } while (0)
    if (a->src_table && a->n_groups > 0 && a->table_steps > 0) {
        size_t bytes = sizeof(double) * (size_t)a->n_groups * a->table_steps;
        CUC(cudaMalloc((void**)&d_table, bytes));
        CUC(cudaMemcpyAsync(d_table, a->src_table, bytes, cudaMemcpyHostToDevice, b->stream));
    }
    if (a->src_group_b) {
        for (int s = 0; s < plan->host.n_sources; ++s) {
            if (!a->src_mode || a->src_mode[s] == CSIM_SRC_CONST) continue;
            for (int64_t i = 0; i < B; ++i) {
                const int32_t gi = a->src_group_b[(size_t)s * B + i];
                if (gi < 0 || gi >= a->n_groups) {
                    cleanup();
                    return fail(CSIM_E_INVALID, "csim_batch_tran: src_group_b entry out of range");
                }
            }
        }
        size_t bytes = sizeof(int32_t) * (size_t)plan->host.n_sources * B;
        CUC(cudaMalloc((void**)&d_group_b, bytes));
        CUC(cudaMemcpyAsync(d_group_b, a->src_group_b, bytes, cudaMemcpyHostToDevice, b->stream));
    }
    if (wave_elems) {
        if (b->wave_bytes < sizeof(double) * wave_elems) {
            cudaFree(b->d_wave);
            b->d_wave = nullptr;
            b->wave_bytes = 0;
            CUC(cudaMalloc((void**)&b->d_wave, sizeof(double) * wave_elems));
            b->wave_bytes = sizeof(double) * wave_elems;
        }
        d_wave = b->d_wave;
        CUC(cudaMemsetAsync(d_wave, 0xFF, sizeof(double) * wave_elems, b->stream));  // NaN = "never written"
    }
    if (iters_step) {
        CUC(cudaMalloc((void**)&d_its, (size_t)nsteps * B));
        CUC(cudaMemsetAsync(d_its, 0, (size_t)nsteps * B, b->stream));
    }
    CUC(cudaMalloc((void**)&d_total, sizeof(int32_t) * B));
    CUC(cudaMemsetAsync(d_total, 0, sizeof(int32_t) * B, b->stream));

    csim_tran_args da = *a;
    da.src_table = d_table;
    da.src_group_b = d_group_b;
    CUC(cudaEventRecord(b->ev0, b->stream));
    if (x_start) {
        CUC(cudaMemcpyAsync(b->d_x, x_start, sizeof(double) * (size_t)N * B, cudaMemcpyHostToDevice, b->stream));
        rc = 0;
    } else {
        // runTransientAnalysisBackwardEuler: DC operating point first (src/tanalisis.cpp:112) ...
        rc = csim_dev_dc(plan, B, b->d_params, b->d_x, nullptr, b->d_hash, b->d_status, b->stream);
    }
    if (rc) {
        cleanup();
        return rc;
    }
    if (x_dc) CUC(cudaMemcpyAsync(x_dc, b->d_x, sizeof(double) * (size_t)N * B, cudaMemcpyDeviceToHost, b->stream));
    // ... then the step loop (255-420).  A long run with a large waveform is cut into chunks of timepoints (the
    // kernels carry every piece of state across a step range: x, iteration totals, pivot hash, status): the rows
    // of `wave` one chunk recorded go to the host on a second stream while the next chunk is computed, so the
    // caller pays for the last chunk's copy only instead of the whole waveform's.
    const int64_t span = (int64_t)a->step_end - a->step_begin;
    int forced = 0;  // CSIM_WAVE_CHUNKS=n: chunk regardless of size (tests exercise the path on small batches)
    if (const char* env = std::getenv("CSIM_WAVE_CHUNKS")) forced = std::atoi(env);
    const bool chunked = wave && wave_elems && a->out_stride > 0 && a->wave_ld > 0 &&
                         (forced > 1 ? span >= forced
                                     : forced == 0 && wave_elems * sizeof(double) >= kChunkMinWaveBytes &&
                                           span >= kChunkMinSteps);
    const int n_chunks = chunked ? (forced > 1 ? forced : kWaveChunks) : 1;
    if (chunked && !b->copy_stream) {
        CUC(cudaStreamCreateWithFlags(&b->copy_stream, cudaStreamNonBlocking));
        CUC(cudaEventCreateWithFlags(&b->ev_chunk, cudaEventDisableTiming));
    }
    for (int c = 0; c < n_chunks; ++c) {
        csim_tran_args ca = da;
        ca.step_begin = a->step_begin + (int32_t)(span * c / n_chunks);
        ca.step_end = a->step_begin + (int32_t)(span * (c + 1) / n_chunks);
        rc = csim_dev_tran(plan, B, b->d_params, &ca, b->d_x, d_wave, d_total, d_its, b->d_hash, b->d_status, b->stream);
        if (rc) {
            if (b->copy_stream) cudaStreamSynchronize(b->copy_stream);
            cleanup();
            return rc;
        }
        if (!chunked) break;
        // rows of `wave` written by steps [step_begin, step_end): step s records row (s+1)/stride - 1 - r0 when
        // (s+1) % stride == 0 and s >= out_begin_step
        const int64_t stride = a->out_stride, r0 = a->out_begin_step / stride;
        int64_t rlo = (ca.step_begin + stride) / stride - 1 - r0, rhi = ca.step_end / stride - r0;
        if (rlo < 0) rlo = 0;
        if (rhi > a->wave_ld) rhi = a->wave_ld;
        if (rhi > rlo) {
            CUC(cudaEventRecord(b->ev_chunk, b->stream));
            CUC(cudaStreamWaitEvent(b->copy_stream, b->ev_chunk, 0));
            CUC(cudaMemcpy2DAsync(wave + rlo, sizeof(double) * (size_t)a->wave_ld, d_wave + rlo,
                                  sizeof(double) * (size_t)a->wave_ld, sizeof(double) * (size_t)(rhi - rlo),
                                  (size_t)B * a->n_probes, cudaMemcpyDeviceToHost, b->copy_stream));
        }
    }
    CUC(cudaEventRecord(b->ev1, b->stream));
    if (x_final) CUC(cudaMemcpyAsync(x_final, b->d_x, sizeof(double) * (size_t)N * B, cudaMemcpyDeviceToHost, b->stream));
    if (wave && wave_elems && !chunked)
        CUC(cudaMemcpyAsync(wave, d_wave, sizeof(double) * wave_elems, cudaMemcpyDeviceToHost, b->stream));
    if (iters_total) CUC(cudaMemcpyAsync(iters_total, d_total, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, b->stream));
    if (iters_step) CUC(cudaMemcpyAsync(iters_step, d_its, (size_t)nsteps * B, cudaMemcpyDeviceToHost, b->stream));
    if (pivot_hash) CUC(cudaMemcpyAsync(pivot_hash, b->d_hash, sizeof(uint64_t) * B, cudaMemcpyDeviceToHost, b->stream));
    if (status) CUC(cudaMemcpyAsync(status, b->d_status, sizeof(uint32_t) * B, cudaMemcpyDeviceToHost, b->stream));
    CUC(cudaStreamSynchronize(b->stream));
    if (chunked) {
        CUC(cudaStreamSynchronize(b->copy_stream));
        // rows no step recorded: "never written" = NaN, as the single-launch path returns them
        const int64_t stride = a->out_stride, r0 = a->out_begin_step / stride;
        int64_t rec_lo = (a->step_begin + stride) / stride - 1 - r0, rec_hi = (int64_t)a->step_end / stride - r0;
        if (rec_lo < 0) rec_lo = 0;
        if (rec_hi > a->wave_ld) rec_hi = a->wave_ld;
        if (rec_hi < rec_lo) rec_hi = rec_lo;
        if (rec_lo > 0 || rec_hi < a->wave_ld)
            for (size_t row = 0; row < (size_t)B * a->n_probes; ++row) {
                double* w = wave + row * (size_t)a->wave_ld;
                std::memset(w, 0xFF, sizeof(double) * (size_t)rec_lo);
                std::memset(w + rec_hi, 0xFF, sizeof(double) * (size_t)(a->wave_ld - rec_hi));
            }
    }
    CUC(cudaEventElapsedTime(&b->last_ms, b->ev0, b->ev1));
    cleanup();
#undef CUC
    return CSIM_OK;
}

int csim_batch_tran(csim_batch* b, const csim_tran_args* a, double* x_dc, double* x_final, double* wave,
                    int32_t* iters_total, uint8_t* iters_step, uint64_t* pivot_hash, uint32_t* status) {
    return batch_tran_impl(b, a, nullptr, x_dc, x_final, wave, iters_total, iters_step, pivot_hash, status);
}

int csim_batch_tran_from(csim_batch* b, const csim_tran_args* a, double* x, double* wave, int32_t* iters_total,
                         uint64_t* pivot_hash, uint32_t* status) {
    if (!x) return fail(CSIM_E_INVALID, "csim_batch_tran_from: null state");
    return batch_tran_impl(b, a, x, nullptr, x, wave, iters_total, nullptr, pivot_hash, status);
}

int csim_batch_sstate(csim_batch* b, const csim_tran_args* a, const csim_sstate_args* ss, double* x_dc, double* x,
                      int32_t* periods, double* delta, int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status) {
    if (!b || !a || !ss || !x) return fail(CSIM_E_INVALID, "csim_batch_sstate: null argument");
    const csim_plan* plan = b->plan;
    CU(cudaSetDevice(plan->device));
    const int N = plan->host.N;
    const int64_t B = b->B;
    int rc = batch_reset_diag(b);
    if (rc) return rc;
    DevBuf d_table, d_group_b, d_periods, d_delta, d_total;
    if (a->src_table && a->n_groups > 0 && a->table_steps > 0)
        CU(d_table.put(a->src_table, sizeof(double) * (size_t)a->n_groups * a->table_steps));
    if (a->src_group_b) CU(d_group_b.put(a->src_group_b, sizeof(int32_t) * (size_t)plan->host.n_sources * B));
    CU(d_periods.put(nullptr, sizeof(int32_t) * B));
    CU(d_delta.put(nullptr, sizeof(double) * B));
    CU(d_total.put(nullptr, sizeof(int32_t) * B));
    CU(cudaMemsetAsync(d_total.p, 0, sizeof(int32_t) * B, b->stream));
    csim_tran_args da = *a;
    da.src_table = a->src_table ? d_table.as<double>() : nullptr;
    da.src_group_b = a->src_group_b ? d_group_b.as<int32_t>() : nullptr;
    CU(cudaEventRecord(b->ev0, b->stream));
    rc = csim_dev_dc(plan, B, b->d_params, b->d_x, nullptr, b->d_hash, b->d_status, b->stream);
    if (rc) return rc;
    if (x_dc) CU(cudaMemcpyAsync(x_dc, b->d_x, sizeof(double) * (size_t)N * B, cudaMemcpyDeviceToHost, b->stream));
    rc = csim_dev_sstate(plan, B, b->d_params, &da, ss, b->d_x, d_periods.as<int32_t>(), d_delta.as<double>(),
                         d_total.as<int32_t>(), b->d_hash, b->d_status, b->stream);
    if (rc) return rc;
    CU(cudaEventRecord(b->ev1, b->stream));
    CU(cudaMemcpyAsync(x, b->d_x, sizeof(double) * (size_t)N * B, cudaMemcpyDeviceToHost, b->stream));
    if (periods) CU(cudaMemcpyAsync(periods, d_periods.p, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, b->stream));
    if (delta) CU(cudaMemcpyAsync(delta, d_delta.p, sizeof(double) * B, cudaMemcpyDeviceToHost, b->stream));
    if (iters_total) CU(cudaMemcpyAsync(iters_total, d_total.p, sizeof(int32_t) * B, cudaMemcpyDeviceToHost, b->stream));
    if (pivot_hash) CU(cudaMemcpyAsync(pivot_hash, b->d_hash, sizeof(uint64_t) * B, cudaMemcpyDeviceToHost, b->stream));
    if (status) CU(cudaMemcpyAsync(status, b->d_status, sizeof(uint32_t) * B, cudaMemcpyDeviceToHost, b->stream));
    CU(cudaStreamSynchronize(b->stream));
    CU(cudaEventElapsedTime(&b->last_ms, b->ev0, b->ev1));
    return CSIM_OK;
}

int csim_host_stamp_csr(const csim_plan* plan, int mode, int64_t B, const double* params, const double* x,
                        const double* x_prev, const double* src_val, double dt, double source_scale, double gmin,
                        double* vals, double* rhs) {
    int rc = need_device(plan, "csim_host_stamp_csr");
    if (rc) return rc;
    if (mode < 0 || mode > 1 || B <= 0 || !params || !x || !vals || !rhs)
        return fail(CSIM_E_INVALID, "csim_host_stamp_csr: bad argument");
    const size_t N = plan->host.N, P = plan->host.n_params, S = plan->host.n_sources;
    const size_t nnz = plan->host.mode[mode].colidx.size();
    DevBuf dp, dx, dxp, dsv, dv, dr;
    CU(dp.put(params, sizeof(double) * P * B));
    CU(dx.put(x, sizeof(double) * N * B));
    if (x_prev) CU(dxp.put(x_prev, sizeof(double) * N * B));
    if (src_val) CU(dsv.put(src_val, sizeof(double) * S * B));
    CU(dv.put(nullptr, sizeof(double) * nnz * B));
    CU(dr.put(nullptr, sizeof(double) * N * B));
    rc = csim_dev_stamp_csr(plan, mode, B, dp.as<double>(), dx.as<double>(), x_prev ? dxp.as<double>() : nullptr,
                            src_val ? dsv.as<double>() : nullptr, dt, source_scale, gmin, dv.as<double>(),
                            dr.as<double>(), nullptr);
    if (rc) return rc;
    CU(cudaMemcpy(vals, dv.p, sizeof(double) * nnz * B, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(rhs, dr.p, sizeof(double) * N * B, cudaMemcpyDeviceToHost));
    return CSIM_OK;
}

int csim_host_lu_solve(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t* perm,
                       int32_t* singular, double lu_eps) {
    if (n <= 0 || n > kBigMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_host_lu_solve: n out of range");
    if (B <= 0 || !A || !b || !x) return fail(CSIM_E_INVALID, "csim_host_lu_solve: bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(CSIM_E_CUDA, "csim_host_lu_solve: no CUDA device (no CPU fallback)");
    DevBuf dA, db, dx, dp, ds;
    CU(dA.put(A, sizeof(double) * (size_t)n * n * B));
    CU(db.put(b, sizeof(double) * (size_t)n * B));
    CU(dx.put(nullptr, sizeof(double) * (size_t)n * B));
    CU(dp.put(nullptr, sizeof(int32_t) * (size_t)n * B));
    CU(ds.put(nullptr, sizeof(int32_t) * (size_t)B));
    int rc = csim_dev_lu_solve(n, B, dA.as<double>(), db.as<double>(), dx.as<double>(), dp.as<int32_t>(),
                               ds.as<int32_t>(), lu_eps, nullptr);
    if (rc) return rc;
    CU(cudaMemcpy(x, dx.p, sizeof(double) * (size_t)n * B, cudaMemcpyDeviceToHost));
    if (perm) CU(cudaMemcpy(perm, dp.p, sizeof(int32_t) * (size_t)n * B, cudaMemcpyDeviceToHost));
    if (singular) CU(cudaMemcpy(singular, ds.p, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost));
    return CSIM_OK;
}

int csim_host_lu_factor(int32_t n, int64_t B, const double* A, double* LU, int32_t* perm, int32_t* singular,
                        double lu_eps) {
    if (n <= 0 || n > kDenseMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_host_lu_factor: n out of range");
    if (B <= 0 || !A || !LU) return fail(CSIM_E_INVALID, "csim_host_lu_factor: bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(CSIM_E_CUDA, "csim_host_lu_factor: no CUDA device (no CPU fallback)");
    DevBuf dA, dLU, dp, ds;
    CU(dA.put(A, sizeof(double) * (size_t)n * n * B));
    CU(dLU.put(nullptr, sizeof(double) * (size_t)n * n * B));
    CU(dp.put(nullptr, sizeof(int32_t) * (size_t)n * B));
    CU(ds.put(nullptr, sizeof(int32_t) * (size_t)B));
    int rc = csim_dev_lu_factor(n, B, dA.as<double>(), dLU.as<double>(), dp.as<int32_t>(), ds.as<int32_t>(), lu_eps, nullptr);
    if (rc) return rc;
    CU(cudaMemcpy(LU, dLU.p, sizeof(double) * (size_t)n * n * B, cudaMemcpyDeviceToHost));
    if (perm) CU(cudaMemcpy(perm, dp.p, sizeof(int32_t) * (size_t)n * B, cudaMemcpyDeviceToHost));
    if (singular) CU(cudaMemcpy(singular, ds.p, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost));
    return CSIM_OK;
}

int csim_dev_gauss_seidel(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t max_iters,
                          double tol, int32_t* sweeps, void* stream) {
    if (n <= 0 || n > kGaussSeidelMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_dev_gauss_seidel: n out of range (1..200)");
    if (B <= 0 || !A || !b || !x || max_iters < 0) return fail(CSIM_E_INVALID, "csim_dev_gauss_seidel: bad argument");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(CSIM_E_CUDA, "csim_dev_gauss_seidel: no CUDA device (no CPU fallback)");
    CU(launch_gauss_seidel(n, B, A, b, x, max_iters, tol, sweeps, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_host_gauss_seidel(int32_t n, int64_t B, const double* A, const double* b, double* x, int32_t max_iters,
                           double tol, int32_t* sweeps) {
    if (n <= 0 || n > kGaussSeidelMaxN) return fail(CSIM_E_UNSUPPORTED, "csim_host_gauss_seidel: n out of range (1..200)");
    if (B <= 0 || !A || !b || !x) return fail(CSIM_E_INVALID, "csim_host_gauss_seidel: bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(CSIM_E_CUDA, "csim_host_gauss_seidel: no CUDA device (no CPU fallback)");
    DevBuf dA, db, dx, ds;
    CU(dA.put(A, sizeof(double) * (size_t)n * n * B));
    CU(db.put(b, sizeof(double) * (size_t)n * B));
    CU(dx.put(x, sizeof(double) * (size_t)n * B));
    CU(ds.put(nullptr, sizeof(int32_t) * (size_t)B));
    int rc = csim_dev_gauss_seidel(n, B, dA.as<double>(), db.as<double>(), dx.as<double>(), max_iters, tol,
                                   ds.as<int32_t>(), nullptr);
    if (rc) return rc;
    CU(cudaMemcpy(x, dx.p, sizeof(double) * (size_t)n * B, cudaMemcpyDeviceToHost));
    if (sweeps) CU(cudaMemcpy(sweeps, ds.p, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost));
    return CSIM_OK;
}

int csim_dev_harmonics(int64_t n_signals, int32_t n_samples, int64_t ld, const double* samples, int32_t n_harm,
                       const double* cos_tab, const double* sin_tab, double* re, double* im, void* stream) {
    if (n_signals <= 0 || n_samples <= 0 || ld < n_samples || !samples || !cos_tab || !sin_tab || !re || !im)
        return fail(CSIM_E_INVALID, "csim_dev_harmonics: bad argument");
    if (n_harm < 0 || n_harm > kHarmonicsMax) return fail(CSIM_E_UNSUPPORTED, "csim_dev_harmonics: n_harm out of range (0..64)");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(CSIM_E_CUDA, "csim_dev_harmonics: no CUDA device (no CPU fallback)");
    CU(launch_harmonics(n_signals, n_samples, ld, samples, n_harm, cos_tab, sin_tab, re, im, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_plan_force_large_path(csim_plan* plan, int on) {
    if (!plan) return fail(CSIM_E_INVALID, "csim_plan_force_large_path: null plan");
    plan->force_big = on < 0 ? 0 : (on > 2 ? 1 : on);
    return CSIM_OK;
}

int csim_plan_jit_mode(csim_plan* plan, int mode) {
    if (!plan || mode < 0 || mode > 3) return fail(CSIM_E_INVALID, "csim_plan_jit_mode: bad argument");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    if (mode != plan->spec.mode) {
        plan->spec.tried = false;
        plan->spec.kind = 0;
        plan->dc_tried = false;
        plan->tpi_dc.ready = false;
    }
    plan->spec.mode = mode;
    return CSIM_OK;
}

int csim_plan_specialize(csim_plan* plan, const int32_t* perm) {
    int rc = need_device(plan, "csim_plan_specialize");
    if (rc) return rc;
    if (!perm) return fail(CSIM_E_INVALID, "csim_plan_specialize: null permutation");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    SpecState& sp = plan->spec;
    sp.tried = true;
    if (sp.mode == 0) sp.mode = 1;
    const std::vector<int32_t> pv(perm, perm + plan->host.N);
    if (sp.mode != 2) {
        rc = tpi_build(plan, pv);
        if (rc == CSIM_OK) sp.kind = 3;
        return rc;
    }
    rc = spec_build(plan, pv);
    if (rc == CSIM_OK) sp.kind = 2;
    return rc;
}

int csim_plan_specialize_variants(csim_plan* plan, int32_t n_variants, const int32_t* perms) {
    int rc = need_device(plan, "csim_plan_specialize_variants");
    if (rc) return rc;
    if (!perms || n_variants < 1 || n_variants > kTpiMaxVariants)
        return fail(CSIM_E_INVALID, "csim_plan_specialize_variants: 1..16 permutations expected");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    SpecState& sp = plan->spec;
    sp.tried = true;
    if (sp.mode == 0 || sp.mode == 2) sp.mode = 1;
    const int N = plan->host.N;
    std::vector<std::vector<int32_t>> pv;
    for (int v = 0; v < n_variants; ++v) pv.emplace_back(perms + (size_t)v * N, perms + (size_t)(v + 1) * N);
    rc = tpi_build(plan, pv);
    if (rc == CSIM_OK) sp.kind = 3;
    return rc;
}

int csim_plan_specialize_dc_variants(csim_plan* plan, int32_t n_variants, const int32_t* perms) {
    int rc = need_device(plan, "csim_plan_specialize_dc_variants");
    if (rc) return rc;
    if (!perms || n_variants < 1 || n_variants > kTpiMaxVariants)
        return fail(CSIM_E_INVALID, "csim_plan_specialize_dc_variants: 1..16 permutations expected");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    if (plan->spec.mode == 0 || plan->spec.mode == 2) plan->spec.mode = 1;
    const int N = plan->host.N;
    std::vector<std::vector<int32_t>> pv;
    for (int v = 0; v < n_variants; ++v) pv.emplace_back(perms + (size_t)v * N, perms + (size_t)(v + 1) * N);
    plan->dc_tried = true;
    return tpi_build(plan, pv, true);
}

int csim_plan_dc_spec_info(const csim_plan* plan, int32_t* active, int32_t* n_variants, uint64_t* fallback_solves,
                           uint64_t* solves, char* log, int32_t log_cap) {
    if (!plan) return fail(CSIM_E_INVALID, "csim_plan_dc_spec_info: null plan");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    const TpiState& tp = plan->tpi_dc;
    if (active) *active = tp.ready ? 1 : 0;
    if (n_variants) *n_variants = tp.ready ? (int32_t)tp.host.perms.size() : 0;
    unsigned long long st[2] = {0, 0};
    if (plan->device >= 0 && tp.ready && tp.d_stats) {
        cudaSetDevice(plan->device);
        cudaDeviceSynchronize();
        cudaMemcpy(st, tp.d_stats, sizeof(st), cudaMemcpyDeviceToHost);
    }
    if (fallback_solves) *fallback_solves = st[0];
    if (solves) *solves = st[1];
    if (log && log_cap > 0) std::snprintf(log, (size_t)log_cap, "%s", tp.log.c_str());
    return CSIM_OK;
}

int csim_plan_generate_source_variants(const csim_plan* plan, int32_t n_variants, const int32_t* perms, char* out,
                                       int32_t cap) {
    if (!plan || !perms || n_variants < 1) return fail(CSIM_E_INVALID, "csim_plan_generate_source_variants: bad argument");
    try {
        const int N = plan->host.N;
        std::vector<std::vector<int32_t>> pv;
        for (int v = 0; v < n_variants; ++v) pv.emplace_back(perms + (size_t)v * N, perms + (size_t)(v + 1) * N);
        TpiHost th = generate_tran_tpi(plan->host, pv, tpi_options_from_env());
        if (out && cap > 0) std::snprintf(out, (size_t)cap, "%s", th.source.c_str());
        return (int)th.source.size();
    } catch (const std::exception& e) {
        return fail(CSIM_E_UNSUPPORTED, e.what());
    }
}

int csim_plan_spec_info(const csim_plan* plan, int32_t* active, int32_t* perm, uint64_t* fallback_solves,
                        uint64_t* solves, char* log, int32_t log_cap) {
    if (!plan) return fail(CSIM_E_INVALID, "csim_plan_spec_info: null plan");
    SpecState& sp = plan->spec;
    TpiState& tp = plan->tpi;
    std::lock_guard<std::recursive_mutex> lock(sp.mu);
    const int kind = sp.mode == 0 ? 0 : (sp.kind == 3 && tp.ready ? 3 : (sp.kind == 2 && sp.ready ? 2 : 0));
    if (active) *active = kind;
    if (perm && kind)
        for (int i = 0; i < plan->host.N; ++i) perm[i] = kind == 3 ? tp.host.perm[i] : sp.host.perm[i];
    unsigned long long st[2] = {0, 0}, st2[2] = {0, 0};
    if (plan->device >= 0 && (tp.d_stats || sp.d_stats)) {
        cudaSetDevice(plan->device);
        cudaDeviceSynchronize();
        if (tp.ready && tp.d_stats) cudaMemcpy(st, tp.d_stats, sizeof(st), cudaMemcpyDeviceToHost);
        if (sp.ready && sp.d_stats) cudaMemcpy(st2, sp.d_stats, sizeof(st2), cudaMemcpyDeviceToHost);
    }
    if (fallback_solves) *fallback_solves = st[0] + st2[0];
    if (solves) *solves = st[1] + st2[1];
    std::string all = sp.log;
    std::string redo_log;  // what the generic redo path met since the kernel was built: per known variant / unknown orders
    if (kind == 3 && plan->device >= 0 && tp.d_miss) {
        std::vector<unsigned long long> tab((size_t)kTpiMissWords, 0);
        if (cudaMemcpy(tab.data(), tp.d_miss, tab.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess) {
            unsigned long long unknown = 0;
            int distinct = 0;
            for (int e = 0; e < kTpiMissEntries; ++e) {
                const unsigned long long hsh = tab[(size_t)e * kTpiMissStride], cnt = tab[(size_t)e * kTpiMissStride + 1];
                if (!hsh) continue;
                const auto it = std::find(tp.host.hashes.begin(), tp.host.hashes.end(), hsh);
                if (it == tp.host.hashes.end()) {
                    unknown += cnt;
                    ++distinct;
                } else {
                    redo_log += " variant " + std::to_string(it - tp.host.hashes.begin()) + ": " + std::to_string(cnt) + ";";
                }
            }
            if (!redo_log.empty() || distinct)
                redo_log = "\nredone solves by the order the reference took:" + redo_log + " other orders (" +
                           std::to_string(distinct) + " distinct logged): " + std::to_string(unknown);
        }
    }
    if (!tp.log.empty()) {
        const size_t nl = tp.log.find('\n');  // summary line, then the redo log, then history and compiler messages
        all = tp.log.substr(0, nl) + redo_log + (nl == std::string::npos ? "" : tp.log.substr(nl)) +
              (all.empty() ? "" : "\n" + all);
    }
    if (log && log_cap > 0) std::snprintf(log, (size_t)log_cap, "%s", all.c_str());
    return CSIM_OK;
}

int csim_plan_spec_source(const csim_plan* plan, char* out, int32_t cap) {
    if (!plan) return fail(CSIM_E_INVALID, "csim_plan_spec_source: null plan");
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    const std::string& src = plan->spec.kind == 3 ? plan->tpi.host.source : plan->spec.host.source;
    if (out && cap > 0) std::snprintf(out, (size_t)cap, "%s", src.c_str());
    return (int)src.size();
}

// Host-only: the CUDA source the generator emits for `perm` (no device, no compilation).
int csim_plan_generate_source(const csim_plan* plan, const int32_t* perm, char* out, int32_t cap) {
    if (!plan || !perm) return fail(CSIM_E_INVALID, "csim_plan_generate_source: null argument");
    try {
        SpecHost sh = generate_tran_spec(plan->host, std::vector<int32_t>(perm, perm + plan->host.N));
        if (out && cap > 0) std::snprintf(out, (size_t)cap, "%s", sh.source.c_str());
        return (int)sh.source.size();
    } catch (const std::exception& e) {
        return fail(CSIM_E_UNSUPPORTED, e.what());
    }
}

// kind: 2 = warp-per-instance kernel (csim_tran_spec), 3 = thread-per-instance kernel (csim_tran_tpi)
int csim_plan_generate_source_kind(const csim_plan* plan, int kind, const int32_t* perm, char* out, int32_t cap) {
    if (!plan || !perm) return fail(CSIM_E_INVALID, "csim_plan_generate_source_kind: null argument");
    if (kind == 2) return csim_plan_generate_source(plan, perm, out, cap);
    if (kind != 3 && kind != 4) return fail(CSIM_E_INVALID, "csim_plan_generate_source_kind: kind must be 2, 3 or 4");
    try {
        const std::vector<std::vector<int32_t>> pv{std::vector<int32_t>(perm, perm + plan->host.N)};
        TpiHost th = kind == 4 ? generate_dc_tpi(plan->host, pv, tpi_options_from_env())
                               : generate_tran_tpi(plan->host, pv, tpi_options_from_env());
        if (out && cap > 0) std::snprintf(out, (size_t)cap, "%s", th.source.c_str());
        return (int)th.source.size();
    } catch (const std::exception& e) {
        return fail(CSIM_E_UNSUPPORTED, e.what());
    }
}

int csim_plan_stamp_info(const csim_plan* plan, int mode, int has_src_val, int32_t* spec_active, int64_t* bytes_read,
                         int64_t* bytes_written, char* log, int32_t log_cap) {
    if (!plan || mode < 0 || mode > 1) return fail(CSIM_E_INVALID, "csim_plan_stamp_info: bad argument");
    const StampTraffic t = stamp_traffic(plan->host, mode, has_src_val != 0);
    if (bytes_read) *bytes_read = 8 * t.doubles_read;
    if (bytes_written) *bytes_written = 8 * t.doubles_written;
    std::lock_guard<std::recursive_mutex> lock(plan->spec.mu);
    const StampSpecState& ss = plan->stamp_spec[mode];
    if (spec_active) *spec_active = ss.ready && plan->spec.mode != 0 ? 1 : 0;
    if (log && log_cap > 0) std::snprintf(log, (size_t)log_cap, "%s", ss.log.c_str());
    return CSIM_OK;
}

int csim_plan_generate_stamp_source(const csim_plan* plan, int mode, char* out, int32_t cap) {
    if (!plan || mode < 0 || mode > 1) return fail(CSIM_E_INVALID, "csim_plan_generate_stamp_source: bad argument");
    try {
        StampSpecHost sh = generate_stamp_spec(plan->host, mode);
        if (out && cap > 0) std::snprintf(out, (size_t)cap, "%s", sh.source.c_str());
        return (int)sh.source.size();
    } catch (const std::exception& e) {
        return fail(CSIM_E_UNSUPPORTED, e.what());
    }
}

int csim_dev_selftest_div(int64_t n, const double* a, const double* b, double* q_split, double* q_ref, void* stream) {
    if (n <= 0 || !a || !b || !q_split || !q_ref) return fail(CSIM_E_INVALID, "csim_dev_selftest_div: bad argument");
    CU(launch_div_selftest(n, a, b, q_split, q_ref, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_dev_selftest_tdiv(int64_t n, const double* a, const double* b, double* q, double* q_ref, int32_t* in_range,
                           void* stream) {
    if (n <= 0 || !a || !b || !q || !q_ref || !in_range) return fail(CSIM_E_INVALID, "csim_dev_selftest_tdiv: bad argument");
    CU(launch_tdiv_selftest(n, a, b, q, q_ref, in_range, (cudaStream_t)stream));
    return CSIM_OK;
}

int csim_dev_margin_probe(double* margin) {
    g_margin = margin;
    return CSIM_OK;
}

int csim_dev_selftest_tdiv_mul(int64_t n, const double* a, const double* b, double* q, double* q_ref, int32_t* flags,
                               void* stream) {
    if (n <= 0 || !a || !b || !q || !q_ref || !flags) return fail(CSIM_E_INVALID, "csim_dev_selftest_tdiv_mul: bad argument");
    CU(launch_tdiv_mul_selftest(n, a, b, q, q_ref, flags, (cudaStream_t)stream));
    return CSIM_OK;
}

float csim_batch_last_kernel_ms(const csim_batch* b) { return b ? b->last_ms : 0.f; }

int csim_dev_fp64_peak(int device, double* tflops_mul_add) {
    if (!tflops_mul_add) return fail(CSIM_E_INVALID, "csim_dev_fp64_peak: null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return fail(CSIM_E_CUDA, "csim_dev_fp64_peak: no such CUDA device");
    CU(cudaSetDevice(device));
    int sms = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    CU(measure_fp64_peak(sms, tflops_mul_add, nullptr));
    return CSIM_OK;
}

}  // extern "C"
