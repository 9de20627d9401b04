// multi.cpp — csim_multi_*: one batch sharded over several GPUs of one box below the C ABI (SURVEY.md §8(b)
// "Threading": one host thread + stream per GPU; §8(e): contiguous shards, no data-path collective, the only
// communication is the final gather — here every device copies its slice straight into the caller's host buffers).
// Built on the single-device entry points (csim_plan_*, csim_batch_*): each device gets its own plan and batch,
// each call runs one host thread per device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/csim_b200.h"

namespace {
struct Shard {
    int device = 0;
    int64_t first = 0, count = 0;
    csim_plan* plan = nullptr;
    csim_batch* batch = nullptr;
    std::vector<double> params;  // [P][count]
    int rc = 0;
    std::string err;
};
thread_local std::string g_multi_err;
}  // namespace

struct csim_multi {
    int32_t N = 0, P = 0, S = 0;
    int64_t B = 0;
    std::vector<Shard> shards;
};

// (capi.cpp) lets this file set the calling thread's csim_last_error()
extern "C" void csim_set_last_error(const char* msg);

namespace {
int fail(int code, const std::string& msg) {
    csim_set_last_error(msg.c_str());
    return code;
}
template <typename F>
int for_each_shard(csim_multi* m, F fn) {
    std::vector<std::thread> th;
    for (Shard& s : m->shards) {
        if (s.count == 0) continue;
        th.emplace_back([&s, &fn]() {
            s.rc = fn(s);
            s.err = s.rc ? csim_last_error() : "";
        });
    }
    for (auto& t : th) t.join();
    for (const Shard& s : m->shards)
        if (s.rc) return fail(s.rc, "device " + std::to_string(s.device) + ": " + s.err);
    return CSIM_OK;
}
// rows [K][B] (global, instance-minor) <-> [K][count] (shard-local)
template <typename T>
void gather_rows(const T* g, int64_t B, int64_t first, int64_t count, int K, T* l) {
    for (int k = 0; k < K; ++k) std::memcpy(l + (size_t)k * count, g + (size_t)k * B + first, sizeof(T) * (size_t)count);
}
template <typename T>
void scatter_rows(const T* l, int64_t B, int64_t first, int64_t count, int K, T* g) {
    for (int k = 0; k < K; ++k) std::memcpy(g + (size_t)k * B + first, l + (size_t)k * count, sizeof(T) * (size_t)count);
}
}  // namespace

extern "C" {

int csim_multi_create(const csim_topology* topo, const csim_constants* consts, int32_t n_devices, const int32_t* devices,
                      int64_t B, csim_multi** out) {
    if (!topo || !out || n_devices <= 0 || B <= 0) return fail(CSIM_E_INVALID, "csim_multi_create: bad argument");
    *out = nullptr;
    const int have = csim_device_count();
    if (have <= 0) return fail(CSIM_E_CUDA, "csim_multi_create: no CUDA device (no CPU fallback)");
    std::unique_ptr<csim_multi> m(new csim_multi());
    m->B = B;
    m->shards.resize(n_devices);
    const int64_t base = B / n_devices, rem = B % n_devices;  // contiguous, remainder to the low ranks
    int64_t first = 0;
    for (int r = 0; r < n_devices; ++r) {
        Shard& s = m->shards[r];
        s.device = devices ? devices[r] : r;
        if (s.device < 0 || s.device >= have) return fail(CSIM_E_INVALID, "csim_multi_create: no such device");
        s.first = first;
        s.count = base + (r < rem ? 1 : 0);
        first += s.count;
    }
    auto cleanup = [&]() {
        for (Shard& s : m->shards) {
            csim_batch_destroy(s.batch);
            csim_plan_destroy(s.plan);
        }
    };
    for (Shard& s : m->shards) {
        if (s.count == 0) continue;
        int rc = csim_plan_create(topo, consts, s.device, &s.plan);
        if (!rc) rc = csim_batch_create(s.plan, s.count, &s.batch);
        if (rc) {
            cleanup();
            return rc;
        }
    }
    csim_plan_dims(m->shards[0].plan, &m->N, &m->P, &m->S);
    *out = m.release();
    return CSIM_OK;
}

void csim_multi_destroy(csim_multi* m) {
    if (!m) return;
    for (Shard& s : m->shards) {
        csim_batch_destroy(s.batch);
        csim_plan_destroy(s.plan);
    }
    delete m;
}

int csim_multi_shard(const csim_multi* m, int32_t rank, int32_t* device, int64_t* first, int64_t* count) {
    if (!m || rank < 0 || rank >= (int)m->shards.size()) return fail(CSIM_E_INVALID, "csim_multi_shard: bad rank");
    if (device) *device = m->shards[rank].device;
    if (first) *first = m->shards[rank].first;
    if (count) *count = m->shards[rank].count;
    return CSIM_OK;
}

int csim_multi_set_params(csim_multi* m, const double* params) {
    if (!m || !params) return fail(CSIM_E_INVALID, "csim_multi_set_params: null argument");
    const int64_t B = m->B;
    const int P = m->P;
    return for_each_shard(m, [&](Shard& s) {
        s.params.resize((size_t)(P > 0 ? P : 1) * s.count);
        gather_rows(params, B, s.first, s.count, P, s.params.data());
        return csim_batch_set_params(s.batch, s.params.data());
    });
}

int csim_multi_dc(csim_multi* m, double* x, int32_t* iters, uint64_t* pivot_hash, uint32_t* status) {
    if (!m) return fail(CSIM_E_INVALID, "csim_multi_dc: null handle");
    const int64_t B = m->B;
    const int N = m->N;
    csim_constants c;
    csim_default_constants(&c);
    return for_each_shard(m, [&](Shard& s) {
        std::vector<double> lx(x ? (size_t)N * s.count : 0);
        std::vector<int32_t> li(iters ? (size_t)c.dc_ramp_steps * s.count : 0);
        int rc = csim_batch_dc(s.batch, x ? lx.data() : nullptr, iters ? li.data() : nullptr,
                               pivot_hash ? pivot_hash + s.first : nullptr, status ? status + s.first : nullptr);
        if (rc) return rc;
        if (x) scatter_rows(lx.data(), B, s.first, s.count, N, x);
        if (iters) scatter_rows(li.data(), B, s.first, s.count, c.dc_ramp_steps, iters);
        return CSIM_OK;
    });
}

int csim_multi_tran(csim_multi* m, const csim_tran_args* a, double* x_dc, double* x_final, double* wave,
                    int32_t* iters_total, uint64_t* pivot_hash, uint32_t* status) {
    if (!m || !a) return fail(CSIM_E_INVALID, "csim_multi_tran: null argument");
    const int64_t B = m->B;
    const int N = m->N, S = m->S;
    const int64_t row = (int64_t)(a->n_probes > 0 ? a->n_probes : 0) * (a->wave_ld > 0 ? a->wave_ld : 0);
    return for_each_shard(m, [&](Shard& s) {
        csim_tran_args la = *a;
        std::vector<int32_t> lg;
        if (a->src_group_b) {
            lg.resize((size_t)S * s.count);
            gather_rows(a->src_group_b, B, s.first, s.count, S, lg.data());
            la.src_group_b = lg.data();
        }
        std::vector<double> ldc(x_dc ? (size_t)N * s.count : 0), lfin(x_final ? (size_t)N * s.count : 0);
        // wave [B][n_probes][wave_ld], per-instance totals, hashes and status are contiguous per shard: written in place
        int rc = csim_batch_tran(s.batch, &la, x_dc ? ldc.data() : nullptr, x_final ? lfin.data() : nullptr,
                                 wave ? wave + (size_t)s.first * row : nullptr, iters_total ? iters_total + s.first : nullptr,
                                 nullptr, pivot_hash ? pivot_hash + s.first : nullptr, status ? status + s.first : nullptr);
        if (rc) return rc;
        if (x_dc) scatter_rows(ldc.data(), B, s.first, s.count, N, x_dc);
        if (x_final) scatter_rows(lfin.data(), B, s.first, s.count, N, x_final);
        return CSIM_OK;
    });
}

float csim_multi_last_kernel_ms(const csim_multi* m) {  // slowest device (they run concurrently)
    float ms = 0.f;
    if (m)
        for (const Shard& s : m->shards)
            if (s.batch) ms = std::max(ms, csim_batch_last_kernel_ms(s.batch));
    return ms;
}

}  // extern "C"
