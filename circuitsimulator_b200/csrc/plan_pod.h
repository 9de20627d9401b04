// plan_pod.h — POD records shared by the host plan builder, the static kernels and the
// NVRTC-compiled specialised kernels.  No standard-library includes (NVRTC-safe).
#pragma once
#include <stdint.h>

namespace csim {

// ---- value ops ------------------------------------------------------------
enum StaticKind : int32_t {
    SK_ONE = 0,       // 1.0
    SK_GMIN = 1,      // gmin (rewritten per Newton iteration in DC)
    SK_RES = 2,       // R==0 ? 0 : 1.0/R                 (src/element.cpp:20-24)
    SK_CAP = 3,       // C<=0||dt<=0 ? 0 : C/dt           (src/tanalisis.cpp:65-67)
    SK_CAPHALF = 4,   // c=0.5*Cj0; same rule             (src/tanalisis.cpp:338-339)
    SK_IND_ONE = 5,   // L<=0 ? 0 : 1.0                   (src/tanalisis.cpp:296, 311-315)
    SK_IND_NREQ = 6,  // L<=0 ? 0 : -(L/dt)               (src/tanalisis.cpp:304, 317)
};
struct StaticOp {  // out = f(kind, params[param])
    int32_t kind, param, out, _pad;
};
enum StepKind : int32_t {
    TK_CAP_HIST = 0,  // out = (-vt[g]) * (xprev[a] - xprev[b])   I_hist, src/tanalisis.cpp:77
    TK_IND_HIST = 1,  // out = vt[g] * xprev[a]                   V_hist = -R_eq*iPrev, src/tanalisis.cpp:305-306
};
struct StepOp {
    int32_t kind, g, a, b, out, _p0, _p1, _p2;  // a, b: equation index or -1 (ground -> 0.0)
};
struct SrcOp {  // one per independent source, element order
    int32_t out, param, is_sin, index;
};
struct MosOp {  // out..out+3 = gd, gg, gs, cst   (src/element.cpp:181-274)
    int32_t out, param, d, g, s, is_p, _p0, _p1;
};

// contribution word: bits 0-5 column (32 = RHS), bit 6 subtract, bits 8.. value index
#ifndef __CUDACC_RTC__
inline uint32_t contrib_word(int col, bool sub, int v) { return (uint32_t)col | (sub ? 64u : 0u) | ((uint32_t)v << 8); }
#endif
constexpr int kRhsCol = 32;

}  // namespace csim
