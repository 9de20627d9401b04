// microbench.cu — FP64 pipe / latency probes on the box (register-resident; no HBM traffic).
// Prints: dependent-chain latency (cycles) of DADD, DMUL, DFMA, double division, SHFL(64-bit),
// REDUX, LDS round trip, and the sustained separate DMUL+DADD throughput of the whole chip
// (the FP64 "peak with FMA disabled" that the fused kernel's utilisation is quoted against).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int OP>
__global__ void chain(double* out, long long* cyc, double a, double b, int iters) {
    double v = a + threadIdx.x * 1e-9;
    __shared__ double sm[64];
    sm[threadIdx.x] = v; sm[threadIdx.x + 32] = b;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            if (OP == 0) v = __dadd_rn(v, b);
            if (OP == 1) v = __dmul_rn(v, b);
            if (OP == 2) v = __fma_rn(v, b, a);
            if (OP == 3) v = v / b;
            if (OP == 4) v = __shfl_xor_sync(0xffffffffu, v, 1);
            if (OP == 5) { unsigned r = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(v)); v = __hiloint2double(r, __double2loint(v)); }
            if (OP == 6) { sm[threadIdx.x] = v; __syncwarp(); v = sm[threadIdx.x ^ 1]; __syncwarp(); }
            if (OP == 7) { v = sm[(__double2loint(v) & 31)]; }
            if (OP == 8) v = sqrt(v);
        }
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// throughput: ILP 8 independent chains per thread, many warps
__global__ void __launch_bounds__(256) tput(double* out, double a, double b, int iters, int mode) {
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = a + k + threadIdx.x * 1e-6;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (mode == 0) v[k] = __dadd_rn(__dmul_rn(v[k], b), a);   // separate mul + add (2 flop, 2 instr)
            else if (mode == 1) v[k] = __fma_rn(v[k], b, a);            // fma (2 flop, 1 instr)
            else v[k] = v[k] / b;
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += v[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    double* out; long long* cyc;
    CK(cudaMalloc(&out, 1 << 24)); CK(cudaMalloc(&cyc, 8));
    const char* names[] = {"DADD", "DMUL", "DFMA", "DDIV", "SHFL.64", "REDUX.u32", "STS+LDS roundtrip", "LDS dependent", "DSQRT"};
    int iters = 2000;
    for (int op = 0; op < 9; ++op) {
        for (int rep = 0; rep < 2; ++rep) {
            switch (op) {
                case 0: chain<0><<<1, 32>>>(out, cyc, 1.0, 1e-9, iters); break;
                case 1: chain<1><<<1, 32>>>(out, cyc, 1.0, 1.0000001, iters); break;
                case 2: chain<2><<<1, 32>>>(out, cyc, 1.0, 0.999999, iters); break;
                case 3: chain<3><<<1, 32>>>(out, cyc, 1.0, 1.0000001, iters); break;
                case 4: chain<4><<<1, 32>>>(out, cyc, 1.0, 1.0, iters); break;
                case 5: chain<5><<<1, 32>>>(out, cyc, 1.0, 1.0, iters); break;
                case 6: chain<6><<<1, 32>>>(out, cyc, 1.0, 1.0, iters); break;
                case 7: chain<7><<<1, 32>>>(out, cyc, 1.0, 1.0, iters); break;
                case 8: chain<8><<<1, 32>>>(out, cyc, 1.5, 1.0, iters); break;
            }
            CK(cudaDeviceSynchronize());
        }
        long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
        printf("latency %-20s %.2f cycles/op\n", names[op], (double)c / (iters * 16.0));
    }
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char* mn[] = {"DMUL+DADD separate", "DFMA", "DDIV"};
    for (int mode = 0; mode < 3; ++mode) {
        int it = mode == 2 ? 2000 : 20000;
        int blocks = sms * 8;
        tput<<<blocks, 256>>>(out, 1.0, 0.9999999, 100, mode);
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        tput<<<blocks, 256>>>(out, 1.0, 0.9999999, it, mode);
        CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        double ops = (double)blocks * 256 * 8 * it;  // per-thread chain steps
        double flop = ops * (mode == 2 ? 1 : 2);
        printf("throughput %-20s %.3f T%s/s  (%d SMs, %.2f ms)\n", mn[mode], flop / ms / 1e9, mode == 2 ? "div" : "flop", sms, ms);
    }
    printf("sm_clock_khz %d\n", prop.clockRate);
    return 0;
}
