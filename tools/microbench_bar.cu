// Latency of hand-overs between the two warps of a 64-thread CTA on sm_100a: named barrier (bar.sync both / bar.arrive +
// bar.sync ping-pong) against a shared-memory flag.   nvcc -arch=sm_100a -o build/microbench_bar tools/microbench_bar.cu
#include <cstdio>
#include <cuda_runtime.h>
#define BAR_SYNC(id) asm volatile("bar.sync %0, 64;" ::"n"(id) : "memory")
#define BAR_ARRIVE(id) asm volatile("bar.arrive %0, 64;" ::"n"(id) : "memory")
__global__ void k_sync(long long* out, int iters) {
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) BAR_SYNC(1);
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
__global__ void k_pingpong(long long* out, int iters) {  // A: arrive 2, sync 3;  B: sync 2, arrive 3  -> one round trip per iteration
    const bool a = threadIdx.x < 32;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (a) { BAR_ARRIVE(2); BAR_SYNC(3); } else { BAR_SYNC(2); BAR_ARRIVE(3); }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
__global__ void k_flag(long long* out, int iters) {  // the same round trip through per-lane shared-memory flags
    __shared__ volatile int fa[32], fb[32];
    const int lane = threadIdx.x & 31;
    const bool a = threadIdx.x < 32;
    if (a) fa[lane] = 0; else fb[lane] = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 1; i <= iters; ++i) {
        if (a) { __threadfence_block(); fa[lane] = i; while (fb[lane] < i) {} }
        else { while (fa[lane] < i) {} __threadfence_block(); fb[lane] = i; }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}
int main() {
    long long* d; cudaMalloc(&d, 8 * 1024);
    const int iters = 10000;
    for (int blocks : {1, 148, 296, 444}) {
        long long h[3];
        k_sync<<<blocks, 64>>>(d, iters); cudaMemcpy(&h[0], d, 8, cudaMemcpyDeviceToHost);
        k_pingpong<<<blocks, 64>>>(d, iters); cudaMemcpy(&h[1], d, 8, cudaMemcpyDeviceToHost);
        k_flag<<<blocks, 64>>>(d, iters); cudaMemcpy(&h[2], d, 8, cudaMemcpyDeviceToHost);
        printf("blocks %3d: bar.sync both %.1f cycles, arrive/sync round trip %.1f cycles, flag round trip %.1f cycles  (%s)\n", blocks,
               (double)h[0] / iters, (double)h[1] / iters, (double)h[2] / iters, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
