// Phase profile of the dense large-matrix LU (dense_kernels.cu built with -DCSIM_DN_PROF): cycles of
// thread 0 per phase, summed over the CTAs, for a batch of random dense systems.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -DCSIM_DN_PROF -Iinclude \
//        -o build/dn_prof tools/dn_prof.cu circuitsimulator_b200/csrc/dense_kernels.cu
//   build/dn_prof [n=404] [B=296]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../circuitsimulator_b200/csrc/kernels.hpp"
namespace csim {
cudaError_t dn_prof_read(unsigned long long* out);
}

int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 404;
    const long B = argc > 2 ? atol(argv[2]) : 296;
    std::vector<double> A((size_t)n * n * B), b((size_t)n * B);
    srand(1);
    for (auto& v : A) v = rand() / (double)RAND_MAX - 0.5;
    for (auto& v : b) v = rand() / (double)RAND_MAX - 0.5;
    double *dA, *db, *dx;
    int *dperm, *dsing;
    cudaMalloc(&dA, A.size() * 8);
    cudaMalloc(&db, b.size() * 8);
    cudaMalloc(&dx, b.size() * 8);
    cudaMalloc(&dperm, b.size() * 4);
    cudaMalloc(&dsing, B * 4);
    cudaMemcpy(dA, A.data(), A.size() * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(db, b.data(), b.size() * 8, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        unsigned long long p0[8], p1[8];
        csim::dn_prof_read(p0);
        cudaEventRecord(e0);
        cudaError_t e = csim::launch_lu_dense(n, B, dA, db, dx, dperm, dsing, 1e-15, 0);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) {
            printf("error %s\n", cudaGetErrorString(e));
            return 1;
        }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        csim::dn_prof_read(p1);
        const char* names[5] = {"panel load", "panel factor", "U11 + strip", "trailing update", "hash + back subst"};
        printf("n=%d B=%ld: %.3f ms, %.0f solves/s;", n, B, ms, B / (ms * 1e-3));
        unsigned long long tot = 0;
        for (int i = 0; i < 5; ++i) tot += p1[i] - p0[i];
        for (int i = 0; i < 5; ++i) printf(" %s %.1f%% (%.0f kcyc/solve)", names[i], 100.0 * (p1[i] - p0[i]) / tot, (p1[i] - p0[i]) / 1e3 / B);
        printf("\n");
    }
    return 0;
}
