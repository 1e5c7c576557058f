// Stand-alone multiplier experiment: modmul/s of fe_mul variants on the device, results cross-checked.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I halo_b200/csrc --expt-relaxed-constexpr -o tools/mulbench/mulbench tools/mulbench/mulbench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "variants.cuh"
using namespace halo;

template <int V, int F> __device__ __forceinline__ Fe<F> mulv(const Fe<F>& a, const Fe<F>& b) {
#if defined(__CUDA_ARCH__)
    if (V == 0) return fe_mul<F>(a, b);
    else return fe_mul_comba<F>(a, b);
#else
    return a;
#endif
}

template <int V, int ILP>
__global__ void __launch_bounds__(128) bench(int iters, Fe<0>* out) {
    u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    Fe<0> a[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[j].l[i] = tid * 2654435761u + 40503u * (i + 8 * j) + 17u;
        a[j].l[7] &= 0x3fffffffu;
    }
    Fe<0> m = a[ILP - 1];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) a[j] = mulv<V, 0>(a[j], m);
    }
    Fe<0> s = a[0];
#pragma unroll
    for (int j = 1; j < ILP; ++j) s = fe_add<0>(s, a[j]);
    out[tid] = s;
}

template <int V, int ILP> double run(int blocks, int iters, Fe<0>* d, Fe<0>* h, size_t n) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        bench<V, ILP><<<blocks, 128>>>(iters, d);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    cudaMemcpy(h, d, n * sizeof(Fe<0>), cudaMemcpyDeviceToHost);
    cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(1); }
    return (double)blocks * 128 * iters * ILP / (best * 1e-3);
}

int main() {
    int sms = 148, blocks = sms * 16, iters = 2000;
    size_t n = (size_t)blocks * 128;
    Fe<0>* d; cudaMalloc(&d, n * sizeof(Fe<0>));
    Fe<0>* h0 = (Fe<0>*)malloc(n * sizeof(Fe<0>)); Fe<0>* h1 = (Fe<0>*)malloc(n * sizeof(Fe<0>));
#define CMP(ILP) { double r0 = run<0, ILP>(blocks, iters, d, h0, n); double r1 = run<1, ILP>(blocks, iters, d, h1, n); \
        size_t bad = 0; for (size_t i = 0; i < n; ++i) for (int k = 0; k < 8; ++k) bad += h0[i].l[k] != h1[i].l[k]; \
        printf("ILP %d: fe_mul %.3e modmul/s  comba %.3e modmul/s  ratio %.3f  mismatches %zu\n", ILP, r0, r1, r1 / r0, bad); }
    CMP(1) CMP(2) CMP(4)
    return 0;
}
