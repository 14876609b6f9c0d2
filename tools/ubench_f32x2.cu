// tools/ubench_f32x2.cu — issue-rate microbenchmark: scalar FFMA against packed fma.rn.f32x2 (FFMA2) on sm_100a.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/ubench_f32x2 tools/ubench_f32x2.cu && gpurun_out/ubench_f32x2
#include <cstdio>
#include <cuda_runtime.h>
// throughput of scalar FFMA vs packed fma.rn.f32x2 on sm_100a: N dependent-chain-free FMAs per thread
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b); rc = *reinterpret_cast<unsigned long long*>(&c);
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
    float a[8], b[8];
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; b[i] = s + i; }
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = fmaf(a[i], b[i], s);
        } else {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                float2 r = ffma2(make_float2(a[i], a[i + 1]), make_float2(b[i], b[i + 1]), make_float2(s, s));
                a[i] = r.x; a[i + 1] = r.y;
            }
        }
    }
    float t = 0; for (int i = 0; i < 8; i++) t += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(d, iters, 1.0001f); else k<1><<<148 * 8, 256>>>(d, iters, 1.0001f);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fma = 148.0 * 8 * 256 * 8 * iters;
            printf("UBENCH mode=%s rep=%d ms=%.3f  FMA/s=%.2f T  (TFLOP/s %.1f)\n", mode ? "f32x2" : "scalar", rep, ms, fma / ms / 1e9, 2 * fma / ms / 1e9);
        }
    }
    return 0;
}
