// view_parallel.cu — per-view densification statistics, one pass over the Gaussians.
//
// Replaces the three boolean-mask-indexed PyTorch updates the reference trainer runs after every backward
// (trainer.py:345 max_radii2D; scene/gaussian_model.py:669-671 add_densification_stats), which cost ~10 small launches
// and three compactions over P.  HBM bound: 12 B (grad xy padded to xyz) + 4 B (radii) read, 3 x 4 B read-modify-write
// for visible Gaussians.  The accumulators stay local to the rank; vcr_view_parallel.py combines them across ranks
// only when the model densifies.
#include "common.cuh"

__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const float* __restrict__ grad_means2D,
                                                            const int* __restrict__ radii, float* __restrict__ accum,
                                                            float* __restrict__ denom, float* __restrict__ max_radii2D) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int r = __ldg(radii + i);
    if (!(r > 0)) return;                                   // visibility_filter = radii > 0 (gaussian_renderer/__init__.py:147)
    const float gx = __ldg(grad_means2D + 3 * (size_t)i), gy = __ldg(grad_means2D + 3 * (size_t)i + 1);
    accum[i] += sqrtf(__fmaf_rn(gx, gx, __fmul_rn(gy, gy)));  // torch.norm(grad[:, :2], dim=-1)
    denom[i] += 1.0f;
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
}

int vcr_launch_densify_stats(int P, const float* grad_means2D, const int* radii, float* xyz_gradient_accum, float* denom,
                             float* max_radii2D, cudaStream_t s) {
    VcrProfScope prof(VCR_K_DENSIFY_STATS, s);
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, grad_means2D, radii, xyz_gradient_accum, denom, max_radii2D);
    VCR_LAUNCH_CHECK("densify_stats_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// Gradient exchange through the NVSwitch: two-shot all-reduce(SUM) over a multicast mapping.
//
// Every rank has its gradient bucket in symmetric memory (same offset on every GPU) and the NVLink multicast object that
// spans all of them is mapped at `mc`.  Rank r owns the r-th 1/world of every range: one `multimem.ld_reduce` pulls a
// 16-byte vector from all ranks and adds it INSIDE the switch, one `multimem.st` pushes the sum back to all ranks.  Per
// GPU that is S bytes out and S bytes in for an S-byte bucket — the all-reduce minimum — with no staging buffer and no
// per-peer loop, and every replica receives bit-identical sums (one reducer per element).
//
// The kernel holds no lock and waits for nobody: the caller orders it after "all ranks have written these rows" and
// orders the consumers after "all ranks have finished this kernel" with the symmetric-memory barrier of its runtime.
// Up to VCR_MM_MAX_RANGES float ranges per launch, so the five parameter-group slices of one Gaussian range go in one
// launch (chunked backward, vcr_rasterize_backward_chunked).
#define VCR_MM_MAX_RANGES 8
struct MmRanges {
    long long first[VCR_MM_MAX_RANGES];   // in floats from the start of the mapping
    long long count[VCR_MM_MAX_RANGES];
    int n;
};

__device__ __forceinline__ float4 mm_ld_reduce4(const float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st4(float* p, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float mm_ld_reduce1(const float* p) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st1(float* p, const float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}

#define MM_UNROLL 4
__global__ void __launch_bounds__(512) multimem_allreduce_kernel(float* __restrict__ mc, const MmRanges rg, const int rank,
                                                                 const int world) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    for (int r = 0; r < rg.n; r++) {
        const long long f0 = rg.first[r], f1 = f0 + rg.count[r];
        const long long a = (f0 + 3) & ~3LL, b = f1 & ~3LL;                 // 16-byte aligned interior [a, b)
        if (b > a) {
            const long long nv = (b - a) >> 2;
            const long long v0 = nv * rank / world, v1 = nv * (rank + 1) / world;   // this rank's share, in float4
            float* base = mc + a;
            long long v = v0 + tid;
            for (; v + (MM_UNROLL - 1) * nthreads < v1; v += MM_UNROLL * nthreads) {
                float4 x[MM_UNROLL];
#pragma unroll
                for (int u = 0; u < MM_UNROLL; u++) x[u] = mm_ld_reduce4(base + 4 * (v + u * nthreads));
#pragma unroll
                for (int u = 0; u < MM_UNROLL; u++) mm_st4(base + 4 * (v + u * nthreads), x[u]);
            }
            for (; v < v1; v += nthreads) mm_st4(base + 4 * v, mm_ld_reduce4(base + 4 * v));
        }
        // unaligned head / tail (at most 3 floats each): rank 0, one thread
        if (rank == 0 && tid == 0) {
            const long long he = a < f1 ? a : f1;
            for (long long i = f0; i < he; i++) mm_st1(mc + i, mm_ld_reduce1(mc + i));
            for (long long i = (b > he ? b : he); i < f1; i++) mm_st1(mc + i, mm_ld_reduce1(mc + i));
        }
    }
}

// The same two-shot reduction of up to 8 ranges plus, in the same launch, an all-gather contribution: this rank's slice
// [bcast_first, +bcast_count) of the symmetric buffer is read locally and multicast to every rank (multimem.st), i.e. S
// bytes out once and (world-1) x S bytes in per GPU with the replication done by the switch.  Used by the compact SH
// exchange: all-reduce the 11 non-SH floats per Gaussian, all-gather the 3 colour-gradient floats.
__global__ void __launch_bounds__(512) multimem_exchange_kernel(float* __restrict__ mc, const float* __restrict__ local,
                                                                const MmRanges rg, const long long bfirst, const long long bcount,
                                                                const int rank, const int world) {
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    // broadcast part first: plain local loads, stores that the switch replicates
    {
        const long long nv = bcount >> 2;
        const float4* src = reinterpret_cast<const float4*>(local + bfirst);
        float* dst = mc + bfirst;
        long long v = tid;
        for (; v + (MM_UNROLL - 1) * nthreads < nv; v += MM_UNROLL * nthreads) {
            float4 x[MM_UNROLL];
#pragma unroll
            for (int u = 0; u < MM_UNROLL; u++) x[u] = src[v + u * nthreads];
#pragma unroll
            for (int u = 0; u < MM_UNROLL; u++) mm_st4(dst + 4 * (v + u * nthreads), x[u]);
        }
        for (; v < nv; v += nthreads) mm_st4(dst + 4 * v, src[v]);
        if (tid == 0) for (long long i = nv << 2; i < bcount; i++) mm_st1(dst + i, local[bfirst + i]);
    }
    for (int r = 0; r < rg.n; r++) {
        const long long f0 = rg.first[r], f1 = f0 + rg.count[r];
        const long long a = (f0 + 3) & ~3LL, b = f1 & ~3LL;
        if (b > a) {
            const long long nv = (b - a) >> 2;
            const long long v0 = nv * rank / world, v1 = nv * (rank + 1) / world;
            float* base = mc + a;
            long long v = v0 + tid;
            for (; v + (MM_UNROLL - 1) * nthreads < v1; v += MM_UNROLL * nthreads) {
                float4 x[MM_UNROLL];
#pragma unroll
                for (int u = 0; u < MM_UNROLL; u++) x[u] = mm_ld_reduce4(base + 4 * (v + u * nthreads));
#pragma unroll
                for (int u = 0; u < MM_UNROLL; u++) mm_st4(base + 4 * (v + u * nthreads), x[u]);
            }
            for (; v < v1; v += nthreads) mm_st4(base + 4 * v, mm_ld_reduce4(base + 4 * v));
        }
        if (rank == 0 && tid == 0) {
            const long long he = a < f1 ? a : f1;
            for (long long i = f0; i < he; i++) mm_st1(mc + i, mm_ld_reduce1(mc + i));
            for (long long i = (b > he ? b : he); i < f1; i++) mm_st1(mc + i, mm_ld_reduce1(mc + i));
        }
    }
}

int vcr_launch_multimem_exchange(float* mc, const float* local, int n_ranges, const long long* first, const long long* count,
                                 long long bcast_first, long long bcast_count, int rank, int world, int blocks, cudaStream_t s) {
    MmRanges rg;
    rg.n = n_ranges;
    for (int i = 0; i < VCR_MM_MAX_RANGES; i++) { rg.first[i] = i < n_ranges ? first[i] : 0; rg.count[i] = i < n_ranges ? count[i] : 0; }
    VcrProfScope prof(VCR_K_MM_ALLREDUCE, s);
    multimem_exchange_kernel<<<blocks, 512, 0, s>>>(mc, local, rg, bcast_first, bcast_count, rank, world);
    VCR_LAUNCH_CHECK("multimem_exchange_kernel");
    return 0;
}

int vcr_launch_multimem_allreduce(float* mc, int n_ranges, const long long* first, const long long* count, int rank, int world,
                                  int blocks, cudaStream_t s) {
    MmRanges rg;
    rg.n = n_ranges;
    for (int i = 0; i < VCR_MM_MAX_RANGES; i++) { rg.first[i] = i < n_ranges ? first[i] : 0; rg.count[i] = i < n_ranges ? count[i] : 0; }
    VcrProfScope prof(VCR_K_MM_ALLREDUCE, s);
    multimem_allreduce_kernel<<<blocks, 512, 0, s>>>(mc, rg, rank, world);
    VCR_LAUNCH_CHECK("multimem_allreduce_kernel");
    return 0;
}
