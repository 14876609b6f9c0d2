// knn.cu — distCUDA2: mean squared distance to the three nearest neighbours of every point.
//
// Replaces SimpleKNN::knn and its kernels (reference submodules/simple-knn/simple_knn.cu:64-222; binding
// spatial.cu:15-26).  Same idea — Morton order, boxes of consecutive points with AABB rejection, exact
// 3-NN — rebuilt without CUB/thrust, without cudaMalloc and without the two device->host copies of the
// scene bounds (the bounds stay on the device):
//   1. knn_bounds_kernel      scene AABB by block reduction + ordered-uint atomics
//   2. knn_morton_kernel      30-bit Morton code per point (+ radix digit histograms)
//   3. radix sort             the library's own (u32 key, u32 value) LSD sort (binning.cu), 4 passes
//   4. knn_gather_box_kernel  gather points into Morton order (float4) + AABB per 256-point box
//   5. knn_search_kernel      per point: seed with the +-3 Morton neighbours, then visit every box whose
//                             AABB distance does not exceed the current third-best distance
// The reference seeds its min/max reductions with {0,0,0} (simple_knn.cu:190-201), which only changes the
// Morton quantisation, not the (exact) result; that quirk is not reproduced.
#include "common.cuh"
#include "../../include/vcr_gaus_b200.h"
#include <float.h>

#define KNN_BOX 256

__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(u);
}

__global__ void knn_bounds_init_kernel(uint32_t* bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xffffffffu;      // min (ordered)
    else if (threadIdx.x < 6) bounds[threadIdx.x] = 0u;          // max (ordered)
}

__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ pts, uint32_t* __restrict__ bounds) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float c = __ldg(pts + 3 * (size_t)i + k);
            mn[k] = fminf(mn[k], c); mx[k] = fmaxf(mx[k], c);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) { atomicMin(bounds + k, f2ord(mn[k])); atomicMax(bounds + 3 + k, f2ord(mx[k])); }
    }
}

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bounds,
                                                         uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                         uint32_t* __restrict__ ghist) {
    __shared__ uint32_t hist[4][256];
    for (int k = threadIdx.x; k < 4 * 256; k += 256) (&hist[0][0])[k] = 0u;
    __syncthreads();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < P) {
        uint32_t code = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float lo = ord2f(bounds[k]), hi = ord2f(bounds[3 + k]);
            const float ext = hi - lo;
            const float c = __ldg(pts + 3 * (size_t)i + k);
            const float nrm = ext > 0.f ? (c - lo) / ext : 0.f;
            const uint32_t q = (uint32_t)fminf(fmaxf(nrm * 1023.0f, 0.f), 1023.f);
            code |= prep_morton(q) << k;
        }
        keys[i] = code;
        vals[i] = (uint32_t)i;
#pragma unroll
        for (int p = 0; p < 4; p++) atomicAdd(&hist[p][(code >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 4; p++) {
        const uint32_t c = hist[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * 256 + threadIdx.x], c);
    }
}

__global__ void __launch_bounds__(KNN_BOX) knn_gather_box_kernel(int P, const float* __restrict__ pts,
                                                                 const uint32_t* __restrict__ order, float4* __restrict__ sorted,
                                                                 float4* __restrict__ boxes) {
    const int i = blockIdx.x * KNN_BOX + threadIdx.x;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t src = order[i];
        const float x = __ldg(pts + 3 * (size_t)src), y = __ldg(pts + 3 * (size_t)src + 1), z = __ldg(pts + 3 * (size_t)src + 2);
        sorted[i] = make_float4(x, y, z, __uint_as_float(src));
        mn[0] = mx[0] = x; mn[1] = mx[1] = y; mn[2] = mx[2] = z;
    }
    __shared__ float smn[3][KNN_BOX / 32], smx[3][KNN_BOX / 32];
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) { smn[k][threadIdx.x >> 5] = mn[k]; smx[k][threadIdx.x >> 5] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            a[k] = smn[k][0]; b[k] = smx[k][0];
            for (int w = 1; w < KNN_BOX / 32; w++) { a[k] = fminf(a[k], smn[k][w]); b[k] = fmaxf(b[k], smx[k][w]); }
        }
        boxes[2 * blockIdx.x] = make_float4(a[0], a[1], a[2], 0.f);
        boxes[2 * blockIdx.x + 1] = make_float4(b[0], b[1], b[2], 0.f);
    }
}

__device__ __forceinline__ void knn_update3(const float4 ref, const float4 q, float (&best)[3]) {
    const float dx = q.x - ref.x, dy = q.y - ref.y, dz = q.z - ref.z;
    float d = dx * dx + dy * dy + dz * dz;  // same expression as updateKBest (simple_knn.cu:132-146)
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (best[j] > d) { const float t = best[j]; best[j] = d; d = t; }
    }
}
__device__ __forceinline__ float knn_box_dist(const float4 mn, const float4 mx, const float4 p) {
    float dx = 0.f, dy = 0.f, dz = 0.f;
    if (p.x < mn.x || p.x > mx.x) dx = fminf(fabsf(p.x - mn.x), fabsf(p.x - mx.x));
    if (p.y < mn.y || p.y > mx.y) dy = fminf(fabsf(p.y - mn.y), fabsf(p.y - mx.y));
    if (p.z < mn.z || p.z > mx.z) dz = fminf(fabsf(p.z - mn.z), fabsf(p.z - mx.z));
    return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(256) knn_search_kernel(int P, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
                                                         int nboxes, float* __restrict__ out) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const bool live = idx < P;
    const float4 me = live ? sorted[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    if (live) {
        for (int i = max(0, idx - 3); i <= min(P - 1, idx + 3); i++) {
            if (i == idx) continue;
            knn_update3(me, sorted[i], best);
        }
    }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    for (int b = 0; b < nboxes; b++) {
        const float4 mn = __ldg(boxes + 2 * b), mx = __ldg(boxes + 2 * b + 1);
        const float d = knn_box_dist(mn, mx, me);
        const bool visit = live && !(d > reject || d > best[2]);
        if (!__any_sync(0xffffffffu, visit)) continue;
        const int lo = b * KNN_BOX, hi = min(P, lo + KNN_BOX);
        if (visit) {
            for (int i = lo; i < hi; i++) {
                if (i == idx) continue;
                knn_update3(me, __ldg(sorted + i), best);
            }
        }
    }
    if (live) out[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

// scratch layout: BinningState(P) (sort buffers) | bounds u32[8] | sorted float4[P] | boxes float4[2*nboxes]
extern "C" size_t vcr_knn_scratch_bytes(int P) {
    const size_t nboxes = ((size_t)P + KNN_BOX - 1) / KNN_BOX;
    return BinningState::bytes(P) + 256 + 64 + 256 + (size_t)P * sizeof(float4) + 256 + 2 * nboxes * sizeof(float4) + 256;
}

extern "C" int vcr_dist2_knn3(int P, const float* points, float* mean_dists, char* scratch, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0) return 0;
    if (!points || !mean_dists || !scratch) { vcr_set_error("dist2_knn3: NULL argument"); return 1; }
    BinningState b = BinningState::carve(scratch, P);
    char* p = scratch + BinningState::bytes(P);
    p = reinterpret_cast<char*>(vcr_align256(reinterpret_cast<size_t>(p)));
    uint32_t* bounds = reinterpret_cast<uint32_t*>(p); p += 64;
    p = reinterpret_cast<char*>(vcr_align256(reinterpret_cast<size_t>(p)));
    float4* sorted = reinterpret_cast<float4*>(p); p += (size_t)P * sizeof(float4);
    p = reinterpret_cast<char*>(vcr_align256(reinterpret_cast<size_t>(p)));
    float4* boxes = reinterpret_cast<float4*>(p);
    const int nboxes = (P + KNN_BOX - 1) / KNN_BOX;
    const int blocks = (P + 255) / 256;

    VcrProfScope prof(VCR_K_KNN, s, 5);
    knn_bounds_init_kernel<<<1, 32, 0, s>>>(bounds);
    VCR_LAUNCH_CHECK("knn_bounds_init_kernel");
    knn_bounds_kernel<<<min(blocks, 148 * 8), 256, 0, s>>>(P, points, bounds);
    VCR_LAUNCH_CHECK("knn_bounds_kernel");
    VCR_CUDA_TRY(cudaMemsetAsync(b.sort_hist, 0, 8 * 256 * sizeof(uint32_t), s));
    knn_morton_kernel<<<blocks, 256, 0, s>>>(P, points, bounds, b.tile_keys, b.point_list, b.sort_hist);
    VCR_LAUNCH_CHECK("knn_morton_kernel");
    if (vcr_launch_radix_sort(b, P, 32, s)) return 1;
    knn_gather_box_kernel<<<nboxes, KNN_BOX, 0, s>>>(P, points, b.point_list, sorted, boxes);
    VCR_LAUNCH_CHECK("knn_gather_box_kernel");
    knn_search_kernel<<<blocks, 256, 0, s>>>(P, sorted, boxes, nboxes, mean_dists);
    VCR_LAUNCH_CHECK("knn_search_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// vcr_dist2_knn3_noalloc — the same computation for callers that cannot obtain device memory at the time of the call.
//
// vcr_dist2_knn3 needs a few megabytes of scratch from its caller.  Some callers have nowhere to get it: a host program
// that has handed the whole GPU to another allocator (a framework's caching pool that already owns all free memory, so
// a fresh cudaMalloc fails), a plug-in running under a "no device allocation after start-up" rule, or code invoked
// from inside an allocator's own callback, where asking for more memory would re-enter it.  For them the scratch lives
// in the fixed device array below, which is part of the library image: every GPU that loads the library sets aside
// VCR_KNN_NOALLOC_BYTES for it until the process exits, whether or not this entry point is ever called.
//
// Limits and rules
//   * at most VCR_KNN_NOALLOC_MAX_POINTS points per call; larger inputs are refused (use vcr_dist2_knn3);
//   * nothing is allocated or freed here, by this function or by anything it calls (kernel launches and
//     cudaMemsetAsync only); inputs and outputs are still the caller's;
//   * the first call on each device looks up the device address of the array (cudaGetSymbolAddress) and remembers it;
//     make that first call outside any stream capture;
//   * all calls on a device share the one array, and it keeps whatever the previous call left in it.  The parts that
//     must start from zero (digit histograms, look-back words, bounds) are cleared on every call, everything else is
//     overwritten before it is read.  Two calls that overlap on the same GPU (different streams, different threads)
//     would trample each other's scratch: serialise them.
// ------------------------------------------------------------------------------------------------------------------
#define VCR_KNN_NOALLOC_MAX_POINTS 65536
#define VCR_KNN_NOALLOC_BYTES (4u << 20)
static __device__ __align__(256) unsigned char g_knn_noalloc_scratch[VCR_KNN_NOALLOC_BYTES];

#include <mutex>
#define VCR_MAX_DEVICES 64
static char* g_knn_noalloc_ptr[VCR_MAX_DEVICES] = {nullptr};
static std::mutex g_knn_noalloc_mutex;

// In host code `g_knn_noalloc_scratch` is only a handle for the symbol, not an address a kernel could use: the device
// address has to be asked for, per device, with that device current.
static int knn_noalloc_scratch_for_current_device(char** out) {
    int dev = -1;
    VCR_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= VCR_MAX_DEVICES) { vcr_set_error("dist2_knn3_noalloc: unsupported device index %d", dev); return 1; }
    std::lock_guard<std::mutex> lock(g_knn_noalloc_mutex);
    if (!g_knn_noalloc_ptr[dev]) {
        char* p = nullptr;
        VCR_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_knn_noalloc_scratch));
        if (!p) { vcr_set_error("dist2_knn3_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev); return 1; }
        g_knn_noalloc_ptr[dev] = p;
    }
    *out = g_knn_noalloc_ptr[dev];
    return 0;
}

extern "C" int vcr_knn_noalloc_max_points(void) { return VCR_KNN_NOALLOC_MAX_POINTS; }

extern "C" int vcr_dist2_knn3_noalloc(int P, const float* points, float* mean_dists, void* stream) {
    if (P <= 0) return 0;
    if (!points || !mean_dists) { vcr_set_error("dist2_knn3_noalloc: NULL argument"); return 1; }
    if (P > VCR_KNN_NOALLOC_MAX_POINTS || vcr_knn_scratch_bytes(P) > (size_t)VCR_KNN_NOALLOC_BYTES) {
        vcr_set_error("dist2_knn3_noalloc: %d points exceed the built-in scratch (at most %d points, %u bytes); "
                      "use vcr_dist2_knn3 with caller-provided scratch", P, VCR_KNN_NOALLOC_MAX_POINTS, VCR_KNN_NOALLOC_BYTES);
        return 1;
    }
    char* scratch = nullptr;
    if (knn_noalloc_scratch_for_current_device(&scratch)) return 1;
    return vcr_dist2_knn3(P, points, mean_dists, scratch, stream);
}
