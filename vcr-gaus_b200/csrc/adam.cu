// adam.cu — Adam step over the flat parameter / gradient bucket of the view-parallel layer.
//
// The reference steps six parameter groups with torch.optim.Adam(l, lr=0.0, eps=1e-15) and per-group learning rates
// (scene/gaussian_model.py:241-258; the xyz rate follows a schedule, :264-270).  With all per-Gaussian parameters and
// gradients in two flat tensors the whole step is one launch: every element is read once (p, g, m, v) and written once
// (p, m, v) — 28 B per float, HBM bound — and the learning rate is looked up from the range the element falls in.
#include "common.cuh"

#define VCR_ADAM_MAX_GROUPS 16
struct AdamGroups {
    long long end[VCR_ADAM_MAX_GROUPS];     // exclusive end of each group's range, ascending; group 0 starts at 0
    float step_size[VCR_ADAM_MAX_GROUPS];   // lr / (1 - beta1^t)
    int n;
};

__global__ void __launch_bounds__(256) adam_step_kernel(const long long n, float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v, const AdamGroups gr,
                                                        const float beta1, const float beta2, const float eps,
                                                        const float inv_sqrt_bc2, const float grad_scale) {
    const long long i4 = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
    if (i4 >= n) return;
    auto one = [&](const long long i, float& pp, const float gg_, float& mm, float& vv) {
        int k = 0;
        while (k + 1 < gr.n && i >= gr.end[k]) k++;
        const float gg = gg_ * grad_scale;
        mm = beta1 * mm + (1.f - beta1) * gg;                  // exp_avg.lerp_(grad, 1 - beta1)
        vv = beta2 * vv + (1.f - beta2) * gg * gg;             // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        const float denom = sqrtf(vv) * inv_sqrt_bc2 + eps;    // (exp_avg_sq.sqrt() / sqrt(bias_correction2)).add_(eps)
        pp -= gr.step_size[k] * (mm / denom);                  // param.addcdiv_(exp_avg, denom, value=-step_size)
    };
    if (i4 + 4 <= n) {
        float4 P4 = reinterpret_cast<float4*>(p)[i4 >> 2];
        const float4 G4 = __ldg(reinterpret_cast<const float4*>(g) + (i4 >> 2));
        float4 M4 = reinterpret_cast<float4*>(m)[i4 >> 2], V4 = reinterpret_cast<float4*>(v)[i4 >> 2];
        one(i4, P4.x, G4.x, M4.x, V4.x); one(i4 + 1, P4.y, G4.y, M4.y, V4.y);
        one(i4 + 2, P4.z, G4.z, M4.z, V4.z); one(i4 + 3, P4.w, G4.w, M4.w, V4.w);
        reinterpret_cast<float4*>(p)[i4 >> 2] = P4; reinterpret_cast<float4*>(m)[i4 >> 2] = M4; reinterpret_cast<float4*>(v)[i4 >> 2] = V4;
    } else {
        for (long long i = i4; i < n; i++) one(i, p[i], g[i], m[i], v[i]);
    }
}

extern "C" int vcr_adam_step(long long numel, float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int n_groups,
                             const long long* group_end_host, const float* lr_host, float beta1, float beta2, float eps, int step,
                             float grad_scale, void* stream) {
    if (numel <= 0) return 0;
    if (!params || !grads || !exp_avg || !exp_avg_sq || !group_end_host || !lr_host) { vcr_set_error("adam_step: NULL argument"); return 1; }
    if (n_groups < 1 || n_groups > VCR_ADAM_MAX_GROUPS) { vcr_set_error("adam_step: 1..16 groups"); return 1; }
    if (step < 1) { vcr_set_error("adam_step: step counts from 1"); return 1; }
    if ((reinterpret_cast<size_t>(params) | reinterpret_cast<size_t>(grads) | reinterpret_cast<size_t>(exp_avg) |
         reinterpret_cast<size_t>(exp_avg_sq)) & 15) { vcr_set_error("adam_step: buffers must be 16-byte aligned"); return 1; }
    AdamGroups gr;
    memset(&gr, 0, sizeof(gr));
    gr.n = n_groups;
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    long long prev = 0;
    for (int k = 0; k < n_groups; k++) {
        if (group_end_host[k] < prev) { vcr_set_error("adam_step: group ends must ascend"); return 1; }
        prev = gr.end[k] = group_end_host[k];
        gr.step_size[k] = (float)((double)lr_host[k] / bc1);
    }
    if (prev != numel) { vcr_set_error("adam_step: the last group must end at numel"); return 1; }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    VcrProfScope prof(VCR_K_ADAM, s);
    const long long quads = (numel + 3) / 4;
    adam_step_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, s>>>(numel, params, grads, exp_avg, exp_avg_sq, gr, beta1, beta2, eps,
                                                                     (float)(1.0 / sqrt(bc2)), grad_scale);
    VCR_LAUNCH_CHECK("adam_step_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Densify / prune as stream compaction.  The reference prunes with ~18 boolean-mask indexing
// operations (every parameter group and both Adam moments: _prune_optimizer, scene/gaussian_model.py:442-460) and grows
// with as many torch.cat (cat_tensors_to_optimizer, :480-502).  Here: one exclusive prefix sum of the keep mask, then ONE
// gather launch that moves the kept rows of up to 8 arrays (of different row widths) into their new places.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) keep_block_sums_kernel(int P, const uint8_t* __restrict__ keep, uint32_t* __restrict__ sums) {
    __shared__ uint32_t warp_tot[8];
    const int i = blockIdx.x * 256 + threadIdx.x;
    const uint32_t v = (i < P && keep[i]) ? 1u : 0u;
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = wsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) t += warp_tot[w];
        sums[blockIdx.x] = t;
    }
}

// one block: exclusive scan of the block sums in place, total -> sums[nb]
__global__ void __launch_bounds__(1024) keep_scan_kernel(uint32_t* __restrict__ sums, int nb) {
    __shared__ uint32_t warp_tot[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int ipt = (nb + 1023) / 1024;
    const int first = t * ipt, last = min(first + ipt, nb);
    uint32_t v = 0;
    for (int i = first; i < last; i++) v += sums[i];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t x = warp_tot[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += n;
        }
        warp_tot[lane] = x;
    }
    __syncthreads();
    uint32_t run = (w ? warp_tot[w - 1] : 0u) + inc - v;
    for (int i = first; i < last; i++) {
        const uint32_t c = sums[i];
        sums[i] = run;
        run += c;
    }
    if (t == 1023) sums[nb] = warp_tot[31];
}

#define VCR_COMPACT_MAX_ARRAYS 8
struct CompactArrays {
    const float* src[VCR_COMPACT_MAX_ARRAYS];
    float* dst[VCR_COMPACT_MAX_ARRAYS];
    int width[VCR_COMPACT_MAX_ARRAYS];      // floats per row
    int n;
};

// block b handles rows [256 b, 256 b + 256): in-block rank of every kept row, then for each array the block's kept rows
// are copied float by float with consecutive threads on consecutive output floats (the kept rows of a block are contiguous
// in the output)
__global__ void __launch_bounds__(256) compact_rows_kernel(int P, const uint8_t* __restrict__ keep, const uint32_t* __restrict__ sums,
                                                           const CompactArrays arr) {
    __shared__ uint32_t s_src[256];          // source row of the block's j-th kept row
    __shared__ uint32_t warp_tot[8];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int i = blockIdx.x * 256 + t;
    const bool k = i < P && keep[i];
    const uint32_t ball = __ballot_sync(0xffffffffu, k);
    if (lane == 0) warp_tot[w] = __popc(ball);
    __syncthreads();
    uint32_t wex = 0, total = 0;
#pragma unroll
    for (int q = 0; q < 8; q++) { wex += (q < w) ? warp_tot[q] : 0u; total += warp_tot[q]; }
    if (k) s_src[wex + __popc(ball & ((1u << lane) - 1u))] = (uint32_t)i;
    __syncthreads();
    const size_t base = sums[blockIdx.x];
    for (int a = 0; a < arr.n; a++) {
        const int wd = arr.width[a];
        const float* __restrict__ src = arr.src[a];
        float* __restrict__ dst = arr.dst[a] + base * wd;
        const uint32_t n = total * (uint32_t)wd;
        for (uint32_t e = t; e < n; e += 256) {
            const uint32_t j = e / (uint32_t)wd, c = e - j * (uint32_t)wd;
            dst[e] = __ldg(src + (size_t)s_src[j] * wd + c);
        }
    }
}

extern "C" size_t vcr_compact_scratch_bytes(int P) { return ((size_t)(P + 255) / 256 + 2) * sizeof(uint32_t) + 256; }

extern "C" int vcr_compact_rows(int P, const uint8_t* keep, int n_arrays, const float* const* src_host, float* const* dst_host,
                                const int* width_host, char* scratch, uint32_t* n_kept_device, void* stream) {
    if (P <= 0) return 0;
    if (!keep || !scratch || n_arrays < 0 || (n_arrays > 0 && (!src_host || !dst_host || !width_host))) { vcr_set_error("compact_rows: NULL argument"); return 1; }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    uint32_t* sums = reinterpret_cast<uint32_t*>(vcr_align256(reinterpret_cast<size_t>(scratch)));
    const int nb = (P + 255) / 256;
    VcrProfScope prof(VCR_K_ADAM, s, 2 + (n_arrays + VCR_COMPACT_MAX_ARRAYS - 1) / VCR_COMPACT_MAX_ARRAYS);
    keep_block_sums_kernel<<<nb, 256, 0, s>>>(P, keep, sums);
    VCR_LAUNCH_CHECK("keep_block_sums_kernel");
    keep_scan_kernel<<<1, 1024, 0, s>>>(sums, nb);
    VCR_LAUNCH_CHECK("keep_scan_kernel");
    if (n_kept_device) VCR_CUDA_TRY(cudaMemcpyAsync(n_kept_device, sums + nb, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
    for (int a0 = 0; a0 < n_arrays; a0 += VCR_COMPACT_MAX_ARRAYS) {
        CompactArrays arr;
        memset(&arr, 0, sizeof(arr));
        arr.n = n_arrays - a0 < VCR_COMPACT_MAX_ARRAYS ? n_arrays - a0 : VCR_COMPACT_MAX_ARRAYS;
        for (int a = 0; a < arr.n; a++) {
            if (!src_host[a0 + a] || !dst_host[a0 + a] || width_host[a0 + a] <= 0) { vcr_set_error("compact_rows: bad array %d", a0 + a); return 1; }
            arr.src[a] = src_host[a0 + a]; arr.dst[a] = dst_host[a0 + a]; arr.width[a] = width_host[a0 + a];
        }
        compact_rows_kernel<<<nb, 256, 0, s>>>(P, keep, sums, arr);
        VCR_LAUNCH_CHECK("compact_rows_kernel");
    }
    return 0;
}
