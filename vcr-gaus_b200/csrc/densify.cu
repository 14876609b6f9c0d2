// densify.cu — the decisions and the new rows of one densification round, on the device.
//
// Replaces, in GaussianModel.densify_and_prune (reference scene/gaussian_model.py:645-659):
//   densify_and_clone                     :628-643   copies of small Gaussians with a large view-space gradient
//   densify_and_split_along_maxscaling    :579-626   two children along the longest axis (RNG-free), parent removed
//   the final opacity / size prune        :651-657
// and the ~60 PyTorch launches (boolean-mask indexing, torch.cat, build_rotation, gather, ...) they take.  The reference
// runs the three stages one after the other, each rebuilding every parameter tensor; the outcome is a pure function of the
// pre-round state, so here ONE kernel decides, per original Gaussian, what survives:
//   bit 0  the Gaussian itself is kept            (= not split and not pruned)
//   bit 1  its clone is appended                  (= cloned and the clone — same values — not pruned)
//   bit 2  its two split children are appended    (= split and the children — same opacity, shrunk longest axis — not pruned)
// Row order of the result, as in the reference: kept originals in index order | clones in index order | first children
// in index order | second children in index order.  vcr_compact_rows moves the kept rows (parameters and Adam moments),
// densify_fill_kernel writes the new rows; their Adam moments are zero (cat_tensors_to_optimizer, :480-502).
//
// Reference quirks kept on purpose: the clones carry gradient 0 into the split stage (padded_grad, :582-584), so a clone is
// never split; max_radii2D has been reset to zero by densification_postfix (:520-523) before the prune reads it, so the
// screen-size test of the prune never fires and only `max scaling > 0.1 extent` is applied when max_screen_size is set.
#include "common.cuh"
#include <math.h>
#include <string.h>

struct DensifyParams {
    int P;
    const float* scaling_raw; const float* opacity_raw; const float* accum; const float* denom;
    float max_grad, min_opacity, dense_limit /* percent_dense * extent */, world_limit /* 0.1 * extent, or +inf */;
};

__device__ __forceinline__ uint32_t densify_decide_one(const DensifyParams& p, int i) {
    float g = p.accum[i] / p.denom[i];                 // grads = xyz_gradient_accum / denom; grads[grads.isnan()] = 0
    if (g != g) g = 0.f;
    const float s0 = expf(p.scaling_raw[3 * (size_t)i]), s1 = expf(p.scaling_raw[3 * (size_t)i + 1]), s2 = expf(p.scaling_raw[3 * (size_t)i + 2]);
    const float smax = fmaxf(s0, fmaxf(s1, s2));
    const bool grad_ok = g >= p.max_grad;
    const bool clone = grad_ok && smax <= p.dense_limit;
    const bool split = grad_ok && smax > p.dense_limit;
    const float opacity = 1.0f / (1.0f + expf(-p.opacity_raw[i]));
    const bool low_opacity = opacity < p.min_opacity;
    const bool prune_self = low_opacity || smax > p.world_limit;
    // children: the longest axis becomes smax / (0.8 * 2); the other two are unchanged
    const float shrunk = smax / 1.6f;
    float c0 = s0, c1 = s1, c2 = s2;
    if (s0 >= s1 && s0 >= s2) c0 = shrunk; else if (s1 >= s2) c1 = shrunk; else c2 = shrunk;   // torch.argmax: first maximum
    const bool prune_child = low_opacity || fmaxf(c0, fmaxf(c1, c2)) > p.world_limit;
    uint32_t f = 0;
    if (!split && !prune_self) f |= 1u;
    if (clone && !prune_self) f |= 2u;
    if (split && !prune_child) f |= 4u;
    return f;
}

// flags + per-256-block counts of the three bits: sums[b], sums[nb1 + b], sums[2 nb1 + b]  (nb1 = nb + 1)
__global__ void __launch_bounds__(256) densify_decide_kernel(const DensifyParams p, uint8_t* __restrict__ flags, uint32_t* __restrict__ sums, int nb1) {
    __shared__ uint32_t wt[3][8];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int i = blockIdx.x * 256 + t;
    uint32_t f = 0;
    if (i < p.P) { f = densify_decide_one(p, i); flags[i] = (uint8_t)f; }
#pragma unroll
    for (int b = 0; b < 3; b++) {
        const uint32_t c = __popc(__ballot_sync(0xffffffffu, (f >> b) & 1u));
        if (lane == 0) wt[b][w] = c;
    }
    __syncthreads();
    if (t < 3) {
        uint32_t s = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) s += wt[t][k];
        sums[(size_t)t * nb1 + blockIdx.x] = s;
    }
}

// one block: exclusive scan of the three count arrays in place, totals -> sums[k * nb1 + nb] and totals[k]
__global__ void __launch_bounds__(1024) densify_scan_kernel(uint32_t* __restrict__ sums, int nb, uint32_t* __restrict__ totals) {
    __shared__ uint32_t warp_tot[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int ipt = (nb + 1023) / 1024;
    const int first = min(t * ipt, nb), last = min(first + ipt, nb);
    for (int k = 0; k < 3; k++) {
        uint32_t* a = sums + (size_t)k * (nb + 1);
        uint32_t v = 0;
        for (int i = first; i < last; i++) v += a[i];
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        __syncthreads();                     // warp_tot of the previous array has been consumed
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
            const uint32_t c = a[i];
            a[i] = run;
            run += c;
        }
        if (t == 1023) { a[nb] = warp_tot[31]; totals[k] = warp_tot[31]; }
    }
}

struct DensifyRows {
    const float* src[6];     // xyz, f_dc, f_rest, opacity, scaling, rotation of the OLD bucket
    float* dst[6];           // the same groups of the NEW bucket
    int width[6];
};

// one thread per original Gaussian: writes its clone row and / or its two child rows into the new bucket
__global__ void __launch_bounds__(256) densify_fill_kernel(int P, const uint8_t* __restrict__ flags, const uint32_t* __restrict__ sums, int nb1,
                                                           uint32_t n_kept, uint32_t n_clone, uint32_t n_split, const DensifyRows r) {
    __shared__ uint32_t wt[2][8];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int i = blockIdx.x * 256 + t;
    const uint32_t f = i < P ? flags[i] : 0u;
    const uint32_t bc = __ballot_sync(0xffffffffu, f & 2u), bs = __ballot_sync(0xffffffffu, f & 4u);
    if (lane == 0) { wt[0][w] = __popc(bc); wt[1][w] = __popc(bs); }
    __syncthreads();
    uint32_t ec = sums[(size_t)nb1 + blockIdx.x], es = sums[(size_t)2 * nb1 + blockIdx.x];
#pragma unroll
    for (int k = 0; k < 8; k++) { ec += (k < w) ? wt[0][k] : 0u; es += (k < w) ? wt[1][k] : 0u; }
    const uint32_t below = (1u << lane) - 1u;
    if (f & 2u) {                                                  // densify_and_clone: a verbatim copy of the raw parameters
        const size_t d = (size_t)n_kept + ec + __popc(bc & below);
#pragma unroll
        for (int a = 0; a < 6; a++)
            for (int c = 0; c < r.width[a]; c++) r.dst[a][d * r.width[a] + c] = __ldg(r.src[a] + (size_t)i * r.width[a] + c);
    }
    if (f & 4u) {                                                  // densify_and_split_along_maxscaling, N = 2, n_std = 2
        const size_t d1 = (size_t)n_kept + n_clone + es + __popc(bs & below), d2 = d1 + n_split;
        const float s[3] = {expf(__ldg(r.src[4] + 3 * (size_t)i)), expf(__ldg(r.src[4] + 3 * (size_t)i + 1)), expf(__ldg(r.src[4] + 3 * (size_t)i + 2))};
        const int axis = (s[0] >= s[1] && s[0] >= s[2]) ? 0 : (s[1] >= s[2] ? 1 : 2);
        const float smax = s[axis];
        // build_rotation (tools/general_utils.py:98-119) of the RAW quaternion: normalised inside; dirs = column `axis` of R
        float q0 = __ldg(r.src[5] + 4 * (size_t)i), q1 = __ldg(r.src[5] + 4 * (size_t)i + 1), q2 = __ldg(r.src[5] + 4 * (size_t)i + 2),
              q3 = __ldg(r.src[5] + 4 * (size_t)i + 3);
        const float nrm = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);
        const float qr = q0 / nrm, x = q1 / nrm, y = q2 / nrm, z = q3 / nrm;
        float dir[3];
        if (axis == 0) { dir[0] = 1 - 2 * (y * y + z * z); dir[1] = 2 * (x * y + qr * z); dir[2] = 2 * (x * z - qr * y); }
        else if (axis == 1) { dir[0] = 2 * (x * y - qr * z); dir[1] = 1 - 2 * (x * x + z * z); dir[2] = 2 * (y * z + qr * x); }
        else { dir[0] = 2 * (x * z + qr * y); dir[1] = 2 * (y * z - qr * x); dir[2] = 1 - 2 * (x * x + y * y); }
        const float radius = 2.f * smax / 3.f;                     // (n_std * max_scaling / 3.)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float m = __ldg(r.src[0] + 3 * (size_t)i + c);
            r.dst[0][d1 * 3 + c] = m + dir[c] * radius;
            r.dst[0][d2 * 3 + c] = m - dir[c] * radius;
            const float ns = logf(c == axis ? smax / 1.6f : s[c]);                      // scaling_inverse_activation
            r.dst[4][d1 * 3 + c] = ns; r.dst[4][d2 * 3 + c] = ns;
        }
#pragma unroll
        for (int a = 1; a < 6; a++) {
            if (a == 4) continue;
            for (int c = 0; c < r.width[a]; c++) {
                const float v = __ldg(r.src[a] + (size_t)i * r.width[a] + c);
                r.dst[a][d1 * r.width[a] + c] = v; r.dst[a][d2 * r.width[a] + c] = v;
            }
        }
    }
}

extern "C" size_t vcr_densify_scratch_bytes(int P) { return 3 * ((size_t)(P + 255) / 256 + 1) * sizeof(uint32_t) + 512; }

// Decide (flags u8[P]) and count.  totals_device u32[3] = {kept originals, clones, split parents whose children are kept}.
extern "C" int vcr_densify_decide(int P, const float* scaling_raw, const float* opacity_raw, const float* xyz_gradient_accum,
                                  const float* denom, float max_grad, float min_opacity, float dense_limit, float world_limit,
                                  uint8_t* flags, char* scratch, uint32_t* totals_device, void* stream) {
    if (P <= 0) return 0;
    if (!scaling_raw || !opacity_raw || !xyz_gradient_accum || !denom || !flags || !scratch || !totals_device) { vcr_set_error("densify_decide: NULL argument"); return 1; }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DensifyParams p;
    p.P = P; p.scaling_raw = scaling_raw; p.opacity_raw = opacity_raw; p.accum = xyz_gradient_accum; p.denom = denom;
    // dense_limit = percent_dense * scene_extent; world_limit = 0.1 * scene_extent when the caller passes a max_screen_size
    // (scene/gaussian_model.py:652-656), +inf otherwise — both formed by the caller exactly as the reference's Python does
    p.max_grad = max_grad; p.min_opacity = min_opacity; p.dense_limit = dense_limit; p.world_limit = world_limit;
    uint32_t* sums = reinterpret_cast<uint32_t*>(vcr_align256(reinterpret_cast<size_t>(scratch)));
    const int nb = (P + 255) / 256;
    VcrProfScope prof(VCR_K_ADAM, s, 2);
    densify_decide_kernel<<<nb, 256, 0, s>>>(p, flags, sums, nb + 1);
    VCR_LAUNCH_CHECK("densify_decide_kernel");
    densify_scan_kernel<<<1, 1024, 0, s>>>(sums, nb, totals_device);
    VCR_LAUNCH_CHECK("densify_scan_kernel");
    return 0;
}

// Write the clone / child rows behind the n_kept compacted rows of the new bucket.  `scratch` is the one vcr_densify_decide filled.
extern "C" int vcr_densify_fill(int P, const uint8_t* flags, const char* scratch, int n_kept, int n_clone, int n_split,
                                const float* const* src_host /* 6 */, float* const* dst_host /* 6 */, const int* width_host /* 6 */,
                                void* stream) {
    if (P <= 0 || (n_clone == 0 && n_split == 0)) return 0;
    if (!flags || !scratch || !src_host || !dst_host || !width_host || n_kept < 0 || n_clone < 0 || n_split < 0) { vcr_set_error("densify_fill: bad argument"); return 1; }
    DensifyRows r;
    memset(&r, 0, sizeof(r));
    static const int expect[6] = {3, 3, -1, 1, 3, 4};       // xyz, f_dc, f_rest (any), opacity, scaling, rotation
    for (int a = 0; a < 6; a++) {
        if (!src_host[a] || !dst_host[a] || width_host[a] <= 0 || (expect[a] > 0 && width_host[a] != expect[a])) {
            if (a == 2 && width_host[a] == 0) { r.src[a] = r.dst[a] = nullptr; r.width[a] = 0; continue; }   // SH degree 0: no f_rest
            vcr_set_error("densify_fill: group %d must be xyz[3], f_dc[3], f_rest[*], opacity[1], scaling[3], rotation[4]", a);
            return 1;
        }
        r.src[a] = src_host[a]; r.dst[a] = dst_host[a]; r.width[a] = width_host[a];
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const uint32_t* sums = reinterpret_cast<const uint32_t*>(vcr_align256(reinterpret_cast<size_t>(scratch)));
    const int nb = (P + 255) / 256;
    VcrProfScope prof(VCR_K_ADAM, s);
    densify_fill_kernel<<<nb, 256, 0, s>>>(P, flags, sums, nb + 1, (uint32_t)n_kept, (uint32_t)n_clone, (uint32_t)n_split, r);
    VCR_LAUNCH_CHECK("densify_fill_kernel");
    return 0;
}
