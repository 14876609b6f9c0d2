// preprocess.cu — per-Gaussian forward preprocessing (EWA projection, SH->RGB, conic, radius, tile
// rectangle) and the frustum test.
//
// Replaces preprocessCUDA<3> (reference DGR/cuda_rasterizer/forward.cu:157-260) with its helpers
// computeCov3D (:120-154), computeCov2D (:76-115), computeColorFromSH (:22-73), in_frustum
// (auxiliary.h:152-177), and checkFrustum (rasterizer_impl.cu:54-66).
//
// Bit-exactness: depth bits, the pixel centre and the radius decide keys / tile lists, which must
// match the reference exactly.  Those chains are written with explicit __fmul_rn/__fmaf_rn in the
// operation order nvcc emits for the reference source (checked against the SASS of the reference
// build, see DESIGN.md "Bit-exact binning"), so they cannot be re-contracted differently here.
//
// The kernel also produces the per-block partial sums of tiles_touched, i.e. the first phase of the
// prefix sum, so the scan never re-reads tiles_touched from HBM for the reduction.
#include "common.cuh"

static __device__ __constant__ float kSH_C0 = 0.28209479177387814f;
static __device__ __constant__ float kSH_C1 = 0.4886025119029199f;
static __device__ __constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                                  -1.0925484305920792f, 0.5462742152960396f};
static __device__ __constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                                  0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                                  -0.5900435899266435f};

// World covariance from scale and (un-normalised) quaternion; operation order of the reference build.
__device__ __forceinline__ void cov3d_from_scale_rot(const float sx, const float sy, const float sz, float mod,
                                                     const float4 q, float* c3) {
    const float s0 = __fmul_rn(mod, sx), s1 = __fmul_rn(mod, sy), s2 = __fmul_rn(mod, sz);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    // R[c][k]: GLM column c, row k
    const float R00 = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(yy, zz)));
    const float R01 = __fmul_rn(2.f, __fmaf_rn(x, y, -rz));
    const float R02 = __fmul_rn(2.f, __fmaf_rn(r, y, xz));
    const float R10 = __fmul_rn(2.f, __fmaf_rn(x, y, rz));
    const float R11 = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, zz)));
    const float R12 = __fmul_rn(2.f, __fmaf_rn(y, z, -rx));
    const float R20 = __fmul_rn(2.f, __fmaf_rn(-r, y, xz));
    const float R21 = __fmul_rn(2.f, __fmaf_rn(y, z, rx));
    const float R22 = __fsub_rn(1.f, __fmul_rn(2.f, __fmaf_rn(x, x, yy)));
    // M = S * R  ->  M[c][k] = s_k * R[c][k]
    const float M00 = __fmul_rn(s0, R00), M01 = __fmul_rn(s1, R01), M02 = __fmul_rn(s2, R02);
    const float M10 = __fmul_rn(s0, R10), M11 = __fmul_rn(s1, R11), M12 = __fmul_rn(s2, R12);
    const float M20 = __fmul_rn(s0, R20), M21 = __fmul_rn(s1, R21), M22 = __fmul_rn(s2, R22);
    // Sigma = transpose(M) * M  ->  Sigma[c][r] = sum_k M[r][k] * M[c][k]
    c3[0] = vcr_dot3(M00, M00, M01, M01, M02, M02);
    c3[1] = vcr_dot3(M10, M00, M11, M01, M12, M02);
    c3[2] = vcr_dot3(M20, M00, M21, M01, M22, M02);
    c3[3] = vcr_dot3(M10, M10, M11, M11, M12, M12);
    c3[4] = vcr_dot3(M20, M10, M21, M11, M22, M12);
    c3[5] = vcr_dot3(M20, M20, M21, M21, M22, M22);
}

// Upper 2x2 of cov = transpose(T) * transpose(Vrk) * T with T = W * J (before the +0.3 dilation).
__device__ __forceinline__ float3 ewa_cov2d(const float3 mean, float fx, float fy, float tanx, float tany,
                                            const float* c3, const float* __restrict__ V) {
    float tx = vcr_affine3(V[0], V[4], V[8], V[12], mean.x, mean.y, mean.z);
    float ty = vcr_affine3(V[1], V[5], V[9], V[13], mean.x, mean.y, mean.z);
    const float tz = vcr_affine3(V[2], V[6], V[10], V[14], mean.x, mean.y, mean.z);
    const float limx = __fmul_rn(1.3f, tanx), limy = __fmul_rn(1.3f, tany);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    tx = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), tz);
    ty = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), tz);
    const float tz2 = __fmul_rn(tz, tz);
    const float J00 = __fdiv_rn(fx, tz), J02 = __fdiv_rn(-__fmul_rn(fx, tx), tz2);
    const float J11 = __fdiv_rn(fy, tz), J12 = __fdiv_rn(-__fmul_rn(fy, ty), tz2);
    // T[c][r] = sum_k W[k][r] * J[c][k], W[k][r] = V[4r + k]; zero terms dropped (they add +-0)
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        T0[r] = __fmaf_rn(V[4 * r + 2], J02, __fmul_rn(V[4 * r], J00));
        T1[r] = __fmaf_rn(V[4 * r + 2], J12, __fmul_rn(V[4 * r + 1], J11));
    }
    const float Vrk[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    // A[c][r] = sum_k T[r][k] * Vrk[c][k]   (r in {0,1}; T[2][*] == 0)
    float A0[3], A1[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        A0[c] = vcr_dot3(T0[0], Vrk[c][0], T0[1], Vrk[c][1], T0[2], Vrk[c][2]);
        A1[c] = vcr_dot3(T1[0], Vrk[c][0], T1[1], Vrk[c][1], T1[2], Vrk[c][2]);
    }
    // cov[c][r] = sum_k A[k][r] * T[c][k]
    float3 cov;
    cov.x = vcr_dot3(A0[0], T0[0], A0[1], T0[1], A0[2], T0[2]);  // cov[0][0]
    cov.y = vcr_dot3(A1[0], T0[0], A1[1], T0[1], A1[2], T0[2]);  // cov[0][1]
    cov.z = vcr_dot3(A1[0], T1[0], A1[1], T1[1], A1[2], T1[2]);  // cov[1][1]
    return cov;
}

__device__ __forceinline__ float3 sh_to_rgb(int deg, const float3 pos, const float3 campos, const float* __restrict__ sh,
                                            uint8_t* clamped) {
    float dx = pos.x - campos.x, dy = pos.y - campos.y, dz = pos.z - campos.z;
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    // 16 coefficients x 3 channels, 192 contiguous bytes per Gaussian (16-byte aligned): float4 loads
    float c[48];
    const int ncoef = (deg + 1) * (deg + 1);
    const float4* sh4 = reinterpret_cast<const float4*>(sh);
#pragma unroll
    for (int k = 0; k < 12; k++) {
        if (k * 4 < ncoef * 3) {
            const float4 v = __ldg(sh4 + k);
            c[4 * k] = v.x; c[4 * k + 1] = v.y; c[4 * k + 2] = v.z; c[4 * k + 3] = v.w;
        } else { c[4 * k] = c[4 * k + 1] = c[4 * k + 2] = c[4 * k + 3] = 0.f; }
    }
    float res[3];
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
#define S(i) c[(i) * 3 + ch]
        float r = kSH_C0 * S(0);
        if (deg > 0) {
            r = r - kSH_C1 * y * S(1) + kSH_C1 * z * S(2) - kSH_C1 * x * S(3);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + kSH_C2[0] * xy * S(4) + kSH_C2[1] * yz * S(5) + kSH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
                    kSH_C2[3] * xz * S(7) + kSH_C2[4] * (xx - yy) * S(8);
                if (deg > 2) {
                    r = r + kSH_C3[0] * y * (3.0f * xx - yy) * S(9) + kSH_C3[1] * xy * z * S(10) +
                        kSH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                        kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                        kSH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + kSH_C3[5] * z * (xx - yy) * S(14) +
                        kSH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        r += 0.5f;
        clamped[ch] = r < 0;
        res[ch] = fmaxf(r, 0.0f);
    }
    return make_float3(res[0], res[1], res[2]);
}

// General-M fallback for SH layouts that are not [P,16,3] (M != 16): scalar loads.
__device__ float3 sh_to_rgb_generic(int deg, int M, const float3 pos, const float3 campos, const float* __restrict__ sh,
                                    uint8_t* clamped) {
    float dx = pos.x - campos.x, dy = pos.y - campos.y, dz = pos.z - campos.z;
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    float res[3];
    for (int ch = 0; ch < 3; ch++) {
#define S(i) __ldg(sh + (i) * 3 + ch)
        float r = kSH_C0 * S(0);
        if (deg > 0 && M >= 4) {
            r = r - kSH_C1 * y * S(1) + kSH_C1 * z * S(2) - kSH_C1 * x * S(3);
            if (deg > 1 && M >= 9) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + kSH_C2[0] * xy * S(4) + kSH_C2[1] * yz * S(5) + kSH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
                    kSH_C2[3] * xz * S(7) + kSH_C2[4] * (xx - yy) * S(8);
                if (deg > 2 && M >= 16) {
                    r = r + kSH_C3[0] * y * (3.0f * xx - yy) * S(9) + kSH_C3[1] * xy * z * S(10) +
                        kSH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                        kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                        kSH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + kSH_C3[5] * z * (xx - yy) * S(14) +
                        kSH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        r += 0.5f;
        clamped[ch] = r < 0;
        res[ch] = fmaxf(r, 0.0f);
    }
    return make_float3(res[0], res[1], res[2]);
}

#ifndef VCR_PRE_MINBLOCKS
#define VCR_PRE_MINBLOCKS 3
#endif
__global__ void __launch_bounds__(256, VCR_PRE_MINBLOCKS) preprocess_kernel(const FwdParams p, const GeomState g) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    __shared__ float sV[16], sP[16];
    __shared__ uint32_t warp_tiles[8];
    __shared__ uint32_t s_dhist[4][256];          // digit histograms of the depth keys (Gaussian-level sort, binning.cu)
    if (threadIdx.x < 16) sV[threadIdx.x] = __ldg(p.viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) sP[threadIdx.x - 16] = __ldg(p.projmatrix + threadIdx.x - 16);
#pragma unroll
    for (int k = 0; k < 4; k++) s_dhist[k][threadIdx.x] = 0u;
    __syncthreads();

    uint32_t tiles = 0;
    int radius = 0;
    uint32_t dkey = 0xFFFFFFFFu;                  // Gaussians that are not rendered sort behind every depth
    if (idx < p.P) {
        const float3 m = make_float3(__ldg(p.means3D + 3 * idx), __ldg(p.means3D + 3 * idx + 1), __ldg(p.means3D + 3 * idx + 2));
        const float vx = vcr_affine3(sV[0], sV[4], sV[8], sV[12], m.x, m.y, m.z);
        const float vy = vcr_affine3(sV[1], sV[5], sV[9], sV[13], m.x, m.y, m.z);
        const float vz = vcr_affine3(sV[2], sV[6], sV[10], sV[14], m.x, m.y, m.z);
        if (vz > 0.2f) {
            const float hx = vcr_affine3(sP[0], sP[4], sP[8], sP[12], m.x, m.y, m.z);
            const float hy = vcr_affine3(sP[1], sP[5], sP[9], sP[13], m.x, m.y, m.z);
            const float hw = vcr_affine3(sP[3], sP[7], sP[11], sP[15], m.x, m.y, m.z);
            const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
            const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);

            float c3[6];
            if (p.cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; k++) c3[k] = __ldg(p.cov3D_precomp + 6 * (size_t)idx + k);
            } else {
                const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
                cov3d_from_scale_rot(__ldg(p.scales + 3 * idx), __ldg(p.scales + 3 * idx + 1), __ldg(p.scales + 3 * idx + 2),
                                     p.scale_modifier, q, c3);
#pragma unroll
                for (int k = 0; k < 6; k++) g.cov3D[6 * (size_t)idx + k] = c3[k];
            }
            float3 cov = ewa_cov2d(m, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, c3, sV);
            cov.x = __fadd_rn(cov.x, 0.3f);
            cov.z = __fadd_rn(cov.z, 0.3f);
            const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
            if (det != 0.0f) {
                const float det_inv = __frcp_rn(det);
                const float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
                const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
                const float lambda1 = __fadd_rn(mid, sq), lambda2 = __fsub_rn(mid, sq);
                const float my_radius = ceilf(__fmul_rn(3.f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
                // ndc2Pix is evaluated in double by the reference (auxiliary.h:54-57)
                const float2 pimg = make_float2((float)((((double)projx + 1.0) * p.W - 1.0) * 0.5),
                                                (float)((((double)projy + 1.0) * p.H - 1.0) * 0.5));
                int2 rmin, rmax;
                vcr_get_rect(pimg, (int)my_radius, p.gx, p.gy, rmin, rmax);
                const int area = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                if (area != 0) {
                    if (p.colors_precomp == nullptr) {
                        const float3 campos = make_float3(__ldg(p.cam_pos), __ldg(p.cam_pos + 1), __ldg(p.cam_pos + 2));
                        uint8_t cl[3];
                        float3 rgb;
                        if (p.M == 16) rgb = sh_to_rgb(p.D, m, campos, p.shs + (size_t)idx * 48, cl);
                        else rgb = sh_to_rgb_generic(p.D, p.M, m, campos, p.shs + (size_t)idx * p.M * 3, cl);
                        g.rgb[3 * (size_t)idx] = rgb.x; g.rgb[3 * (size_t)idx + 1] = rgb.y; g.rgb[3 * (size_t)idx + 2] = rgb.z;
                        g.clamped[3 * (size_t)idx] = cl[0]; g.clamped[3 * (size_t)idx + 1] = cl[1]; g.clamped[3 * (size_t)idx + 2] = cl[2];
                    }
                    g.depths[idx] = vz;
                    g.xyz_view[3 * (size_t)idx] = vx; g.xyz_view[3 * (size_t)idx + 1] = vy; g.xyz_view[3 * (size_t)idx + 2] = vz;
                    g.means2D[idx] = pimg;
                    g.conic_opacity[idx] = make_float4(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv),
                                                       __fmul_rn(cov.x, det_inv), __ldg(p.opacities + idx));
                    radius = (int)my_radius;
                    tiles = (uint32_t)area;
                    if (tiles) dkey = __float_as_uint(vz);
                }
            }
        } else if (p.prefiltered) {
            // reference: printf + __trap() (auxiliary.h:169-173).  Here: raise a flag the host turns into an error.
            g.num_rendered[1] = 1u;
        }
        p.radii[idx] = radius;
        g.tiles_touched[idx] = tiles;
    }
    // input of the Gaussian-level depth sort: (depth bits, index) + the digit histograms of its four passes
    if (idx < p.P) {
        g.gs_keys[0][idx] = dkey;
        g.gs_vals[0][idx] = (uint32_t)idx;
#pragma unroll
        for (int k = 0; k < 4; k++) atomicAdd(&s_dhist[k][(dkey >> (8 * k)) & 255u], 1u);
    }
    // phase 1 of the prefix sum: per-block total
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, tiles);
    if ((threadIdx.x & 31) == 0) warp_tiles[threadIdx.x >> 5] = wsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) t += warp_tiles[w];
        g.block_sums[blockIdx.x] = t;
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t c = s_dhist[k][threadIdx.x];
        if (c) atomicAdd(&g.gs_hist[k * 256 + threadIdx.x], c);
    }
}

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ V, uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= P) return;
    const float z = vcr_affine3(__ldg(V + 2), __ldg(V + 6), __ldg(V + 10), __ldg(V + 14), __ldg(means3D + 3 * idx),
                                __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2));
    present[idx] = z > 0.2f;
}

int vcr_launch_preprocess(const FwdParams& p, GeomState& g, cudaStream_t s) {
    const int blocks = (p.P + 255) / 256;
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(g.num_rendered, 0, 4 * sizeof(uint32_t), s)); }
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(g.gs_hist, 0, 4 * 256 * sizeof(uint32_t), s)); }
    VcrProfScope prof(VCR_K_PREPROCESS, s);
    preprocess_kernel<<<blocks, 256, 0, s>>>(p, g);
    VCR_LAUNCH_CHECK("preprocess_kernel");
    return 0;
}

int vcr_launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s) {
    VcrProfScope prof(VCR_K_MARK_VISIBLE, s);
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
    VCR_LAUNCH_CHECK("mark_visible_kernel");
    return 0;
}
