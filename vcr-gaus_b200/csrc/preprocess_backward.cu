// preprocess_backward.cu — per-Gaussian tail of the backward pass, one kernel.
//
// Replaces computeCov2DCUDA (reference DGR/cuda_rasterizer/backward.cu:144-274), preprocessCUDA<3>
// backward (:346-419) with computeColorFromSH backward (:20-139) and computeCov3D backward (:278-341),
// and the 14 torch::zeros allocations of RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:463-476):
// every output row is written exactly once here (zeros for Gaussians that were not rendered), so no
// output tensor needs a memset.
#include "common.cuh"

static __device__ __constant__ float kC0 = 0.28209479177387814f;
static __device__ __constant__ float kC1 = 0.4886025119029199f;
static __device__ __constant__ float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                               -1.0925484305920792f, 0.5462742152960396f};
static __device__ __constant__ float kC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                               0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                               -0.5900435899266435f};

// Per-Gaussian work.  With COOP (M == 16) the SH gradient is not written here: the thread leaves its 16 basis values
// and its 3 clamp-masked colour gradients in shared memory (both pre-zeroed by the caller) and the block writes the
// [256][16][3] gradient tile with fully coalesced stores afterwards.
// dRGB/dsh_k for the unit view direction (x, y, z): the SH basis of computeColorFromSH (forward.cu:22-74), shared by the
// three channels.  Used by the backward below and by the view-parallel reconstruction of dL_dsh at the end of this file.
__device__ __forceinline__ void vcr_sh_basis(const float x, const float y, const float z, float (&basis)[16]) {
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    basis[0] = kC0;
    basis[1] = -kC1 * y; basis[2] = kC1 * z; basis[3] = -kC1 * x;
    basis[4] = kC2[0] * xy; basis[5] = kC2[1] * yz; basis[6] = kC2[2] * (2.f * zz - xx - yy);
    basis[7] = kC2[3] * xz; basis[8] = kC2[4] * (xx - yy);
    basis[9] = kC3[0] * y * (3.f * xx - yy); basis[10] = kC3[1] * xy * z; basis[11] = kC3[2] * y * (4.f * zz - xx - yy);
    basis[12] = kC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); basis[13] = kC3[4] * x * (4.f * zz - xx - yy);
    basis[14] = kC3[5] * z * (xx - yy); basis[15] = kC3[6] * x * (xx - 3.f * yy);
}

template <bool COOP>
__device__ __forceinline__ void preprocess_backward_body(const BwdParams& p, const GeomState& g, const int idx, const float* V,
                                                         const float* Pm, float* s_basis, float* s_gg) {
    if (idx >= p.P) return;
    const size_t i = (size_t)idx;
    const int M = p.M;

    if (!(p.radii[idx] > 0)) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            p.dL_dmean2D[3 * i + k] = 0.f; p.dL_dmean2D_densify[3 * i + k] = 0.f;
            if (p.dL_dcolor) p.dL_dcolor[3 * i + k] = 0.f;
            p.dL_dnormal[3 * i + k] = 0.f; p.dL_dmean3D[3 * i + k] = 0.f; p.dL_dscale[3 * i + k] = 0.f;
        }
        p.dL_dopacity[i] = 0.f;
        if (p.dL_dobjects) { p.dL_dobjects[2 * i] = 0.f; p.dL_dobjects[2 * i + 1] = 0.f; }
#pragma unroll
        for (int k = 0; k < 6; k++) if (p.dL_dcov3D) p.dL_dcov3D[6 * i + k] = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) p.dL_drot[4 * i + k] = 0.f;
        if (p.dL_dcolor_masked) { p.dL_dcolor_masked[3 * i] = 0.f; p.dL_dcolor_masked[3 * i + 1] = 0.f; p.dL_dcolor_masked[3 * i + 2] = 0.f; }
        else if (!COOP && p.dL_dsh) for (int k = 0; k < 3 * M; k++) p.dL_dsh[i * 3 * M + k] = 0.f;
        return;
    }

    float a[VCR_ACC_STRIDE];
    {
        const float4* rec = reinterpret_cast<const float4*>(p.acc + i * VCR_ACC_STRIDE);
#pragma unroll
        for (int k = 0; k < VCR_ACC_STRIDE / 4; k++) {
            const float4 q = rec[k];
            a[4 * k] = q.x; a[4 * k + 1] = q.y; a[4 * k + 2] = q.z; a[4 * k + 3] = q.w;
        }
    }
    // outputs of the compositing backward: raw moments -> the reference's per-Gaussian gradients (common.cuh ACC_*)
    const float4 co = g.conic_opacity[i];
    const float half_w = 0.5f * p.W, half_h = 0.5f * p.H;
    const float g2x = -half_w * co.w * (co.x * a[ACC_MX] + co.y * a[ACC_MY]);
    const float g2y = -half_h * co.w * (co.z * a[ACC_MY] + co.y * a[ACC_MX]);
    p.dL_dmean2D[3 * i] = g2x; p.dL_dmean2D[3 * i + 1] = g2y; p.dL_dmean2D[3 * i + 2] = 0.f;
    p.dL_dmean2D_densify[3 * i] = half_w * fabsf(co.w) * a[ACC_ABS_X];
    p.dL_dmean2D_densify[3 * i + 1] = half_h * fabsf(co.w) * a[ACC_ABS_Y];
    p.dL_dmean2D_densify[3 * i + 2] = 0.f;
    p.dL_dopacity[i] = a[ACC_M0];
#pragma unroll
    for (int k = 0; k < 3; k++) { if (p.dL_dcolor) p.dL_dcolor[3 * i + k] = a[ACC_COLOR_R + k]; p.dL_dnormal[3 * i + k] = a[ACC_NORMAL_X + k]; }
    if (p.dL_dobjects) { p.dL_dobjects[2 * i] = a[ACC_OBJ_0]; p.dL_dobjects[2 * i + 1] = a[ACC_OBJ_1]; }
    const float gcx = -0.5f * co.w * a[ACC_MXX], gcy = -0.5f * co.w * a[ACC_MXY], gcz = -0.5f * co.w * a[ACC_MYY];

    const float3 mean = make_float3(__ldg(p.means3D + 3 * i), __ldg(p.means3D + 3 * i + 1), __ldg(p.means3D + 3 * i + 2));
    const float* c3 = p.cov3D_precomp ? p.cov3D_precomp + 6 * i : g.cov3D + 6 * i;
    const float h_x = p.focal_x, h_y = p.focal_y;

    // ---------------- computeCov2DCUDA ----------------
    float3 t = make_float3(V[0] * mean.x + V[4] * mean.y + V[8] * mean.z + V[12],
                           V[1] * mean.x + V[5] * mean.y + V[9] * mean.z + V[13],
                           V[2] * mean.x + V[6] * mean.y + V[10] * mean.z + V[14]);
    const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    const float J00 = h_x / t.z, J02 = -(h_x * t.x) / (t.z * t.z), J11 = h_y / t.z, J12 = -(h_y * t.y) / (t.z * t.z);
    // T[c][r] (GLM column c): T0[r] = T[0][r], T1[r] = T[1][r]; W[k][r] = V[4r+k]
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        T0[r] = V[4 * r] * J00 + V[4 * r + 2] * J02;
        T1[r] = V[4 * r + 1] * J11 + V[4 * r + 2] * J12;
    }
    const float Vrk[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    // s0[k] = sum_m T0[m] Vrk[k][m], s1[k] = sum_m T1[m] Vrk[k][m]
    float s0[3], s1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        s0[k] = T0[0] * Vrk[k][0] + T0[1] * Vrk[k][1] + T0[2] * Vrk[k][2];
        s1[k] = T1[0] * Vrk[k][0] + T1[1] * Vrk[k][1] + T1[2] * Vrk[k][2];
    }
    const float ca = s0[0] * T0[0] + s0[1] * T0[1] + s0[2] * T0[2] + 0.3f;
    const float cb = s1[0] * T0[0] + s1[1] * T0[1] + s1[2] * T0[2];
    const float cc = s1[0] * T1[0] + s1[1] * T1[1] + s1[2] * T1[2] + 0.3f;
    const float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float gcov[6];
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * gcx + 2 * cb * cc * gcy + (denom - ca * cc) * gcz);
        dL_dc = denom2inv * (-ca * ca * gcz + 2 * ca * cb * gcy + (denom - ca * cc) * gcx);
        dL_db = denom2inv * 2 * (cb * cc * gcx - (denom + 2 * cb * cb) * gcy + ca * cb * gcz);
        gcov[0] = (T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc);
        gcov[3] = (T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc);
        gcov[5] = (T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc);
        gcov[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
        gcov[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
        gcov[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
    } else {
#pragma unroll
        for (int k = 0; k < 6; k++) gcov[k] = 0.f;
    }
#pragma unroll
    for (int k = 0; k < 6; k++) if (p.dL_dcov3D) p.dL_dcov3D[6 * i + k] = gcov[k];
    float dL_dT0[3], dL_dT1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        dL_dT0[k] = 2 * s0[k] * dL_da + s1[k] * dL_db;
        dL_dT1[k] = 2 * s1[k] * dL_dc + s0[k] * dL_db;
    }
    // W[c][k] = V[4k + c]
    const float dL_dJ00 = V[0] * dL_dT0[0] + V[4] * dL_dT0[1] + V[8] * dL_dT0[2];
    const float dL_dJ02 = V[2] * dL_dT0[0] + V[6] * dL_dT0[1] + V[10] * dL_dT0[2];
    const float dL_dJ11 = V[1] * dL_dT1[0] + V[5] * dL_dT1[1] + V[9] * dL_dT1[2];
    const float dL_dJ12 = V[2] * dL_dT1[0] + V[6] * dL_dT1[1] + V[10] * dL_dT1[2];
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    float gm[3] = {V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz, V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz,
                   V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz};

    // ---------------- preprocessCUDA backward ----------------
    const float m_hw = Pm[3] * mean.x + Pm[7] * mean.y + Pm[11] * mean.z + Pm[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (Pm[0] * mean.x + Pm[4] * mean.y + Pm[8] * mean.z + Pm[12]) * m_w * m_w;
    const float mul2 = (Pm[1] * mean.x + Pm[5] * mean.y + Pm[9] * mean.z + Pm[13]) * m_w * m_w;
    gm[0] += (Pm[0] * m_w - Pm[3] * mul1) * g2x + (Pm[1] * m_w - Pm[3] * mul2) * g2y;
    gm[1] += (Pm[4] * m_w - Pm[7] * mul1) * g2x + (Pm[5] * m_w - Pm[7] * mul2) * g2y;
    gm[2] += (Pm[8] * m_w - Pm[11] * mul1) * g2x + (Pm[9] * m_w - Pm[11] * mul2) * g2y;
    if (p.dirs == nullptr) {
        const float gd = a[ACC_SV_OR_DEPTH];
        gm[0] += V[2] * gd; gm[1] += V[6] * gd; gm[2] += V[10] * gd;
    } else {
        const float sv = a[ACC_SV_OR_DEPTH];
        const float gvx = sv * __ldg(p.normals + 3 * i), gvy = sv * __ldg(p.normals + 3 * i + 1), gvz = sv * __ldg(p.normals + 3 * i + 2);
        gm[0] += V[0] * gvx + V[1] * gvy + V[2] * gvz;
        gm[1] += V[4] * gvx + V[5] * gvy + V[6] * gvz;
        gm[2] += V[8] * gvx + V[9] * gvy + V[10] * gvz;
    }

    if (p.shs) {
        const float* sh = p.shs + i * 3 * M;
        float* gsh = p.dL_dsh + i * 3 * M;
        const float3 campos = make_float3(__ldg(p.campos), __ldg(p.campos + 1), __ldg(p.campos + 2));
        const float dox = mean.x - campos.x, doy = mean.y - campos.y, doz = mean.z - campos.z;
        const float len = sqrtf(dox * dox + doy * doy + doz * doz);
        const float x = dox / len, y = doy / len, z = doz / len;
        const int D = p.D;
        float gg[3];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) gg[ch] = a[ACC_COLOR_R + ch] * (g.clamped[3 * i + ch] ? 0.f : 1.f);
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        float basis[16];
        vcr_sh_basis(x, y, z, basis);
        const int nact = (D + 1) * (D + 1);
        if (p.dL_dcolor_masked) {
            // compact mode: hand out the masked colour gradient; dL_dsh = basis (x) this is rebuilt after the exchange
#pragma unroll
            for (int ch = 0; ch < 3; ch++) p.dL_dcolor_masked[3 * i + ch] = gg[ch];
        } else if (COOP) {
#pragma unroll
            for (int k = 0; k < 16; k++) s_basis[threadIdx.x * 17 + k] = k < nact ? basis[k] : 0.f;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) s_gg[threadIdx.x * 3 + ch] = gg[ch];
        } else {
            for (int k = 0; k < M; k++) {
                const float bk = (k < nact && k < 16) ? basis[k] : 0.f;
                gsh[3 * k] = bk * gg[0]; gsh[3 * k + 1] = bk * gg[1]; gsh[3 * k + 2] = bk * gg[2];
            }
        }
        if (M == 16) {
            // Colour is linear in the coefficients, so d colour / d dir . gg = sum_k (d basis_k / d dir) t_k with
            // t_k = gg . sh[k]: one dot product per coefficient instead of the three per-channel polynomials, and the row is
            // read as twelve 16-byte loads (the per-channel form issued 45 scalar loads 192 bytes apart across the warp:
            // 32 lines per request, the L1 pipe was the limit of this kernel).
            const float4* sh4 = reinterpret_cast<const float4*>(sh);
            float t[16];
#pragma unroll
            for (int q = 0; q < 4; q++) {           // floats 12q .. 12q+11 = coefficients 4q .. 4q+3
                const float4 a0 = __ldg(sh4 + 3 * q), a1 = __ldg(sh4 + 3 * q + 1), a2 = __ldg(sh4 + 3 * q + 2);
                t[4 * q + 0] = a0.x * gg[0] + a0.y * gg[1] + a0.z * gg[2];
                t[4 * q + 1] = a0.w * gg[0] + a1.x * gg[1] + a1.y * gg[2];
                t[4 * q + 2] = a1.z * gg[0] + a1.w * gg[1] + a2.x * gg[2];
                t[4 * q + 3] = a2.y * gg[0] + a2.z * gg[1] + a2.w * gg[2];
            }
            if (D > 0) {
                ddx = -kC1 * t[3]; ddy = -kC1 * t[1]; ddz = kC1 * t[2];
                if (D > 1) {
                    ddx += kC2[0] * y * t[4] + kC2[2] * 2.f * -x * t[6] + kC2[3] * z * t[7] + kC2[4] * 2.f * x * t[8];
                    ddy += kC2[0] * x * t[4] + kC2[1] * z * t[5] + kC2[2] * 2.f * -y * t[6] + kC2[4] * 2.f * -y * t[8];
                    ddz += kC2[1] * y * t[5] + kC2[2] * 2.f * 2.f * z * t[6] + kC2[3] * x * t[7];
                    if (D > 2) {
                        ddx += kC3[0] * t[9] * 3.f * 2.f * xy + kC3[1] * t[10] * yz + kC3[2] * t[11] * -2.f * xy +
                               kC3[3] * t[12] * -3.f * 2.f * xz + kC3[4] * t[13] * (-3.f * xx + 4.f * zz - yy) +
                               kC3[5] * t[14] * 2.f * xz + kC3[6] * t[15] * 3.f * (xx - yy);
                        ddy += kC3[0] * t[9] * 3.f * (xx - yy) + kC3[1] * t[10] * xz + kC3[2] * t[11] * (-3.f * yy + 4.f * zz - xx) +
                               kC3[3] * t[12] * -3.f * 2.f * yz + kC3[4] * t[13] * -2.f * xy + kC3[5] * t[14] * -2.f * yz +
                               kC3[6] * t[15] * -3.f * 2.f * xy;
                        ddz += kC3[1] * t[10] * xy + kC3[2] * t[11] * 4.f * 2.f * yz + kC3[3] * t[12] * 3.f * (2.f * zz - xx - yy) +
                               kC3[4] * t[13] * 4.f * 2.f * xz + kC3[5] * t[14] * (xx - yy);
                    }
                }
            }
        } else {
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
#define S(k) __ldg(sh + (k) * 3 + ch)
                float dx_ = 0.f, dy_ = 0.f, dz_ = 0.f;
                if (D > 0) {
                    dx_ = -kC1 * S(3); dy_ = -kC1 * S(1); dz_ = kC1 * S(2);
                    if (D > 1) {
                        dx_ += kC2[0] * y * S(4) + kC2[2] * 2.f * -x * S(6) + kC2[3] * z * S(7) + kC2[4] * 2.f * x * S(8);
                        dy_ += kC2[0] * x * S(4) + kC2[1] * z * S(5) + kC2[2] * 2.f * -y * S(6) + kC2[4] * 2.f * -y * S(8);
                        dz_ += kC2[1] * y * S(5) + kC2[2] * 2.f * 2.f * z * S(6) + kC2[3] * x * S(7);
                        if (D > 2) {
                            dx_ += kC3[0] * S(9) * 3.f * 2.f * xy + kC3[1] * S(10) * yz + kC3[2] * S(11) * -2.f * xy +
                                   kC3[3] * S(12) * -3.f * 2.f * xz + kC3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                                   kC3[5] * S(14) * 2.f * xz + kC3[6] * S(15) * 3.f * (xx - yy);
                            dy_ += kC3[0] * S(9) * 3.f * (xx - yy) + kC3[1] * S(10) * xz + kC3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
                                   kC3[3] * S(12) * -3.f * 2.f * yz + kC3[4] * S(13) * -2.f * xy + kC3[5] * S(14) * -2.f * yz +
                                   kC3[6] * S(15) * -3.f * 2.f * xy;
                            dz_ += kC3[1] * S(10) * xy + kC3[2] * S(11) * 4.f * 2.f * yz + kC3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
                                   kC3[4] * S(13) * 4.f * 2.f * xz + kC3[5] * S(14) * (xx - yy);
                        }
                    }
                }
#undef S
                ddx += dx_ * gg[ch]; ddy += dy_ * gg[ch]; ddz += dz_ * gg[ch];
            }
        }
        // dnormvdv (auxiliary.h:102-113)
        const float sum2 = dox * dox + doy * doy + doz * doz;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        gm[0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
        gm[1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
        gm[2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
    } else if (p.dL_dcolor_masked) {
        p.dL_dcolor_masked[3 * i] = 0.f; p.dL_dcolor_masked[3 * i + 1] = 0.f; p.dL_dcolor_masked[3 * i + 2] = 0.f;
    } else if (!COOP && p.dL_dsh) {
        for (int k = 0; k < 3 * M; k++) p.dL_dsh[i * 3 * M + k] = 0.f;
    }
    p.dL_dmean3D[3 * i] = gm[0]; p.dL_dmean3D[3 * i + 1] = gm[1]; p.dL_dmean3D[3 * i + 2] = gm[2];

    if (p.scales) {  // computeCov3D backward
        const float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        // R[c][k]: GLM column c, row k
        const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                               {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                               {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        const float s[3] = {p.scale_modifier * __ldg(p.scales + 3 * i), p.scale_modifier * __ldg(p.scales + 3 * i + 1),
                            p.scale_modifier * __ldg(p.scales + 3 * i + 2)};
        float Mm[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int k = 0; k < 3; k++) Mm[c][k] = s[k] * R[c][k];
        const float dS[3][3] = {{gcov[0], 0.5f * gcov[1], 0.5f * gcov[2]},
                                {0.5f * gcov[1], gcov[3], 0.5f * gcov[4]},
                                {0.5f * gcov[2], 0.5f * gcov[4], gcov[5]}};
        // dL_dM = 2 * M * dL_dSigma ; dMt[k][j] = dL_dM[j][k]
        float dMt[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++)
                dMt[rr][c] = 2.0f * (Mm[0][rr] * dS[c][0] + Mm[1][rr] * dS[c][1] + Mm[2][rr] * dS[c][2]);
#pragma unroll
        for (int k = 0; k < 3; k++)
            p.dL_dscale[3 * i + k] = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int j = 0; j < 3; j++) dMt[k][j] *= s[k];
        float4 gq;
        gq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
        gq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
        gq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
        gq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
        reinterpret_cast<float4*>(p.dL_drot)[idx] = gq;  // no quaternion-normalisation Jacobian (backward.cu:340)
    } else {
#pragma unroll
        for (int k = 0; k < 3; k++) p.dL_dscale[3 * i + k] = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) p.dL_drot[4 * i + k] = 0.f;
    }
}

#ifndef VCR_PB_MINBLOCKS
#define VCR_PB_MINBLOCKS 4
#endif
template <bool COOP>
__global__ void __launch_bounds__(256, VCR_PB_MINBLOCKS) preprocess_backward_kernel(const BwdParams p, const GeomState g, const int first_block) {
    const int block = first_block + (int)blockIdx.x;        // a launch covers blocks [first_block, first_block + gridDim.x)
    const int idx = block * 256 + threadIdx.x;
    __shared__ float V[16], Pm[16];
    __shared__ float s_basis[COOP ? 256 * 17 : 1];  // stride 17: conflict-free per-thread rows
    __shared__ float s_gg[COOP ? 256 * 3 : 1];
    if (threadIdx.x < 16) V[threadIdx.x] = __ldg(p.viewmatrix + threadIdx.x);
    else if (threadIdx.x < 32) Pm[threadIdx.x - 16] = __ldg(p.projmatrix + threadIdx.x - 16);
    if (COOP) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++) s_gg[threadIdx.x * 3 + ch] = 0.f;  // zero colour gradient => zero SH gradient row
#pragma unroll
        for (int k = 0; k < 16; k++) s_basis[threadIdx.x * 17 + k] = 0.f;  // (uninitialised shared memory could hold NaNs)
    }
    __syncthreads();
    preprocess_backward_body<COOP>(p, g, idx, V, Pm, s_basis, s_gg);
    if (COOP) {
        __syncthreads();
        // dL_dsh tile of this block: 256 Gaussians x 48 floats = 3072 float4, written as coalesced 16-byte stores
        const size_t base = (size_t)block * 256 * 48;
        const size_t total = (size_t)p.P * 48;
        float4* out4 = reinterpret_cast<float4*>(p.dL_dsh + base);
#pragma unroll 4
        for (int r = 0; r < 12; r++) {
            const int e4 = r * 256 + threadIdx.x;   // float4 index inside the tile
            const int e = e4 * 4;
            if (base + e >= total) continue;
            const int gi = e / 48, k0 = e - gi * 48;
            float o[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int k = k0 + q;                // never crosses a Gaussian: 48 is a multiple of 4
                o[q] = s_basis[gi * 17 + k / 3] * s_gg[gi * 3 + k % 3];
            }
            out4[e4] = make_float4(o[0], o[1], o[2], o[3]);
        }
    }
}

int vcr_launch_preprocess_backward(const BwdParams& p, GeomState& g, int first_gaussian, int n_gaussians, cudaStream_t s) {
    // Gaussians [first_gaussian, first_gaussian + n_gaussians); first_gaussian must be a multiple of 256 (one block's worth)
    if (n_gaussians <= 0) return 0;
    VcrProfScope prof(VCR_K_PREPROCESS_BWD, s);
    const bool coop = p.shs != nullptr && p.M == 16 && p.dL_dsh != nullptr && p.dL_dcolor_masked == nullptr;
    const int first_block = first_gaussian / 256, blocks = (n_gaussians + 255) / 256;
    if (coop) preprocess_backward_kernel<true><<<blocks, 256, 0, s>>>(p, g, first_block);
    else preprocess_backward_kernel<false><<<blocks, 256, 0, s>>>(p, g, first_block);
    VCR_LAUNCH_CHECK("preprocess_backward_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// View-parallel reconstruction of the SH gradient.  On every rank the SH gradient of a view is the outer product of the
// clamp-masked colour gradient (3 floats) with the SH basis of the direction camera -> Gaussian, i.e. rank one per
// Gaussian and view.  Instead of all-reducing 48 floats per Gaussian the ranks therefore all-gather 3 (+ their camera
// position) and each rebuilds   dL_dsh[g] = sum_r basis(normalize(mean[g] - campos_r)) (x) dcolor_r[g]   locally, in rank
// order (so all replicas get the same bits).  `slices` holds one slice per rank: [3P colour gradients | campos xyz | pad].
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sh_grad_from_colors_kernel(const int P, const int D, const int M, const int world,
                                                                  const float* __restrict__ means3D, const float* __restrict__ slices,
                                                                  const long long slice_stride, float* __restrict__ dL_dsh) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    const bool live = idx < P;
    const size_t i = (size_t)(live ? idx : 0);
    const float3 mean = make_float3(__ldg(means3D + 3 * i), __ldg(means3D + 3 * i + 1), __ldg(means3D + 3 * i + 2));
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; k++) acc[k] = 0.f;
    const int nact = min((D + 1) * (D + 1), 16);
    for (int r = 0; r < (live ? world : 0); r++) {
        const float* sl = slices + (size_t)r * slice_stride;
        const float g0 = __ldg(sl + 3 * i), g1 = __ldg(sl + 3 * i + 1), g2 = __ldg(sl + 3 * i + 2);
        if (g0 == 0.f && g1 == 0.f && g2 == 0.f) continue;              // not rendered / fully clamped on that rank
        const float dox = mean.x - __ldg(sl + 3 * (size_t)P), doy = mean.y - __ldg(sl + 3 * (size_t)P + 1),
                    doz = mean.z - __ldg(sl + 3 * (size_t)P + 2);
        const float len = sqrtf(dox * dox + doy * doy + doz * doz);
        float basis[16];
        vcr_sh_basis(dox / len, doy / len, doz / len, basis);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const float bk = k < nact ? basis[k] : 0.f;
            acc[3 * k] += bk * g0; acc[3 * k + 1] += bk * g1; acc[3 * k + 2] += bk * g2;
        }
    }
    if (M == 16) {
        // coalesced store of the block's [256][48] tile through shared memory, in two halves of 24 floats per Gaussian
        // (row stride 25: conflict-free per-thread rows); 96-byte runs per Gaussian and half, i.e. whole sectors
        __shared__ float s_t[256 * 25];
        const size_t tile0 = (size_t)blockIdx.x * 256;
        const int n_g = min(256, P - (int)tile0);
#pragma unroll
        for (int h = 0; h < 2; h++) {
            __syncthreads();
#pragma unroll
            for (int e = 0; e < 24; e++) s_t[threadIdx.x * 25 + e] = acc[24 * h + e];
            __syncthreads();
            for (int q = threadIdx.x; q < n_g * 6; q += 256) {
                const int gi = q / 6, f4 = q - gi * 6;
                const float* src = s_t + gi * 25 + 4 * f4;
                reinterpret_cast<float4*>(dL_dsh + (tile0 + gi) * 48 + 24 * h)[f4] = make_float4(src[0], src[1], src[2], src[3]);
            }
        }
    } else if (live) {
        float* out = dL_dsh + i * 3 * (size_t)M;
#pragma unroll
        for (int e = 0; e < 48; e++) if (e < 3 * M) out[e] = acc[e];
        for (int e = 48; e < 3 * M; e++) out[e] = 0.f;
    }
}

int vcr_launch_sh_grad_from_colors(int P, int D, int M, int world, const float* means3D, const float* slices, long long slice_stride,
                                   float* dL_dsh, cudaStream_t s) {
    VcrProfScope prof(VCR_K_PREPROCESS_BWD, s);
    sh_grad_from_colors_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, D, M, world, means3D, slices, slice_stride, dL_dsh);
    VCR_LAUNCH_CHECK("sh_grad_from_colors_kernel");
    return 0;
}
