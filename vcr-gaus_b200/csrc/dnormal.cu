// dnormal.cu — the D-Normal regularizer as one fused kernel per direction.
//
// Replaces the PyTorch op chain of the reference (~15 launches + torch.inverse with a host sync, and
// >= 10 full-image temporaries per view):
//   depth -> camera points      tools/graphics_utils.py:111-131 (depth2point_cam, ndc_2_cam)
//   finite differences, cross, normalise   tools/normal_utils.py:24-41 (compute_gradient, compute_normals)
//   F.normalize of the rendered normal     gaussian_renderer/__init__.py:133-134
//   mask = cam.mask & depth < extent*thr   gaussian_renderer/__init__.py:125-131
//   cos_weight, monosdf_normal_loss        tools/loss_utils.py:122-142
//   masked means + assembly                trainer.py:266-280
//
// forward : est_normal (optional), per-pixel confidence weight (optional), and the three sums
//           { sum_M w*|n-g|_1, sum_M w*(1-n.g), |M| };  loss = (s0 + s1) / s2   (0 when M is empty).
// backward: dL/ddepth by a 5-point gather — each pixel re-evaluates the loss gradient of the (at most
//           five) pixels whose normal depends on its depth; no saved intermediates, no atomics.
//
// K = [[fx,0,cx],[0,fy,cy],[0,0,1]] so K^-1 [u+.5, v+.5, 1]^T z = ((u+.5-cx)/fx, (v+.5-cy)/fy, 1) z.
#include "common.cuh"

struct DnParams {
    const float* depth; const float* gt; const float* rn; const uint8_t* mask;
    int W, H;
    float inv_fx, inv_fy, cx, cy, exp_t, mask_depth_max;
    size_t rn_ps, rn_cs;   // element strides of the rendered normal: per pixel / per channel (HWC: 3,1 ; CHW planes: 1,H*W)
};

// torch.gradient (spacing 1, edge_order 1): central differences inside, one-sided at the borders.
// X(u,v) = (a(u) z, b(v) z, z) with a(u) = (u + .5 - cx)/fx, so along u
//   X(u1) - X(u0) = (a(u0) dz + (u1-u0)/fx z1,  b dz,  dz),  dz = z1 - z0,
// which avoids the cancellation of subtracting two O(1) coordinates that differ by O(1/fx) (the reference chain
// differences the float32 coordinates themselves; both agree with the exact value, this form just carries less noise).
__device__ __forceinline__ float3 dn_diff_u(const DnParams& p, int u, int v) {
    if (p.W == 1) return make_float3(0.f, 0.f, 0.f);
    const int u0 = max(u - 1, 0), u1 = min(u + 1, p.W - 1);
    const float s = (u1 - u0 == 2) ? 0.5f : 1.0f;
    const float z0 = __ldg(p.depth + (size_t)v * p.W + u0), z1 = __ldg(p.depth + (size_t)v * p.W + u1);
    const float dz = z1 - z0;
    const float a0 = ((u0 + 0.5f) - p.cx) * p.inv_fx, b = ((v + 0.5f) - p.cy) * p.inv_fy;
    return make_float3((a0 * dz + (float)(u1 - u0) * p.inv_fx * z1) * s, b * dz * s, dz * s);
}
__device__ __forceinline__ float3 dn_diff_v(const DnParams& p, int u, int v) {
    if (p.H == 1) return make_float3(0.f, 0.f, 0.f);
    const int v0 = max(v - 1, 0), v1 = min(v + 1, p.H - 1);
    const float s = (v1 - v0 == 2) ? 0.5f : 1.0f;
    const float z0 = __ldg(p.depth + (size_t)v0 * p.W + u), z1 = __ldg(p.depth + (size_t)v1 * p.W + u);
    const float dz = z1 - z0;
    const float a = ((u + 0.5f) - p.cx) * p.inv_fx, b0 = ((v0 + 0.5f) - p.cy) * p.inv_fy;
    return make_float3(a * dz * s, (b0 * dz + (float)(v1 - v0) * p.inv_fy * z1) * s, dz * s);
}
__device__ __forceinline__ float3 cross3(const float3 a, const float3 b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ bool dn_masked_in(const DnParams& p, size_t pix) {
    bool m = p.mask ? (__ldg(p.mask + pix) != 0) : true;
    if (p.mask_depth_max > 0.f && !(__ldg(p.depth + pix) < p.mask_depth_max)) m = false;
    return m;
}
__device__ __forceinline__ float dn_weight(const DnParams& p, size_t pix, const float3 g) {
    if (!(p.exp_t > 0.f)) return 1.0f;
    const float rx = __ldg(p.rn + p.rn_ps * pix), ry = __ldg(p.rn + p.rn_ps * pix + p.rn_cs), rz = __ldg(p.rn + p.rn_ps * pix + 2 * p.rn_cs);
    const float rl = fmaxf(sqrtf(rx * rx + ry * ry + rz * rz), 1e-12f);  // F.normalize eps
    const float cs = (rx * g.x + ry * g.y + rz * g.z) / rl;
    return expf((cs - 1.0f) / p.exp_t);
}

__global__ void __launch_bounds__(256) dnormal_forward_kernel(const DnParams p, float* __restrict__ est_normal,
                                                              float* __restrict__ weight_out, float* __restrict__ sums) {
    const int u = blockIdx.x * 32 + (threadIdx.x & 31);
    const int v = blockIdx.y * 8 + (threadIdx.x >> 5);
    float l1 = 0.f, lc = 0.f, cnt = 0.f;
    if (u < p.W && v < p.H) {
        const size_t pix = (size_t)v * p.W + u;
        const float3 c = cross3(dn_diff_u(p, u, v), dn_diff_v(p, u, v));
        const float len = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
        const float inv = 1.0f / fmaxf(len, 1e-12f);
        const float3 n = make_float3(c.x * inv, c.y * inv, c.z * inv);
        if (est_normal) { est_normal[3 * pix] = n.x; est_normal[3 * pix + 1] = n.y; est_normal[3 * pix + 2] = n.z; }
        float w = 0.f;
        if (dn_masked_in(p, pix)) {
            const float3 g = make_float3(__ldg(p.gt + 3 * pix), __ldg(p.gt + 3 * pix + 1), __ldg(p.gt + 3 * pix + 2));
            w = dn_weight(p, pix, g);
            l1 = w * (fabsf(n.x - g.x) + fabsf(n.y - g.y) + fabsf(n.z - g.z));
            lc = w * (1.0f - (n.x * g.x + n.y * g.y + n.z * g.z));
            cnt = 1.f;
        }
        if (weight_out) weight_out[pix] = w;
    }
    __shared__ float red[3][8];
    l1 = vcr_warp_sum(l1); lc = vcr_warp_sum(lc); cnt = vcr_warp_sum(cnt);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = l1; red[1][threadIdx.x >> 5] = lc; red[2][threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x < 3) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) t += red[threadIdx.x][k];
        if (t != 0.f) atomicAdd(sums + threadIdx.x, t);
    }
}

// dL/dc at pixel (u,v): loss gradient w.r.t. the un-normalised cross product; also returns dx, dy there.
__device__ __forceinline__ float3 dn_grad_c(const DnParams& p, int u, int v, float scale, const float* __restrict__ up,
                                            float3& dxo, float3& dyo) {
    const size_t pix = (size_t)v * p.W + u;
    const float3 dx = dn_diff_u(p, u, v), dy = dn_diff_v(p, u, v);
    dxo = dx; dyo = dy;
    const float3 c = cross3(dx, dy);
    const float len = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
    const float den = fmaxf(len, 1e-12f);
    const float inv = 1.0f / den;
    const float3 n = make_float3(c.x * inv, c.y * inv, c.z * inv);
    float3 dn = make_float3(0.f, 0.f, 0.f);
    if (dn_masked_in(p, pix)) {
        const float3 g = make_float3(__ldg(p.gt + 3 * pix), __ldg(p.gt + 3 * pix + 1), __ldg(p.gt + 3 * pix + 2));
        const float w = dn_weight(p, pix, g) * scale;
        const float ex = n.x - g.x, ey = n.y - g.y, ez = n.z - g.z;
        dn.x = w * ((float)((ex > 0.f) - (ex < 0.f)) - g.x);
        dn.y = w * ((float)((ey > 0.f) - (ey < 0.f)) - g.y);
        dn.z = w * ((float)((ez > 0.f) - (ez < 0.f)) - g.z);
    }
    if (up) { dn.x += __ldg(up + 3 * pix); dn.y += __ldg(up + 3 * pix + 1); dn.z += __ldg(up + 3 * pix + 2); }
    if (len > 1e-12f) {
        const float d = dn.x * n.x + dn.y * n.y + dn.z * n.z;
        return make_float3((dn.x - n.x * d) * inv, (dn.y - n.y * d) * inv, (dn.z - n.z * d) * inv);
    }
    return make_float3(dn.x * inv, dn.y * inv, dn.z * inv);
}

// Tile of VL_TW x VL_TH pixels per 256-thread block (two rows per thread).  The backward is a 5-point gather: the depth
// gradient of a pixel collects cross(dy, gc) from its left/right neighbours and cross(gc, dx) from its upper/lower ones,
// where gc = dL/d(cross product) of THAT neighbour.  gc costs four depth loads, the ground-truth and rendered normals, a
// square root, a division and an exp, so it is evaluated once per pixel of the tile + a one-pixel halo into shared memory
// (1.2 evaluations per output) instead of once per (pixel, neighbour) pair (4-5 per output).
#define VL_TW 32
#define VL_TH 16
__global__ void __launch_bounds__(256) dnormal_backward_kernel(const DnParams p, const float* __restrict__ sums,
                                                               const float* __restrict__ dL_dloss,
                                                               const float* __restrict__ dL_dest, float* __restrict__ dL_ddepth) {
    constexpr int HWD = VL_TW + 2, HHT = VL_TH + 2;
    __shared__ float s_tx[3][HHT][HWD];      // dL/ddx(q) = dy(q) x gc(q) at the halo tile
    __shared__ float s_ty[3][HHT][HWD];      // dL/ddy(q) = gc(q) x dx(q)
    const float cnt = __ldg(sums + 2);
    const float scale = cnt > 0.f ? __ldg(dL_dloss) / cnt : 0.f;
    const int u_base = blockIdx.x * VL_TW, v_base = blockIdx.y * VL_TH;
    for (int idx = threadIdx.x; idx < HWD * HHT; idx += 256) {
        const int ly = idx / HWD, lx = idx - ly * HWD;
        const int qu = u_base + lx - 1, qv = v_base + ly - 1;
        float3 tx = make_float3(0.f, 0.f, 0.f), ty = tx;
        if (qu >= 0 && qu < p.W && qv >= 0 && qv < p.H) {
            float3 dx, dy;
            const float3 gc = dn_grad_c(p, qu, qv, scale, dL_dest, dx, dy);
            tx = cross3(dy, gc);
            ty = cross3(gc, dx);
        }
        s_tx[0][ly][lx] = tx.x; s_tx[1][ly][lx] = tx.y; s_tx[2][ly][lx] = tx.z;
        s_ty[0][ly][lx] = ty.x; s_ty[1][ly][lx] = ty.y; s_ty[2][ly][lx] = ty.z;
    }
    __syncthreads();
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const int lx = threadIdx.x & 31, ly = (threadIdx.x >> 5) + 8 * half;
        const int u = u_base + lx, v = v_base + ly;
        if (u >= p.W || v >= p.H) continue;
        float3 gX = make_float3(0.f, 0.f, 0.f);
        // contributions through dx(q): q = (u-1,v) [+s], q = (u+1,v) [-s], q = (u,v) at a border [-1 / +1]
        auto via_dx = [&](int qu, float coef) {
            const int hx = qu - u_base + 1, hy = ly + 1;
            gX.x += coef * s_tx[0][hy][hx]; gX.y += coef * s_tx[1][hy][hx]; gX.z += coef * s_tx[2][hy][hx];
        };
        auto via_dy = [&](int qv, float coef) {
            const int hx = lx + 1, hy = qv - v_base + 1;
            gX.x += coef * s_ty[0][hy][hx]; gX.y += coef * s_ty[1][hy][hx]; gX.z += coef * s_ty[2][hy][hx];
        };
        if (p.W > 1) {
            if (u > 0) { const int q = u - 1; via_dx(q, (q > 0) ? 0.5f : 1.0f); }               // X(p) is the upper sample of q
            if (u < p.W - 1) { const int q = u + 1; via_dx(q, (q < p.W - 1) ? -0.5f : -1.0f); } // X(p) is the lower sample of q
            if (u == 0) via_dx(0, -1.0f);
            if (u == p.W - 1) via_dx(u, 1.0f);
        }
        if (p.H > 1) {
            if (v > 0) { const int q = v - 1; via_dy(q, (q > 0) ? 0.5f : 1.0f); }
            if (v < p.H - 1) { const int q = v + 1; via_dy(q, (q < p.H - 1) ? -0.5f : -1.0f); }
            if (v == 0) via_dy(0, -1.0f);
            if (v == p.H - 1) via_dy(v, 1.0f);
        }
        const float ax = ((u + 0.5f) - p.cx) * p.inv_fx, ay = ((v + 0.5f) - p.cy) * p.inv_fy;
        dL_ddepth[(size_t)v * p.W + u] = gX.x * ax + gX.y * ay + gX.z;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Fused per-view image loss: everything trainer.py computes from the rasterizer's output map in image space, except
// SSIM, in one kernel per direction:
//   l1          = mean |render - gt_image|                                   tools/loss_utils.py l1_loss, trainer.py:236
//   mono_normal = monosdf_normal_loss(normalize(render_normal), gt_normal)    trainer.py:261-264, loss_utils.py:122-132
//   depth_normal (D-Normal, as above)                                         trainer.py:266-280
// forward : sums[6] = { sum |dRGB|, sum |rn-g|_1, sum (1 - rn.g), sum_M w|n-g|_1, sum_M w(1-n.g), |M| }
//           loss = w_l1 s0/(3N) + w_mono (s1+s2)/N + w_dn (s3+s4)/s5
// backward: dL/d outmap [10,H,W] written in full (zeros in the alpha / semantic planes).
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) view_loss_forward_kernel(const DnParams p, const float* __restrict__ out,
                                                                const float* __restrict__ gt_image, float* __restrict__ sums) {
    const int u = blockIdx.x * 32 + (threadIdx.x & 31);
    const int v = blockIdx.y * 8 + (threadIdx.x >> 5);
    const size_t HW = (size_t)p.W * p.H;
    float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (u < p.W && v < p.H) {
        const size_t pix = (size_t)v * p.W + u;
#pragma unroll
        for (int c = 0; c < 3; c++) acc[0] += fabsf(__ldg(out + c * HW + pix) - __ldg(gt_image + c * HW + pix));
        const float3 g = make_float3(__ldg(p.gt + 3 * pix), __ldg(p.gt + 3 * pix + 1), __ldg(p.gt + 3 * pix + 2));
        const float rx = __ldg(p.rn + pix), ry = __ldg(p.rn + HW + pix), rz = __ldg(p.rn + 2 * HW + pix);
        const float rinv = 1.0f / fmaxf(sqrtf(rx * rx + ry * ry + rz * rz), 1e-12f);
        const float3 rn = make_float3(rx * rinv, ry * rinv, rz * rinv);
        const float rcos = rn.x * g.x + rn.y * g.y + rn.z * g.z;
        acc[1] = fabsf(rn.x - g.x) + fabsf(rn.y - g.y) + fabsf(rn.z - g.z);
        acc[2] = 1.0f - rcos;
        if (dn_masked_in(p, pix)) {
            const float3 c = cross3(dn_diff_u(p, u, v), dn_diff_v(p, u, v));
            const float inv = 1.0f / fmaxf(sqrtf(c.x * c.x + c.y * c.y + c.z * c.z), 1e-12f);
            const float3 n = make_float3(c.x * inv, c.y * inv, c.z * inv);
            const float w = p.exp_t > 0.f ? expf((rcos - 1.0f) / p.exp_t) : 1.0f;
            acc[3] = w * (fabsf(n.x - g.x) + fabsf(n.y - g.y) + fabsf(n.z - g.z));
            acc[4] = w * (1.0f - (n.x * g.x + n.y * g.y + n.z * g.z));
            acc[5] = 1.f;
        }
    }
    __shared__ float red[6][8];
#pragma unroll
    for (int k = 0; k < 6; k++) {
        const float t = vcr_warp_sum(acc[k]);
        if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = t;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; k++) t += red[threadIdx.x][k];
        if (t != 0.f) atomicAdd(sums + threadIdx.x, t);
    }
}

// Same halo-tile scheme as dnormal_backward_kernel (VL_TW x VL_TH pixels per block) for the D-Normal part.
__global__ void __launch_bounds__(256) view_loss_backward_kernel(const DnParams p, const float* __restrict__ out,
                                                                 const float* __restrict__ gt_image, const float* __restrict__ sums,
                                                                 const float* __restrict__ dL_dloss, float w_l1, float w_mono,
                                                                 float w_dn, float* __restrict__ dL_dout) {
    constexpr int HWD = VL_TW + 2, HHT = VL_TH + 2;
    __shared__ float s_tx[3][HHT][HWD];      // cross(dy, gc) at the halo tile
    __shared__ float s_ty[3][HHT][HWD];      // cross(gc, dx)
    const size_t HW = (size_t)p.W * p.H;
    const float gl = __ldg(dL_dloss);
    const float cnt = __ldg(sums + 5);
    const float a_l1 = gl * w_l1 / (3.0f * (float)HW), a_mono = gl * w_mono / (float)HW;
    const float a_dn = cnt > 0.f ? gl * w_dn / cnt : 0.f;
    const int u_base = blockIdx.x * VL_TW, v_base = blockIdx.y * VL_TH;
    for (int idx = threadIdx.x; idx < HWD * HHT; idx += 256) {
        const int ly = idx / HWD, lx = idx - ly * HWD;
        const int qu = u_base + lx - 1, qv = v_base + ly - 1;
        float3 tx = make_float3(0.f, 0.f, 0.f), ty = tx;
        if (a_dn != 0.f && qu >= 0 && qu < p.W && qv >= 0 && qv < p.H) {
            float3 dx, dy;
            const float3 gc = dn_grad_c(p, qu, qv, a_dn, nullptr, dx, dy);
            tx = cross3(dy, gc);
            ty = cross3(gc, dx);
        }
        s_tx[0][ly][lx] = tx.x; s_tx[1][ly][lx] = tx.y; s_tx[2][ly][lx] = tx.z;
        s_ty[0][ly][lx] = ty.x; s_ty[1][ly][lx] = ty.y; s_ty[2][ly][lx] = ty.z;
    }
    __syncthreads();
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const int lx = threadIdx.x & 31, ly = (threadIdx.x >> 5) + 8 * half;
        const int u = u_base + lx, v = v_base + ly;
        if (u >= p.W || v >= p.H) continue;
        const size_t pix = (size_t)v * p.W + u;
        // RGB: d|x|/dx = sign(x)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float e = __ldg(out + c * HW + pix) - __ldg(gt_image + c * HW + pix);
            dL_dout[c * HW + pix] = a_l1 * (float)((e > 0.f) - (e < 0.f));
        }
        // rendered normal: gradient of |rn-g|_1 + 1 - rn.g through rn = n / max(|n|, eps)
        {
            const float3 g = make_float3(__ldg(p.gt + 3 * pix), __ldg(p.gt + 3 * pix + 1), __ldg(p.gt + 3 * pix + 2));
            const float rx = __ldg(p.rn + pix), ry = __ldg(p.rn + HW + pix), rz = __ldg(p.rn + 2 * HW + pix);
            const float len = sqrtf(rx * rx + ry * ry + rz * rz);
            const float rinv = 1.0f / fmaxf(len, 1e-12f);
            const float3 rn = make_float3(rx * rinv, ry * rinv, rz * rinv);
            const float ex = rn.x - g.x, ey = rn.y - g.y, ez = rn.z - g.z;
            float3 d = make_float3(a_mono * ((float)((ex > 0.f) - (ex < 0.f)) - g.x), a_mono * ((float)((ey > 0.f) - (ey < 0.f)) - g.y),
                                   a_mono * ((float)((ez > 0.f) - (ez < 0.f)) - g.z));
            if (len > 1e-12f) {
                const float dd = d.x * rn.x + d.y * rn.y + d.z * rn.z;
                d = make_float3((d.x - rn.x * dd) * rinv, (d.y - rn.y * dd) * rinv, (d.z - rn.z * dd) * rinv);
            } else {
                d = make_float3(d.x * rinv, d.y * rinv, d.z * rinv);
            }
            dL_dout[4 * HW + pix] = d.x; dL_dout[5 * HW + pix] = d.y; dL_dout[6 * HW + pix] = d.z;
        }
        // depth: the D-Normal 5-point gather (same terms, in the same order, as dnormal_backward_kernel)
        float3 gX = make_float3(0.f, 0.f, 0.f);
        auto via_dx = [&](int qu, float coef) {
            const int hx = qu - u_base + 1, hy = ly + 1;
            gX.x += coef * s_tx[0][hy][hx]; gX.y += coef * s_tx[1][hy][hx]; gX.z += coef * s_tx[2][hy][hx];
        };
        auto via_dy = [&](int qv, float coef) {
            const int hx = lx + 1, hy = qv - v_base + 1;
            gX.x += coef * s_ty[0][hy][hx]; gX.y += coef * s_ty[1][hy][hx]; gX.z += coef * s_ty[2][hy][hx];
        };
        if (a_dn != 0.f) {
            if (p.W > 1) {
                if (u > 0) { const int q = u - 1; via_dx(q, (q > 0) ? 0.5f : 1.0f); }
                if (u < p.W - 1) { const int q = u + 1; via_dx(q, (q < p.W - 1) ? -0.5f : -1.0f); }
                if (u == 0) via_dx(0, -1.0f);
                if (u == p.W - 1) via_dx(u, 1.0f);
            }
            if (p.H > 1) {
                if (v > 0) { const int q = v - 1; via_dy(q, (q > 0) ? 0.5f : 1.0f); }
                if (v < p.H - 1) { const int q = v + 1; via_dy(q, (q < p.H - 1) ? -0.5f : -1.0f); }
                if (v == 0) via_dy(0, -1.0f);
                if (v == p.H - 1) via_dy(v, 1.0f);
            }
        }
        const float ax = ((u + 0.5f) - p.cx) * p.inv_fx, ay = ((v + 0.5f) - p.cy) * p.inv_fy;
        dL_dout[3 * HW + pix] = gX.x * ax + gX.y * ay + gX.z;
        dL_dout[7 * HW + pix] = 0.f; dL_dout[8 * HW + pix] = 0.f; dL_dout[9 * HW + pix] = 0.f;
    }
}

// the scalar loss from the six sums: one thread instead of a dozen single-element PyTorch launches
__global__ void view_loss_total_kernel(const float* __restrict__ sums, const float N, const float w_l1, const float w_mono,
                                       const float w_dn, float* __restrict__ loss) {
    if (threadIdx.x != 0) return;
    const float s5 = sums[5];
    const float dn = s5 > 0.f ? (sums[3] + sums[4]) / fmaxf(s5, 1.0f) : 0.f;
    *loss = __fadd_rn(__fadd_rn(__fdiv_rn(__fmul_rn(w_l1, sums[0]), 3.0f * N), __fdiv_rn(__fmul_rn(w_mono, sums[1] + sums[2]), N)),
                      __fmul_rn(w_dn, dn));
}

extern "C" {

int vcr_view_loss_forward(const float* outmap, const float* gt_image, const float* gt_normal, const uint8_t* mask, int width,
                          int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max, float* sums,
                          void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!outmap || !gt_image || !gt_normal || !sums || width <= 0 || height <= 0) { vcr_set_error("view_loss_forward: bad argument"); return 1; }
    const size_t HW = (size_t)width * height;
    DnParams p{outmap + 3 * HW, gt_normal, outmap + 4 * HW, mask, width, height, 1.0f / fx, 1.0f / fy, cx, cy, exp_t, mask_depth_max, 1, HW};
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(sums, 0, 8 * sizeof(float), s)); }
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    VcrProfScope prof(VCR_K_DNORMAL_FWD, s);
    view_loss_forward_kernel<<<grid, 256, 0, s>>>(p, outmap, gt_image, sums);
    VCR_LAUNCH_CHECK("view_loss_forward_kernel");
    return 0;
}

int vcr_view_loss_backward(const float* outmap, const float* gt_image, const float* gt_normal, const uint8_t* mask, int width,
                           int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max, const float* sums,
                           const float* dL_dloss, float w_l1, float w_mono, float w_dn, float* dL_doutmap, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!outmap || !gt_image || !gt_normal || !sums || !dL_dloss || !dL_doutmap || width <= 0 || height <= 0) { vcr_set_error("view_loss_backward: bad argument"); return 1; }
    const size_t HW = (size_t)width * height;
    DnParams p{outmap + 3 * HW, gt_normal, outmap + 4 * HW, mask, width, height, 1.0f / fx, 1.0f / fy, cx, cy, exp_t, mask_depth_max, 1, HW};
    const dim3 grid((width + VL_TW - 1) / VL_TW, (height + VL_TH - 1) / VL_TH);
    VcrProfScope prof(VCR_K_DNORMAL_BWD, s);
    view_loss_backward_kernel<<<grid, 256, 0, s>>>(p, outmap, gt_image, sums, dL_dloss, w_l1, w_mono, w_dn, dL_doutmap);
    VCR_LAUNCH_CHECK("view_loss_backward_kernel");
    return 0;
}

int vcr_view_loss_total(const float* sums, int width, int height, float w_l1, float w_mono, float w_dn, float* loss, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!sums || !loss || width <= 0 || height <= 0) { vcr_set_error("view_loss_total: bad argument"); return 1; }
    VcrProfScope prof(VCR_K_DNORMAL_FWD, s);
    view_loss_total_kernel<<<1, 32, 0, s>>>(sums, (float)((double)width * (double)height), w_l1, w_mono, w_dn, loss);
    VCR_LAUNCH_CHECK("view_loss_total_kernel");
    return 0;
}

int vcr_dnormal_forward(const float* depth, const float* gt_normal, const float* render_normal, const uint8_t* mask, int width,
                        int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max, float* est_normal,
                        float* weight, float* sums, int zero_sums, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!depth || !gt_normal || !sums || width <= 0 || height <= 0) { vcr_set_error("dnormal_forward: bad argument"); return 1; }
    if (exp_t > 0.f && !render_normal) { vcr_set_error("dnormal_forward: render_normal is required when exp_t > 0"); return 1; }
    DnParams p{depth, gt_normal, render_normal, mask, width, height, 1.0f / fx, 1.0f / fy, cx, cy, exp_t, mask_depth_max, 3, 1};
    if (zero_sums) VCR_CUDA_TRY(cudaMemsetAsync(sums, 0, 3 * sizeof(float), s));
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    VcrProfScope prof(VCR_K_DNORMAL_FWD, s);
    dnormal_forward_kernel<<<grid, 256, 0, s>>>(p, est_normal, weight, sums);
    VCR_LAUNCH_CHECK("dnormal_forward_kernel");
    return 0;
}

int vcr_dnormal_backward(const float* depth, const float* gt_normal, const float* render_normal, const uint8_t* mask, int width,
                         int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max, const float* sums,
                         const float* dL_dloss, const float* dL_dest_normal, float* dL_ddepth, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!depth || !gt_normal || !sums || !dL_dloss || !dL_ddepth || width <= 0 || height <= 0) { vcr_set_error("dnormal_backward: bad argument"); return 1; }
    if (exp_t > 0.f && !render_normal) { vcr_set_error("dnormal_backward: render_normal is required when exp_t > 0"); return 1; }
    DnParams p{depth, gt_normal, render_normal, mask, width, height, 1.0f / fx, 1.0f / fy, cx, cy, exp_t, mask_depth_max, 3, 1};
    const dim3 grid((width + VL_TW - 1) / VL_TW, (height + VL_TH - 1) / VL_TH);
    VcrProfScope prof(VCR_K_DNORMAL_BWD, s);
    dnormal_backward_kernel<<<grid, 256, 0, s>>>(p, sums, dL_dloss, dL_dest_normal, dL_ddepth);
    VCR_LAUNCH_CHECK("dnormal_backward_kernel");
    return 0;
}

}  // extern "C"

// Ground-truth images cross the bus as the 8-bit images they are; dst = (float)src / 255 (IEEE division: the value the reference's
// loader computes, utils/general_utils.py PILtoTorch).  Sixteen pixels per thread.
__global__ void __launch_bounds__(256) unpack_u8_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, const long long n) {
    const long long i = 16ll * ((long long)blockIdx.x * 256 + threadIdx.x);
    if (i + 16 <= n) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(src + i));
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; k++)
            reinterpret_cast<float4*>(dst + i)[k] = make_float4(__fdiv_rn((float)(w[k] & 255u), 255.0f), __fdiv_rn((float)((w[k] >> 8) & 255u), 255.0f),
                                                                __fdiv_rn((float)((w[k] >> 16) & 255u), 255.0f), __fdiv_rn((float)(w[k] >> 24), 255.0f));
    } else {
        for (long long j = i; j < n; j++) dst[j] = __fdiv_rn((float)src[j], 255.0f);
    }
}
extern "C" int vcr_unpack_u8(const uint8_t* src, float* dst, long long n, void* stream) {
    if (n <= 0) return 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 15)) { vcr_set_error("vcr_unpack_u8: src and dst must be 16-byte aligned"); return 1; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long threads = (n + 15) / 16;
    unpack_u8_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(src, dst, n);
    VCR_LAUNCH_CHECK("unpack_u8_kernel");
    return 0;
}
