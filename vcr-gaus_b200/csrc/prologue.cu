// prologue.cu — everything render() does per Gaussian BEFORE the rasterizer call, in one pass each way.
//
// Replaces, per view, the PyTorch chain
//   pc.get_opacity / get_scaling / get_rotation / get_features      scene/gaussian_model.py:125-162 (sigmoid, exp, normalize, cat)
//   pc.get_normal(...)                                               scene/gaussian_model.py:168-192 (build_rotation, argmin, gather)
//   flip towards the camera + rotate into camera space               gaussian_renderer/__init__.py:95-101
// (~40 elementwise launches over P forward and as many backward, including a 192 B/Gaussian concat copy), by one kernel
// per direction.  HBM bound: forward reads 59 raw floats + 3 (xyz) and writes 62 activated floats + 1 byte per Gaussian.
#include "common.cuh"

struct PrologueParams {
    int P, n_rest;                      // n_rest = SH coefficients beyond DC (15 at degree 3)
    const float* xyz;                   // [P,3]
    const float* features_dc;           // [P,1,3]
    const float* features_rest;         // [P,n_rest,3]
    const float* opacity_raw;           // [P,1]   (logit)
    const float* scaling_raw;           // [P,3]   (log)
    const float* rotation_raw;          // [P,4]   (unnormalised quaternion r,x,y,z)
    const uint8_t* valid;               // [P] or NULL: Gaussians outside get a zero normal (get_normal(is_all=False))
    const float* cam_center;            // [3]
    const float* R_w2c;                 // [3,3] row-major (torch.tensor(camera.R.T))
    float* opacities; float* scales; float* rotations; float* shs; float* normals;
    uint8_t* aux;                       // [P]: bits 0-1 = argmin axis, bit 2 = flipped, bit 3 = has normal
    // backward
    const float* dL_dopacities; const float* dL_dscales; const float* dL_drotations; const float* dL_dshs; const float* dL_dnormals;
    float* dL_dfeatures_dc; float* dL_dfeatures_rest; float* dL_dopacity_raw; float* dL_dscaling_raw; float* dL_drotation_raw;
};

// column `axis` of build_rotation(q) (tools/general_utils.py:98-119), q = (r, x, y, z) already of unit length
__device__ __forceinline__ float3 rot_column(const float r, const float x, const float y, const float z, const int axis) {
    if (axis == 0) return make_float3(1.f - 2.f * (y * y + z * z), 2.f * (x * y + r * z), 2.f * (x * z - r * y));
    if (axis == 1) return make_float3(2.f * (x * y - r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + r * x));
    return make_float3(2.f * (x * z + r * y), 2.f * (y * z - r * x), 1.f - 2.f * (x * x + y * y));
}

// copy the SH coefficients of 256 Gaussians between the split layout (dc | rest) and the joint [P,1+n_rest,3] layout,
// coalesced on both sides: `joint` rows are (1+n_rest)*3 floats, the first 3 of which belong to dc
template <bool TO_JOINT>
__device__ __forceinline__ void sh_copy_block(const int P, const int n_rest, float* joint, float* dc, float* rest, const int block) {
    const int row = (1 + n_rest) * 3;
    const size_t g0 = (size_t)block * 256;
    const size_t n_g = min((size_t)256, (size_t)P - g0);
    const size_t total = n_g * row;
    for (size_t e = threadIdx.x; e < total; e += 256) {
        const size_t gi = e / row;
        const int k = (int)(e - gi * row);
        float* split = k < 3 ? dc + (g0 + gi) * 3 + k : rest + (g0 + gi) * (size_t)(n_rest * 3) + (k - 3);
        if (TO_JOINT) joint[g0 * row + e] = *split;
        else *split = joint[g0 * row + e];
    }
}

__global__ void __launch_bounds__(256) prologue_forward_kernel(const PrologueParams p) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    sh_copy_block<true>(p.P, p.n_rest, p.shs, const_cast<float*>(p.features_dc), const_cast<float*>(p.features_rest), blockIdx.x);
    if (i >= p.P) return;
    // opacity_activation = sigmoid, scaling_activation = exp, rotation_activation = F.normalize (eps 1e-12)
    p.opacities[i] = 1.0f / (1.0f + expf(-__ldg(p.opacity_raw + i)));
    const float s0 = expf(__ldg(p.scaling_raw + 3 * (size_t)i)), s1 = expf(__ldg(p.scaling_raw + 3 * (size_t)i + 1)),
                s2 = expf(__ldg(p.scaling_raw + 3 * (size_t)i + 2));
    p.scales[3 * (size_t)i] = s0; p.scales[3 * (size_t)i + 1] = s1; p.scales[3 * (size_t)i + 2] = s2;
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(p.rotation_raw) + i);
    const float n1 = fmaxf(sqrtf(q4.x * q4.x + q4.y * q4.y + q4.z * q4.z + q4.w * q4.w), 1e-12f);
    const float4 qn = make_float4(q4.x / n1, q4.y / n1, q4.z / n1, q4.w / n1);
    reinterpret_cast<float4*>(p.rotations)[i] = qn;
    // normal = column argmin(scaling) of build_rotation(get_rotation), which normalises once more
    const bool has = p.valid == nullptr || p.valid[i] != 0;
    const int axis = (s1 < s0) ? ((s2 < s1) ? 2 : 1) : ((s2 < s0) ? 2 : 0);        // torch.argmin: first minimum
    float3 nc = make_float3(0.f, 0.f, 0.f);
    uint32_t flags = (uint32_t)axis;
    if (has) {
        const float n2 = sqrtf(qn.x * qn.x + qn.y * qn.y + qn.z * qn.z + qn.w * qn.w);
        const float3 nw = rot_column(qn.x / n2, qn.y / n2, qn.z / n2, qn.w / n2, axis);
        const float vx = __ldg(p.xyz + 3 * (size_t)i) - __ldg(p.cam_center), vy = __ldg(p.xyz + 3 * (size_t)i + 1) - __ldg(p.cam_center + 1),
                    vz = __ldg(p.xyz + 3 * (size_t)i + 2) - __ldg(p.cam_center + 2);
        // normal * (((view_dir . normal) > 0) * 1 - 0.5) * 2: +1 when the normal points away from the camera, else -1
        const bool pos = (vx * nw.x + vy * nw.y + vz * nw.z) > 0.f;
        const float sg = pos ? 1.f : -1.f;
        const float* M = p.R_w2c;      // normals_precomp = normal @ R_w2c^T  ->  n_cam[j] = sum_i n[i] * M[j][i]
        nc.x = sg * (nw.x * __ldg(M + 0) + nw.y * __ldg(M + 1) + nw.z * __ldg(M + 2));
        nc.y = sg * (nw.x * __ldg(M + 3) + nw.y * __ldg(M + 4) + nw.z * __ldg(M + 5));
        nc.z = sg * (nw.x * __ldg(M + 6) + nw.y * __ldg(M + 7) + nw.z * __ldg(M + 8));
        flags |= (pos ? 0u : 4u) | 8u;
    }
    p.normals[3 * (size_t)i] = nc.x; p.normals[3 * (size_t)i + 1] = nc.y; p.normals[3 * (size_t)i + 2] = nc.z;
    p.aux[i] = (uint8_t)flags;
}

__global__ void __launch_bounds__(256) prologue_backward_kernel(const PrologueParams p) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    sh_copy_block<false>(p.P, p.n_rest, const_cast<float*>(p.dL_dshs), p.dL_dfeatures_dc, p.dL_dfeatures_rest, blockIdx.x);
    if (i >= p.P) return;
    const float op = p.opacities[i];
    p.dL_dopacity_raw[i] = __ldg(p.dL_dopacities + i) * op * (1.f - op);
#pragma unroll
    for (int k = 0; k < 3; k++)
        p.dL_dscaling_raw[3 * (size_t)i + k] = __ldg(p.dL_dscales + 3 * (size_t)i + k) * p.scales[3 * (size_t)i + k];
    const float4 qn = reinterpret_cast<const float4*>(p.rotations)[i];
    float4 g = __ldg(reinterpret_cast<const float4*>(p.dL_drotations) + i);       // gradient w.r.t. the normalised quaternion
    const uint32_t flags = p.aux[i];
    if (flags & 8u) {
        const float sg = (flags & 4u) ? -1.f : 1.f;
        const float* M = p.R_w2c;
        const float gx = __ldg(p.dL_dnormals + 3 * (size_t)i), gy = __ldg(p.dL_dnormals + 3 * (size_t)i + 1),
                    gz = __ldg(p.dL_dnormals + 3 * (size_t)i + 2);
        // d n_cam / d n_world (times the sign)
        const float3 gw = make_float3(sg * (gx * __ldg(M + 0) + gy * __ldg(M + 3) + gz * __ldg(M + 6)),
                                      sg * (gx * __ldg(M + 1) + gy * __ldg(M + 4) + gz * __ldg(M + 7)),
                                      sg * (gx * __ldg(M + 2) + gy * __ldg(M + 5) + gz * __ldg(M + 8)));
        const float n2 = sqrtf(qn.x * qn.x + qn.y * qn.y + qn.z * qn.z + qn.w * qn.w);
        const float r = qn.x / n2, x = qn.y / n2, y = qn.z / n2, z = qn.w / n2;
        float4 gq;                      // d (column axis of R) / d (r, x, y, z), contracted with gw
        const int axis = (int)(flags & 3u);
        if (axis == 0) {
            gq.x = 2.f * (z * gw.y - y * gw.z);
            gq.y = 2.f * (y * gw.y + z * gw.z);
            gq.z = 2.f * (-2.f * y * gw.x + x * gw.y - r * gw.z);
            gq.w = 2.f * (-2.f * z * gw.x + r * gw.y + x * gw.z);
        } else if (axis == 1) {
            gq.x = 2.f * (-z * gw.x + x * gw.z);
            gq.y = 2.f * (y * gw.x - 2.f * x * gw.y + r * gw.z);
            gq.z = 2.f * (x * gw.x + z * gw.z);
            gq.w = 2.f * (-r * gw.x - 2.f * z * gw.y + y * gw.z);
        } else {
            gq.x = 2.f * (y * gw.x - x * gw.y);
            gq.y = 2.f * (z * gw.x - r * gw.y - 2.f * x * gw.z);
            gq.z = 2.f * (r * gw.x + z * gw.y - 2.f * y * gw.z);
            gq.w = 2.f * (x * gw.x + y * gw.y);
        }
        // through build_rotation's own normalisation q = qn / |qn|
        const float d2 = r * gq.x + x * gq.y + y * gq.z + z * gq.w;
        g.x += (gq.x - r * d2) / n2; g.y += (gq.y - x * d2) / n2; g.z += (gq.z - y * d2) / n2; g.w += (gq.w - z * d2) / n2;
    }
    // through rotation_activation: qn = raw / max(|raw|, eps)
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(p.rotation_raw) + i);
    const float n1 = fmaxf(sqrtf(q4.x * q4.x + q4.y * q4.y + q4.z * q4.z + q4.w * q4.w), 1e-12f);
    const float d1 = qn.x * g.x + qn.y * g.y + qn.z * g.z + qn.w * g.w;
    reinterpret_cast<float4*>(p.dL_drotation_raw)[i] =
        make_float4((g.x - qn.x * d1) / n1, (g.y - qn.y * d1) / n1, (g.z - qn.z * d1) / n1, (g.w - qn.w * d1) / n1);
}

int vcr_launch_prologue(const PrologueParams& p, bool backward, cudaStream_t s) {
    VcrProfScope prof(VCR_K_PROLOGUE, s);
    const int blocks = (p.P + 255) / 256;
    if (backward) prologue_backward_kernel<<<blocks, 256, 0, s>>>(p);
    else prologue_forward_kernel<<<blocks, 256, 0, s>>>(p);
    VCR_LAUNCH_CHECK(backward ? "prologue_backward_kernel" : "prologue_forward_kernel");
    return 0;
}

extern "C" int vcr_prologue_forward(int P, int n_rest, const float* xyz, const float* features_dc, const float* features_rest,
                                    const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                                    const uint8_t* valid, const float* cam_center, const float* R_w2c, float* opacities,
                                    float* scales, float* rotations, float* shs, float* normals, uint8_t* aux, void* stream) {
    if (P <= 0) return 0;
    if (n_rest < 0 || !xyz || !features_dc || (n_rest > 0 && !features_rest) || !opacity_raw || !scaling_raw || !rotation_raw ||
        !cam_center || !R_w2c || !opacities || !scales || !rotations || !shs || !normals || !aux) {
        vcr_set_error("prologue_forward: NULL argument");
        return 1;
    }
    if ((reinterpret_cast<size_t>(rotation_raw) | reinterpret_cast<size_t>(rotations)) & 15) {
        vcr_set_error("prologue_forward: rotation_raw and rotations must be 16-byte aligned (read / written as float4)");
        return 1;
    }
    PrologueParams p;
    memset(&p, 0, sizeof(p));
    p.P = P; p.n_rest = n_rest; p.xyz = xyz; p.features_dc = features_dc; p.features_rest = features_rest;
    p.opacity_raw = opacity_raw; p.scaling_raw = scaling_raw; p.rotation_raw = rotation_raw; p.valid = valid;
    p.cam_center = cam_center; p.R_w2c = R_w2c; p.opacities = opacities; p.scales = scales; p.rotations = rotations;
    p.shs = shs; p.normals = normals; p.aux = aux;
    return vcr_launch_prologue(p, false, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vcr_prologue_backward(int P, int n_rest, const float* rotation_raw, const float* opacities, const float* scales,
                                     const float* rotations, const uint8_t* aux, const float* R_w2c, const float* dL_dopacities,
                                     const float* dL_dscales, const float* dL_drotations, const float* dL_dshs,
                                     const float* dL_dnormals, float* dL_dfeatures_dc, float* dL_dfeatures_rest,
                                     float* dL_dopacity_raw, float* dL_dscaling_raw, float* dL_drotation_raw, void* stream) {
    if (P <= 0) return 0;
    if (n_rest < 0 || !rotation_raw || !opacities || !scales || !rotations || !aux || !R_w2c || !dL_dopacities || !dL_dscales ||
        !dL_drotations || !dL_dshs || !dL_dnormals || !dL_dfeatures_dc || (n_rest > 0 && !dL_dfeatures_rest) || !dL_dopacity_raw ||
        !dL_dscaling_raw || !dL_drotation_raw) {
        vcr_set_error("prologue_backward: NULL argument");
        return 1;
    }
    if ((reinterpret_cast<size_t>(rotation_raw) | reinterpret_cast<size_t>(rotations) | reinterpret_cast<size_t>(dL_drotations) |
         reinterpret_cast<size_t>(dL_drotation_raw)) & 15) {
        vcr_set_error("prologue_backward: rotation_raw, rotations, dL_drotations and dL_drotation_raw must be 16-byte aligned (float4 accesses)");
        return 1;
    }
    PrologueParams p;
    memset(&p, 0, sizeof(p));
    p.P = P; p.n_rest = n_rest; p.rotation_raw = rotation_raw; p.opacities = const_cast<float*>(opacities);
    p.scales = const_cast<float*>(scales); p.rotations = const_cast<float*>(rotations); p.aux = const_cast<uint8_t*>(aux);
    p.R_w2c = R_w2c; p.dL_dopacities = dL_dopacities; p.dL_dscales = dL_dscales; p.dL_drotations = dL_drotations;
    p.dL_dshs = dL_dshs; p.dL_dnormals = dL_dnormals; p.dL_dfeatures_dc = dL_dfeatures_dc; p.dL_dfeatures_rest = dL_dfeatures_rest;
    p.dL_dopacity_raw = dL_dopacity_raw; p.dL_dscaling_raw = dL_dscaling_raw; p.dL_drotation_raw = dL_drotation_raw;
    return vcr_launch_prologue(p, true, reinterpret_cast<cudaStream_t>(stream));
}
