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
