// render_backward.cu — back-to-front replay of the compositing with per-Gaussian gradient accumulation.
//
// Replaces renderCUDA<3,2> backward (reference DGR/cuda_rasterizer/backward.cu:422-812).
//
// Same tiling as the forward (one 256-thread block per 16x16 tile, one pixel per thread), Gaussians
// staged in shared memory in reverse list order.  Where the reference issues ~20 atomicAdd per
// contributing (pixel, Gaussian) pair, this kernel
//   0. drops, per block and per warp, the tail of the list behind the last contributor of any of its pixels, and
//      culls (warp, Gaussian) pairs with a conservative footprint box against the warp's 16x2 pixel strip
//      (8 ballots per 256 staged Gaussians); only candidates enter the per-pixel code,
//   1. skips a candidate for the whole warp when no lane contributes (__any_sync),
//   2. reduces the 16 main gradient components across the warp with a transposed butterfly
//      (15 + 1 shuffles instead of 16 x 5), leaving component k on lane k, and
//   3. lands them with ONE coalesced 64-byte atomic request per (warp, Gaussian) into a per-Gaussian
//      AoS record (common.cuh ACC_*), from which preprocess_backward reads them back once.
// The formulas (including the reference's quirks) are kept term by term; only the
// division by the final weight is replaced by a multiplication with its reciprocal.
#include "common.cuh"
#include <atomic>

#define RB_BATCH 256
#define RB_SPARSE_LANES 2

struct __align__(16) StagedGeoB {
    float x, y, ca, cb;
    float cc, op, nx;    // nx: dot(normal, xyz_view) (intersection mode) or view depth
    uint32_t id_inside;  // (id << 1) | inside
};

template <bool HAS_DIRS>
struct RbStage {                     // one staging buffer (24 KB with intersection depth, 20 KB without)
    StagedGeoB geo[RB_BATCH];
    float4 box[RB_BATCH];            // conservative footprint box {xlo, xhi, ylo, yhi}
    float2 ex[RB_BATCH];             // exact strip test: { -(tau + slack), 1 / conic.x }  (vcr_alpha_rowtest)
    float4 nrm[RB_BATCH];            // normal.xyz, semantics[0]
    float4 col[RB_BATCH];            // rgb, semantics[1]
    float4 xv[HAS_DIRS ? RB_BATCH : 1];  // view-space position (intersection mode)
};

// Transposed warp reduction: every lane passes 16 values; on return lane L (and lane L+16) holds the
// warp-wide sum of value (L & 15).  8+4+2+1 exchange shuffles + 1 plain butterfly step.
__device__ __forceinline__ float warp_reduce16_transposed(float (&v)[16], int lane) {
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float send = hi ? v[i] : v[i + 8];
            const float keep = hi ? v[i + 8] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
    }
    {
        const bool hi = lane & 4;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float send = hi ? v[i] : v[i + 4];
            const float keep = hi ? v[i + 4] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
    }
    {
        const bool hi = lane & 2;
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const float send = hi ? v[i] : v[i + 2];
            const float keep = hi ? v[i + 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        }
    }
    {
        const bool hi = lane & 1;
        const float send = hi ? v[0] : v[1];
        const float keep = hi ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

template <bool HAS_DIRS, bool HAS_OBJ>
__global__ void __launch_bounds__(256) render_backward_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, int gx, const uint32_t* __restrict__ point_list, int W, int H, const float* __restrict__ bg,
    const float2* __restrict__ means2D, const float* __restrict__ xyz_view, const float4* __restrict__ conic_opacity,
    const float* __restrict__ colors, const float* __restrict__ depths, const float* __restrict__ normals,
    const float* __restrict__ objects, const float* __restrict__ dirs, const float* __restrict__ alpha_img,
    const float* __restrict__ depth_img, const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpixels, const uint8_t* __restrict__ inside_gs, float* __restrict__ acc,
    const unsigned int* __restrict__ abort_flag) {
    extern __shared__ __align__(16) unsigned char rb_smem[];
    if (__ldg(abort_flag)) return;      // the sort of this view failed (binning.cu): the lists are garbage, accumulate nothing
    RbStage<HAS_DIRS>* s_stage = reinterpret_cast<RbStage<HAS_DIRS>*>(rb_smem);   // [2]: double-buffered staging
    __shared__ uint32_t s_block_last;

    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t tile = __ldg(tile_order + blockIdx.x);      // blocks take tiles longest-list first
    const uint32_t tile_x = tile % (uint32_t)gx, tile_y = tile / (uint32_t)gx;
    const uint32_t px = tile_x * VCR_TILE_X + (tid & 15);
    const uint32_t py = tile_y * VCR_TILE_Y + (tid >> 4);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const size_t HW = (size_t)H * W;
    const size_t pix_id = (size_t)W * py + px;
    const float pfx = (float)px, pfy = (float)py;
    const float strip_x0 = (float)(tile_x * VCR_TILE_X), strip_x1 = strip_x0 + (float)(VCR_TILE_X - 1);
    const float strip_y0 = (float)(tile_y * VCR_TILE_Y + 2 * (tid >> 5)), strip_y1 = strip_y0 + 1.0f;

    const uint2 range = ranges[tile];
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    // Entries behind the last contributor of every pixel of the warp / block can never pass `contributor <
    // last_contributor` (backward.cu:589-591): drop them without staging or evaluating them.
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    if (tid == 0) s_block_last = 0u;
    __syncthreads();
    if (lane == 0 && warp_last) atomicMax(&s_block_last, warp_last);
    __syncthreads();
    const uint32_t block_last = s_block_last;          // <= range.y - range.x
    const uint32_t list_end = range.x + block_last;    // effective list: [range.x, list_end)
    const int rounds = (int)((block_last + RB_BATCH - 1) / RB_BATCH);
    int toDo = (int)block_last;

    const float T_final = inside ? final_Ts[pix_id] : 0.f;
    float T = T_final;
    const float inv_weight = inside ? 1.0f / alpha_img[pix_id] : 1.f;
    const float depth_final = inside ? depth_img[pix_id] : 1.f;

    float accum_rec[3] = {0.f, 0.f, 0.f}, dL_dpixel[3] = {0.f, 0.f, 0.f};
    float accum_depth_rec = 0.f, accum_rec2 = 0.f, dL_dpixel_depth = 0.f;
    float accum_normal_rec[3] = {0.f, 0.f, 0.f}, dL_dpixel_normal[3] = {0.f, 0.f, 0.f};
    float accum_rec_obj[2] = {0.f, 0.f}, dL_dpixel_obj[2] = {0.f, 0.f};
    float accum_alpha_rec = 0.f, dL_dalpha = 0.f;
    float3 dir = make_float3(0.f, 0.f, 1.f);
    if (inside) {
#pragma unroll
        for (int i = 0; i < 3; i++) dL_dpixel[i] = __ldg(dL_dpixels + i * HW + pix_id);
        dL_dpixel_depth = __ldg(dL_dpixels + VCR_DEPTH_OFFSET * HW + pix_id);
#pragma unroll
        for (int i = 0; i < 3; i++) dL_dpixel_normal[i] = __ldg(dL_dpixels + (VCR_NORMAL_OFFSET + i) * HW + pix_id);
        if (HAS_OBJ) {
#pragma unroll
            for (int i = 0; i < 2; i++) dL_dpixel_obj[i] = __ldg(dL_dpixels + (VCR_OBJECT_OFFSET + i) * HW + pix_id);
        }
        dL_dalpha = __ldg(dL_dpixels + VCR_ALPHA_OFFSET * HW + pix_id);
        if (HAS_DIRS) dir = make_float3(__ldg(dirs + pix_id), __ldg(dirs + HW + pix_id), __ldg(dirs + 2 * HW + pix_id));
    }
    float bg_dot_dpixel = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) bg_dot_dpixel += __ldg(bg + i) * dL_dpixel[i];

    float last_alpha = 0.f, last_color[3] = {0.f, 0.f, 0.f}, last_depth = 0.f, last_normal[3] = {0.f, 0.f, 0.f};
    float last_object[2] = {0.f, 0.f};
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

    // gather reversed-list entry (batch, tid) into staging buffer `buf`
    auto stage = [&](int batch, int buf) {
        const uint32_t progress = (uint32_t)batch * RB_BATCH + tid;
        if (range.x + progress >= list_end) return;
        RbStage<HAS_DIRS>& st = s_stage[buf];
        const uint32_t id = __ldg(point_list + (list_end - progress - 1));
        const float2 xy = __ldg(means2D + id);
        const float4 co = __ldg(conic_opacity + id);
        const float3 n = make_float3(__ldg(normals + 3 * (size_t)id), __ldg(normals + 3 * (size_t)id + 1),
                                     __ldg(normals + 3 * (size_t)id + 2));
        StagedGeoB sg;
        sg.x = xy.x; sg.y = xy.y; sg.ca = co.x; sg.cb = co.y; sg.cc = co.z; sg.op = co.w;
        st.box[tid] = vcr_alpha_bbox(xy, co);
        st.ex[tid] = vcr_alpha_rowtest(co);
        if (HAS_DIRS) {
            const float3 xv = make_float3(__ldg(xyz_view + 3 * (size_t)id), __ldg(xyz_view + 3 * (size_t)id + 1),
                                          __ldg(xyz_view + 3 * (size_t)id + 2));
            sg.nx = n.x * xv.x + n.y * xv.y + n.z * xv.z;
            st.xv[tid] = make_float4(xv.x, xv.y, xv.z, 0.f);
        } else {
            sg.nx = __ldg(depths + id);
        }
        sg.id_inside = (id << 1) | (__ldg(inside_gs + id) ? 1u : 0u);
        st.geo[tid] = sg;
        float o0 = 0.f, o1 = 0.f;
        if (HAS_OBJ) { o0 = __ldg(objects + 2 * (size_t)id); o1 = __ldg(objects + 2 * (size_t)id + 1); }
        st.nrm[tid] = make_float4(n.x, n.y, n.z, o0);
        st.col[tid] = make_float4(__ldg(colors + 3 * (size_t)id), __ldg(colors + 3 * (size_t)id + 1),
                                  __ldg(colors + 3 * (size_t)id + 2), o1);
    };

    if (rounds > 0) stage(0, 0);
    __syncthreads();
    for (int i = 0; i < rounds; i++, toDo -= RB_BATCH) {
        const int buf = i & 1;
        if (i + 1 < rounds) stage(i + 1, buf ^ 1);   // overlaps with the replay of batch i
        const RbStage<HAS_DIRS>& st = s_stage[buf];

        const int cnt = min(RB_BATCH, toDo);
        // entry j of this batch has list position (contributor index) block_last - 1 - (i*RB_BATCH + j)
        const int first_pos = (int)block_last - 1 - i * RB_BATCH;
#pragma unroll 1
        for (int k = 0; k < RB_BATCH / 32; k++) {
            if (k * 32 >= cnt) break;
            const int jj = k * 32 + lane;
            const float4 bx = st.box[jj];
            bool hit = jj < cnt && (uint32_t)(first_pos - jj) < warp_last && bx.x <= strip_x1 && bx.y >= strip_x0 &&
                       bx.z <= strip_y1 && bx.w >= strip_y0;
            if (hit) {
                const float4 gq = *reinterpret_cast<const float4*>(&st.geo[jj]);        // x, y, ca, cb
                hit = vcr_strip_may_contribute(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], strip_x0, strip_x1, strip_y0);
            }
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int j = k * 32 + __ffs(m) - 1;
                m &= m - 1;
                const uint32_t contributor = (uint32_t)(first_pos - j);
                const float4 g0 = *reinterpret_cast<const float4*>(&st.geo[j]);
                const float4 g1 = *reinterpret_cast<const float4*>(&st.geo[j].cc);
                const float dx = g0.x - pfx, dy = g0.y - pfy;
                bool contrib = inside && contributor < last_contributor;
                const float q = __fmaf_rn(__fmul_rn(g0.z, dx), dx, __fmul_rn(__fmul_rn(g1.x, dy), dy));
                const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(g0.w, dx), dy));
                contrib = contrib && !(power > 0.0f);
                if (!__any_sync(0xffffffffu, contrib)) continue;
                // Fast intrinsics from here on: the backward only has to meet 1e-3 relative L2 (the forward, which must
                // match the reference bit for bit, keeps expf and IEEE division).
                const float G = __expf(power);
                const float alpha = fminf(0.99f, g1.y * G);
                contrib = contrib && !(alpha < 1.0f / 255.0f);
                const float4 nrm = st.nrm[j];
                float c_d, inv_pd = 1.f;
                if (HAS_DIRS) {
                    inv_pd = __fdividef(1.0f, nrm.x * dir.x + nrm.y * dir.y + nrm.z * dir.z);
                    c_d = (g1.z * inv_pd) * dir.z;
                } else {
                    c_d = g1.z;
                }
                contrib = contrib && !(c_d < 0.2f);  // == (double)c_d < 0.2, backward.cu:616
                const uint32_t cmask = __ballot_sync(0xffffffffu, contrib);
                if (cmask == 0u) continue;

                float v[16];
#pragma unroll
                for (int q_ = 0; q_ < 16; q_++) v[q_] = 0.f;
                float ex_obj1 = 0.f;
                const uint32_t idi = __float_as_uint(g1.w);
                if (contrib) {
                    const float inv1ma = __fdividef(1.0f, 1.f - alpha);
                    T = T * inv1ma;
                    const float aT = alpha * T;
                    float dL_dopa = 0.0f;
                    const float4 col = st.col[j];
                    const float cc[3] = {col.x, col.y, col.z};
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = cc[ch];
                        dL_dopa += (cc[ch] - accum_rec[ch]) * dL_dpixel[ch];
                        v[ACC_COLOR_R + ch] = aT * dL_dpixel[ch];
                    }
                    if (idi & 1u) {
                        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
                        accum_rec2 = last_alpha + (1.f - last_alpha) * accum_rec2;  // == accum_normal_rec2[ch], same recurrence
                        dL_dopa += (((c_d - accum_depth_rec) - depth_final * (1 - accum_rec2)) * inv_weight) * dL_dpixel_depth;
                        const float aTw = aT * inv_weight;
                        const float dL_dd_j = aTw * dL_dpixel_depth;
                        float dnd[3] = {0.f, 0.f, 0.f};
                        if (HAS_DIRS) {  // backward.cu:722-728
                            const float4 xv = st.xv[j];
                            const float sv = dL_dd_j * dir.z * inv_pd;      // dL/dxyz_view = normal * sum(sv)
                            const float s2 = sv * inv_pd;
                            const float pro_dot = nrm.x * dir.x + nrm.y * dir.y + nrm.z * dir.z;
                            dnd[0] = s2 * (pro_dot * xv.x - dir.x * g1.z);
                            dnd[1] = s2 * (pro_dot * xv.y - dir.y * g1.z);
                            dnd[2] = s2 * (pro_dot * xv.z - dir.z * g1.z);
                            v[ACC_SV_OR_DEPTH] = sv;
                        } else {
                            v[ACC_SV_OR_DEPTH] = dL_dd_j;
                        }
                        const float nn[3] = {nrm.x, nrm.y, nrm.z};
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            const float n = nn[ch];
                            accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
                            last_normal[ch] = n;
                            // reference quirk: last_normal was just overwritten with n (backward.cu:741-745)
                            dL_dopa += (((n - accum_normal_rec[ch]) - n * (1 - accum_rec2)) * inv_weight) * dL_dpixel_normal[ch];
                            v[ACC_NORMAL_X + ch] = aTw * dL_dpixel_normal[ch] + dnd[ch];
                        }
                    }
                    if (HAS_OBJ) {
                        const float oo[2] = {nrm.w, col.w};
                        float eo[2];
#pragma unroll
                        for (int ch = 0; ch < 2; ch++) {
                            accum_rec_obj[ch] = last_alpha * last_object[ch] + (1.f - last_alpha) * accum_rec_obj[ch];
                            last_object[ch] = oo[ch];
                            dL_dalpha += (oo[ch] - accum_rec_obj[ch]) * dL_dpixel_obj[ch];  // persists across Gaussians (backward.cu:766)
                            eo[ch] = aT * dL_dpixel_obj[ch];
                        }
                        v[ACC_OBJ_0] = eo[0]; ex_obj1 = eo[1];
                    }
                    accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
                    dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
                    dL_dopa *= T;
                    last_alpha = alpha;
                    last_depth = c_d;
                    dL_dopa += (-T_final * inv1ma) * bg_dot_dpixel;

                    // raw moments; the per-Gaussian factors (opacity, conic, 0.5 W) are applied once in preprocess_backward
                    const float b = G * dL_dopa;
                    const float bx = b * dx, by = b * dy;
                    v[ACC_M0] = b; v[ACC_MX] = bx; v[ACC_MY] = by;
                    v[ACC_MXX] = bx * dx; v[ACC_MXY] = bx * dy; v[ACC_MYY] = by * dy;
                    v[ACC_ABS_X] = fabsf(bx * g0.z + by * g0.w);
                    v[ACC_ABS_Y] = fabsf(by * g1.x + bx * g0.w);
                }
                float* rec = acc + (size_t)(idi >> 1) * VCR_ACC_STRIDE;
                if (__popc(cmask) <= RB_SPARSE_LANES) {
                    // few contributing pixels (footprint edges): cheaper to let each of them issue its own atomics than
                    // to run the 32-lane reduction
                    if (contrib) {
#pragma unroll
                        for (int q_ = 0; q_ < (HAS_OBJ ? 16 : 15); q_++) atomicAdd(rec + q_, v[q_]);
                        if (HAS_OBJ) atomicAdd(rec + ACC_OBJ_1, ex_obj1);
                    }
                    continue;
                }
                const float red = warp_reduce16_transposed(v, lane);
                if (lane < (HAS_OBJ ? 16 : 15)) atomicAdd(rec + lane, red);
                if (HAS_OBJ) {
                    const float o1 = vcr_warp_sum(ex_obj1);
                    if (lane == 16) atomicAdd(rec + ACC_OBJ_1, o1);
                }
            }
        }
        // one barrier per batch: publishes batch i+1 and frees buffer `buf` for batch i+2
        __syncthreads();
    }
}

int vcr_launch_render_backward(const BwdParams& p, GeomState& g, BinningState& b, ImageState& img, cudaStream_t s) {
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(p.acc, 0, (size_t)p.P * VCR_ACC_STRIDE * sizeof(float), s)); }
    const unsigned int* abort_flag = vcr_radix_abort_flag();
    if (!abort_flag) { vcr_set_error("cannot resolve the sort status flag on this device"); return 1; }
    VcrProfScope prof(VCR_K_RENDER_BWD, s);
    const dim3 grid(p.gx * p.gy, 1, 1);
    const float* col = p.colors_precomp ? p.colors_precomp : g.rgb;
#define RB_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, p.background, g.means2D, g.xyz_view, g.conic_opacity, col, g.depths, \
                p.normals, p.sh_objs, p.dirs, p.alpha, p.depth, img.accum_alpha, img.n_contrib, p.dL_dpix, p.inside, p.acc, abort_flag
    // the opt-in shared-memory size is a per-device function attribute
    // (atomic flag: the forward thread and autograd's backward thread may both get here; setting the attribute twice is harmless)
    static std::atomic<bool> attr_set_dev[64];
    int dev_ = 0;
    VCR_CUDA_TRY(cudaGetDevice(&dev_));
    if (dev_ >= 0 && dev_ < 64 && !attr_set_dev[dev_].load(std::memory_order_acquire)) {
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<true>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<true>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<false>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<false>))));
        attr_set_dev[dev_].store(true, std::memory_order_release);
    }
    if (p.dirs) {
        const size_t sm = 2 * sizeof(RbStage<true>);
        if (p.sh_objs) render_backward_kernel<true, true><<<grid, 256, sm, s>>>(RB_ARGS);
        else render_backward_kernel<true, false><<<grid, 256, sm, s>>>(RB_ARGS);
    } else {
        const size_t sm = 2 * sizeof(RbStage<false>);
        if (p.sh_objs) render_backward_kernel<false, true><<<grid, 256, sm, s>>>(RB_ARGS);
        else render_backward_kernel<false, false><<<grid, 256, sm, s>>>(RB_ARGS);
    }
#undef RB_ARGS
    VCR_LAUNCH_CHECK("render_backward_kernel");
    return 0;
}
