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
#ifndef RB_SPARSE_LANES
#define RB_SPARSE_LANES 2
#endif

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

// DEPTH2: the reference built with NUM_DEPTH2 = 2 (config.h:18): dL/dpixel has two more channels (sum of NDC depth and of its
// square, DEPTH2_OFFSET = 10) whose gradient reaches the Gaussians through the depth only — DETACH_WEIGHT = 1 (auxiliary.h:32)
// removes the opacity term (backward.cu:656-680).  The NUM_DIST = 1 channel has no gradient in the reference as written: the
// values it would use are declared a second time inside `if (NUM_DIST) { ... }` (backward.cu:494-499) and the outer ones stay 0.
template <bool HAS_DIRS, bool HAS_OBJ, bool DEPTH2 = false>
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
    float dL_dd1 = 0.f, dL_dd2 = 0.f;
    float3 dir = make_float3(0.f, 0.f, 1.f);
    if (inside && DEPTH2) {
        dL_dd1 = __ldg(dL_dpixels + VCR_OUT_CH * HW + pix_id);
        dL_dd2 = __ldg(dL_dpixels + (VCR_OUT_CH + 1) * HW + pix_id);
    }
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
        sg.id_inside = (id << 1) | ((inside_gs == nullptr || __ldg(inside_gs + id)) ? 1u : 0u);
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
                        float dL_dd_j = aTw * dL_dpixel_depth;
                        if (DEPTH2) {        // backward.cu:656-664
                            const float dn = (float)((100.0 * (double)c_d - 100.0 * 0.2) / ((100.0 - 0.2) * (double)c_d));
                            const float dmax_t_dd = (float)((100.0 * 0.2) / ((100.0 - 0.2) * (double)c_d * (double)c_d));
                            dL_dd_j += aT * dmax_t_dd * dL_dd1;
                            dL_dd_j += 2.0f * aT * dn * dmax_t_dd * dL_dd2;
                        }
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

// ------------------------------------------------------------------------------------------------------------------
// Two pixels per thread (the production kernel when every Gaussian is `inside`, i.e. the caller passed inside = NULL — what
// render() does, gaussian_renderer/__init__.py:103; an explicit inside mask takes the one-pixel kernel above, which keeps the
// reference's behaviour for mixed masks term by term).
//
// Same tile decomposition as render_forward2_kernel: 128 threads per 16x16 tile, a thread owns (x, y) and (x, y + 1), a
// warp a 16x4 strip.  What that buys here:
//   * warp-uniform work (staged-record reads, candidate bookkeeping) and the 16-component warp reduction are spent once
//     per two pixels — the two pixels' contributions to a Gaussian are added in registers before the reduction;
//   * the per-pixel arithmetic of the two pixels runs as packed FFMA2 / FMUL2 / FADD2;
//   * the running "accum_rec" recurrences are kept in their folded form S <- S + alpha (x - S) (the value the reference's
//     `last_alpha * last_x + (1 - last_alpha) * accum_rec` takes at the next contributor), which needs one state variable
//     per channel instead of three and makes a non-contributing pixel a no-op by setting its alpha to 0 — no per-variable
//     selects.  Algebraically identical to backward.cu:556-812 (with all Gaussians inside, accum_rec2 and accum_alpha_rec
//     are the same sequence); rounding differs at the 1e-7 level, far inside the 1e-3 gradient gate.
// ------------------------------------------------------------------------------------------------------------------
#define RB2_THREADS 128
#define RB2_BATCH 128
#ifndef VCR_RING_B
#define VCR_RING_B 1          // 1: staging buffer as a ring of 32-entry chunks with ready flags, no block barrier (see render_forward.cu)
#endif
#ifndef RB2_CHUNKS
#define RB2_CHUNKS 8
#endif
#define RB2_SLOTS (VCR_RING_B ? 32 * RB2_CHUNKS : RB2_BATCH)
#ifndef VCR_RING_B_AHEAD
#define VCR_RING_B_AHEAD 4
#endif
#ifndef VCR_RB2_MINBLOCKS
#define VCR_RB2_MINBLOCKS 5
#endif

template <bool HAS_DIRS>
struct Rb2Stage {
    StagedGeoB geo[RB2_SLOTS];
    float4 box[RB2_SLOTS];
    float2 ex[RB2_SLOTS];
#if VCR_QUAD
    float exc[RB2_SLOTS];             // 1 / conic.z for the column tests
#endif
    float4 nrm[RB2_SLOTS];            // normal.xyz, semantics[0]
    float4 col[RB2_SLOTS];            // rgb, semantics[1]
    float4 xv[HAS_DIRS ? RB2_SLOTS : 1];
};

__device__ __forceinline__ float2 vcr_sub2(const float2 a, const float2 b) { return vcr_add2(a, make_float2(-b.x, -b.y)); }

template <bool HAS_DIRS, bool HAS_OBJ>
__global__ void __launch_bounds__(RB2_THREADS, VCR_RB2_MINBLOCKS) render_backward2_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, int gx, const uint32_t* __restrict__ point_list, int W, int H, const float* __restrict__ bg,
    const float2* __restrict__ means2D, const float* __restrict__ xyz_view, const float4* __restrict__ conic_opacity,
    const float* __restrict__ colors, const float* __restrict__ depths, const float* __restrict__ normals,
    const float* __restrict__ objects, const float* __restrict__ dirs, const float* __restrict__ alpha_img,
    const float* __restrict__ depth_img, const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpixels, float* __restrict__ acc, const unsigned int* __restrict__ abort_flag) {
    __shared__ Rb2Stage<HAS_DIRS> s_stage[VCR_RING_B ? 1 : 2];
    __shared__ uint32_t s_block_last;
    if (__ldg(abort_flag)) return;      // the sort of this view failed (binning.cu): the lists are garbage, accumulate nothing

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = __ldg(tile_order + blockIdx.x);      // blocks take tiles longest-list first
    const uint32_t tile_x = tile % (uint32_t)gx, tile_y = tile / (uint32_t)gx;
#if VCR_QUAD
    // warp w owns the 8x8 quadrant (w & 1, w >> 1) of the tile; lane l the pixels (l & 7, 2 (l >> 3)) and the one below
    const uint32_t strip_c = tile_x * VCR_TILE_X + 8 * (warp & 1), strip_r = tile_y * VCR_TILE_Y + 8 * (warp >> 1);
    const uint32_t px = strip_c + (lane & 7);
    const uint32_t py0 = strip_r + 2 * (lane >> 3), py1 = py0 + 1;
#else
    const uint32_t px = tile_x * VCR_TILE_X + (lane & 15);
    const uint32_t py0 = tile_y * VCR_TILE_Y + 4 * warp + 2 * (lane >> 4), py1 = py0 + 1;
#endif
    const bool in0 = px < (uint32_t)W && py0 < (uint32_t)H, in1 = px < (uint32_t)W && py1 < (uint32_t)H;
    const size_t HW = (size_t)H * W;
    const size_t pix0 = (size_t)W * py0 + px, pix1 = pix0 + (size_t)W;
    const float pfx = (float)px;
    const float2 npfy = make_float2(-(float)py0, -(float)py1);
#if VCR_QUAD
    const float strip_x0 = (float)strip_c, strip_x1 = strip_x0 + 7.0f, strip_y0 = (float)strip_r, strip_y1 = strip_y0 + 7.0f;
#else
    const float strip_x0 = (float)(tile_x * VCR_TILE_X), strip_x1 = strip_x0 + (float)(VCR_TILE_X - 1);
    const float strip_y0 = (float)(tile_y * VCR_TILE_Y + 4 * warp), strip_y1 = strip_y0 + 3.0f;
#endif

    const uint2 range = ranges[tile];
    const uint32_t lastc0 = in0 ? n_contrib[pix0] : 0u, lastc1 = in1 ? n_contrib[pix1] : 0u;
    // Entries behind the last contributor of every pixel of the warp / block can never pass `contributor <
    // last_contributor` (backward.cu:589-591): drop them without staging or evaluating them.
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, max(lastc0, lastc1));
    if (tid == 0) s_block_last = 0u;
    __syncthreads();
    if (lane == 0 && warp_last) atomicMax(&s_block_last, warp_last);
    __syncthreads();
    const uint32_t block_last = s_block_last;          // <= range.y - range.x
    const uint32_t list_end = range.x + block_last;    // effective list: [range.x, list_end)
    const int rounds = (int)((block_last + RB2_BATCH - 1) / RB2_BATCH);
    int toDo = (int)block_last;

    auto ld2 = [&](const float* base, float dflt) {
        return make_float2(in0 ? __ldg(base + pix0) : dflt, in1 ? __ldg(base + pix1) : dflt);
    };
    const float2 T_final = ld2(final_Ts, 0.f);
    float2 T = T_final;
    const float2 wimg = ld2(alpha_img, 1.f);
    const float2 inv_weight = make_float2(1.0f / wimg.x, 1.0f / wimg.y);
    const float2 depth_final = ld2(depth_img, 1.f);
    const float2 gC0 = ld2(dL_dpixels, 0.f), gC1 = ld2(dL_dpixels + HW, 0.f), gC2 = ld2(dL_dpixels + 2 * HW, 0.f);
    const float2 gD = ld2(dL_dpixels + VCR_DEPTH_OFFSET * HW, 0.f);
    const float2 gN0 = ld2(dL_dpixels + (VCR_NORMAL_OFFSET + 0) * HW, 0.f), gN1 = ld2(dL_dpixels + (VCR_NORMAL_OFFSET + 1) * HW, 0.f),
                 gN2 = ld2(dL_dpixels + (VCR_NORMAL_OFFSET + 2) * HW, 0.f);
    float2 gO0 = make_float2(0.f, 0.f), gO1 = gO0;
    if (HAS_OBJ) { gO0 = ld2(dL_dpixels + (VCR_OBJECT_OFFSET + 0) * HW, 0.f); gO1 = ld2(dL_dpixels + (VCR_OBJECT_OFFSET + 1) * HW, 0.f); }
    float2 gA = ld2(dL_dpixels + VCR_ALPHA_OFFSET * HW, 0.f);             // grows with the semantics terms (backward.cu:766)
    float2 dirx = make_float2(0.f, 0.f), diry = dirx, dirz = make_float2(1.f, 1.f);
    if (HAS_DIRS) { dirx = ld2(dirs, 0.f); diry = ld2(dirs + HW, 0.f); dirz = ld2(dirs + 2 * HW, 1.f); }
    const float bgr = __ldg(bg), bgg = __ldg(bg + 1), bgb = __ldg(bg + 2);
    const float2 bg_dot = make_float2(bgr * gC0.x + bgg * gC1.x + bgb * gC2.x, bgr * gC0.y + bgg * gC1.y + bgb * gC2.y);
    // per-pixel constants folded once: (inv_weight * dL/dpixel) for the depth and normal channels
    const float2 kD = vcr_mul2(inv_weight, gD);
    const float2 kN0 = vcr_mul2(inv_weight, gN0), kN1 = vcr_mul2(inv_weight, gN1), kN2 = vcr_mul2(inv_weight, gN2);
    const float2 nTf_bg = make_float2(-T_final.x * bg_dot.x, -T_final.y * bg_dot.y);

    // folded recurrences (see the header): colour, depth, the constant 1, normal, semantics
    const float2 Z = make_float2(0.f, 0.f);
    float2 Sc0 = Z, Sc1 = Z, Sc2 = Z, Sd = Z, S1 = Z, Sn0 = Z, Sn1 = Z, Sn2 = Z, So0 = Z, So1 = Z;

    auto stage_one = [&](const uint32_t progress, Rb2Stage<HAS_DIRS>& st, const int tid) {    // entry `progress` of the reversed list into slot `tid`
        if (range.x + progress >= list_end) return;
        const uint32_t id = __ldg(point_list + (list_end - progress - 1));
        const float2 xy = __ldg(means2D + id);
        const float4 co = __ldg(conic_opacity + id);
        const float3 n = make_float3(__ldg(normals + 3 * (size_t)id), __ldg(normals + 3 * (size_t)id + 1),
                                     __ldg(normals + 3 * (size_t)id + 2));
        StagedGeoB sg;
        sg.x = xy.x; sg.y = xy.y; sg.ca = co.x; sg.cb = co.y; sg.cc = co.z; sg.op = co.w;
        st.box[tid] = vcr_alpha_bbox(xy, co);
        st.ex[tid] = vcr_alpha_rowtest(co);
#if VCR_QUAD
        st.exc[tid] = 1.0f / co.z;      // only read where ex says the conic is sound
#endif
        if (HAS_DIRS) {
            const float3 xv = make_float3(__ldg(xyz_view + 3 * (size_t)id), __ldg(xyz_view + 3 * (size_t)id + 1),
                                          __ldg(xyz_view + 3 * (size_t)id + 2));
            sg.nx = n.x * xv.x + n.y * xv.y + n.z * xv.z;
            st.xv[tid] = make_float4(xv.x, xv.y, xv.z, 0.f);
        } else {
            sg.nx = __ldg(depths + id);
        }
        sg.id_inside = id;
        st.geo[tid] = sg;
        float o0 = 0.f, o1 = 0.f;
        if (HAS_OBJ) { o0 = __ldg(objects + 2 * (size_t)id); o1 = __ldg(objects + 2 * (size_t)id + 1); }
        st.nrm[tid] = make_float4(n.x, n.y, n.z, o0);
        st.col[tid] = make_float4(__ldg(colors + 3 * (size_t)id), __ldg(colors + 3 * (size_t)id + 1),
                                  __ldg(colors + 3 * (size_t)id + 2), o1);
    };

    // One chunk of 32 staged entries, slots slot0 .. slot0+31 of `st`, the first n valid; the entry in slot j is contributor
    // cbase - j of the tile's list.  Footprint culling on the lanes, then the surviving candidates one at a time.
    auto consume = [&](const Rb2Stage<HAS_DIRS>& st, const int slot0, const int n, const int cbase) {
        const int jj = slot0 + lane;
        const float4 bx = st.box[jj];
        bool hit = lane < n && (uint32_t)(cbase - jj) < warp_last && bx.x <= strip_x1 && bx.y >= strip_x0 &&
                   bx.z <= strip_y1 && bx.w >= strip_y0;
        if (hit) {
            const float4 gq = *reinterpret_cast<const float4*>(&st.geo[jj]);        // x, y, ca, cb
            #if VCR_QUAD
            hit = vcr_rect_may_contribute(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], st.exc[jj], strip_x0, strip_x1, strip_y0, strip_y1);
#else
            hit = vcr_strip_may_contribute_rows<4>(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], strip_x0, strip_x1, strip_y0);
#endif
        }
        uint32_t m = __ballot_sync(0xffffffffu, hit);
        // One candidate in two halves, like the forward: eval() is everything that does not depend on the running
        // transmittance and the recurrences (exponent, alpha, depth along the ray, the tests) and ends in the warp's vote;
        // process() replays the pixel's state and accumulates.  (Issuing the evals of two candidates together, which pays in
        // the forward, costs registers here: 877 us against 836.)
        struct Ev { float2 Gr, ipd; bool c0, c1; uint32_t cmask; };      // the rest is cheap to redo in process()
        auto eval = [&](const int j) -> Ev {
            Ev e;
            e.cmask = 0u; e.c0 = false; e.c1 = false;
            e.Gr = e.ipd = make_float2(0.f, 0.f);
            const uint32_t contributor = (uint32_t)(cbase - j);
            const float4 g0 = *reinterpret_cast<const float4*>(&st.geo[j]);        // x, y, ca, cb
            const float4 g1 = *reinterpret_cast<const float4*>(&st.geo[j].cc);     // cc, op, nx, id
            const float dx = g0.x - pfx;
            const float2 dy = vcr_add2(vcr_bcast2(g0.y), npfy);
            const float adx = __fmul_rn(g0.z, dx), nbdx = -__fmul_rn(g0.w, dx);
            const float2 q = vcr_fma2(vcr_bcast2(adx), vcr_bcast2(dx), vcr_mul2(vcr_mul2(vcr_bcast2(g1.x), dy), dy));
            const float2 power = vcr_fma2(q, vcr_bcast2(-0.5f), vcr_mul2(vcr_bcast2(nbdx), dy));
            bool c0 = contributor < lastc0 && !(power.x > 0.0f);
            bool c1 = contributor < lastc1 && !(power.y > 0.0f);
            if (!__any_sync(0xffffffffu, c0 || c1)) return e;
            // Fast intrinsics from here on: the backward only has to meet 1e-3 relative L2 (the forward, which must
            // match the reference bit for bit, keeps expf and IEEE division).
            e.Gr = make_float2(__expf(power.x), __expf(power.y));
            const float2 ar = make_float2(fminf(0.99f, g1.y * e.Gr.x), fminf(0.99f, g1.y * e.Gr.y));
            c0 = c0 && !(ar.x < 1.0f / 255.0f);
            c1 = c1 && !(ar.y < 1.0f / 255.0f);
            const float4 nrm = st.nrm[j];
            float2 pd = make_float2(1.f, 1.f), cdr; e.ipd = pd;
            if (HAS_DIRS) {
            pd = vcr_fma2(vcr_bcast2(nrm.z), dirz, vcr_fma2(vcr_bcast2(nrm.x), dirx, vcr_mul2(vcr_bcast2(nrm.y), diry)));
            e.ipd = make_float2(__fdividef(1.0f, pd.x), __fdividef(1.0f, pd.y));
            cdr = vcr_mul2(vcr_mul2(vcr_bcast2(g1.z), e.ipd), dirz);
            } else {
            cdr = vcr_bcast2(g1.z);
            }
            c0 = c0 && !(cdr.x < 0.2f);  // == (double)c_d < 0.2, backward.cu:616
            c1 = c1 && !(cdr.y < 0.2f);
            e.c0 = c0; e.c1 = c1;
            e.cmask = __ballot_sync(0xffffffffu, c0 || c1);
            return e;
        };
        auto process = [&](const int j, const Ev& e) {
            const float4 g0 = *reinterpret_cast<const float4*>(&st.geo[j]);        // x, y, ca, cb
            const float4 g1 = *reinterpret_cast<const float4*>(&st.geo[j].cc);     // cc, op, nx, id
            const float4 nrm = st.nrm[j];
            const float dx = g0.x - pfx;
            const float2 dy = vcr_add2(vcr_bcast2(g0.y), npfy);
            const bool c0 = e.c0, c1 = e.c1;
            const uint32_t cmask = e.cmask;
            const float2 Gr = e.Gr;
            const float2 ar = make_float2(fminf(0.99f, g1.y * Gr.x), fminf(0.99f, g1.y * Gr.y));
            float2 ipd = e.ipd, pd = make_float2(1.f, 1.f), cdr = vcr_bcast2(g1.z);
            if (HAS_DIRS) {
                pd = vcr_fma2(vcr_bcast2(nrm.z), dirz, vcr_fma2(vcr_bcast2(nrm.x), dirx, vcr_mul2(vcr_bcast2(nrm.y), diry)));
                cdr = vcr_mul2(vcr_mul2(vcr_bcast2(g1.z), ipd), dirz);
            }

            // a pixel that does not take this Gaussian: alpha = G = 0 (and finite stand-ins where it may hold inf / NaN)
            const float2 alpha = make_float2(c0 ? ar.x : 0.f, c1 ? ar.y : 0.f);
            const float2 G = make_float2(c0 ? Gr.x : 0.f, c1 ? Gr.y : 0.f);
            const float2 c_d = make_float2(c0 ? cdr.x : 0.f, c1 ? cdr.y : 0.f);
            if (HAS_DIRS) ipd = make_float2(c0 ? ipd.x : 0.f, c1 ? ipd.y : 0.f);
            const float2 one = make_float2(1.f, 1.f);
            const float2 oma = vcr_sub2(one, alpha);
            const float2 inv1ma = make_float2(__fdividef(1.0f, oma.x), __fdividef(1.0f, oma.y));
            T = vcr_mul2(T, inv1ma);
            const float2 aT = vcr_mul2(alpha, T);
            const float4 col = st.col[j];
            float2 v[16];
            // colour (backward.cu:632-641): dL_dopa += (c - accum_rec) * dL_dpixel;  dL_dcolor += alpha T dL_dpixel
            float2 d0 = vcr_sub2(vcr_bcast2(col.x), Sc0), d1 = vcr_sub2(vcr_bcast2(col.y), Sc1), d2 = vcr_sub2(vcr_bcast2(col.z), Sc2);
            float2 dopa = vcr_fma2(d2, gC2, vcr_fma2(d1, gC1, vcr_mul2(d0, gC0)));
            Sc0 = vcr_fma2(alpha, d0, Sc0); Sc1 = vcr_fma2(alpha, d1, Sc1); Sc2 = vcr_fma2(alpha, d2, Sc2);
            v[ACC_COLOR_R] = vcr_mul2(aT, gC0); v[ACC_COLOR_R + 1] = vcr_mul2(aT, gC1); v[ACC_COLOR_R + 2] = vcr_mul2(aT, gC2);
            // depth and normal (backward.cu:643-754); om1 = 1 - accum_rec2 = 1 - accum_alpha_rec
            const float2 om1 = vcr_sub2(one, S1);
            {
            const float2 dd = vcr_sub2(c_d, Sd);
            const float2 t = vcr_fma2(make_float2(-depth_final.x, -depth_final.y), om1, dd);
            dopa = vcr_fma2(t, kD, dopa);
            Sd = vcr_fma2(alpha, dd, Sd);
            }
            const float2 dL_dd_j = vcr_mul2(aT, kD);          // alpha T / weight * dL/ddepth, with kD = dL/ddepth / weight folded per pixel
            float2 dn0 = Z, dn1 = Z, dn2 = Z;
            if (HAS_DIRS) {  // backward.cu:722-728
            const float4 xv = st.xv[j];
            const float2 sv = vcr_mul2(vcr_mul2(dL_dd_j, dirz), ipd);      // dL/dxyz_view = normal * sum(sv)
            const float2 s2 = vcr_mul2(sv, ipd);
            const float2 nnx = vcr_bcast2(-g1.z);
            dn0 = vcr_mul2(s2, vcr_fma2(pd, vcr_bcast2(xv.x), vcr_mul2(dirx, nnx)));
            dn1 = vcr_mul2(s2, vcr_fma2(pd, vcr_bcast2(xv.y), vcr_mul2(diry, nnx)));
            dn2 = vcr_mul2(s2, vcr_fma2(pd, vcr_bcast2(xv.z), vcr_mul2(dirz, nnx)));
            v[ACC_SV_OR_DEPTH] = sv;
            } else {
            v[ACC_SV_OR_DEPTH] = dL_dd_j;
            }
            {
            // reference quirk kept: the "final normal" of the normalisation term is this Gaussian's own n (backward.cu:741-745)
            const float2 n0 = vcr_bcast2(nrm.x), n1 = vcr_bcast2(nrm.y), n2 = vcr_bcast2(nrm.z);
            const float2 e0 = vcr_sub2(n0, Sn0), e1 = vcr_sub2(n1, Sn1), e2 = vcr_sub2(n2, Sn2);
            const float2 nom1 = make_float2(-om1.x, -om1.y);
            dopa = vcr_fma2(vcr_fma2(n0, nom1, e0), kN0, dopa);
            dopa = vcr_fma2(vcr_fma2(n1, nom1, e1), kN1, dopa);
            dopa = vcr_fma2(vcr_fma2(n2, nom1, e2), kN2, dopa);
            Sn0 = vcr_fma2(alpha, e0, Sn0); Sn1 = vcr_fma2(alpha, e1, Sn1); Sn2 = vcr_fma2(alpha, e2, Sn2);
            v[ACC_NORMAL_X] = vcr_fma2(aT, kN0, dn0); v[ACC_NORMAL_X + 1] = vcr_fma2(aT, kN1, dn1); v[ACC_NORMAL_X + 2] = vcr_fma2(aT, kN2, dn2);
            }
            float2 ex_obj1 = Z;
            v[ACC_OBJ_0] = Z;
            if (HAS_OBJ) {
            // dL_dalpha += (o - accum_rec_obj) * dL_dpixel_obj persists across Gaussians (backward.cu:766); contributing pixels only
            const float2 f0 = vcr_sub2(vcr_bcast2(nrm.w), So0), f1 = vcr_sub2(vcr_bcast2(col.w), So1);
            const float2 add = vcr_fma2(f1, gO1, vcr_mul2(f0, gO0));
            gA = make_float2(c0 ? gA.x + add.x : gA.x, c1 ? gA.y + add.y : gA.y);
            So0 = vcr_fma2(alpha, f0, So0); So1 = vcr_fma2(alpha, f1, So1);
            v[ACC_OBJ_0] = vcr_mul2(aT, gO0); ex_obj1 = vcr_mul2(aT, gO1);
            }
            dopa = vcr_fma2(om1, gA, dopa);                                   // (1 - accum_alpha_rec) * dL_dalpha
            S1 = vcr_fma2(alpha, om1, S1);
            dopa = vcr_fma2(inv1ma, nTf_bg, vcr_mul2(dopa, T));               // * T, + (-T_final / (1 - alpha)) * bg . dL_dpixel
            // raw moments; the per-Gaussian factors (opacity, conic, 0.5 W) are applied once in preprocess_backward
            const float2 bb = vcr_mul2(G, dopa);
            const float2 bxx = vcr_mul2(bb, vcr_bcast2(dx)), byy = vcr_mul2(bb, dy);
            v[ACC_M0] = bb; v[ACC_MX] = bxx; v[ACC_MY] = byy;
            v[ACC_MXX] = vcr_mul2(bxx, vcr_bcast2(dx)); v[ACC_MXY] = vcr_mul2(bxx, dy); v[ACC_MYY] = vcr_mul2(byy, dy);
            const float2 ax = vcr_fma2(byy, vcr_bcast2(g0.w), vcr_mul2(bxx, vcr_bcast2(g0.z)));
            const float2 ay = vcr_fma2(bxx, vcr_bcast2(g0.w), vcr_mul2(byy, vcr_bcast2(g1.x)));
            v[ACC_ABS_X] = make_float2(fabsf(ax.x), fabsf(ax.y)); v[ACC_ABS_Y] = make_float2(fabsf(ay.x), fabsf(ay.y));

            // the thread's two pixels first, then the warp
            float s[16];
#pragma unroll
            for (int q_ = 0; q_ < 16; q_++) s[q_] = v[q_].x + v[q_].y;
            const float eo1 = ex_obj1.x + ex_obj1.y;
            float* rec = acc + (size_t)__float_as_uint(g1.w) * VCR_ACC_STRIDE;
            if (__popc(cmask) <= RB_SPARSE_LANES) {
            if (c0 || c1) {
#pragma unroll
                for (int q_ = 0; q_ < (HAS_OBJ ? 16 : 15); q_++) atomicAdd(rec + q_, s[q_]);
                if (HAS_OBJ) atomicAdd(rec + ACC_OBJ_1, eo1);
            }
            return;
            }
            const float red = warp_reduce16_transposed(s, lane);
            if (lane < (HAS_OBJ ? 16 : 15)) atomicAdd(rec + lane, red);
            if (HAS_OBJ) {
            const float o1 = vcr_warp_sum(eo1);
            if (lane == 16) atomicAdd(rec + ACC_OBJ_1, o1);
            }
        };
        while (m) {
            const int ja = slot0 + __ffs(m) - 1;
            m &= m - 1;
            const Ev ea = eval(ja);
            if (ea.cmask) process(ja, ea);
        }
    };

#if VCR_RING_B
    // Ring protocol of render_forward2_kernel: chunk c of the reversed list in slots 32 (c mod RB2_CHUNKS) .., staged by the
    // warp that claims it, published through s_ready, slots reused once every warp has left the chunk that was there.  A warp
    // starts at the first chunk that holds a contributor below its warp_last; until then it only stages.
    __shared__ uint32_t s_ready[RB2_CHUNKS];
    __shared__ uint32_t s_tail[RB2_THREADS / 32];
    __shared__ uint32_t s_next;
    const uint32_t nchunks = (block_last + 31u) >> 5;
    const uint32_t first_chunk = warp_last ? (block_last - warp_last) >> 5 : nchunks;      // chunks before it hold nothing for this warp
    if (tid < RB2_CHUNKS) s_ready[tid] = 0u;
    if (lane == 0) s_tail[warp] = first_chunk;
    if (tid == 0) s_next = 0u;
    __syncthreads();
    {
        Rb2Stage<HAS_DIRS>& st = s_stage[0];
        volatile uint32_t* v_ready = s_ready;
        volatile uint32_t* v_tail = s_tail;
        volatile uint32_t* v_next = &s_next;
        auto try_stage = [&](const uint32_t limit) -> bool {
            const uint32_t c = *v_next;
            if (c >= limit) return false;
            const uint32_t behind = min(min(v_tail[0], v_tail[1]), min(v_tail[2], v_tail[3]));
            if (c >= RB2_CHUNKS && c - RB2_CHUNKS >= behind) return false;
            uint32_t got = c + 1u;
            if (lane == 0) got = atomicCAS(&s_next, c, c + 1u);
            got = __shfl_sync(0xffffffffu, got, 0);
            if (got == c) {
                stage_one(32u * c + lane, st, (int)((32u * c + lane) % RB2_SLOTS));
                __threadfence_block();
                __syncwarp();
                if (lane == 0) v_ready[c % RB2_CHUNKS] = c + 1u;
            }
            return true;
        };
        bool broken = false;
        for (uint32_t my = first_chunk; my < nchunks && !broken; my++) {
            try_stage(min(nchunks, my + VCR_RING_B_AHEAD));
            uint32_t spins = 0;
            while (v_ready[my % RB2_CHUNKS] != my + 1u) {
                if (!try_stage(nchunks)) {
                    __nanosleep(VCR_RING_NAP);
                    // never reached unless the protocol is broken: leave through the error flag of the sort instead of hanging
                    if (++spins > (VCR_RING_SPIN_LIMIT)) { atomicExch(const_cast<unsigned int*>(abort_flag), 1u); broken = true; break; }
                }
            }
            if (broken) break;
            __threadfence_block();
            const int slot0 = (int)((32u * my) % RB2_SLOTS);
            consume(st, slot0, min(32, (int)(block_last - 32u * my)), (int)block_last - 1 - (int)(32u * my) + slot0);
            __threadfence_block();
            __syncwarp();
            if (lane == 0) v_tail[warp] = my + 1u;
        }
        if (lane == 0) v_tail[warp] = 0xffffffffu;
        for (uint32_t spins = 0; spins < (VCR_RING_SPIN_LIMIT); ) {        // finished (or nothing to do): keep staging for the others
            if (*v_next >= nchunks) break;
            if ((v_tail[0] & v_tail[1] & v_tail[2] & v_tail[3]) == 0xffffffffu) break;
            if (!try_stage(nchunks)) { __nanosleep(2 * VCR_RING_NAP); spins++; }
        }
    }
#else
    if (rounds > 0) stage_one((uint32_t)tid, s_stage[0], tid);
    __syncthreads();
    for (int i = 0; i < rounds; i++, toDo -= RB2_BATCH) {
        const int buf = i & 1;
        if (i + 1 < rounds) stage_one((uint32_t)(i + 1) * RB2_BATCH + tid, s_stage[buf ^ 1], tid);   // overlaps with the replay of batch i
        const Rb2Stage<HAS_DIRS>& st = s_stage[buf];
        const int cnt = min(RB2_BATCH, toDo);
        // entry j of this batch has list position (contributor index) block_last - 1 - (i*RB2_BATCH + j)
        const int first_pos = (int)block_last - 1 - i * RB2_BATCH;
#pragma unroll 1
        for (int k = 0; k < RB2_BATCH / 32; k++) {
            if (k * 32 >= cnt) break;
            consume(st, k * 32, cnt - k * 32, first_pos);
        }
        // one barrier per batch: publishes batch i+1 and frees buffer `buf` for batch i+2
        __syncthreads();
    }
#endif
}

int vcr_render_version();      // render_forward.cu

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
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<true>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<true>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<false>))));
        VCR_CUDA_TRY(cudaFuncSetAttribute(render_backward_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * sizeof(RbStage<false>))));
        attr_set_dev[dev_].store(true, std::memory_order_release);
    }
    if (p.inside == nullptr && vcr_render_version() == 2 && !(p.variant & 2)) {
        // every Gaussian inside: the two-pixels-per-thread kernel (static shared memory)
#define RB2_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, p.background, g.means2D, g.xyz_view, g.conic_opacity, col, g.depths, \
                 p.normals, p.sh_objs, p.dirs, p.alpha, p.depth, img.accum_alpha, img.n_contrib, p.dL_dpix, p.acc, abort_flag
        if (p.dirs) {
            if (p.sh_objs) render_backward2_kernel<true, true><<<grid, RB2_THREADS, 0, s>>>(RB2_ARGS);
            else render_backward2_kernel<true, false><<<grid, RB2_THREADS, 0, s>>>(RB2_ARGS);
        } else {
            if (p.sh_objs) render_backward2_kernel<false, true><<<grid, RB2_THREADS, 0, s>>>(RB2_ARGS);
            else render_backward2_kernel<false, false><<<grid, RB2_THREADS, 0, s>>>(RB2_ARGS);
        }
#undef RB2_ARGS
        VCR_LAUNCH_CHECK("render_backward2_kernel");
        return 0;
    }
    if (p.variant & 2) {
        if (p.dirs) {
            const size_t sm = 2 * sizeof(RbStage<true>);
            if (p.sh_objs) render_backward_kernel<true, true, true><<<grid, 256, sm, s>>>(RB_ARGS);
            else render_backward_kernel<true, false, true><<<grid, 256, sm, s>>>(RB_ARGS);
        } else {
            const size_t sm = 2 * sizeof(RbStage<false>);
            if (p.sh_objs) render_backward_kernel<false, true, true><<<grid, 256, sm, s>>>(RB_ARGS);
            else render_backward_kernel<false, false, true><<<grid, 256, sm, s>>>(RB_ARGS);
        }
        VCR_LAUNCH_CHECK("render_backward_kernel (depth2)");
        return 0;
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
