// render_forward.cu — per-tile front-to-back compositing of colour / depth / normal / alpha / semantics.
//
// Replaces renderCUDA<3,2> (reference DGR/cuda_rasterizer/forward.cu:265-478) and, as MODE 1 / 3, the
// pruning variants renderCUDA_count (:480-603) and renderCUDA_visi_acc (:735-835).
//
// One 256-thread block per 16x16 tile, one pixel per thread, FP32 on the CUDA cores (this is not a
// dense contraction; no tensor cores).  Differences to the reference, none of which changes a result:
//   * everything the inner loop touches is staged in shared memory per batch of 256 list entries (the
//     reference reloads normal / colour / semantics from global memory for every contributing pair);
//   * staging is double-buffered: batch i+1 is gathered while batch i is composited, one barrier per batch;
//   * per warp and batch, 8 ballots test the conservative footprint boxes of the 256 staged Gaussians
//     (vcr_alpha_bbox) against the warp's 16x2 pixel strip, and only candidates enter the per-pixel code;
//   * the near-plane test is done in FP32 (x <= 0.2 in double is the same predicate as x < 0.2f for a
//     float x — no F2F.F64/DSETP per pair), and n.x_view of the ray/plane intersection is computed once
//     per staged Gaussian instead of once per pixel.
// The per-pixel arithmetic keeps the reference's operation order, so images are bit-identical to the
// reference build on identical inputs.
#include "common.cuh"

#define RF_BATCH 256

struct __align__(16) StagedGeo {  // 32 bytes
    float x, y;        // pixel-space centre
    float ca, cb;      // conic xx, xy
    float cc, op;      // conic yy, opacity
    float nx;          // dot(normal, xyz_view)  (intersection mode)  |  view depth (plain mode)
    uint32_t id_inside;  // (id << 1) | inside
};

template <int MODE>
struct RfStage {                       // one staging buffer
    StagedGeo geo[RF_BATCH];
    float4 box[RF_BATCH];              // conservative {xlo, xhi, ylo, yhi} of the alpha >= 1/255 footprint
    float2 ex[RF_BATCH];               // exact strip test: { -(tau + slack), 1 / conic.x }  (vcr_alpha_rowtest)
    float4 nrm[MODE == 0 ? RF_BATCH : 1];                              // normal.xyz, semantics[0]
    float4 col[MODE != VCR_MODE_VISI_ACC ? RF_BATCH : 1];              // rgb, semantics[1]
};

// Instrumentation (bench.py `scene`): when a counter array has been registered with vcr_pair_counters() the forward runs
// its STATS instantiation, which adds up K_e (pixel x Gaussian pairs whose exponent is evaluated), K_c (pairs that
// contribute), sum(n_contrib) and the number of (warp, Gaussian) candidates.  The production instantiation has none of this.
static unsigned long long* g_pair_counters = nullptr;
extern "C" int vcr_pair_counters(unsigned long long* counters_device) { g_pair_counters = counters_device; return 0; }
// Block timeline (bench.py VCR_BENCH_TIMELINE): while a device array of 16 u64 per tile is registered, every block of the
// two-pixel forward writes [0] start, [1] end (globaltimer, ns), [2] tile << 32 | SM id, [3] list length and, in the
// instrumented instantiation, [4..7] the candidates each warp evaluated, [8..11] the cycles each warp waited at the per-batch barrier, [12] the block's cycles. One uniform test per block when switched off.
static unsigned long long* g_block_timeline = nullptr;
extern "C" int vcr_debug_block_timeline(unsigned long long* timeline_device) { g_block_timeline = timeline_device; return 0; }
__device__ __forceinline__ unsigned long long vcr_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// VAR: the reference's compile-time variants (DGR/cuda_rasterizer/config.h:17-18) as a template parameter — bit 0: NUM_DIST = 1
// (depth-distortion channel, forward.cu:400-412, 464-472), bit 1: NUM_DEPTH2 = 2 (sums of NDC depth and its square over the
// inside Gaussians, forward.cu:426-430, 473-476).  Both write from channel 10 on (DISTORTION_OFFSET == DEPTH2_OFFSET,
// auxiliary.h:25-26), so with both set the distortion value is overwritten, as in the reference.
__device__ __forceinline__ float vcr_depth_ndc(const float depth) {
    // (FAR_PLANE * depth - FAR_PLANE * NEAR_PLANE) / ((FAR_PLANE - NEAR_PLANE) * depth): the macros are double literals (auxiliary.h:30-31)
    return (float)((100.0 * (double)depth - 100.0 * 0.2) / ((100.0 - 0.2) * (double)depth));
}

template <int MODE, bool HAS_DIRS, bool HAS_OBJ, bool STATS = false, int VAR = 0>
__global__ void __launch_bounds__(256, VAR ? 3 : 4) render_forward_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, int gx, const uint32_t* __restrict__ point_list, int W, int H,
    const float2* __restrict__ means2D, const float* __restrict__ xyz_view, const float* __restrict__ features,
    const float* __restrict__ depths, const float* __restrict__ normals, const float* __restrict__ obj,
    const float* __restrict__ dirs, const float4* __restrict__ conic_opacity, const uint8_t* __restrict__ inside_gs,
    const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
    float* __restrict__ out_color, int* __restrict__ gaussian_count, float* __restrict__ important_score,
    const unsigned int* __restrict__ abort_flag, unsigned long long* __restrict__ pair_counters = nullptr) {
    __shared__ RfStage<MODE> s_stage[2];
    unsigned int st_ke = 0, st_kc = 0, st_cand = 0;
    if (__ldg(abort_flag)) return;      // the sort of this view failed (binning.cu): the lists are garbage, leave the outputs unwritten

    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t tile = __ldg(tile_order + blockIdx.x);      // blocks take tiles longest-list first
    const uint32_t tile_x = tile % (uint32_t)gx, tile_y = tile / (uint32_t)gx;
    const uint32_t px = tile_x * VCR_TILE_X + (tid & 15);
    const uint32_t py = tile_y * VCR_TILE_Y + (tid >> 4);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const size_t HW = (size_t)H * W;
    const size_t pix_id = (size_t)W * py + px;
    const float pfx = (float)px, pfy = (float)py;
    bool done = !inside;
    // the warp's pixel strip: the full tile width, two rows
    const float strip_x0 = (float)(tile_x * VCR_TILE_X), strip_x1 = strip_x0 + (float)(VCR_TILE_X - 1);
    const float strip_y0 = (float)(tile_y * VCR_TILE_Y + 2 * (tid >> 5)), strip_y1 = strip_y0 + 1.0f;

    float3 dir = make_float3(0.f, 0.f, 0.f);
    if (HAS_DIRS && inside) dir = make_float3(__ldg(dirs + pix_id), __ldg(dirs + HW + pix_id), __ldg(dirs + 2 * HW + pix_id));

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + RF_BATCH - 1) / RF_BATCH;

    // gather list entry (batch, tid) into staging buffer `buf`
    auto stage = [&](int batch, int buf) {
        const uint32_t pos = range.x + (uint32_t)batch * RF_BATCH + tid;
        if (pos >= range.y) return;
        RfStage<MODE>& st = s_stage[buf];
        const uint32_t id = __ldg(point_list + pos);
        const float2 xy = __ldg(means2D + id);
        const float4 co = __ldg(conic_opacity + id);
        StagedGeo sg;
        sg.x = xy.x; sg.y = xy.y; sg.ca = co.x; sg.cb = co.y; sg.cc = co.z; sg.op = co.w;
        st.box[tid] = vcr_alpha_bbox(xy, co);
        st.ex[tid] = vcr_alpha_rowtest(co);
        uint32_t ins = 1u;
        float o0 = 0.f, o1 = 0.f;
        if (MODE == 0) {
            const float3 n = make_float3(__ldg(normals + 3 * (size_t)id), __ldg(normals + 3 * (size_t)id + 1),
                                         __ldg(normals + 3 * (size_t)id + 2));
            if (HAS_DIRS) {
                const float3 xv = make_float3(__ldg(xyz_view + 3 * (size_t)id), __ldg(xyz_view + 3 * (size_t)id + 1),
                                              __ldg(xyz_view + 3 * (size_t)id + 2));
                sg.nx = n.x * xv.x + n.y * xv.y + n.z * xv.z;  // dot(n, x) of intersect(), auxiliary.h:262-271
            } else {
                sg.nx = __ldg(depths + id);
            }
            ins = (inside_gs == nullptr || __ldg(inside_gs + id)) ? 1u : 0u;      // NULL: every Gaussian inside
            if (HAS_OBJ) { o0 = __ldg(obj + 2 * (size_t)id); o1 = __ldg(obj + 2 * (size_t)id + 1); }
            st.nrm[tid] = make_float4(n.x, n.y, n.z, o0);
        } else {
            sg.nx = 0.f;
        }
        sg.id_inside = (id << 1) | ins;
        st.geo[tid] = sg;
        if (MODE != VCR_MODE_VISI_ACC)
            st.col[tid] = make_float4(__ldg(features + 3 * (size_t)id), __ldg(features + 3 * (size_t)id + 1),
                                      __ldg(features + 3 * (size_t)id + 2), o1);
    };

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float weight = 1e-10f;  // reference quirk: alpha output starts at 1e-10 (forward.cu:326)
    float D = 0.f, N[3] = {0.f, 0.f, 0.f}, O[2] = {0.f, 0.f};
    float dist1 = 0.f, dist2 = 0.f, distortion = 0.f;       // VAR only

    if (rounds > 0) stage(0, 0);
    __syncthreads();
    for (int i = 0; i < rounds; i++) {
        const int buf = i & 1;
        if (i + 1 < rounds) stage(i + 1, buf ^ 1);   // overlaps with compositing batch i
        const RfStage<MODE>& st = s_stage[buf];
        const int cnt = min(RF_BATCH, total - i * RF_BATCH);
        // Warp-level culling: each lane tests 8 staged Gaussians against the warp's 16x2 pixel strip; the ballots
        // form the candidate masks and only candidates enter the per-pixel loop (in list order).
        bool warp_done = __all_sync(0xffffffffu, done);
#pragma unroll 1
        for (int k = 0; k < RF_BATCH / 32 && !warp_done; k++) {
            if (k * 32 >= cnt) break;
            const int jj = k * 32 + lane;
            const float4 bx = st.box[jj];
            bool hit = jj < cnt && bx.x <= strip_x1 && bx.y >= strip_x0 && bx.z <= strip_y1 && bx.w >= strip_y0;
            if (hit) {
                const float4 gq = *reinterpret_cast<const float4*>(&st.geo[jj]);        // x, y, ca, cb
                hit = vcr_strip_may_contribute(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], strip_x0, strip_x1, strip_y0);
            }
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                // m is warp-uniform, so every lane passes here once per candidate: a safe place to vote
                if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
                const int j = k * 32 + __ffs(m) - 1;
                m &= m - 1;
                if (STATS && lane == 0) st_cand++;
                if (done) continue;
                if (STATS) st_ke++;
                const float4 g0 = *reinterpret_cast<const float4*>(&st.geo[j]);        // x, y, ca, cb
                const float4 g1 = *reinterpret_cast<const float4*>(&st.geo[j].cc);     // cc, op, nx, id_inside
                const float dx = g0.x - pfx, dy = g0.y - pfy;
                // power = -0.5f*(ca*dx*dx + cc*dy*dy) - cb*dx*dy, contracted exactly like the reference build
                const float q = __fmaf_rn(__fmul_rn(g0.z, dx), dx, __fmul_rn(__fmul_rn(g1.x, dy), dy));
                const float power = __fmaf_rn(q, -0.5f, -__fmul_rn(__fmul_rn(g0.w, dx), dy));
                if (power > 0.0f) continue;
                // the reference tests depth before alpha (forward.cu:382-391); both are side-effect-free `continue`s,
                // so alpha goes first here: most rejected pairs fail it, and it spares them the IEEE division
                const float alpha = fminf(0.99f, g1.y * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                float depth = 0.f;
                float4 nrm = make_float4(0.f, 0.f, 0.f, 0.f);
                if (MODE == 0) {
                    nrm = st.nrm[j];
                    if (HAS_DIRS) {
                        const float t = g1.z / (nrm.x * dir.x + nrm.y * dir.y + nrm.z * dir.z);
                        depth = t * dir.z;
                    } else {
                        depth = g1.z;
                    }
                    if (depth < 0.2f) continue;  // == (double)depth <= 0.2, forward.cu:382
                }
                const uint32_t idi = __float_as_uint(g1.w);
                if (MODE == VCR_MODE_VISI_ACC) {  // first contributor only (forward.cu:809-826)
                    atomicAdd(gaussian_count + (idi >> 1), 1);  // reference: racy ++ (forward.cu:822)
                    done = true;
                    continue;
                }
                const float test_T = T * (1 - alpha);
                if (test_T < 0.0001f) { done = true; continue; }
                const float4 col = st.col[j];
                if (MODE == VCR_MODE_COUNT) {
                    // the reference's plain ++ / += race between pixels (forward.cu:575-576); atomics give the exact totals
                    atomicAdd(gaussian_count + (idi >> 1), 1);
                    atomicAdd(important_score + (idi >> 1), g1.y);
                    C[0] += col.x * alpha * T; C[1] += col.y * alpha * T; C[2] += col.z * alpha * T;
                } else {
                    const float w = alpha * T;
                    if (VAR & 1) {          // distortion loss of 2DGS (forward.cu:400-412)
                        const float dn = vcr_depth_ndc(depth);
                        const float A = 1 - T;
                        const float error = dn * dn * A + dist2 - 2 * dn * dist1;
                        distortion += error * w;
                        dist1 += dn * w;
                        dist2 += dn * dn * w;
                    }
                    C[0] = __fmaf_rn(col.x, w, C[0]); C[1] = __fmaf_rn(col.y, w, C[1]); C[2] = __fmaf_rn(col.z, w, C[2]);
                    if (HAS_OBJ) { O[0] = __fmaf_rn(nrm.w, w, O[0]); O[1] = __fmaf_rn(col.w, w, O[1]); }
                    weight += w;
                    if (idi & 1u) {
                        D = __fmaf_rn(depth, w, D);
                        if (VAR & 2) { const float dn = vcr_depth_ndc(depth); dist1 += dn * w; dist2 += dn * dn * w; }
                        N[0] = __fmaf_rn(nrm.x, w, N[0]); N[1] = __fmaf_rn(nrm.y, w, N[1]); N[2] = __fmaf_rn(nrm.z, w, N[2]);
                    }
                }
                T = test_T;
                if (STATS) st_kc++;
                last_contributor = (uint32_t)(i * RF_BATCH + j + 1);
            }
        }
        // one barrier per batch: publishes batch i+1, frees buffer `buf` for batch i+2, and votes on completion
        if (__syncthreads_and(done)) break;
    }

    if (STATS) {
        const unsigned int ke = __reduce_add_sync(0xffffffffu, st_ke), kc = __reduce_add_sync(0xffffffffu, st_kc);
        const unsigned int nc = __reduce_add_sync(0xffffffffu, inside ? last_contributor : 0u);
        if (lane == 0) {
            atomicAdd(pair_counters + 0, (unsigned long long)ke); atomicAdd(pair_counters + 1, (unsigned long long)kc);
            atomicAdd(pair_counters + 2, (unsigned long long)nc); atomicAdd(pair_counters + 3, (unsigned long long)st_cand);
        }
    }
    if (inside && MODE != VCR_MODE_VISI_ACC) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = __fmaf_rn(T, __ldg(bg + ch), C[ch]);
        if (MODE == 0) {
            out_color[VCR_DEPTH_OFFSET * HW + pix_id] = D / weight;
            out_color[VCR_ALPHA_OFFSET * HW + pix_id] = weight;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) out_color[(VCR_NORMAL_OFFSET + ch) * HW + pix_id] = N[ch] / weight;
            // the reference leaves the semantic channels at their zero initialisation when there are none
            out_color[(VCR_OBJECT_OFFSET + 0) * HW + pix_id] = HAS_OBJ ? O[0] : 0.f;
            out_color[(VCR_OBJECT_OFFSET + 1) * HW + pix_id] = HAS_OBJ ? O[1] : 0.f;
            if (VAR & 1) out_color[VCR_OUT_CH * HW + pix_id] = (float)((double)distortion / ((double)((1 - T) * (1 - T)) + 1e-7));
            if (VAR & 2) { out_color[VCR_OUT_CH * HW + pix_id] = dist1; out_color[(VCR_OUT_CH + 1) * HW + pix_id] = dist2; }
            if (VAR == 3) out_color[(VCR_OUT_CH + 2) * HW + pix_id] = 0.f;     // the reference's 13th channel stays at its zero fill
        }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Two pixels per thread (the production kernel of f_count = 0).
//
// ncu of the one-pixel kernel above: instruction-issue bound (70 % issue-active, 3.6 % DRAM), 85 warp-instructions per
// (warp, Gaussian) candidate, of which only the FMA chains are per-pixel work — the shared-memory reads of the staged
// record, the candidate bookkeeping and the loop control are warp-uniform.  Here a 16x16 tile is drawn by 128 threads; a
// thread owns the vertically adjacent pixels (x, y) and (x, y + 1), a warp a 16x4 strip.  Warp-uniform work is spent once
// per two pixels, the number of (warp, Gaussian) candidates drops (a footprint that touches the upper 16x2 strip nearly
// always touches the lower one), and the paired FMA chains go through the packed FFMA2 / FMUL2 / FADD2 instructions of
// sm_100 (fma.rn.f32x2: one issue slot, two IEEE-rounded results; 1.5x the FMA throughput of scalar FFMA measured with
// tools/ubench_f32x2.cu).  Every lane of a packed operation is rounded exactly like the scalar operation it replaces and
// the per-pixel operation order is the reference's, so the images stay bit-identical to the reference build.
// A pixel that does not take a Gaussian adds w = 0: fma(c, 0, acc) == acc for every finite c, which is the one assumption
// (finite colours / normals / semantics) this formulation adds.
// ------------------------------------------------------------------------------------------------------------------
#define RF2_THREADS 128
#define RF2_BATCH 128
#ifndef VCR_RING
#define VCR_RING 1            // 1: the staging buffer is a ring of 32-entry chunks with per-chunk ready flags (no block barrier)
#endif
#ifndef RF2_CHUNKS
#define RF2_CHUNKS 8                      // ring length in chunks
#endif
#ifndef VCR_RING_AHEAD
#define VCR_RING_AHEAD 2                  // a warp stages a chunk before consuming if the ring is less than this far ahead of it
#endif
#define RF2_SLOTS (VCR_RING ? 32 * RF2_CHUNKS : RF2_BATCH)
#ifndef VCR_RF2_MINBLOCKS
#define VCR_RF2_MINBLOCKS 7
#endif
#ifndef VCR_RF2_WAY
#define VCR_RF2_WAY 2          // candidates evaluated together per loop iteration
#endif

struct Rf2Stage {
    StagedGeo geo[RF2_SLOTS];
    float4 box[RF2_SLOTS];
    float2 ex[RF2_SLOTS];
#if VCR_QUAD
    float exc[RF2_SLOTS];              // 1 / conic.z for the column tests
#endif
    float4 nrm[RF2_SLOTS];             // normal.xyz, semantics[0]
    float4 col[RF2_SLOTS];             // rgb, semantics[1]
};

template <bool HAS_DIRS, bool HAS_OBJ, bool STATS>
__global__ void __launch_bounds__(RF2_THREADS, VCR_RF2_MINBLOCKS) render_forward2_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, int gx, const uint32_t* __restrict__ point_list, int W, int H,
    const float2* __restrict__ means2D, const float* __restrict__ xyz_view, const float* __restrict__ features,
    const float* __restrict__ depths, const float* __restrict__ normals, const float* __restrict__ obj,
    const float* __restrict__ dirs, const float4* __restrict__ conic_opacity, const uint8_t* __restrict__ inside_gs,
    const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
    float* __restrict__ out_color, const unsigned int* __restrict__ abort_flag, unsigned long long* __restrict__ pair_counters,
    unsigned long long* __restrict__ timeline) {
    __shared__ Rf2Stage s_stage[VCR_RING ? 1 : 2];
    if (timeline && threadIdx.x == 0) timeline[16 * (size_t)blockIdx.x] = vcr_globaltimer();
    unsigned int st_ke = 0, st_kc = 0, st_cand = 0;
    long long st_wait = 0;             // cycles this warp spent at the per-batch barrier (instrumented runs)
    const long long st_t0 = STATS ? clock64() : 0;
    if (__ldg(abort_flag)) return;      // the sort of this view failed (binning.cu): the lists are garbage, leave the outputs unwritten

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t tile = __ldg(tile_order + blockIdx.x);      // blocks take tiles longest-list first
    const uint32_t tile_x = tile % (uint32_t)gx, tile_y = tile / (uint32_t)gx;
#if VCR_QUAD
    // warp w owns the 8x8 quadrant (w & 1, w >> 1) of the tile; lane l the pixels (l & 7, 2 (l >> 3)) and the one below
    const uint32_t strip_c = tile_x * VCR_TILE_X + 8 * (warp & 1), strip_r = tile_y * VCR_TILE_Y + 8 * (warp >> 1);
    const uint32_t px = strip_c + (lane & 7);
    const uint32_t py0 = strip_r + 2 * (lane >> 3), py1 = py0 + 1;
#else
    // lanes 0-15: rows 4w, 4w+1; lanes 16-31: rows 4w+2, 4w+3 — the warp's strip is 16 columns x 4 rows
    const uint32_t px = tile_x * VCR_TILE_X + (lane & 15);
    const uint32_t py0 = tile_y * VCR_TILE_Y + 4 * warp + 2 * (lane >> 4), py1 = py0 + 1;
#endif
    const bool in0 = px < (uint32_t)W && py0 < (uint32_t)H, in1 = px < (uint32_t)W && py1 < (uint32_t)H;
    const size_t HW = (size_t)H * W;
    const size_t pix0 = (size_t)W * py0 + px, pix1 = pix0 + (size_t)W;
    const float pfx = (float)px;
    const float2 npfy = make_float2(-(float)py0, -(float)py1);       // dy = g.y - pfy as g.y + (-pfy): the same rounding
    uint32_t alive = (in0 ? 1u : 0u) | (in1 ? 2u : 0u);               // bit k: pixel k still composites
#if VCR_QUAD
    const float strip_x0 = (float)strip_c, strip_x1 = strip_x0 + 7.0f, strip_y0 = (float)strip_r, strip_y1 = strip_y0 + 7.0f;
#else
    const float strip_x0 = (float)(tile_x * VCR_TILE_X), strip_x1 = strip_x0 + (float)(VCR_TILE_X - 1);
    const float strip_y0 = (float)(tile_y * VCR_TILE_Y + 4 * warp), strip_y1 = strip_y0 + 3.0f;
#endif

    float2 dirx = make_float2(0.f, 0.f), diry = dirx, dirz = dirx;
    if (HAS_DIRS) {
        if (in0) { dirx.x = __ldg(dirs + pix0); diry.x = __ldg(dirs + HW + pix0); dirz.x = __ldg(dirs + 2 * HW + pix0); }
        if (in1) { dirx.y = __ldg(dirs + pix1); diry.y = __ldg(dirs + HW + pix1); dirz.y = __ldg(dirs + 2 * HW + pix1); }
    }

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + RF2_BATCH - 1) / RF2_BATCH;

    auto stage_one = [&](const uint32_t pos, Rf2Stage& st, const int tid) {      // list entry `pos` into slot `tid`
        if (pos >= range.y) return;
        const uint32_t id = __ldg(point_list + pos);
        const float2 xy = __ldg(means2D + id);
        const float4 co = __ldg(conic_opacity + id);
        StagedGeo sg;
        sg.x = xy.x; sg.y = xy.y; sg.ca = co.x; sg.cb = co.y; sg.cc = co.z; sg.op = co.w;
        st.box[tid] = vcr_alpha_bbox(xy, co);
        st.ex[tid] = vcr_alpha_rowtest(co);
#if VCR_QUAD
        st.exc[tid] = 1.0f / co.z;      // only read where ex says the conic is sound
#endif
        float o0 = 0.f, o1 = 0.f;
        const float3 n = make_float3(__ldg(normals + 3 * (size_t)id), __ldg(normals + 3 * (size_t)id + 1), __ldg(normals + 3 * (size_t)id + 2));
        if (HAS_DIRS) {
            const float3 xv = make_float3(__ldg(xyz_view + 3 * (size_t)id), __ldg(xyz_view + 3 * (size_t)id + 1),
                                          __ldg(xyz_view + 3 * (size_t)id + 2));
            sg.nx = n.x * xv.x + n.y * xv.y + n.z * xv.z;  // dot(n, x) of intersect(), auxiliary.h:262-271
        } else {
            sg.nx = __ldg(depths + id);
        }
        const uint32_t ins = (inside_gs == nullptr || __ldg(inside_gs + id)) ? 1u : 0u;      // NULL: every Gaussian inside
        if (HAS_OBJ) { o0 = __ldg(obj + 2 * (size_t)id); o1 = __ldg(obj + 2 * (size_t)id + 1); }
        st.nrm[tid] = make_float4(n.x, n.y, n.z, o0);
        sg.id_inside = (id << 1) | ins;
        st.geo[tid] = sg;
        st.col[tid] = make_float4(__ldg(features + 3 * (size_t)id), __ldg(features + 3 * (size_t)id + 1),
                                  __ldg(features + 3 * (size_t)id + 2), o1);
    };

    float2 T = make_float2(1.0f, 1.0f);
    uint32_t last0 = 0, last1 = 0;
    float2 C0 = make_float2(0.f, 0.f), C1 = C0, C2 = C0;
    float2 weight = make_float2(1e-10f, 1e-10f);  // reference quirk: alpha output starts at 1e-10 (forward.cu:326)
    float2 D = C0, N0 = C0, N1 = C0, N2 = C0, O0 = C0, O1 = C0;

    // One candidate in two halves. eval: everything that does not depend on the running transmittance (two shared-memory
    // reads, the quadratic form, expf). commit: the tests against `alive` and T, and the accumulation, in list order.
    auto eval = [&](const Rf2Stage& st, const int j, uint32_t& pos) -> float2 {
        const float4 g0 = *reinterpret_cast<const float4*>(&st.geo[j]);        // x, y, ca, cb
        const float2 g1 = *reinterpret_cast<const float2*>(&st.geo[j].cc);     // cc, op
        const float dx = g0.x - pfx;
        const float2 dy = vcr_add2(vcr_bcast2(g0.y), npfy);
        // power = -0.5f*(ca*dx*dx + cc*dy*dy) - cb*dx*dy, contracted exactly like the reference build:
        //   q = fma(fmul(ca,dx), dx, fmul(fmul(cc,dy), dy));  power = fma(q, -0.5, -fmul(fmul(cb,dx), dy))
        const float adx = __fmul_rn(g0.z, dx), nbdx = -__fmul_rn(g0.w, dx);
        const float2 q = vcr_fma2(vcr_bcast2(adx), vcr_bcast2(dx), vcr_mul2(vcr_mul2(vcr_bcast2(g1.x), dy), dy));
        const float2 power = vcr_fma2(q, vcr_bcast2(-0.5f), vcr_mul2(vcr_bcast2(nbdx), dy));
        // the reference tests depth before alpha (forward.cu:382-391); both are side-effect-free `continue`s,
        // so alpha goes first here: most rejected pairs fail it, and it spares them the IEEE division
        // power > 0 rejects the pair (forward.cu:386): two bits of `pos` carry that to commit, the alpha chain stays branch-free
        pos = (power.x > 0.0f ? 1u : 0u) | (power.y > 0.0f ? 2u : 0u);
        const float2 e = vcr_expf2(power);                      // == expf on both lanes, exhaustively (tools/expf2_check.cu)
        return make_float2(fminf(0.99f, g1.y * e.x), fminf(0.99f, g1.y * e.y));
    };
    auto commit = [&](const Rf2Stage& st, const int list_base, const int j, const float2 al, const uint32_t pos) {
        const float a0 = al.x, a1 = al.y;
        if (STATS) st_ke += __popc(alive);
        bool ok0 = (alive & ~pos & 1u) && !(a0 < 1.0f / 255.0f);
        bool ok1 = (alive & ~pos & 2u) && !(a1 < 1.0f / 255.0f);
        if (!(ok0 || ok1)) return;
        const float2 g1 = *reinterpret_cast<const float2*>(&st.geo[j].nx);     // nx, id_inside
        const float4 nrm = st.nrm[j];
        float2 depth;
        if (HAS_DIRS) {
            // t = nx / dot(n, dir) with the dot product in the reference's order: n.y*d.y, then n.x*d.x, then n.z*d.z
            const float2 pd = vcr_fma2(vcr_bcast2(nrm.z), dirz, vcr_fma2(vcr_bcast2(nrm.x), dirx, vcr_mul2(vcr_bcast2(nrm.y), diry)));
            depth = vcr_mul2(make_float2(__fdiv_rn(g1.x, pd.x), __fdiv_rn(g1.x, pd.y)), dirz);
        } else {
            depth = vcr_bcast2(g1.x);
        }
        ok0 = ok0 && !(depth.x < 0.2f);  // == (double)depth <= 0.2, forward.cu:382
        ok1 = ok1 && !(depth.y < 0.2f);
        const float2 test_T = vcr_mul2(T, make_float2(1 - a0, 1 - a1));
        if (ok0 && test_T.x < 0.0001f) { alive &= ~1u; ok0 = false; }
        if (ok1 && test_T.y < 0.0001f) { alive &= ~2u; ok1 = false; }
        if (!(ok0 || ok1)) return;
        const float4 col = st.col[j];
        const uint32_t idi = __float_as_uint(g1.y);
        const float2 w = make_float2(ok0 ? __fmul_rn(a0, T.x) : 0.f, ok1 ? __fmul_rn(a1, T.y) : 0.f);
        C0 = vcr_fma2(vcr_bcast2(col.x), w, C0); C1 = vcr_fma2(vcr_bcast2(col.y), w, C1); C2 = vcr_fma2(vcr_bcast2(col.z), w, C2);
        if (HAS_OBJ) { O0 = vcr_fma2(vcr_bcast2(nrm.w), w, O0); O1 = vcr_fma2(vcr_bcast2(col.w), w, O1); }
        weight = vcr_add2(weight, w);
        if (idi & 1u) {
            D = vcr_fma2(make_float2(ok0 ? depth.x : 0.f, ok1 ? depth.y : 0.f), w, D);
            N0 = vcr_fma2(vcr_bcast2(nrm.x), w, N0); N1 = vcr_fma2(vcr_bcast2(nrm.y), w, N1); N2 = vcr_fma2(vcr_bcast2(nrm.z), w, N2);
        }
        const uint32_t here = (uint32_t)(list_base + j + 1);       // list position + 1 of the entry in slot j
        if (ok0) { T.x = test_T.x; last0 = here; }
        if (ok1) { T.y = test_T.y; last1 = here; }
        if (STATS) st_kc += (ok0 ? 1u : 0u) + (ok1 ? 1u : 0u);
    };

    // One chunk of 32 staged entries, slots slot0 .. slot0+31 of `st`, the first n of them valid; list_base + slot = list
    // position.  Footprint culling on the lanes, then the surviving candidates one pair at a time.  Returns true when every
    // pixel of the warp has finished.
    auto consume = [&](const Rf2Stage& st, const int slot0, const int n, const int list_base) -> bool {
        const int jj = slot0 + lane;
        const float4 bx = st.box[jj];
        bool hit = lane < n && bx.x <= strip_x1 && bx.y >= strip_x0 && bx.z <= strip_y1 && bx.w >= strip_y0;
        if (hit) {
            const float4 gq = *reinterpret_cast<const float4*>(&st.geo[jj]);        // x, y, ca, cb
#if VCR_QUAD
            hit = vcr_rect_may_contribute(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], st.exc[jj], strip_x0, strip_x1, strip_y0, strip_y1);
#else
            hit = vcr_strip_may_contribute_rows<4>(gq.x, gq.y, gq.z, gq.w, st.geo[jj].cc, st.ex[jj], strip_x0, strip_x1, strip_y0);
#endif
        }
        uint32_t m = __ballot_sync(0xffffffffu, hit);
        while (m) {
            // m is warp-uniform, so every lane passes here once per candidate group: a safe place to vote
            if (__all_sync(0xffffffffu, alive == 0u)) return true;
            // VCR_RF2_WAY candidates per iteration: their alpha chains are independent of T and of each other, so they are
            // issued together (the kernel is as long as its longest tile, whose four warps are latency-bound), then
            // committed one after the other, in list order.
            int js[VCR_RF2_WAY];
            float2 als[VCR_RF2_WAY];
            uint32_t pos[VCR_RF2_WAY];
            int nc = 0;
#pragma unroll
            for (int w = 0; w < VCR_RF2_WAY; w++) {
                if (m) { js[w] = slot0 + __ffs(m) - 1; m &= m - 1; nc = w + 1; }
                else js[w] = js[0];
            }
            if (STATS && lane == 0) st_cand += (unsigned int)nc;
            if (alive == 0u) continue;
#pragma unroll
            for (int w = 0; w < VCR_RF2_WAY; w++) als[w] = eval(st, js[w], pos[w]);
#pragma unroll
            for (int w = 0; w < VCR_RF2_WAY; w++)
                if (w < nc && alive != 0u) commit(st, list_base, js[w], als[w], pos[w]);
        }
        return false;
    };

#if VCR_RING
    // Ring of RF2_CHUNKS chunks of 32 entries, shared by the four warps, no block barrier: chunk c lives in slots
    // 32 (c mod RF2_CHUNKS) .., is staged by whichever warp claims it (s_next, compare-and-swap) and published through
    // s_ready[c mod RF2_CHUNKS] = c + 1; a slot may be overwritten once every warp has left the chunk that was there
    // (s_tail).  A warp that finds its next chunk missing stages chunks itself, so the warps that are ahead do the staging
    // and the warp with the most candidates — the one that decides when the block ends — finds its chunks ready.
    __shared__ uint32_t s_ready[RF2_CHUNKS];
    __shared__ uint32_t s_tail[RF2_THREADS / 32];    // chunks each warp has finished with; 0xffffffff once it has left the loop
    __shared__ uint32_t s_next;                      // first chunk nobody has claimed
    if (tid < RF2_CHUNKS) s_ready[tid] = 0u;
    if (tid < RF2_THREADS / 32) s_tail[tid] = 0u;
    if (tid == 0) s_next = 0u;
    __syncthreads();
    {
        Rf2Stage& st = s_stage[0];
        volatile uint32_t* v_ready = s_ready;
        volatile uint32_t* v_tail = s_tail;
        volatile uint32_t* v_next = &s_next;
        const uint32_t nchunks = ((uint32_t)total + 31u) >> 5;
        // stage the first unclaimed chunk if it lies below `limit` and its slots are free; false if there was nothing to do
        auto try_stage = [&](const uint32_t limit) -> bool {
            const uint32_t c = *v_next;
            if (c >= limit) return false;
            const uint32_t behind = min(min(v_tail[0], v_tail[1]), min(v_tail[2], v_tail[3]));
            if (c >= RF2_CHUNKS && c - RF2_CHUNKS >= behind) return false;
            uint32_t got = c + 1u;
            if (lane == 0) got = atomicCAS(&s_next, c, c + 1u);
            got = __shfl_sync(0xffffffffu, got, 0);
            if (got == c) {
                stage_one(range.x + 32u * c + lane, st, (int)((32u * c + lane) % RF2_SLOTS));
                __threadfence_block();
                __syncwarp();
                if (lane == 0) v_ready[c % RF2_CHUNKS] = c + 1u;
            }
            return true;
        };
        for (uint32_t my = 0; my < nchunks; my++) {
            if (__all_sync(0xffffffffu, alive == 0u)) break;
            try_stage(min(nchunks, my + VCR_RING_AHEAD));            // keep the ring ahead of the warps that lead
            const long long st_b0 = STATS ? clock64() : 0;
            uint32_t spins = 0;
            while (v_ready[my % RF2_CHUNKS] != my + 1u) {
                if (!try_stage(nchunks)) {
                    __nanosleep(VCR_RING_NAP);
                    // never reached unless the protocol is broken: leave through the error flag of the sort (reported by the
                    // next call) instead of hanging the device
                    if (++spins > (VCR_RING_SPIN_LIMIT)) { atomicExch(const_cast<unsigned int*>(abort_flag), 1u); break; }
                }
            }
            if (spins > (VCR_RING_SPIN_LIMIT)) break;
            __threadfence_block();
            if (STATS) st_wait += clock64() - st_b0;
            const bool done = consume(st, (int)((32u * my) % RF2_SLOTS), min(32, total - (int)(32u * my)), (int)(32u * my - (32u * my) % RF2_SLOTS));
            __threadfence_block();
            __syncwarp();
            if (done) break;
            if (lane == 0) v_tail[warp] = my + 1u;
        }
        if (lane == 0) v_tail[warp] = 0xffffffffu;
        // A warp that has finished keeps staging for the others (they would otherwise pay the load latency of every chunk
        // themselves — in the longest tiles one warp is often the only one left for most of the list).
        for (uint32_t spins = 0; spins < (VCR_RING_SPIN_LIMIT); ) {
            if (*v_next >= nchunks) break;                                                              // everything is staged
            if ((v_tail[0] & v_tail[1] & v_tail[2] & v_tail[3]) == 0xffffffffu) break;                  // nobody is left to read it
            if (!try_stage(nchunks)) { __nanosleep(2 * VCR_RING_NAP); spins++; }
        }
    }
#else
    if (rounds > 0) stage_one(range.x + tid, s_stage[0], tid);
    __syncthreads();
    for (int i = 0; i < rounds; i++) {
        const int buf = i & 1;
        if (i + 1 < rounds) stage_one(range.x + (uint32_t)(i + 1) * RF2_BATCH + tid, s_stage[buf ^ 1], tid);   // overlaps with compositing batch i
        const Rf2Stage& st = s_stage[buf];
        const int cnt = min(RF2_BATCH, total - i * RF2_BATCH);
        bool warp_done = __all_sync(0xffffffffu, alive == 0u);
#pragma unroll 1
        for (int k = 0; k < RF2_BATCH / 32 && !warp_done; k++) {
            if (k * 32 >= cnt) break;
            warp_done = consume(st, k * 32, cnt - k * 32, i * RF2_BATCH);
        }
        // one barrier per batch: publishes batch i+1, frees buffer `buf` for batch i+2, and votes on completion
        const long long st_b0 = STATS ? clock64() : 0;
        const bool all_done = __syncthreads_and(alive == 0u);
        if (STATS) st_wait += clock64() - st_b0;
        if (all_done) break;
    }
#endif

    if (STATS) {
        const unsigned int ke = __reduce_add_sync(0xffffffffu, st_ke), kc = __reduce_add_sync(0xffffffffu, st_kc);
        const unsigned int nc = __reduce_add_sync(0xffffffffu, (in0 ? last0 : 0u) + (in1 ? last1 : 0u));
        if (lane == 0) {
            atomicAdd(pair_counters + 0, (unsigned long long)ke); atomicAdd(pair_counters + 1, (unsigned long long)kc);
            atomicAdd(pair_counters + 2, (unsigned long long)nc); atomicAdd(pair_counters + 3, (unsigned long long)st_cand);
        }
        if (tid == 0) {         // [4] longest block in SM cycles, [5] sum over blocks: is the kernel bound by its longest tile?
            const unsigned long long dt = (unsigned long long)(clock64() - st_t0);
            atomicMax(pair_counters + 4, dt); atomicAdd(pair_counters + 5, dt);
        }
    }
    const float bg0 = __ldg(bg), bg1 = __ldg(bg + 1), bg2 = __ldg(bg + 2);
    auto write = [&](const size_t pix, const float Tv, const uint32_t last, const float c0, const float c1, const float c2, const float wt,
                     const float d, const float n0, const float n1, const float n2, const float o0, const float o1) {
        final_T[pix] = Tv;
        n_contrib[pix] = last;
        out_color[pix] = __fmaf_rn(Tv, bg0, c0);
        out_color[HW + pix] = __fmaf_rn(Tv, bg1, c1);
        out_color[2 * HW + pix] = __fmaf_rn(Tv, bg2, c2);
        out_color[VCR_DEPTH_OFFSET * HW + pix] = d / wt;
        out_color[VCR_ALPHA_OFFSET * HW + pix] = wt;
        out_color[(VCR_NORMAL_OFFSET + 0) * HW + pix] = n0 / wt;
        out_color[(VCR_NORMAL_OFFSET + 1) * HW + pix] = n1 / wt;
        out_color[(VCR_NORMAL_OFFSET + 2) * HW + pix] = n2 / wt;
        // the reference leaves the semantic channels at their zero initialisation when there are none
        out_color[(VCR_OBJECT_OFFSET + 0) * HW + pix] = HAS_OBJ ? o0 : 0.f;
        out_color[(VCR_OBJECT_OFFSET + 1) * HW + pix] = HAS_OBJ ? o1 : 0.f;
    };
    if (in0) write(pix0, T.x, last0, C0.x, C1.x, C2.x, weight.x, D.x, N0.x, N1.x, N2.x, O0.x, O1.x);
    if (in1) write(pix1, T.y, last1, C0.y, C1.y, C2.y, weight.y, D.y, N0.y, N1.y, N2.y, O0.y, O1.y);
    if (timeline) {
        __syncthreads();
        unsigned long long* rec = timeline + 16 * (size_t)blockIdx.x;
        if (tid == 0) {
            unsigned int smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            rec[1] = vcr_globaltimer();
            rec[2] = ((unsigned long long)tile << 32) | smid;
            rec[3] = (unsigned long long)(uint32_t)total;
        }
        if (STATS && lane == 0) {
            rec[4 + warp] = st_cand;                           // (warp, Gaussian) candidates after culling, per warp
            rec[8 + warp] = (unsigned long long)st_wait;       // cycles at the per-batch barrier, per warp
            if (warp == 0) rec[12] = (unsigned long long)(clock64() - st_t0);
        }
    }
}

// vcr_debug_render_version(1) selects the one-pixel-per-thread kernels (A/B measurements, tools/); 2 (default) = two pixels.
static int g_render_version = 2;
extern "C" int vcr_debug_render_version(int v) { if (v == 1 || v == 2) g_render_version = v; return g_render_version; }
int vcr_render_version() { return g_render_version; }

int vcr_launch_render_forward(const FwdParams& p, GeomState& g, BinningState& b, ImageState& img, cudaStream_t s) {
    const dim3 grid(p.gx * p.gy, 1, 1);
    const float* feat = p.colors_precomp ? p.colors_precomp : g.rgb;
    const unsigned int* abort_flag = vcr_radix_abort_flag();
    if (!abort_flag) { vcr_set_error("cannot resolve the sort status flag on this device"); return 1; }
    VcrProfScope prof(VCR_K_RENDER_FWD, s);
    if (p.variant) {          // distortion / depth-variance channels: the one-pixel kernel carries the variants
#define RFV_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, g.means2D, g.xyz_view, feat, g.depths, p.normals, p.sh_objs, p.dirs, \
                 g.conic_opacity, p.inside, p.background, img.accum_alpha, img.n_contrib, p.out_color, nullptr, nullptr, abort_flag
#define RFV_LAUNCH(V_)                                                                                              \
        do {                                                                                                        \
            if (p.dirs) { if (p.sh_objs) render_forward_kernel<0, true, true, false, V_><<<grid, 256, 0, s>>>(RFV_ARGS);   \
                          else render_forward_kernel<0, true, false, false, V_><<<grid, 256, 0, s>>>(RFV_ARGS); }          \
            else { if (p.sh_objs) render_forward_kernel<0, false, true, false, V_><<<grid, 256, 0, s>>>(RFV_ARGS);         \
                   else render_forward_kernel<0, false, false, false, V_><<<grid, 256, 0, s>>>(RFV_ARGS); }                \
        } while (0)
        if (p.variant == 1) RFV_LAUNCH(1); else if (p.variant == 2) RFV_LAUNCH(2); else RFV_LAUNCH(3);
#undef RFV_LAUNCH
#undef RFV_ARGS
        VCR_LAUNCH_CHECK("render_forward_kernel (variant)");
        return 0;
    }
    if (g_render_version == 2) {
#define RF2_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, g.means2D, g.xyz_view, feat, g.depths, p.normals, p.sh_objs, p.dirs, \
                 g.conic_opacity, p.inside, p.background, img.accum_alpha, img.n_contrib, p.out_color, abort_flag, g_pair_counters, g_block_timeline
#define RF2_LAUNCH(D_, O_)                                                                                   \
        do {                                                                                                 \
            if (g_pair_counters) render_forward2_kernel<D_, O_, true><<<grid, RF2_THREADS, 0, s>>>(RF2_ARGS); \
            else render_forward2_kernel<D_, O_, false><<<grid, RF2_THREADS, 0, s>>>(RF2_ARGS);                \
        } while (0)
        if (p.dirs) { if (p.sh_objs) RF2_LAUNCH(true, true); else RF2_LAUNCH(true, false); }
        else { if (p.sh_objs) RF2_LAUNCH(false, true); else RF2_LAUNCH(false, false); }
#undef RF2_LAUNCH
#undef RF2_ARGS
        VCR_LAUNCH_CHECK("render_forward2_kernel");
        return 0;
    }
#define RF_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, g.means2D, g.xyz_view, feat, g.depths, p.normals, p.sh_objs, p.dirs, \
                g.conic_opacity, p.inside, p.background, img.accum_alpha, img.n_contrib, p.out_color, nullptr, nullptr, abort_flag
    if (g_pair_counters) {          // instrumented run (never inside a timed region)
        if (p.dirs) {
            if (p.sh_objs) render_forward_kernel<0, true, true, true><<<grid, 256, 0, s>>>(RF_ARGS, g_pair_counters);
            else render_forward_kernel<0, true, false, true><<<grid, 256, 0, s>>>(RF_ARGS, g_pair_counters);
        } else {
            if (p.sh_objs) render_forward_kernel<0, false, true, true><<<grid, 256, 0, s>>>(RF_ARGS, g_pair_counters);
            else render_forward_kernel<0, false, false, true><<<grid, 256, 0, s>>>(RF_ARGS, g_pair_counters);
        }
    } else if (p.dirs) {
        if (p.sh_objs) render_forward_kernel<0, true, true><<<grid, 256, 0, s>>>(RF_ARGS);
        else render_forward_kernel<0, true, false><<<grid, 256, 0, s>>>(RF_ARGS);
    } else {
        if (p.sh_objs) render_forward_kernel<0, false, true><<<grid, 256, 0, s>>>(RF_ARGS);
        else render_forward_kernel<0, false, false><<<grid, 256, 0, s>>>(RF_ARGS);
    }
#undef RF_ARGS
    VCR_LAUNCH_CHECK("render_forward_kernel");
    return 0;
}

int vcr_launch_render_aux(int mode, const FwdParams& p, GeomState& g, BinningState& b, ImageState& img,
                          int* gaussians_count, float* important_score, cudaStream_t s) {
    const dim3 grid(p.gx * p.gy, 1, 1);
    const float* feat = p.colors_precomp ? p.colors_precomp : g.rgb;
    const unsigned int* abort_flag = vcr_radix_abort_flag();
    if (!abort_flag) { vcr_set_error("cannot resolve the sort status flag on this device"); return 1; }
    VcrProfScope prof(VCR_K_RENDER_AUX, s);
#define RA_ARGS img.ranges, img.tile_order, p.gx, b.point_list, p.W, p.H, g.means2D, g.xyz_view, feat, g.depths, nullptr, nullptr, nullptr, \
                g.conic_opacity, nullptr, p.background, img.accum_alpha, img.n_contrib, p.out_color, gaussians_count, important_score, abort_flag
    if (mode == VCR_MODE_COUNT) render_forward_kernel<VCR_MODE_COUNT, false, false><<<grid, 256, 0, s>>>(RA_ARGS);
    else render_forward_kernel<VCR_MODE_VISI_ACC, false, false><<<grid, 256, 0, s>>>(RA_ARGS);
#undef RA_ARGS
    VCR_LAUNCH_CHECK("render_aux_kernel");
    return 0;
}
