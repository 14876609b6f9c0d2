// capi.cu — the extern "C" surface declared in include/vcr_gaus_b200.h: argument checking, opaque
// buffer layouts, and the launch sequences of the forward / backward passes.
//
// Host orchestration replaces CudaRasterizer::Rasterizer::forward / backward / markVisible /
// forwardCount / forwardVisiAcc (reference DGR/cuda_rasterizer/rasterizer_impl.cu:141-153, 199-348,
// 352-476, 482-626, 782-917).  All launches go to the caller's stream.
#include "common.cuh"
#include "../../include/vcr_gaus_b200.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

static thread_local char g_err[512] = "";
void vcr_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---------------- per-kernel profiling (not thread-safe; meant for bench.py) ----------------
#include <vector>
struct ProfRec { int id; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;
static long long g_launches[VCR_K_COUNT] = {0};
static const char* kKernelNames[VCR_K_COUNT] = {
    "preprocess_kernel", "scan_block_sums_kernel", "duplicate_kernel", "radix_scan_hist_kernel", "radix_pass_kernel",
    "tile_ranges_kernel", "render_forward_kernel", "render_backward_kernel", "preprocess_backward_kernel",
    "dnormal_forward_kernel", "dnormal_backward_kernel", "render_aux_kernel", "knn_kernels", "mark_visible_kernel", "memset", "densify_stats_kernel", "multimem_allreduce_kernel", "prologue_kernels", "ssim_kernels", "adam_step_kernel"};
static cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
VcrProfScope::VcrProfScope(int id_, cudaStream_t s_, int launches) : id(id_), s(s_), slot(-1) {
    g_launches[id] += launches;
    if (g_prof_on) {
        ProfRec r{id, prof_event(), prof_event()};
        cudaEventRecord(r.a, s);
        slot = (int)g_prof_recs.size();
        g_prof_recs.push_back(r);
    }
}
VcrProfScope::~VcrProfScope() { if (slot >= 0) cudaEventRecord(g_prof_recs[slot].b, s); }

// ---------------- layouts ----------------
template <typename T>
static inline void carve_one(char*& p, T*& out, size_t count) {
    p = reinterpret_cast<char*>(vcr_align256(reinterpret_cast<size_t>(p)));
    out = reinterpret_cast<T*>(p);
    p += count * sizeof(T);
}
static inline size_t radix_blocks(size_t R) { return (R + VCR_RADIX_TILE - 1) / VCR_RADIX_TILE; }
GeomState GeomState::carve(char* chunk, int P) {
    GeomState g;
    const size_t n = (size_t)P;
    carve_one(chunk, g.depths, n);
    carve_one(chunk, g.clamped, 3 * n);
    carve_one(chunk, g.means2D, n);
    carve_one(chunk, g.xyz_view, 3 * n);
    carve_one(chunk, g.cov3D, 6 * n);
    carve_one(chunk, g.conic_opacity, n);
    carve_one(chunk, g.rgb, 3 * n);
    carve_one(chunk, g.tiles_touched, n);
    carve_one(chunk, g.point_offsets, n);
    carve_one(chunk, g.block_sums, (n + 255) / 256 + 1);
    carve_one(chunk, g.num_rendered, (size_t)4);
    carve_one(chunk, g.gs_keys[0], n);
    carve_one(chunk, g.gs_keys[1], n);
    carve_one(chunk, g.gs_vals[0], n);
    carve_one(chunk, g.gs_vals[1], n);
    carve_one(chunk, g.gs_hist, (size_t)4 * 256);
    carve_one(chunk, g.gs_status, (size_t)4 * radix_blocks(n) * 256 + 16);
    carve_one(chunk, g.sorted_sums, (n + 255) / 256 + 1);
    return g;
}
size_t GeomState::bytes(int P) {
    char* base = nullptr;
    GeomState g = carve(base, P);
    return reinterpret_cast<size_t>(g.sorted_sums + ((size_t)P + 255) / 256 + 1) + 256;
}
ImageState ImageState::carve(char* chunk, int W, int H) {
    ImageState s;
    const size_t N = (size_t)W * H;
    const size_t tiles = (size_t)((W + VCR_TILE_X - 1) / VCR_TILE_X) * ((H + VCR_TILE_Y - 1) / VCR_TILE_Y);
    carve_one(chunk, s.accum_alpha, N);
    carve_one(chunk, s.n_contrib, N);
    carve_one(chunk, s.ranges, tiles);
    carve_one(chunk, s.tile_order, tiles);
    return s;
}
size_t ImageState::bytes(int W, int H) {
    char* base = nullptr;
    ImageState s = carve(base, W, H);
    const size_t tiles = (size_t)((W + VCR_TILE_X - 1) / VCR_TILE_X) * ((H + VCR_TILE_Y - 1) / VCR_TILE_Y);
    return reinterpret_cast<size_t>(s.tile_order + tiles) + 256;
}
BinningState BinningState::carve(char* chunk, int R) {
    BinningState b;
    const size_t n = (size_t)R;
    carve_one(chunk, b.point_list, n);
    carve_one(chunk, b.tile_keys, n);
    carve_one(chunk, b.alt_vals, n);
    carve_one(chunk, b.alt_keys, n);
    carve_one(chunk, b.sort_hist, (size_t)8 * 256);
    carve_one(chunk, b.sort_status, (size_t)4 * radix_blocks(n) * 256 + 16);
    carve_one(chunk, b.point_list_keys, n);
    carve_one(chunk, b.point_list_unsorted, n);
    carve_one(chunk, b.point_list_keys_unsorted, n);
    return b;
}
size_t BinningState::bytes(int R) {
    char* base = nullptr;
    BinningState b = carve(base, R);
    return reinterpret_cast<size_t>(b.point_list_keys_unsorted + (size_t)R) + 256;
}

static uint32_t higher_msb(uint32_t n) {  // rasterizer_impl.cu:35-50
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// The device copy of R is read back through a pinned word (one per host thread).  The copy is followed by an event, not
// by a stream synchronisation: work queued behind it (the Gaussian-level depth sort) keeps the GPU busy while the host
// waits for R, sizes the binning buffer and issues the rest of the forward.
struct ReadBack {
    uint32_t* pinned = nullptr;
    cudaEvent_t ev = nullptr;
};
static int read_back_begin(const uint32_t* dev, ReadBack*& rb_out, cudaStream_t s) {
    // one pinned word block + one event per host thread AND device: an event belongs to the device that was current when it
    // was created, so a thread that drives several GPUs needs one each (they live until the process exits)
    static thread_local ReadBack per_device[64];
    int d = 0;
    VCR_CUDA_TRY(cudaGetDevice(&d));
    if (d < 0 || d >= 64) { vcr_set_error("unsupported device index %d", d); return 1; }
    ReadBack& rb = per_device[d];
    if (!rb.pinned) VCR_CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&rb.pinned), 4 * sizeof(uint32_t), cudaHostAllocDefault));
    if (!rb.ev) VCR_CUDA_TRY(cudaEventCreateWithFlags(&rb.ev, cudaEventDisableTiming));
    VCR_CUDA_TRY(cudaMemcpyAsync(rb.pinned, dev, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    VCR_CUDA_TRY(cudaEventRecord(rb.ev, s));
    rb_out = &rb;
    return 0;
}
static int read_back_end(ReadBack* rb, uint32_t out[4]) {
    VCR_CUDA_TRY(cudaEventSynchronize(rb->ev));
    memcpy(out, rb->pinned, 4 * sizeof(uint32_t));
    return 0;
}

static int debug_sync(int debug, cudaStream_t s, const char* what) {  // CHECK_CUDA of the reference (auxiliary.h:181-188)
    if (!debug) return 0;
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { vcr_set_error("[CUDA ERROR] after %s: %s", what, cudaGetErrorString(e)); return 1; }
    return 0;
}

// Shared front half of every forward variant: preprocess -> scan -> R read-back (event) || depth sort of the Gaussians ->
// emission in depth order -> sort by tile -> ranges + tile schedule.
static int forward_front(vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* user,
                         FwdParams& p, GeomState& g, BinningState& b, ImageState& img, int debug, int* num_rendered_host,
                         cudaStream_t s) {
    if (p.P <= 0) { vcr_set_error("P must be positive (the P == 0 short-circuit lives in the caller)"); return 1; }
    if (p.W <= 0 || p.H <= 0) { vcr_set_error("bad image size %dx%d", p.W, p.H); return 1; }
    if (!geometry_buffer || !binning_buffer || !image_buffer) { vcr_set_error("NULL buffer callback"); return 1; }
    if (!p.means3D || !p.opacities || !p.viewmatrix || !p.projmatrix || !p.radii) { vcr_set_error("NULL required input"); return 1; }
    if (!p.colors_precomp && !p.shs) { vcr_set_error("either shs or colors_precomp is required"); return 1; }
    if (!p.cov3D_precomp && (!p.scales || !p.rotations)) { vcr_set_error("either cov3D_precomp or scales+rotations is required"); return 1; }
    if (!p.colors_precomp && !p.cam_pos) { vcr_set_error("cam_pos is required for SH colours"); return 1; }
    // float4 accesses (include/vcr_gaus_b200.h "Alignment"): a misaligned pointer would fault inside the kernel
    if (!p.cov3D_precomp && (reinterpret_cast<size_t>(p.rotations) & 15)) { vcr_set_error("rotations must be 16-byte aligned (read as float4)"); return 1; }
    if (!p.colors_precomp && p.M == 16 && (reinterpret_cast<size_t>(p.shs) & 15)) { vcr_set_error("shs must be 16-byte aligned when M == 16 (read as float4)"); return 1; }
    p.gx = (p.W + VCR_TILE_X - 1) / VCR_TILE_X;
    p.gy = (p.H + VCR_TILE_Y - 1) / VCR_TILE_Y;
    p.focal_y = p.H / (2.0f * p.tan_fovy);  // rasterizer_impl.cu:227-228
    p.focal_x = p.W / (2.0f * p.tan_fovx);

    char* gchunk = geometry_buffer(user, GeomState::bytes(p.P));
    char* ichunk = image_buffer(user, ImageState::bytes(p.W, p.H));
    if (!gchunk || !ichunk) { vcr_set_error("buffer callback returned NULL"); return 1; }
    g = GeomState::carve(gchunk, p.P);
    img = ImageState::carve(ichunk, p.W, p.H);

    if (vcr_launch_preprocess(p, g, s)) return 1;
    if (debug_sync(debug, s, "preprocess")) return 1;
    if (vcr_launch_scan_block_sums(p.P, g, s)) return 1;
    uint32_t host[4];
    ReadBack* rb = nullptr;
    if (read_back_begin(g.num_rendered, rb, s)) return 1;
    if (vcr_launch_gaussian_sort(p.P, g, s)) return 1;          // runs while the host waits for R
    if (read_back_end(rb, host)) return 1;
    if (host[1]) { vcr_set_error("Point is filtered although prefiltered is set. This shouldn't happen!"); return 1; }
    if (host[3]) {
        vcr_set_error("radix sort look-back timed out in an earlier call on this device (or a staging-ring wait of its compositing kernels): the instance lists (and every result) of that "
                      "call are invalid");
        return 1;
    }
    if (host[0] >= (1u << 30)) { vcr_set_error("%u (Gaussian, tile) instances: at most 2^30 - 1 are supported", host[0]); return 1; }
    const int R = (int)host[0];
    *num_rendered_host = R;

    char* bchunk = binning_buffer(user, BinningState::bytes(R));
    if (!bchunk) { vcr_set_error("binning buffer callback returned NULL"); return 1; }
    b = BinningState::carve(bchunk, R);
    const int tiles = p.gx * p.gy;
    // bits of the largest tile id (the reference sorts bits [0, 32 + getHigherMsb(tiles)), rasterizer_impl.cu:305-314, i.e. one
    // more bit whenever the tile count is a power of two; any bit count that covers every tile id gives the same order)
    int tile_bits = 1;
    while (tile_bits < 32 && ((uint32_t)(tiles - 1) >> tile_bits)) tile_bits++;
    if (R > 0) {
        if (debug_sync(debug, s, "depth sort")) return 1;
        if (vcr_launch_duplicate_sorted(p, g, b, R, tile_bits, s)) return 1;
        if (debug_sync(debug, s, "duplicate")) return 1;
        if (vcr_launch_tile_sort(b, R, tile_bits, s)) return 1;
        if (debug_sync(debug, s, "tile sort")) return 1;
    }
    if (vcr_launch_tile_ranges(b, img, R, tiles, s)) return 1;
    if (debug_sync(debug, s, "tile ranges")) return 1;
    if (debug) {                                  // debug mode reports a sort failure of THIS call before returning
        if (vcr_radix_take_error(g.num_rendered + 3, s)) return 1;
        uint32_t flag = 0;
        VCR_CUDA_TRY(cudaMemcpyAsync(&flag, g.num_rendered + 3, sizeof(flag), cudaMemcpyDeviceToHost, s));
        VCR_CUDA_TRY(cudaStreamSynchronize(s));
        if (flag) { vcr_set_error("radix sort look-back timed out: the instance lists of this call are invalid"); return 1; }
    }
    return 0;
}

extern "C" {

int vcr_abi_version(void) { return VCR_ABI_VERSION; }
const char* vcr_last_error(void) { return g_err; }
uint32_t vcr_higher_msb(uint32_t n) { return higher_msb(n); }
size_t vcr_geom_bytes(int P) { return GeomState::bytes(P); }
size_t vcr_image_bytes(int width, int height) { return ImageState::bytes(width, height); }
size_t vcr_binning_bytes(int num_rendered) { return BinningState::bytes(num_rendered); }
size_t vcr_backward_scratch_bytes(int P) { return (size_t)P * VCR_ACC_STRIDE * sizeof(float) + 256; }

static int rasterize_forward_impl(int variant, vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
                          int P, int D, int M, const float* background, int width, int height, const float* means3D,
                          const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                          const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix, const float* dirs,
                          const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color, int* radii,
                          const uint8_t* inside, int debug, int* num_rendered_host, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (!normals_precomp) { vcr_set_error("normals_precomp is required (the reference kernel dereferences it unconditionally, forward.cu:374)"); return 1; }
    if (!background || !out_color || !num_rendered_host) { vcr_set_error("NULL required argument"); return 1; }   // inside == NULL: all inside
    FwdParams p;
    memset(&p, 0, sizeof(p));
    p.P = P; p.D = D; p.M = M; p.W = width; p.H = height;
    p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.normals = normals_precomp; p.sh_objs = sh_objs;
    p.opacities = opacities; p.scales = scales; p.scale_modifier = scale_modifier; p.rotations = rotations;
    p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix; p.dirs = dirs; p.cam_pos = cam_pos;
    p.background = background; p.tan_fovx = tan_fovx; p.tan_fovy = tan_fovy; p.prefiltered = prefiltered;
    p.out_color = out_color; p.radii = radii; p.inside = inside; p.variant = variant;
    GeomState g; BinningState b; ImageState img;
    if (forward_front(geometry_buffer, binning_buffer, image_buffer, alloc_user, p, g, b, img, debug, num_rendered_host, s)) return 1;
    if (vcr_launch_render_forward(p, g, b, img, s)) return 1;
    if (debug_sync(debug, s, "render forward")) return 1;
    return 0;
}

int vcr_rasterize_forward(vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
                          int P, int D, int M, const float* background, int width, int height, const float* means3D,
                          const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                          const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix, const float* dirs,
                          const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color, int* radii,
                          const uint8_t* inside, int debug, int* num_rendered_host, void* stream) {
    return rasterize_forward_impl(0, geometry_buffer, binning_buffer, image_buffer, alloc_user, P, D, M, background, width, height, means3D, shs, colors_precomp, normals_precomp, sh_objs, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, dirs, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color, radii, inside, debug, num_rendered_host, stream);
}

static int variant_code(int num_dist, int num_depth2, int* code) {
    if ((num_dist != 0 && num_dist != 1) || (num_depth2 != 0 && num_depth2 != 2)) {
        vcr_set_error("NUM_DIST must be 0 or 1 and NUM_DEPTH2 0 or 2 (DGR/cuda_rasterizer/config.h:17-18), got %d / %d", num_dist, num_depth2);
        return 1;
    }
    *code = (num_dist ? 1 : 0) | (num_depth2 ? 2 : 0);
    return 0;
}

int vcr_variant_channels(int num_dist, int num_depth2) { return VCR_OUT_CH + (num_dist ? 1 : 0) + (num_depth2 ? 2 : 0); }

int vcr_rasterize_forward_variant(int num_dist, int num_depth2, vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
                          int P, int D, int M, const float* background, int width, int height, const float* means3D,
                          const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                          const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix, const float* dirs,
                          const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color, int* radii,
                          const uint8_t* inside, int debug, int* num_rendered_host, void* stream) {
    int code = 0;
    if (variant_code(num_dist, num_depth2, &code)) return 1;
    return rasterize_forward_impl(code, geometry_buffer, binning_buffer, image_buffer, alloc_user, P, D, M, background, width, height, means3D, shs, colors_precomp, normals_precomp, sh_objs, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, dirs, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color, radii, inside, debug, num_rendered_host, stream);
}

static int rasterize_backward_impl(int variant, int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
                           const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                           const float* alpha, const float* depth, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* dirs, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                           char* geom_buffer, char* binning_buffer, char* img_buffer, const float* dL_dpix, float* dL_dmean2D,
                           float* dL_dmean2D_densify, float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_dobjects,
                           float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                           const uint8_t* inside, char* scratch, int debug, void* stream, int n_chunks,
                           vcr_chunk_fn chunk_enqueued, void* chunk_user, float* dL_dcolor_masked) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0) { vcr_set_error("P must be positive"); return 1; }
    if (!geom_buffer || !binning_buffer || !img_buffer || !scratch) { vcr_set_error("NULL state/scratch buffer"); return 1; }
    // dL_dcolor / dL_dobjects / dL_dcov3D may be NULL: they are the gradients of optional INPUTS (colors_precomp, sh_objs,
    // cov3D_precomp); a caller that did not pass the input has no use for the array and saves its 44 B per Gaussian of writes
    if (!dL_dpix || !dL_dmean2D || !dL_dmean2D_densify || !dL_dopacity || !dL_dnormal ||
        !dL_dmean3D || !dL_dscale || !dL_drot || (colors_precomp && !dL_dcolor) || (sh_objs && !dL_dobjects) ||
        (cov3D_precomp && !dL_dcov3D) || (M > 0 && shs && !dL_dsh && !dL_dcolor_masked)) {
        vcr_set_error("NULL gradient output");
        return 1;
    }
    if (!normals_precomp || !radii || !alpha || !depth) { vcr_set_error("NULL required input"); return 1; }   // inside == NULL: all inside
    if (scales && ((reinterpret_cast<size_t>(rotations) | reinterpret_cast<size_t>(dL_drot)) & 15)) {
        vcr_set_error("rotations and dL_drot must be 16-byte aligned (float4 accesses)");
        return 1;
    }
    if (shs && M == 16 && dL_dsh && !dL_dcolor_masked && (reinterpret_cast<size_t>(dL_dsh) & 15)) {
        vcr_set_error("dL_dsh must be 16-byte aligned when M == 16 (written as float4)");
        return 1;
    }
    BwdParams p;
    memset(&p, 0, sizeof(p));
    p.P = P; p.D = D; p.M = M; p.R = R; p.W = width; p.H = height;
    p.gx = (width + VCR_TILE_X - 1) / VCR_TILE_X; p.gy = (height + VCR_TILE_Y - 1) / VCR_TILE_Y;
    p.background = background; p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.normals = normals_precomp;
    p.sh_objs = sh_objs; p.alpha = alpha; p.depth = depth; p.scales = scales; p.scale_modifier = scale_modifier;
    p.rotations = rotations; p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix; p.projmatrix = projmatrix;
    p.dirs = dirs; p.campos = campos; p.tan_fovx = tan_fovx; p.tan_fovy = tan_fovy;
    p.focal_y = height / (2.0f * tan_fovy); p.focal_x = width / (2.0f * tan_fovx);
    p.radii = radii; p.dL_dpix = dL_dpix;
    p.dL_dmean2D = dL_dmean2D; p.dL_dmean2D_densify = dL_dmean2D_densify; p.dL_dopacity = dL_dopacity; p.dL_dcolor = dL_dcolor;
    p.dL_dnormal = dL_dnormal; p.dL_dobjects = dL_dobjects; p.dL_dmean3D = dL_dmean3D; p.dL_dcov3D = dL_dcov3D; p.dL_dsh = dL_dsh;
    p.dL_dscale = dL_dscale; p.dL_drot = dL_drot; p.inside = inside;
    p.dL_dcolor_masked = shs ? dL_dcolor_masked : nullptr; p.variant = variant;
    p.acc = reinterpret_cast<float*>(vcr_align256(reinterpret_cast<size_t>(scratch)));
    GeomState g = GeomState::carve(geom_buffer, P);
    BinningState b = BinningState::carve(binning_buffer, R);
    ImageState img = ImageState::carve(img_buffer, width, height);
    if (vcr_launch_render_backward(p, g, b, img, s)) return 1;
    if (debug_sync(debug, s, "render backward")) return 1;
    // The per-Gaussian stage runs as n_chunks launches over consecutive ranges (multiples of 256 Gaussians).  After each
    // launch the caller is told which rows of the gradient outputs that launch completes, so it can queue the exchange of
    // those rows (an all-reduce ordered after this stream's work so far) while the next range is still being computed.
    if (n_chunks < 1) n_chunks = 1;
    const int blocks = (P + 255) / 256;
    if (n_chunks > blocks) n_chunks = blocks;
    for (int c = 0; c < n_chunks; c++) {
        const int b0 = (int)((long long)blocks * c / n_chunks), b1 = (int)((long long)blocks * (c + 1) / n_chunks);
        const int first = b0 * 256, count = (b1 * 256 < P ? b1 * 256 : P) - first;
        if (vcr_launch_preprocess_backward(p, g, first, count, s)) return 1;
        if (chunk_enqueued) chunk_enqueued(chunk_user, c, first, count);
    }
    if (debug_sync(debug, s, "preprocess backward")) return 1;
    return 0;
}

int vcr_rasterize_backward_chunked(int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
                           const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                           const float* alpha, const float* depth, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* dirs, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                           char* geom_buffer, char* binning_buffer, char* img_buffer, const float* dL_dpix, float* dL_dmean2D,
                           float* dL_dmean2D_densify, float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_dobjects,
                           float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                           const uint8_t* inside, char* scratch, int debug, void* stream, int n_chunks,
                           vcr_chunk_fn chunk_enqueued, void* chunk_user, float* dL_dcolor_masked) {
    return rasterize_backward_impl(0, P, D, M, R, background, width, height, means3D, shs, colors_precomp, normals_precomp, sh_objs, alpha, depth, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, dirs, campos, tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, img_buffer, dL_dpix, dL_dmean2D, dL_dmean2D_densify, dL_dopacity, dL_dcolor, dL_dnormal, dL_dobjects, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, inside, scratch, debug, stream, n_chunks, chunk_enqueued, chunk_user, dL_dcolor_masked);
}

int vcr_rasterize_backward_variant(int num_dist, int num_depth2, int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
                           const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                           const float* alpha, const float* depth, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* dirs, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                           char* geom_buffer, char* binning_buffer, char* img_buffer, const float* dL_dpix, float* dL_dmean2D,
                           float* dL_dmean2D_densify, float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_dobjects,
                           float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                           const uint8_t* inside, char* scratch, int debug, void* stream) {
    int code = 0;
    if (variant_code(num_dist, num_depth2, &code)) return 1;
    return rasterize_backward_impl(code, P, D, M, R, background, width, height, means3D, shs, colors_precomp, normals_precomp, sh_objs, alpha, depth, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, dirs, campos, tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, img_buffer, dL_dpix, dL_dmean2D, dL_dmean2D_densify, dL_dopacity, dL_dcolor, dL_dnormal, dL_dobjects, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, inside, scratch, debug, stream, 1, nullptr, nullptr, nullptr);
}

int vcr_rasterize_backward(int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
                           const float* shs, const float* colors_precomp, const float* normals_precomp, const float* sh_objs,
                           const float* alpha, const float* depth, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* dirs, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                           char* geom_buffer, char* binning_buffer, char* img_buffer, const float* dL_dpix, float* dL_dmean2D,
                           float* dL_dmean2D_densify, float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_dobjects,
                           float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                           const uint8_t* inside, char* scratch, int debug, void* stream) {
    return vcr_rasterize_backward_chunked(P, D, M, R, background, width, height, means3D, shs, colors_precomp, normals_precomp, sh_objs,
                                          alpha, depth, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, dirs,
                                          campos, tan_fovx, tan_fovy, radii, geom_buffer, binning_buffer, img_buffer, dL_dpix,
                                          dL_dmean2D, dL_dmean2D_densify, dL_dopacity, dL_dcolor, dL_dnormal, dL_dobjects, dL_dmean3D,
                                          dL_dcov3D, dL_dsh, dL_dscale, dL_drot, inside, scratch, debug, stream, 1, nullptr, nullptr, nullptr);
}

int vcr_sh_grad_from_colors(int P, int D, int M, int world, const float* means3D, const float* slices, long long slice_stride,
                            float* dL_dsh, void* stream) {
    if (P <= 0) return 0;
    if (!means3D || !slices || !dL_dsh || world < 1 || M < 1 || slice_stride < 3LL * P + 3) { vcr_set_error("sh_grad_from_colors: bad argument"); return 1; }
    if (M == 16 && (reinterpret_cast<size_t>(dL_dsh) & 15)) { vcr_set_error("sh_grad_from_colors: dL_dsh must be 16-byte aligned when M == 16 (written as float4)"); return 1; }
    return vcr_launch_sh_grad_from_colors(P, D, M, world, means3D, slices, slice_stride, dL_dsh, reinterpret_cast<cudaStream_t>(stream));
}

int vcr_multimem_exchange(void* multicast_base, const float* local_base, int n_ranges, const long long* first_float_host,
                          const long long* n_floats_host, long long bcast_first, long long bcast_count, int rank, int world,
                          int blocks, void* stream) {
    if (!multicast_base) { vcr_set_error("NULL multicast address (no NVLink multicast mapping for this buffer)"); return 1; }
    if (n_ranges < 0 || n_ranges > 8) { vcr_set_error("at most 8 ranges per launch"); return 1; }
    if (world < 1 || rank < 0 || rank >= world) { vcr_set_error("bad rank / world"); return 1; }
    if (n_ranges > 0 && (!first_float_host || !n_floats_host)) { vcr_set_error("NULL range arrays"); return 1; }
    for (int i = 0; i < n_ranges; i++)
        if (first_float_host[i] < 0 || n_floats_host[i] < 0) { vcr_set_error("negative range"); return 1; }
    if (bcast_count < 0 || bcast_first < 0 || (bcast_count > 0 && (!local_base || (bcast_first & 3)))) { vcr_set_error("bad broadcast range (16-byte aligned start, local base required)"); return 1; }
    if (n_ranges == 0 && bcast_count == 0) return 0;
    if (blocks <= 0) blocks = 37;
    return vcr_launch_multimem_exchange(static_cast<float*>(multicast_base), local_base, n_ranges, first_float_host, n_floats_host,
                                        bcast_first, bcast_count, rank, world, blocks, reinterpret_cast<cudaStream_t>(stream));
}

int vcr_multimem_allreduce_sum(void* multicast_base, int n_ranges, const long long* first_float_host,
                               const long long* n_floats_host, int rank, int world, int blocks, void* stream) {
    if (!multicast_base) { vcr_set_error("NULL multicast address (no NVLink multicast mapping for this buffer)"); return 1; }
    if (n_ranges < 0 || n_ranges > 8) { vcr_set_error("at most 8 ranges per launch"); return 1; }
    if (world < 1 || rank < 0 || rank >= world) { vcr_set_error("bad rank / world"); return 1; }
    if (n_ranges == 0) return 0;
    if (!first_float_host || !n_floats_host) { vcr_set_error("NULL range arrays"); return 1; }
    for (int i = 0; i < n_ranges; i++)
        if (first_float_host[i] < 0 || n_floats_host[i] < 0) { vcr_set_error("negative range"); return 1; }
    if (blocks <= 0) blocks = 37;   // 1.2 MB in flight already saturates the links (tools/mm_bench.py); 3/4 of the SMs stay free
    return vcr_launch_multimem_allreduce(static_cast<float*>(multicast_base), n_ranges, first_float_host, n_floats_host, rank, world,
                                         blocks, reinterpret_cast<cudaStream_t>(stream));
}

int vcr_binning_reference_lists(int P, int R, int width, int height, const int* radii, char* geom_buffer, char* binning_buffer,
                                void* stream) {
    if (P <= 0 || R <= 0) return 0;
    if (!radii || !geom_buffer || !binning_buffer || width <= 0 || height <= 0) { vcr_set_error("binning_reference_lists: bad argument"); return 1; }
    GeomState g = GeomState::carve(geom_buffer, P);
    BinningState b = BinningState::carve(binning_buffer, R);
    return vcr_launch_reference_lists(P, R, (width + VCR_TILE_X - 1) / VCR_TILE_X, (height + VCR_TILE_Y - 1) / VCR_TILE_Y, radii, g, b,
                                      reinterpret_cast<cudaStream_t>(stream));
}

int vcr_densify_stats_accumulate(int P, const float* grad_means2D_densify, const int* radii, float* xyz_gradient_accum,
                                 float* denom, float* max_radii2D, void* stream) {
    if (P <= 0) return 0;
    if (!grad_means2D_densify || !radii || !xyz_gradient_accum || !denom || !max_radii2D) { vcr_set_error("NULL argument"); return 1; }
    return vcr_launch_densify_stats(P, grad_means2D_densify, radii, xyz_gradient_accum, denom, max_radii2D,
                                    reinterpret_cast<cudaStream_t>(stream));
}

int vcr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present, void* stream) {
    (void)projmatrix;  // the reference computes the projection too but only tests view-space z (auxiliary.h:160-166)
    if (P <= 0) return 0;
    if (!means3D || !viewmatrix || !present) { vcr_set_error("NULL argument"); return 1; }
    return vcr_launch_mark_visible(P, means3D, viewmatrix, present, reinterpret_cast<cudaStream_t>(stream));
}

static int aux_forward(int mode, vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
                       int P, int D, int M, const float* background, int width, int height, const float* means3D, const float* shs,
                       const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                       const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                       const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color,
                       int* gaussians_count, float* important_score, int* radii, int debug, int* num_rendered_host, cudaStream_t s) {
    if (!gaussians_count || !num_rendered_host) { vcr_set_error("NULL required argument"); return 1; }
    FwdParams p;
    memset(&p, 0, sizeof(p));
    p.P = P; p.D = D; p.M = M; p.W = width; p.H = height;
    p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.opacities = opacities; p.scales = scales;
    p.scale_modifier = scale_modifier; p.rotations = rotations; p.cov3D_precomp = cov3D_precomp; p.viewmatrix = viewmatrix;
    p.projmatrix = projmatrix; p.cam_pos = cam_pos; p.background = background; p.tan_fovx = tan_fovx; p.tan_fovy = tan_fovy;
    p.prefiltered = prefiltered; p.out_color = out_color; p.radii = radii;
    GeomState g; BinningState b; ImageState img;
    if (forward_front(geometry_buffer, binning_buffer, image_buffer, alloc_user, p, g, b, img, debug, num_rendered_host, s)) return 1;
    VCR_CUDA_TRY(cudaMemsetAsync(gaussians_count, 0, (size_t)P * sizeof(int), s));
    if (important_score) VCR_CUDA_TRY(cudaMemsetAsync(important_score, 0, (size_t)P * sizeof(float), s));
    if (vcr_launch_render_aux(mode, p, g, b, img, gaussians_count, important_score, s)) return 1;
    if (debug_sync(debug, s, "render aux")) return 1;
    return 0;
}

int vcr_rasterize_count(vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user, int P,
                        int D, int M, const float* background, int width, int height, const float* means3D, const float* shs,
                        const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                        const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                        const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color,
                        int* gaussians_count, float* important_score, int* radii, int debug, int* num_rendered_host, void* stream) {
    if (!out_color || !important_score || !background) { vcr_set_error("NULL required argument"); return 1; }
    return aux_forward(VCR_MODE_COUNT, geometry_buffer, binning_buffer, image_buffer, alloc_user, P, D, M, background, width, height,
                       means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
                       projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color, gaussians_count, important_score, radii,
                       debug, num_rendered_host, reinterpret_cast<cudaStream_t>(stream));
}

int vcr_rasterize_visi_acc(vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
                           int P, int D, int M, int width, int height, const float* means3D, const float* shs,
                           const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                           const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, int* gaussians_count, int* radii,
                           int debug, int* num_rendered_host, void* stream) {
    return aux_forward(VCR_MODE_VISI_ACC, geometry_buffer, binning_buffer, image_buffer, alloc_user, P, D, M, nullptr, width, height,
                       means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
                       projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, nullptr, gaussians_count, nullptr, radii, debug,
                       num_rendered_host, reinterpret_cast<cudaStream_t>(stream));
}

int vcr_profile_enable(int enable) { g_prof_on = enable != 0; return 0; }
int vcr_profile_kernel_count(void) { return VCR_K_COUNT; }
const char* vcr_profile_kernel_name(int id) { return (id >= 0 && id < VCR_K_COUNT) ? kKernelNames[id] : ""; }
int vcr_profile_read(float* ms, long long* timed, long long* launches, int reset) {
    for (int k = 0; k < VCR_K_COUNT; k++) { if (ms) ms[k] = 0.f; if (timed) timed[k] = 0; if (launches) launches[k] = g_launches[k]; }
    for (ProfRec& r : g_prof_recs) {
        VCR_CUDA_TRY(cudaEventSynchronize(r.b));
        float t = 0.f;
        VCR_CUDA_TRY(cudaEventElapsedTime(&t, r.a, r.b));
        if (ms) ms[r.id] += t;
        if (timed) timed[r.id] += 1;
    }
    if (reset) {
        for (ProfRec& r : g_prof_recs) { g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b); }
        g_prof_recs.clear();
        for (int k = 0; k < VCR_K_COUNT; k++) g_launches[k] = 0;
    }
    return 0;
}

void vcr_geom_layout(char* chunk, int P, vcr_geom_view* out) {
    GeomState g = GeomState::carve(chunk, P);
    out->depths = g.depths; out->clamped = g.clamped; out->means2D = reinterpret_cast<float*>(g.means2D);
    out->xyz_view = g.xyz_view; out->cov3D = g.cov3D; out->conic_opacity = reinterpret_cast<float*>(g.conic_opacity);
    out->rgb = g.rgb; out->tiles_touched = g.tiles_touched; out->point_offsets = g.point_offsets;
}
void vcr_binning_layout(char* chunk, int num_rendered, vcr_binning_view* out) {
    BinningState b = BinningState::carve(chunk, num_rendered);
    out->point_list = b.point_list; out->point_list_keys = b.point_list_keys;
    out->point_list_unsorted = b.point_list_unsorted; out->point_list_keys_unsorted = b.point_list_keys_unsorted;
}
void vcr_image_layout(char* chunk, int width, int height, vcr_image_view* out) {
    ImageState s = ImageState::carve(chunk, width, height);
    out->accum_alpha = s.accum_alpha; out->n_contrib = s.n_contrib; out->ranges = reinterpret_cast<uint32_t*>(s.ranges);
}

}  // extern "C"
