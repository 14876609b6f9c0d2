/*
 * vcr_gaus_b200.h — C ABI of libvcr_gaus_b200.so, the sm_100a rasterizer / D-Normal / kNN library.
 *
 * This is the drop-in boundary of the hot path.  Every entry point replaces one native interface
 * of the reference (paths relative to the reference root; DGR = submodules/diff-gaussian-rasterization):
 *
 *   vcr_rasterize_forward        <- CudaRasterizer::Rasterizer::forward       DGR/cuda_rasterizer/rasterizer.h:31-59
 *                                   (called from RasterizeGaussiansCUDA,        DGR/rasterize_points.cu:49-152)
 *   vcr_rasterize_backward       <- CudaRasterizer::Rasterizer::backward      DGR/cuda_rasterizer/rasterizer.h:137-177
 *                                   (called from RasterizeGaussiansBackwardCUDA, DGR/rasterize_points.cu:418-529)
 *   vcr_mark_visible             <- CudaRasterizer::Rasterizer::markVisible   DGR/cuda_rasterizer/rasterizer.h:24-29
 *   vcr_rasterize_count          <- CudaRasterizer::Rasterizer::forwardCount  DGR/cuda_rasterizer/rasterizer.h:62-86
 *   vcr_rasterize_visi_acc       <- CudaRasterizer::Rasterizer::forwardVisiAcc DGR/cuda_rasterizer/rasterizer.h (f_count=3)
 *   vcr_dist2_knn3               <- SimpleKNN::knn                            submodules/simple-knn/simple_knn.h:17
 *   vcr_dnormal_forward/backward <- tools/normal_utils.py:30-41 + tools/loss_utils.py:122-142 + trainer.py:266-280
 *                                   (a PyTorch op chain in the reference; one fused kernel each way here)
 *   vcr_view_loss_forward/backward/total <- the image-space losses of trainer.py:236-280 (L1 + mono-normal + D-Normal)
 *   vcr_ssim_forward/backward    <- tools/loss_utils.py:47-92
 *   vcr_prologue_forward/backward <- the per-Gaussian prologue of render(), gaussian_renderer/__init__.py:60-101
 *   vcr_adam_step                <- torch.optim.Adam over the six groups of scene/gaussian_model.py:241-258
 *   vcr_densify_stats_accumulate <- trainer.py:345 + scene/gaussian_model.py:669-671
 *   vcr_densify_decide / _fill   <- GaussianModel.densify_and_prune, scene/gaussian_model.py:579-659
 *   vcr_compact_rows             <- prune_points / _prune_optimizer, scene/gaussian_model.py:442-478
 * and, with no counterpart in the (single-GPU) reference, the view-parallel pieces: vcr_rasterize_backward_chunked,
 * vcr_multimem_allreduce_sum, vcr_multimem_exchange, vcr_sh_grad_from_colors; plus the test-only parity view
 * vcr_binning_reference_lists and vcr_dist2_knn3_noalloc.
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless the name ends in _host.  The caller owns all memory;
 *     the library never frees, and never keeps a pointer past the call.
 *   - Absent optional inputs are NULL (the reference passes zero-sized tensors whose data_ptr is
 *     nullptr, DGR/cuda_rasterizer/forward.cu:208,244,302,418).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream, which is where the
 *     reference launches everything).
 *   - Opaque per-view state (geometry / binning / image buffers) is obtained through caller-supplied
 *     grow callbacks, exactly like the reference's std::function<char*(size_t)> resize lambdas
 *     (DGR/rasterize_points.cu:28-34, rasterizer.h:32-34).  The byte layout is private to this
 *     library (documented in DESIGN.md); forward and backward agree on it.
 *   - Alignment: `rotations` (f32[P,4]), `shs` and `dL_dsh` when M == 16 (f32[P,16,3]), `dL_drot`, and the
 *     prologue's rotation_raw / rotations / dL_drotations / dL_drotation_raw are accessed as float4 and must be
 *     16-byte aligned; the buffers of vcr_adam_step likewise.  A tensor that starts a torch allocation always is;
 *     a slice of a flat parameter bucket is only if its float offset is a multiple of 4
 *     (vcr_view_parallel.GaussianBucket pads its groups accordingly).  Misaligned pointers are rejected with an
 *     error before anything is launched.  All other arrays only need their natural 4-byte alignment.
 *   - Return value: 0 = success, non-zero = error; vcr_last_error() returns a thread-local message.
 *     Negative num_rendered is never returned.
 */
#ifndef VCR_GAUS_B200_H
#define VCR_GAUS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VCR_ABI_VERSION 1
#define VCR_OUT_CHANNELS 10 /* RGB 0-2, depth 3, normal 4-6, alpha 7, semantics 8-9 (DGR auxiliary.h:21-28) */
#define VCR_NUM_OBJECTS 2   /* DGR cuda_rasterizer/config.h:16 */
#define VCR_TILE 16         /* DGR cuda_rasterizer/config.h:19-20 */

/* Grow callback: return a device pointer to at least `bytes` bytes (256-byte aligned), valid until the
 * matching backward call has finished.  `user` is passed through untouched. */
typedef char* (*vcr_alloc_fn)(void* user, size_t bytes);

int vcr_abi_version(void);
const char* vcr_last_error(void);

/* Number of key bits used for tile ids (reference: getHigherMsb, DGR/cuda_rasterizer/rasterizer_impl.cu:35-50). */
uint32_t vcr_higher_msb(uint32_t n);

/* Buffer sizes the grow callbacks will be asked for (exposed for tests and pre-allocation). */
size_t vcr_geom_bytes(int P);
size_t vcr_image_bytes(int width, int height);
size_t vcr_binning_bytes(int num_rendered);

/* Full forward pass of one view.  On success *num_rendered_host receives R (the number of
 * (Gaussian, tile) instances).  out_color is f32[10,H,W] and radii is i32[P]; both are fully
 * written (no pre-zeroing needed).  The host waits once inside (to learn R before sizing the binning
 * buffer, as in the reference, rasterizer_impl.cu:287) — on an event recorded after the copy of R, while
 * work already queued behind it on `stream` keeps running. */
int vcr_rasterize_forward(
    vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
    int P, int D, int M,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* normals_precomp,
    const float* sh_objs,
    const float* opacities,
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* dirs,
    const float* cam_pos,
    float tan_fovx, float tan_fovy,
    int prefiltered,
    float* out_color,
    int* radii,
    const uint8_t* inside,       /* u8/bool[P]: which Gaussians enter the depth / normal maps (the fork's `inside` mask); NULL = all,
                                    what the reference wrapper substitutes for inside=None — and the fast path of the backward */
    int debug,
    int* num_rendered_host,
    void* stream);

/* Backward pass of one view.  Every output gradient array is fully written by the library
 * (zeros for Gaussians that were not rendered), so the caller may pass uninitialised memory.
 * `scratch` must hold vcr_backward_scratch_bytes(P) bytes. */
size_t vcr_backward_scratch_bytes(int P);
int vcr_rasterize_backward(
    int P, int D, int M, int R,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* normals_precomp,
    const float* sh_objs,
    const float* alpha,          /* out_color channel 7 of the forward, f32[H*W] */
    const float* depth,          /* out_color channel 3 of the forward, f32[H*W] */
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* dirs,
    const float* campos,
    float tan_fovx, float tan_fovy,
    const int* radii,
    char* geom_buffer, char* binning_buffer, char* img_buffer,
    const float* dL_dpix,        /* f32[10,H,W] */
    float* dL_dmean2D,           /* f32[P,3] */
    float* dL_dmean2D_densify,   /* f32[P,3] */
    float* dL_dopacity,          /* f32[P,1] */
    float* dL_dcolor,            /* f32[P,3]; may be NULL when colors_precomp is NULL */
    float* dL_dnormal,           /* f32[P,3] */
    float* dL_dobjects,          /* f32[P,2]; may be NULL when sh_objs is NULL */
    float* dL_dmean3D,           /* f32[P,3] */
    float* dL_dcov3D,            /* f32[P,6]; may be NULL when cov3D_precomp is NULL */
    float* dL_dsh,               /* f32[P,M,3] (may be NULL when M == 0) */
    float* dL_dscale,            /* f32[P,3] */
    float* dL_drot,              /* f32[P,4] */
    const uint8_t* inside,
    char* scratch,
    int debug,
    void* stream);

/* The same backward with its last, per-Gaussian stage (computeCov2DCUDA + preprocessCUDA of the backward,
 * DGR/cuda_rasterizer/backward.cu:144-419) split into `n_chunks` launches over consecutive ranges of Gaussians
 * (range starts are multiples of 256).  Right after the launch of chunk c has been enqueued on `stream`, the library
 * calls chunk_enqueued(chunk_user, c, first_gaussian, n_gaussians) on the calling host thread: rows
 * [first_gaussian, first_gaussian + n_gaussians) of every gradient output are final once the work enqueued so far has
 * run.  A view-parallel caller queues the all-reduce of those rows from the callback (ordered after `stream`), so the
 * exchange of chunk c overlaps the computation of chunk c+1 ("overlappable with bwd preprocess").
 * n_chunks <= 1 with a NULL callback is exactly vcr_rasterize_backward.  The reference has no counterpart: it is
 * single-GPU (tools/distributed.py is never called). */
typedef void (*vcr_chunk_fn)(void* user, int chunk, int first_gaussian, int n_gaussians);
int vcr_rasterize_backward_chunked(
    int P, int D, int M, int R,
    const float* background,
    int width, int height,
    const float* means3D,
    const float* shs,
    const float* colors_precomp,
    const float* normals_precomp,
    const float* sh_objs,
    const float* alpha,
    const float* depth,
    const float* scales,
    float scale_modifier,
    const float* rotations,
    const float* cov3D_precomp,
    const float* viewmatrix,
    const float* projmatrix,
    const float* dirs,
    const float* campos,
    float tan_fovx, float tan_fovy,
    const int* radii,
    char* geom_buffer, char* binning_buffer, char* img_buffer,
    const float* dL_dpix,
    float* dL_dmean2D,
    float* dL_dmean2D_densify,
    float* dL_dopacity,
    float* dL_dcolor,
    float* dL_dnormal,
    float* dL_dobjects,
    float* dL_dmean3D,
    float* dL_dcov3D,
    float* dL_dsh,
    float* dL_dscale,
    float* dL_drot,
    const uint8_t* inside,
    char* scratch,
    int debug,
    void* stream,
    int n_chunks,
    vcr_chunk_fn chunk_enqueued,
    void* chunk_user,
    float* dL_dcolor_masked);    /* optional f32[P,3], see vcr_sh_grad_from_colors; when non-NULL dL_dsh is NOT written */

/* Per-view densification statistics, in place, for Gaussians with radii > 0:
 *   xyz_gradient_accum[i] += |grad_means2D_densify[i, 0:2]|;  denom[i] += 1     (scene/gaussian_model.py:669-671)
 *   max_radii2D[i] = max(max_radii2D[i], radii[i])                              (trainer.py:345)
 * One pass instead of the reference's three boolean-mask-indexed PyTorch updates.  All arrays f32[P] except radii i32[P]
 * and grad_means2D_densify f32[P,3]. */
int vcr_densify_stats_accumulate(int P, const float* grad_means2D_densify, const int* radii, float* xyz_gradient_accum,
                                 float* denom, float* max_radii2D, void* stream);

/* Parity view of the binning state (tests; no caller in the reference).  The production forward sorts the Gaussians by
 * depth first, emits their (tile, Gaussian) instances in that order and then sorts by tile id only, which yields the
 * reference's sorted point list and tile ranges without ever materialising the reference's intermediate lists.  This
 * call writes them, bit for bit in the reference's formats, into the binning / geometry buffers of a finished forward:
 * the duplicated keys and values in Gaussian-index order (duplicateWithKeys, rasterizer_impl.cu:70-111), point_offsets
 * (the inclusive scan, :283) and the sorted 64-bit keys (tile << 32 | depth bits, :309-314).  Idempotent. */
int vcr_binning_reference_lists(int P, int R, int width, int height, const int* radii, char* geom_buffer, char* binning_buffer,
                                void* stream);

/* *loss (device f32) = w_l1*s0/(3N) + w_mono*(s1+s2)/N + w_dn*(s3+s4)/s5 (last term 0 when s5 == 0) from the sums of
 * vcr_view_loss_forward, N = width*height: one launch instead of a dozen single-element PyTorch launches. */
int vcr_view_loss_total(const float* sums, int width, int height, float w_l1, float w_mono, float w_dn, float* loss, void* stream);
/* dst[i] = (float)src[i] / 255 for n 8-bit samples (IEEE division: the float image the reference's loader forms,
 * utils/general_utils.py PILtoTorch), so that ground-truth images can cross the bus in their file format.  No counterpart in
 * the reference (it keeps float images on the device).  src and dst 16-byte aligned. */
int vcr_unpack_u8(const uint8_t* src, float* dst, long long n, void* stream);

/* Everything render() computes per Gaussian before it calls the rasterizer, fused:
 *   opacities = sigmoid(opacity_raw), scales = exp(scaling_raw), rotations = normalize(rotation_raw),
 *   shs = cat(features_dc, features_rest)                              (scene/gaussian_model.py:125-162)
 *   normals = column argmin(scales) of build_rotation(rotations)       (scene/gaussian_model.py:168-192,
 *             flipped to point away from the camera and rotated         tools/general_utils.py:98-119,
 *             into camera space by R_w2c                                 gaussian_renderer/__init__.py:95-101)
 * n_rest = SH coefficients beyond DC (15 at degree 3); features_dc f32[P,1,3], features_rest f32[P,n_rest,3],
 * shs f32[P,1+n_rest,3].  valid (u8[P], may be NULL = all) is the inside mask of get_normal(is_all=False): Gaussians
 * outside get a zero normal and no normal gradient.  cam_center f32[3] and R_w2c f32[3,3] (row-major, the reference's
 * torch.tensor(camera.R.T)) are DEVICE arrays.  aux u8[P] carries the argmin axis / flip sign to the backward.
 * The backward takes the gradients of the five outputs and returns those of the raw parameters (xyz only enters
 * the non-differentiable flip test). */
int vcr_prologue_forward(int P, int n_rest, const float* xyz, const float* features_dc, const float* features_rest,
                         const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                         const uint8_t* valid, const float* cam_center, const float* R_w2c, float* opacities,
                         float* scales, float* rotations, float* shs, float* normals, uint8_t* aux, void* stream);
int vcr_prologue_backward(int P, int n_rest, const float* rotation_raw, const float* opacities, const float* scales,
                          const float* rotations, const uint8_t* aux, const float* R_w2c, const float* dL_dopacities,
                          const float* dL_dscales, const float* dL_drotations, const float* dL_dshs,
                          const float* dL_dnormals, float* dL_dfeatures_dc, float* dL_dfeatures_rest,
                          float* dL_dopacity_raw, float* dL_dscaling_raw, float* dL_drotation_raw, void* stream);

/* SSIM of tools/loss_utils.py:47-92 (11x11 Gaussian window, sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2) over
 * `planes` independent height x width images (planes = batch x channels), fused.
 * forward: *ssim_sum (device double) = sum of the SSIM map over all elements (the caller divides by planes*H*W for
 * size_average=True); dmaps f32[3,planes,H,W] receives the map's partial derivatives for the backward.
 * backward: dL_dimg1 = dL_dssim (device f32 scalar, gradient of the MEAN) * d mean(ssim) / d img1.  img2 is the
 * ground truth and gets no gradient. */
int vcr_ssim_forward(int planes, int height, int width, const float* img1, const float* img2, float* dmaps,
                     double* ssim_sum, void* stream);
int vcr_ssim_backward(int planes, int height, int width, const float* img1, const float* img2, const float* dmaps,
                      const float* dL_dssim, float* dL_dimg1, void* stream);

/* One Adam step (torch.optim.Adam semantics: no weight decay, no amsgrad) over a flat parameter tensor and its flat
 * gradient, with a learning rate per contiguous range — the reference's six per-Gaussian parameter groups
 * (scene/gaussian_model.py:241-258, eps = 1e-15) become six ranges of the view-parallel bucket and ONE launch
 * (SURVEY.md 8f rank 3).  group_end_host[k] is the exclusive end (in floats) of range k, ascending, the last one equal
 * to numel; lr_host[k] its learning rate; `step` counts from 1 (bias correction); every gradient is multiplied by
 * grad_scale first (1/world after a SUM all-reduce).  All four device buffers must be 16-byte aligned. */
int vcr_adam_step(long long numel, float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int n_groups,
                  const long long* group_end_host, const float* lr_host, float beta1, float beta2, float eps, int step,
                  float grad_scale, void* stream);

/* Compact exchange of the SH gradient (view-parallel; no reference counterpart).  For one view the gradient of the
 * loss w.r.t. the SH coefficients of a Gaussian is the outer product of its clamp-masked colour gradient (3 floats) with
 * the SH basis of the direction camera -> Gaussian (computeColorFromSH backward, DGR/cuda_rasterizer/backward.cu:20-139),
 * so the ranks need not all-reduce 48 floats per Gaussian: vcr_rasterize_backward_chunked(..., dL_dcolor_masked) hands out
 * the 3 floats, the ranks all-gather them together with their camera positions, and this call rebuilds
 *   dL_dsh[g] = sum over ranks r of basis(normalize(means3D[g] - campos_r)) (x) dcolor_r[g]
 * on every rank, in rank order.  `slices`: world slices of slice_stride floats, slice r = [ dcolor_r f32[P,3] | campos_r
 * f32[3] | padding ].  dL_dsh f32[P,M,3] is fully written. */
int vcr_sh_grad_from_colors(int P, int D, int M, int world, const float* means3D, const float* slices, long long slice_stride,
                            float* dL_dsh, void* stream);

/* vcr_multimem_allreduce_sum plus, in the same launch, this rank's contribution to an all-gather: the floats
 * [bcast_first, +bcast_count) of the symmetric buffer (local_base = this rank's own mapping of it, bcast_first a
 * multiple of 4) are read locally and multicast to every rank.  Same ordering contract as vcr_multimem_allreduce_sum. */
int vcr_multimem_exchange(void* multicast_base, const float* local_base, int n_ranges, const long long* first_float_host,
                          const long long* n_floats_host, long long bcast_first, long long bcast_count, int rank, int world,
                          int blocks, void* stream);

/* Prune as stream compaction (replaces the ~18 boolean-mask indexing operations of _prune_optimizer,
 * scene/gaussian_model.py:442-460): rows i with keep[i] != 0 of every array a (f32[P, width[a]]) are copied, in order, to
 * dst[a] (row j = the j-th kept row).  src / dst / width are HOST arrays of n_arrays entries (device pointers inside);
 * src and dst must not overlap.  scratch: vcr_compact_scratch_bytes(P) device bytes.  The number of kept rows is written
 * to *n_kept_device (device u32, may be NULL); a caller that needs it on the host before sizing dst counts the mask
 * itself.  One prefix sum + one gather launch per 8 arrays. */
size_t vcr_compact_scratch_bytes(int P);
int vcr_compact_rows(int P, const uint8_t* keep, int n_arrays, const float* const* src_host, float* const* dst_host,
                     const int* width_host, char* scratch, uint32_t* n_kept_device, void* stream);

/* One densification round on the raw parameters (replaces GaussianModel.densify_and_prune, scene/gaussian_model.py:645-659:
 * densify_and_clone :628-643, densify_and_split_along_maxscaling :579-626 and the final prune :651-657).
 * vcr_densify_decide writes one flag byte per Gaussian — bit 0: the Gaussian itself survives (not split, not pruned),
 * bit 1: a clone is appended, bit 2: two children split along the longest axis are appended — from the accumulated
 * statistic grads = xyz_gradient_accum / denom (NaN -> 0) compared with max_grad, max(exp(scaling_raw)) compared with
 * dense_limit (= percent_dense * scene_extent), sigmoid(opacity_raw) < min_opacity, and max scaling > world_limit
 * (= 0.1 * scene_extent when the reference is called with a max_screen_size, +inf otherwise; the reference's screen-size
 * test itself never fires because max_radii2D has been reset by then).  totals_device u32[3] receives the number of
 * surviving originals, clones and split parents.  vcr_densify_fill then writes the n_clone + 2 n_split new rows of the six
 * groups (xyz[3], f_dc[3], f_rest[*], opacity[1], scaling[3], rotation[4]; HOST arrays of device pointers) behind the
 * n_kept rows that vcr_compact_rows(keep = flags & 1) has moved: clones verbatim; children at mean +- dir * 2 smax / 3 with the
 * longest axis shrunk to smax / 1.6, everything else copied.  Row order as in the reference (kept | clones | first
 * children | second children).  scratch: vcr_densify_scratch_bytes(P), shared by the two calls. */
size_t vcr_densify_scratch_bytes(int P);
int vcr_densify_decide(int P, const float* scaling_raw, const float* opacity_raw, const float* xyz_gradient_accum, const float* denom,
                       float max_grad, float min_opacity, float dense_limit, float world_limit, uint8_t* flags, char* scratch,
                       uint32_t* totals_device, void* stream);
int vcr_densify_fill(int P, const uint8_t* flags, const char* scratch, int n_kept, int n_clone, int n_split,
                     const float* const* src_host, float* const* dst_host, const int* width_host, void* stream);

/* Gradient exchange of the view-parallel layer through the NVSwitch (the reference is single-GPU and
 * has no counterpart).  `multicast_base` is the multicast mapping (NVLink SHARP, e.g. torch symmetric memory's
 * multicast_ptr) of a float buffer that every rank holds at the same offset.  For each of the n_ranges (<= 8) float
 * ranges [first_float_host[i], +n_floats_host[i]) this rank reduces its 1/world share with multimem.ld_reduce (the sum is
 * formed inside the switch) and writes it to every rank with multimem.st: a two-shot all-reduce(SUM) in ONE launch, S
 * bytes in and S bytes out per GPU.  The range arrays are HOST arrays, consumed before the call returns.  `blocks` <= 0
 * picks 37.  The launch waits for nobody: the caller must order it after every rank has written the ranges and
 * order consumers after every rank's launch has completed (symmetric-memory barrier on `stream`, see
 * vcr_view_parallel.py).  All ranks must pass identical ranges. */
int vcr_multimem_allreduce_sum(void* multicast_base, int n_ranges, const long long* first_float_host,
                               const long long* n_floats_host, int rank, int world, int blocks, void* stream);

/* present[i] = (view-space z of means3D[i]) > 0.2  (checkFrustum, rasterizer_impl.cu:54-66). */
int vcr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* f_count = 1: RGB image + per-Gaussian contribution count and summed opacity
 * (forwardCount / renderCUDA_count, forward.cu:480-603).  The reference's increments are racy
 * non-atomic read-modify-writes; this library uses atomics, i.e. returns the exact counts. */
int vcr_rasterize_count(
    vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
    int P, int D, int M,
    const float* background, int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy, int prefiltered,
    float* out_color /* f32[3,H,W] */, int* gaussians_count /* i32[P] */, float* important_score /* f32[P] */,
    int* radii, int debug, int* num_rendered_host, void* stream);

/* f_count = 3: mark the first contributing Gaussian of every pixel (forwardVisiAcc /
 * renderCUDA_visi_acc, forward.cu:735-835). */
int vcr_rasterize_visi_acc(
    vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
    int P, int D, int M,
    int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy, int prefiltered,
    int* gaussians_count /* i32[P] */, int* radii, int debug, int* num_rendered_host, void* stream);

/* D-Normal regularizer, fused.  depth f32[H*W]; gt_normal and render_normal f32[H,W,3]
 * (render_normal is the UN-normalised rasterizer output permuted to HWC, or NULL with exp_t <= 0);
 * mask u8[H*W] or NULL (all valid); K is fx, fy, cx, cy.  mask_depth_max > 0 additionally masks
 * pixels with depth >= mask_depth_max (gaussian_renderer/__init__.py:125-131).
 * forward writes est_normal f32[H,W,3] (may be NULL), per-pixel weight f32[H*W] (may be NULL) and
 * sums[3] = { sum_M w*|n-g|_1, sum_M w*(1-n.g), |M| } (sums must be zeroed by the caller or pass
 * zero_sums = 1).  loss = (sums[0] + sums[1]) / sums[2]  (0 if sums[2] == 0).
 * backward writes dL_ddepth f32[H*W] given dL_dloss (DEVICE scalar, so no host sync is needed between
 * autograd nodes), the forward's sums and the same inputs; dL_dest_normal is an optional extra upstream
 * gradient w.r.t. est_normal (e.g. from a consistency or curvature loss). */
int vcr_dnormal_forward(const float* depth, const float* gt_normal, const float* render_normal,
                        const uint8_t* mask, int width, int height,
                        float fx, float fy, float cx, float cy,
                        float exp_t, float mask_depth_max,
                        float* est_normal, float* weight, float* sums, int zero_sums, void* stream);
int vcr_dnormal_backward(const float* depth, const float* gt_normal, const float* render_normal,
                         const uint8_t* mask, int width, int height,
                         float fx, float fy, float cx, float cy,
                         float exp_t, float mask_depth_max,
                         const float* sums, const float* dL_dloss, const float* dL_dest_normal /* optional f32[H,W,3] */,
                         float* dL_ddepth, void* stream);

/* Fused per-view image loss on the rasterizer's output map (outmap f32[10,H,W], planes as in vcr_rasterize_forward):
 *   loss = w_l1 * mean|RGB - gt_image| + w_mono * monosdf_normal_loss(normalize(normal planes), gt_normal)
 *        + w_dn * D-Normal(depth plane, gt_normal, confidence from the normal planes)
 * i.e. l1_loss (tools/loss_utils.py:22-23, trainer.py:236), the mono-normal term (trainer.py:261-264 with
 * gaussian_renderer/__init__.py:133-134) and the depth-normal term (trainer.py:266-280), ~60 PyTorch launches per view in
 * the reference, as one kernel per direction.  gt_image f32[3,H,W], gt_normal f32[H,W,3], mask u8[H*W] or NULL.
 * forward fills sums[8] (zeroed inside): { sum|dRGB|, sum|rn-g|_1, sum(1-rn.g), sum_M w|n-g|_1, sum_M w(1-n.g), |M|, -, - };
 * loss = w_l1*s0/(3N) + w_mono*(s1+s2)/N + w_dn*(s3+s4)/s5 (the last term 0 when s5 == 0).
 * backward writes dL/d outmap f32[10,H,W] in full, given dL_dloss (DEVICE scalar) and the forward's sums. */
int vcr_view_loss_forward(const float* outmap, const float* gt_image, const float* gt_normal, const uint8_t* mask,
                          int width, int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max,
                          float* sums, void* stream);
int vcr_view_loss_backward(const float* outmap, const float* gt_image, const float* gt_normal, const uint8_t* mask,
                           int width, int height, float fx, float fy, float cx, float cy, float exp_t, float mask_depth_max,
                           const float* sums, const float* dL_dloss, float w_l1, float w_mono, float w_dn,
                           float* dL_doutmap, void* stream);

/* distCUDA2: mean squared distance to the 3 nearest neighbours.  scratch must hold
 * vcr_knn_scratch_bytes(P) bytes. */
size_t vcr_knn_scratch_bytes(int P);
int vcr_dist2_knn3(int P, const float* points /* f32[P,3] */, float* mean_dists /* f32[P] */,
                   char* scratch, void* stream);

/* distCUDA2 without caller scratch.  No counterpart in the reference (its SimpleKNN::knn allocates with cudaMalloc and
 * thrust on every call, simple_knn.cu:186-222).  For hosts that cannot hand over device memory at call time — the GPU's
 * free memory is already owned by another allocator, allocation is forbidden after start-up, or the call comes from
 * inside an allocator callback — the scratch is a fixed array inside the library (4 MiB per GPU that loads the library,
 * held until exit).  At most vcr_knn_noalloc_max_points() (65536) points; more is refused with an error.  Allocates and
 * frees nothing.  The first call on each device resolves the array's device address: make it before any stream
 * capture.  Calls share the array, so calls that overlap in time on one GPU must be serialised by the caller. */
int vcr_knn_noalloc_max_points(void);
int vcr_dist2_knn3_noalloc(int P, const float* points /* f32[P,3] */, float* mean_dists /* f32[P] */, void* stream);

/* ---- the reference's compile-time variants (DGR/cuda_rasterizer/config.h:17-18, README.md:150-155) as run-time arguments ----
 * num_dist in {0, 1} = NUM_DIST (a depth-distortion channel, forward.cu:400-412, 464-472), num_depth2 in {0, 2} = NUM_DEPTH2
 * (the sums of NDC depth and of its square over the inside Gaussians, forward.cu:426-430, 473-476; what render() turns into
 * `depth_var`, gaussian_renderer/__init__.py:154-158).  out_color and dL_dpix then have vcr_variant_channels() channels
 * (OUTPUT_CHANNELS, auxiliary.h:28); both variants write from channel 10 on, as in the reference (DISTORTION_OFFSET ==
 * DEPTH2_OFFSET), so with both set channel 10 holds the first depth sum and channel 12 stays zero.  Gradients: the depth sums
 * reach the Gaussians through the depth only (DETACH_WEIGHT = 1, backward.cu:656-664); the distortion channel has none — in
 * the reference as written its backward terms multiply variables that are re-declared in an inner scope and stay zero
 * (backward.cu:489-499).  Not reproduced: with NUM_DIST = 1 the reference also stores two values per pixel behind the end
 * of its final_T array (forward.cu:469-470 against rasterizer_impl.cu:176), over its own n_contrib.  All other arguments as
 * vcr_rasterize_forward / vcr_rasterize_backward.  The variants run on the one-pixel-per-thread kernels. */
int vcr_variant_channels(int num_dist, int num_depth2);
int vcr_rasterize_forward_variant(
    int num_dist, int num_depth2,
    vcr_alloc_fn geometry_buffer, vcr_alloc_fn binning_buffer, vcr_alloc_fn image_buffer, void* alloc_user,
    int P, int D, int M, const float* background, int width, int height, const float* means3D, const float* shs,
    const float* colors_precomp, const float* normals_precomp, const float* sh_objs, const float* opacities, const float* scales,
    float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
    const float* dirs, const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color, int* radii,
    const uint8_t* inside, int debug, int* num_rendered_host, void* stream);
int vcr_rasterize_backward_variant(
    int num_dist, int num_depth2,
    int P, int D, int M, int R, const float* background, int width, int height, const float* means3D, const float* shs,
    const float* colors_precomp, const float* normals_precomp, const float* sh_objs, const float* alpha, const float* depth,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
    const float* projmatrix, const float* dirs, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
    char* geom_buffer, char* binning_buffer, char* img_buffer, const float* dL_dpix, float* dL_dmean2D, float* dL_dmean2D_densify,
    float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
    float* dL_dscale, float* dL_drot, const uint8_t* inside, char* scratch, int debug, void* stream);

/* ---- measurement hooks (bench.py; no counterpart in the reference, which has no per-kernel timers — SURVEY.md §5) ----
 * With profiling enabled every kernel launch of this library is bracketed by CUDA events on its own stream.
 * vcr_profile_read sums the elapsed milliseconds and the number of timed launches per kernel id since the last reset
 * (it synchronises on the recorded events) and reports the launch counters, which are kept even when profiling is off.
 * Not thread-safe. */
int vcr_profile_enable(int enable);
int vcr_profile_kernel_count(void);
const char* vcr_profile_kernel_name(int id);
int vcr_profile_read(float* ms, long long* timed, long long* launches, int reset);
/* Pair counters (SURVEY.md §8d asks for K_e, K_c and sum n_contrib with every result; no counterpart in the reference).
 * While a device array of at least 4 u64 is registered, vcr_rasterize_forward runs an instrumented instantiation of its
 * compositing kernel that ADDS to it: [0] K_e = (pixel, Gaussian) pairs whose exponent was evaluated, [1] K_c = pairs that
 * contributed, [2] sum over pixels of n_contrib, [3] (warp, Gaussian) candidates after culling.  NULL switches it off.
 * Process-wide, not thread-safe; meant for bench.py outside its timed regions. */
int vcr_pair_counters(unsigned long long* counters_device);
/* Block timeline of the two-pixel forward compositing kernel (bench.py VCR_BENCH_TIMELINE; no counterpart in the reference).
 * While a device array of 16 u64 per tile is registered, block b writes at 16b: [0] its start and [1] its end
 * (globaltimer, ns), [2] = (tile << 32) | SM id, [3] = its list length and, while vcr_pair_counters is registered too,
 * [4..7] = the candidates each of its warps evaluated, [8..11] = the cycles each warp waited at the per-batch barrier,
 * [12] = the block's cycles.  NULL switches it off.  Process-wide, not
 * thread-safe, never inside a timed region. */
int vcr_debug_block_timeline(unsigned long long* timeline_device);
/* Test hook for the sort's failure path (no counterpart in the reference, whose CUB sort cannot time out): with on != 0 every
 * multi-block radix pass is made to time out in its look-back (block 0 publishes nothing, 2000 polls).  The affected call
 * produces invalid instance lists; the NEXT vcr_rasterize_forward / _count / _visi_acc on the device (or the same call when
 * debug != 0) returns an error that says so.  Process-wide; for tests only. */
int vcr_debug_radix_fault(int on);
/* Development A/B: 1 selects the one-pixel-per-thread compositing kernels for every call, 2 (the default) the
 * two-pixels-per-thread ones (forward always; backward when inside == NULL).  Returns the version in force. */
int vcr_debug_render_version(int version);

/* ---- test/parity accessors: decode the opaque buffers (used by tests/, never by the product path) ---- */
typedef struct vcr_geom_view {
    float* depths; uint8_t* clamped; float* means2D; float* xyz_view; float* cov3D;
    float* conic_opacity; float* rgb; uint32_t* tiles_touched; uint32_t* point_offsets;
} vcr_geom_view;
typedef struct vcr_binning_view {
    uint32_t* point_list; uint64_t* point_list_keys; uint32_t* point_list_unsorted; uint64_t* point_list_keys_unsorted;
} vcr_binning_view;
typedef struct vcr_image_view { float* accum_alpha; uint32_t* n_contrib; uint32_t* ranges; } vcr_image_view;
void vcr_geom_layout(char* chunk, int P, vcr_geom_view* out);
void vcr_binning_layout(char* chunk, int num_rendered, vcr_binning_view* out);
void vcr_image_layout(char* chunk, int width, int height, vcr_image_view* out);

#ifdef __cplusplus
}
#endif
#endif /* VCR_GAUS_B200_H */
