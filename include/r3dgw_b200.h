/*
 * r3dgw_b200.h - C-ABI of the B200-native (sm_100a) Gaussian-splat rasterizer + SH relighting path.
 *
 * This is the drop-in boundary for the reference's native layer.  Every entry point takes plain
 * device pointers, sizes and a CUDA stream (passed as void* == cudaStream_t); no torch types.
 * All functions return 0 on success, a positive cudaError_t value on a CUDA failure, or a negative
 * R3DGW_ERR_* code on misuse.  Nothing is retained across calls except what the caller-owned state
 * buffers hold (same ownership model as the reference: the caller allocates through callbacks and
 * hands the opaque buffers back to the backward call).
 *
 * Reference interfaces replaced (paths relative to /root/reference/submodules):
 *   r3dgw_rasterize_forward   <- CudaRasterizer::Rasterizer::forward   diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:34-57
 *   r3dgw_rasterize_backward  <- CudaRasterizer::Rasterizer::backward  diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:59-86
 *   r3dgw_mark_visible        <- CudaRasterizer::Rasterizer::markVisible diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:27-32
 *   r3dgw_knn_dist2           <- SimpleKNN::knn                         simple-knn/simple_knn.h:18 (simple_knn.cu:185-221)
 *   r3dgw_shade_forward/backward <- EnvironmentLight.shade + sky eval_sh  scene/NVDIFFREC/light.py:131-193,
 *                                   gaussian_renderer/__init__.py:132-156 (torch graph in the reference)
 * The staged entry points (geometry / blend) are the same pipeline split at the point where the
 * reference's six per-render calls repeat identical work (gaussian_renderer/__init__.py:160-264).
 */
#ifndef R3DGW_B200_H_INCLUDED
#define R3DGW_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define R3DGW_ERR_BAD_ARG      (-1)  /* null pointer / negative size / bad combination            */
#define R3DGW_ERR_UNSUPPORTED  (-2)  /* channel count not instantiated, SH degree > 3, ...        */
#define R3DGW_ERR_ALLOC        (-3)  /* an allocation callback returned NULL                      */
#define R3DGW_ERR_OVERFLOW     (-4)  /* num_rendered would exceed 2^31-1 (reference cap, 32-bit)  */
#define R3DGW_ERR_TOO_LARGE    (-5)  /* size beyond the static scratch of an allocation-free entry point */

#define R3DGW_MAX_CHANNELS 32

/* Allocation callback: must return a device pointer to >= bytes bytes, 256-B aligned, that stays
 * valid until the matching backward call has returned (reference: resizeFunctional,
 * diff-gaussian-rasterization/rasterize_points.cu:27-33). */
typedef char* (*r3dgw_alloc_fn)(void* ctx, size_t bytes);

/* Camera / raster settings shared by every stage (reference: the scalar tail of
 * Rasterizer::forward's argument list; GaussianRasterizationSettings,
 * diff_gaussian_rasterization/__init__.py:133-144).  All pointers are DEVICE pointers. */
typedef struct r3dgw_camera {
    int32_t width, height;
    float tan_fovx, tan_fovy;
    float scale_modifier;
    int32_t prefiltered;
    const float* viewmatrix;   /* 16 floats, column-major (already transposed by the caller)  */
    const float* projmatrix;   /* 16 floats, column-major                                     */
    const float* cam_pos;      /* 3 floats                                                    */
} r3dgw_camera;

/* Byte offsets of the sub-arrays inside the opaque state buffers, for tests and tooling. */
typedef struct r3dgw_geom_layout {
    size_t depths;         /* f32[P]     view-space z                                   */
    size_t means2D;        /* float2[P]  pixel-space centre                             */
    size_t conic_opacity;  /* float4[P]  (conic.x, conic.y, conic.z, opacity)           */
    size_t cov3D;          /* f32[6P]                                                   */
    size_t tiles_touched;  /* u32[P]                                                    */
    size_t point_offsets;  /* u32[P]     inclusive prefix sum of tiles_touched          */
    size_t rgb;            /* f32[3P]    SH->RGB result (only when shs are given)       */
    size_t clamped;        /* u8[3P]                                                    */
    size_t scan_state;     /* u64[blocks] decoupled look-back tile states               */
    size_t depth_sort;     /* per-Gaussian depth sort of the depth-first binning: u32 keys[P], keys_alt[P], ids[P],
                              ids_alt[P], radix workspace, K2' look-back states            */
    size_t total;
} r3dgw_geom_layout;

typedef struct r3dgw_binning_layout {
    size_t point_list;          /* u32[R]   sorted Gaussian ids                          */
    size_t point_list_keys;     /* u64[R]   sorted (tile<<32 | depth bits) keys - filled on demand by
                                             r3dgw_materialize_sorted_keys (the pipeline sorts 32-bit tile ids) */
    size_t point_list_alt;      /* u32[R]   ping-pong value buffer                       */
    size_t point_list_keys_alt; /* u32[R] x 2 tile-id keys (sorted result in the first half) */
    size_t ranges;              /* uint2[tiles]                                          */
    size_t splats;              /* 2 x float4[R] sorted-order splat stream */
    size_t sort_hist;           /* u32 radix histograms + look-back state                */
    size_t total;
} r3dgw_binning_layout;

typedef struct r3dgw_image_layout {
    size_t final_T;    /* f32[N] */
    size_t n_contrib;  /* u32[N] */
    size_t total;
} r3dgw_image_layout;

int r3dgw_get_geom_layout(int P, int has_sh, r3dgw_geom_layout* out);
int r3dgw_get_binning_layout(int R, int width, int height, r3dgw_binning_layout* out);
int r3dgw_get_image_layout(int width, int height, r3dgw_image_layout* out);

/* Version / build info: returns a static string such as "r3dgw_b200 0.1 sm_100a". */
const char* r3dgw_version(void);
/* Number of kernel launches issued by this library in this process so far. */
uint64_t r3dgw_launch_count(void);

/* Optional per-kernel device timing (CUDA events recorded on the caller's stream around every launch).
 * Off by default.  r3dgw_profile_read synchronises the recorded events, adds them to per-kernel totals and
 * copies totals (ms) / launch counts into arrays of r3dgw_profile_kernel_count() entries. */
void r3dgw_profile_enable(int on);
int r3dgw_profile_kernel_count(void);
const char* r3dgw_profile_kernel_name(int id);
int r3dgw_profile_read(double* total_ms, uint64_t* launches, int reset);

/* ---- staged pipeline ------------------------------------------------------------------------ */

/* K1 preprocess (+fused prefix sum) -> per-Gaussian depth sort (P 32-bit keys) -> D2H of num_rendered -> K2'
 * duplicates in depth order (tile-id keys) -> stable onesweep radix sort on the tile bits only -> K3 tile ranges +
 * splat-stream gather.  Same sorted order as the reference's single (tile|depth) sort.
 * Reference: rasterizer_impl.cu:216-318. */
int r3dgw_geometry_forward(
    r3dgw_alloc_fn geom_alloc, void* geom_ctx,
    r3dgw_alloc_fn binning_alloc, void* binning_ctx,
    int P, int D, int M,
    const r3dgw_camera* cam,
    const float* means3D, const float* shs, const float* opacities,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    int* radii,            /* out, i32[P] */
    int* num_rendered,     /* out, host   */
    void* stream);

/* Rebuilds the reference's sorted 64-bit keys (tile << 32 | depth bits, rasterizer_impl.cu:70-111, 303-308) in
 * binning.point_list_keys from the sorted tile ids, the sorted point list and the depths (tests / tooling). */
int r3dgw_materialize_sorted_keys(int P, int R, int width, int height, const char* geom_buffer,
                                  char* binning_buffer, void* stream);

/* K4: front-to-back blend of C-channel per-Gaussian features over the cached geometry.
 * features == NULL selects the SH->RGB colours in the geometry buffer (C must be 3).
 * Reference: FORWARD::render, forward.cu:261-374. */
int r3dgw_blend_forward(
    int P, int R, int C,
    const r3dgw_camera* cam,
    const char* geom_buffer, const char* binning_buffer, char* image_buffer,
    const float* features,   /* f32[P,C] */
    const float* background, /* f32[C]   */
    float* out_color,        /* f32[C,H,W] */
    void* stream);

/* K4 with an alpha cache, for repeated blends over ONE geometry (the six raster calls of render(),
 * gaussian_renderer/__init__.py:160-264; one view x many environment maps in the relit-eval grid).
 *   cache_mode 0: plain K4 (alpha_cache ignored) == r3dgw_blend_forward
 *   cache_mode 1: plain K4 that also records, per (sorted list entry, 8x4-pixel warp footprint) that survives the
 *                 footprint cull, the 32 pixels' effective alphas (0 = did not blend) into alpha_cache
 *   cache_mode 2: replays alpha_cache instead of re-evaluating conic / exp / thresholds; image_buffer must be
 *                 the one a mode-0/1 call on the same geometry filled (final_T, n_contrib are read, not written).
 * Output is bit-identical in all three modes.  alpha_cache: r3dgw_alpha_cache_bytes(R) bytes (R * 1 KiB, only the
 * surviving rows are ever touched), caller-owned device memory. */
size_t r3dgw_alpha_cache_bytes(int R);
int r3dgw_blend_forward_cached(
    int P, int R, int C,
    const r3dgw_camera* cam,
    const char* geom_buffer, const char* binning_buffer, char* image_buffer,
    const float* features, const float* background, float* out_color,
    float* alpha_cache, int cache_mode,
    void* stream);

/* K5: backward of the blend; ACCUMULATES (+=) into dL_dmean2D[P,3], dL_dconic[P,4],
 * dL_dopacity[P], dL_dfeatures[P,C] (caller zeroes them).  Reference: backward.cu:399-557.
 * dL_dfeatures may be NULL: geometry-only pass (see r3dgw_blend_backward_color_cached). */
int r3dgw_blend_backward(
    int P, int R, int C,
    const r3dgw_camera* cam,
    const char* geom_buffer, const char* binning_buffer, const char* image_buffer,
    const float* features, const float* background,
    const float* dL_dpixels,  /* f32[C,H,W] */
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dfeatures,
    void* stream);

/* Deferred backward for repeated blends over one geometry (the 5-6 raster backwards of one render()):
 *   per call   r3dgw_blend_backward_color_cached : dL/dfeatures[P,C] (+=) only, replayed front to back from the alpha
 *              cache a mode-1 r3dgw_blend_forward_cached call recorded (C <= 4);
 *   once       r3dgw_blend_backward with dL_dfeatures == NULL over the CONCATENATED features [P, sum C], background
 *              [sum C] and pixel gradients [sum C, H, W] of all calls: the geometry gradients (mean2D, conic, opacity)
 *              of the whole frame in one walk of the tile lists instead of one per call.
 * The sums are the ones the reference accumulates call by call (backward.cu:399-557); only the atomic order differs. */
int r3dgw_blend_backward_color_cached(
    int P, int R, int C,
    const r3dgw_camera* cam,
    const char* binning_buffer, const char* image_buffer,
    const float* alpha_cache,
    const float* dL_dpixels,   /* f32[C,H,W] */
    float* dL_dfeatures,       /* f32[P,grad_stride], accumulated */
    int grad_stride,           /* floats per Gaussian row of dL_dfeatures: C, or 4 (16-B aligned rows: the C sums of a
                                * (warp, Gaussian) pair leave as ONE 16-byte reduction instead of C 4-byte ones) */
    void* stream);

/* Geometry gradients (dL/dmean2D [P,3], dL/dconic [P,4], dL/dopacity [P]; ACCUMULATED, zero them first) of one or
 * several blends over one geometry, from the alpha cache, walking FRONT TO BACK: the reference's suffix sums
 * (backward.cu:464-520) are the pixel's <forward output, dL/dpixel> minus a running prefix, so no conic / exp / T
 * division is re-evaluated.  features f32[P,C], out_images f32[C,H,W] = the forward outputs of those blends
 * (background included), background f32[C], dL_dpixels f32[C,H,W]; C in {3,4,6,8,9,12,15,16} (several calls
 * concatenated along C). */
int r3dgw_blend_backward_geometry_cached(int P, int R, int C, const r3dgw_camera* cam, const char* binning_buffer,
                                         const char* image_buffer, const float* alpha_cache, const float* features,
                                         const float* out_images, const float* background, const float* dL_dpixels,
                                         float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, void* stream);
/* The same pass over n_calls blend calls whose tensors stay where they are (no concatenation copies): channels[i]
 * (host array; their sum must be one of the totals above, <= 16), and HOST arrays of n_calls DEVICE pointers:
 * features[i] f32[P, channels[i]], out_images[i] / dL_dpixels[i] f32[channels[i], H, W], backgrounds[i]
 * f32[channels[i]].  This is how the six-call pattern of render() (gaussian_renderer/__init__.py:160-264) reaches
 * the pass: one geometry-gradient walk for the five differentiated calls. */
int r3dgw_blend_backward_geometry_cached_calls(int P, int R, int n_calls, const int* channels, const r3dgw_camera* cam,
                                               const char* binning_buffer, const char* image_buffer,
                                               const float* alpha_cache, const float* const* features,
                                               const float* const* out_images, const float* const* backgrounds,
                                               const float* const* dL_dpixels, float* dL_dmean2D, float* dL_dconic,
                                               float* dL_dopacity, void* stream);

/* K6+K7 fused: conic/mean2D gradients -> means3D, cov3D, scales, rotations (and SH).
 * Overwrites dL_dmean3D/dL_dcov3D/dL_dscale/dL_drot rows of visible Gaussians (caller zeroes).
 * Reference: BACKWARD::preprocess, backward.cu:144-396, 560-624. */
int r3dgw_geometry_backward(
    int P, int D, int M,
    const r3dgw_camera* cam,
    const float* means3D, const float* shs, const float* scales, const float* rotations,
    const float* cov3D_precomp, const int* radii,
    const char* geom_buffer,
    const float* dL_dmean2D, const float* dL_dconic,
    float* dL_dmean3D, float* dL_dcolor, float* dL_dcov3D, float* dL_dsh,
    float* dL_dscale, float* dL_drot,
    void* stream);

/* ---- reference-shaped one-call entry points ------------------------------------------------- */

int r3dgw_rasterize_forward(
    r3dgw_alloc_fn geom_alloc, void* geom_ctx,
    r3dgw_alloc_fn binning_alloc, void* binning_ctx,
    r3dgw_alloc_fn image_alloc, void* image_ctx,
    int P, int D, int M,
    const float* background, const r3dgw_camera* cam,
    const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    float* out_color, int* radii, int* num_rendered, void* stream);

int r3dgw_rasterize_backward(
    int P, int D, int M, int R,
    const float* background, const r3dgw_camera* cam,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, const float* rotations, const float* cov3D_precomp,
    const int* radii,
    const char* geom_buffer, const char* binning_buffer, const char* image_buffer,
    const float* dL_dpix,
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
    float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
    void* stream);

int r3dgw_mark_visible(int P, const float* means3D, const float* viewmatrix,
                       const float* projmatrix, uint8_t* present, void* stream);

/* ---- simple-knn ------------------------------------------------------------------------------ */
/* workspace: query bytes with r3dgw_knn_workspace_bytes(P); device memory owned by the caller. */
size_t r3dgw_knn_workspace_bytes(int P);
int r3dgw_knn_dist2(int P, const float* points /* f32[P,3] */, float* mean_dist2 /* f32[P] */,
                    char* workspace, void* stream);

/* ---- fused per-Gaussian relit shading ---------------------------------------------------------
 * One thread per Gaussian.  For foreground Gaussians (is_sky[i]==0):
 *   EnvironmentLight.shade (light.py:131-193): order-2 SH irradiance, optional SH-domain blurred
 *   specular with the split-sum FG LUT, Fresnel mix, gamma.
 * For sky Gaussians: clamp_min(eval_sh(1, sky_sh, dir) + 0.5, 0) or ones (fix_sky)
 *   (gaussian_renderer/__init__.py:143-148).
 * Material arrays (albedo/roughness/metalness) are indexed by FOREGROUND rank via fg_index[i]
 * (the reference stores them for foreground Gaussians only).  fg_index may be NULL (identity).
 * Outputs f32[P,3] each: rgb, diffuse, specular (sky rows of diffuse/specular are zero). */
typedef struct r3dgw_shade_args {
    int32_t P;
    int32_t specular;      /* 0: Lambertian only (light.py:160-163) */
    int32_t fix_sky;
    int32_t sh_degree;     /* env SH degree (<=4 on the fused path) */
    const float* positions;   /* f32[P,3] */
    const float* normals;     /* f32[P,3] */
    const float* albedo;      /* f32[Pfg,3] */
    const float* roughness;   /* f32[Pfg]   */
    const float* metalness;   /* f32[Pfg] or NULL */
    const uint8_t* is_sky;    /* u8[P] or NULL */
    const int32_t* fg_index;  /* i32[P] or NULL */
    const float* cam_pos;     /* f32[3] device */
    const float* env_sh;      /* f32[(deg+1)^2,3] device */
    const float* sky_sh;      /* f32[4,3] device ([coef][channel]) or NULL */
    const float* fg_lut;      /* f32[256,256,2] device */
} r3dgw_shade_args;

int r3dgw_shade_forward(const r3dgw_shade_args* a, float* rgb, float* diffuse, float* specular,
                        void* stream);
/* Gradients w.r.t. normals, albedo, roughness, metalness, positions (through wo / sky dir),
 * env_sh and sky_sh.  env/sky SH gradients are ACCUMULATED (+=) with atomics; the per-Gaussian
 * ones are overwritten. Any output pointer may be NULL to skip it. */
int r3dgw_shade_backward(const r3dgw_shade_args* a,
                         const float* dL_drgb, const float* dL_ddiffuse, const float* dL_dspecular,
                         float* dL_dpositions, float* dL_dnormals, float* dL_dalbedo,
                         float* dL_droughness, float* dL_dmetalness,
                         float* dL_denv_sh, float* dL_dsky_sh, void* stream);

/* ---- pre-shading fusion ------------------------------------------------
 * The per-Gaussian caller graph of render() in front of the shading, fused into the shading kernels:
 *   GaussianModel.get_normal      scene/gaussian_model.py:115-122 (build_rotation / get_minimum_axis /
 *                                 flip_align_view, utils/general_utils.py:98-119, 158-170)
 *   dir_pp_normalized             gaussian_renderer/__init__.py:127-129 (safe_normalize(xyz - campos))
 *   GaussianModel.get_depth       scene/gaussian_model.py:125-130 (+ .repeat(1,3), render():182-183)
 *   normal "colour" 0.5 n + 0.5   gaussian_renderer/__init__.py:187-188
 * shade.normals is ignored (may be NULL): the normal is the column of R(q / |q|) at argmin(scale), flipped
 * towards the camera.  Outputs (and upstream gradients) are addressed as base[stride * i + c]: five dense
 * [P,3] arrays (stride 3) for the six-call drop-in pattern, or columns of ONE [P,16] feature matrix
 * (rgb 0-2, diffuse 3-5, specular 6-8, depth 9-11, normal 12-14, ones 15; stride 16) for the single-pass
 * r3dgw_blend_forward with C = 16.  depth3 / normal01 / ones / normals may be NULL.  normals: dense [P,3]. */
typedef struct r3dgw_preshade_args {
    r3dgw_shade_args shade;
    const float* scales;      /* f32[P,3] post-activation */
    const float* rotations;   /* f32[P,4] (r,x,y,z) */
    const float* viewmatrix;  /* f32[16] world_view_transform as stored (transposed), device */
} r3dgw_preshade_args;
int r3dgw_preshade_forward(const r3dgw_preshade_args* a, int out_stride, float* rgb, float* diffuse,
                           float* specular, float* depth3, float* normal01, float* ones, float* normals,
                           void* stream);
/* dL_drotations includes the Jacobian of build_rotation's normalisation; scales receive no gradient (argmin). */
int r3dgw_preshade_backward(const r3dgw_preshade_args* a, int g_stride, const float* dL_drgb,
                            const float* dL_ddiffuse, const float* dL_dspecular, const float* dL_ddepth3,
                            const float* dL_dnormal01, float* dL_dpositions, float* dL_drotations,
                            float* dL_dalbedo, float* dL_droughness, float* dL_dmetalness,
                            float* dL_denv_sh, float* dL_dsky_sh, void* stream);
/* GaussianModel.get_normal alone (cam_pos == NULL: no flip, as get_normal(dir_pp_normalized=None)). */
int r3dgw_gaussian_normal_forward(int P, const float* scales, const float* rotations, const float* positions,
                                  const float* cam_pos, float* normals, void* stream);
int r3dgw_gaussian_normal_backward(int P, const float* scales, const float* rotations, const float* positions,
                                   const float* cam_pos, const float* dL_dnormals, float* dL_drotations,
                                   void* stream);

/* ---- per-Gaussian regularisers of the training step ------------------------------
 *   min_scale_loss        utils/loss_utils.py:210-220 (mean over visible foreground Gaussians of min(scale))
 *   depth_loss_gaussians  utils/loss_utils.py:140-148 (exp(-gamma (mean depth visible sky - mean depth
 *                         visible foreground [detached])))
 * visible = radii > 0.  out: device f32[8] = { scale_loss, depth_loss, count fg, count sky, mean depth sky,
 * mean depth fg, count visible, 0 }.  Fixed-order reduction (bit-reproducible), no host round trip.
 * backward: grads = device f32[2] { dL/dscale_loss, dL/ddepth_loss }; both outputs are fully overwritten. */
size_t r3dgw_gaussian_reg_workspace_bytes(int P);
int r3dgw_gaussian_regularisers_forward(int P, const int* radii, const uint8_t* is_sky, const float* scales,
                                        const float* means3D, const float* viewmatrix, float gamma,
                                        char* workspace, float* out, void* stream);
int r3dgw_gaussian_regularisers_backward(int P, const int* radii, const uint8_t* is_sky, const float* scales,
                                         const float* viewmatrix, float gamma, const float* out,
                                         const float* grads, float* dL_dscales, float* dL_dmeans3D, void* stream);
/* Allocation-free variant of r3dgw_gaussian_regularisers_forward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs, same `out`, same kernels and reduction order (results are
 * bit-identical to the workspace entry point); the per-CTA partial sums live in a file-scope
 * `static __device__ __align__(256)` array of 28 416 bytes that every GPU loading the library holds until exit,
 * resolved once per device with cudaGetSymbolAddress (that device must be current at the call, as for every entry
 * point here).  Calls no cudaMalloc / cudaFree / allocating library routine and ignores r3dgw_profile_enable.
 * Limit: P <= r3dgw_gaussian_regularisers_noalloc_max_gaussians() = 16 777 216 Gaussians (2^24, up to which the f32 counts
 * in out[2], out[3], out[6] are exact); a larger P returns R3DGW_ERR_TOO_LARGE (and says so on stderr) without
 * launching anything.  Every call on a device shares the one array, which keeps the last call's contents (nothing
 * in it has to start at zero: each call overwrites every slot it later reads): calls that OVERLAP on the GPU
 * (different streams without an ordering between them) RACE - serialise them. */
int r3dgw_gaussian_regularisers_noalloc_max_gaussians(void);
int r3dgw_gaussian_regularisers_forward_noalloc(int P, const int* radii, const uint8_t* is_sky, const float* scales,
                                                const float* means3D, const float* viewmatrix, float gamma,
                                                float* out, void* stream);

/* ---- image-space post-ops of the training step ---------------------------------
 * Replace torch graphs in the reference (paths relative to /root/reference):
 *   r3dgw_ssim_l1_forward/backward   <- utils/loss_utils.py:27-36 (l1_loss) + :41-99 (ssim / _ssim, 11x11 Gaussian
 *                                       window, sigma 1.5, zero padding, mask-weighted mean over mask == 1)
 *   r3dgw_masked_l1_forward/backward <- utils/loss_utils.py:27-36 with mask or (1 - mask) (train.py:86-87)
 *   r3dgw_depth_to_normal_forward/backward <- utils/graphics_utils.py:141-169 (depths_to_points + depth_to_normal)
 * Images are f32[C,H,W], contiguous.  mask is f32 with channel stride mask_channel_stride elements (0 for a
 * [1,H,W] mask broadcast over channels, H*W for a full [C,H,W] mask) or NULL (all ones).
 * out / sums: device f32[4] = { mean of ssim_map over mask==1 (1 if none), masked L1 mean (0 if none),
 * count(mask == 1), 1/count (0 if none) }.  workspace: r3dgw_image_loss_workspace_bytes(C,H,W) bytes.
 * The per-block partial sums are reduced in a fixed order: results are bit-reproducible. */
size_t r3dgw_image_loss_workspace_bytes(int C, int H, int W);
int r3dgw_ssim_l1_forward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                          size_t mask_channel_stride,
                          float* dmap,      /* f32[3,C,H,W] saved for backward, or NULL when no gradient is needed */
                          char* workspace, float* out, void* stream);
/* dL/dimg1 = coef[0] * d(ssim mean)/dimg1 + coef[1] * d(l1 mean)/dimg1; coef: device f32[2] (upstream grads). */
int r3dgw_ssim_l1_backward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                           size_t mask_channel_stride, const float* dmap, const float* sums, const float* coef,
                           float* dL_dimg1, void* stream);
/* target may be NULL (zeros); complement != 0 uses (1 - mask).  out[0] is unused (0). */
int r3dgw_masked_l1_forward(int C, int H, int W, const float* x, const float* target, const float* mask,
                            size_t mask_channel_stride, int complement, char* workspace, float* out, void* stream);
/* dL/dx = coef[0] * sign(x*m - t*m) * m / count; coef: device f32[1]. */
int r3dgw_masked_l1_backward(int C, int H, int W, const float* x, const float* target, const float* mask,
                             size_t mask_channel_stride, int complement, const float* sums, const float* coef,
                             float* dL_dx, void* stream);
/* depth f32[H,W]; rays: device f32[21] = intrins.inverse().T (3x3 row-major), c2w[:3,:3].T (3x3 row-major),
 * rays_o[3] (graphics_utils.py:142-155); normal f32[H,W,3] (zero on the 1-pixel border). */
int r3dgw_depth_to_normal_forward(int H, int W, const float* depth, const float* rays, float* normal, void* stream);
int r3dgw_depth_to_normal_backward(int H, int W, const float* depth, const float* rays, const float* dL_dnormal,
                                   float* dL_ddepth, void* stream);

/* image-space tail of render() (gaussian_renderer/__init__.py:231-237, 266-271), one kernel each way:
 *   normal     = ((n01 - 0.5) * 2) * sky + (1 - sky)            n01: rasterized 0.5 n + 0.5 image, f32[3,H,W]
 *   normal_ref = depth_to_normal(depth * sky) * alpha + (1 - sky)  depth: f32[H,W] (channel 0 of the depth image);
 *                                                                  alpha: f32 planes at alpha_channel_stride elements
 *                                                                  (H*W for a [3,H,W] image, 0 for one [H,W] plane)
 * rays: the 21 floats of r3dgw_depth_to_normal_*.  alpha is treated as a constant (the reference detaches it).
 * backward: any of dL_dnormal / dL_dnormal_ref may be NULL (zero); outputs dL_dn01 f32[3,H,W], dL_ddepth f32[H,W]. */
int r3dgw_normal_tail_forward(int H, int W, const float* n01, const float* depth, const float* alpha,
                              size_t alpha_channel_stride, const float* sky, const float* rays, float* normal,
                              float* normal_ref, void* stream);
int r3dgw_normal_tail_backward(int H, int W, const float* depth, const float* alpha, size_t alpha_channel_stride,
                               const float* sky, const float* rays, const float* dL_dnormal, const float* dL_dnormal_ref,
                               float* dL_dn01, float* dL_ddepth, void* stream);
/* normal-consistency term (train.py:92-96): out[0] = mean over the H*W pixels of
 * 1 - sum_c (normal_c occl sky)(normal_ref_c occl sky); occl / sky f32[H,W] or NULL (ones); fixed-order reduction.
 * workspace: r3dgw_image_loss_workspace_bytes(1,H,W).  backward: coef = device f32[1] upstream gradient. */
int r3dgw_normal_loss_forward(int H, int W, const float* normal, const float* normal_ref, const float* occl,
                              const float* sky, char* workspace, float* out, void* stream);
int r3dgw_normal_loss_backward(int H, int W, const float* normal, const float* normal_ref, const float* occl,
                               const float* sky, const float* coef, float* dL_dnormal, float* dL_dnormal_ref,
                               void* stream);

/* ---- host side of the eval grid on the device -----------------------------------
 * Replace numpy / cv2 / spaudiopy host code in the reference (paths relative to /root/reference):
 *   r3dgw_envmap_resize_cubic <- utils/sh_additional_utils.py:48-70 (resize_image -> cv2.resize INTER_CUBIC, the
 *                                (1000, 500) down-scale of :181-184)
 *   r3dgw_envmap_sh_project   <- utils/sh_additional_utils.py:176-206 (get_coefficients_from_image: real SH basis
 *                                of utils/spherical_harmonics.py:148-216 x per-row solid angle, summed per channel)
 *   r3dgw_sh_rotate_sweep     <- spaudiopy.sph.rotate_sh(F, yaw, pitch, roll, 'real') applied twice per sun angle,
 *                                eval_with_gt_envmaps.py:95-100 / relit_novel_view.py:72-74 (spaudiopy==0.1.6,
 *                                environment.yml:239; Ivanic-Ruedenberg recursion)
 * The source image is EITHER f32 (f32 != NULL) OR u8 (u8 != NULL), [H,W,stride] with stride >= 3 channels per pixel
 * (extra channels ignored).  For u8 the value is (float)u8 / 255.f and, when boost != 0, values > threshold are
 * multiplied by scale in f32 (process_environment_map_image, eval_with_gt_envmaps.py:42-49).
 * Projection: H must equal W/2 (the reference's yres = int(xres/2)); l_max <= 4; coeffs: device f64[(l_max+1)^2, 3];
 * workspace: r3dgw_envmap_sh_workspace_bytes(H) bytes.  Row partials are reduced in a fixed order.
 * Rotation sweep: coeffs device f64[n, C] (n = (l_max+1)^2, C <= 4); first_ypr: device f64[3] (yaw, pitch, roll) of
 * a rotation applied to every entry first, or NULL; angles_ypr: device f64[B,3]; out32 f32[B,n,C] and/or out64
 * f64[B,n,C] (either may be NULL). */
size_t r3dgw_envmap_sh_workspace_bytes(int H);
int r3dgw_envmap_resize_cubic(int H, int W, int stride, const float* f32, const uint8_t* u8, int boost, float threshold,
                              float scale, int Hd, int Wd, float* dst /* f32[Hd,Wd,3] */, void* stream);
int r3dgw_envmap_sh_project(int H, int W, int stride, const float* f32, const uint8_t* u8, int boost, float threshold,
                            float scale, int l_max, char* workspace, double* coeffs, void* stream);
/* EnvironmentLight.render_sh's sh_render (utils/sh_additional_utils.py:288-299): coeffs f64[(l_max+1)^2,3] ->
 * out f64[W/2, W, 3], the band-weighted (get_diffuse_coefficients, :267-280) SH reconstruction, linear (no gamma). */
int r3dgw_envmap_sh_render(int l_max, const double* coeffs, int W, double* out, void* stream);
int r3dgw_sh_rotate_sweep(int l_max, int C, const double* coeffs, const double* first_ypr, int B,
                          const double* angles_ypr, float* out32, double* out64, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* R3DGW_B200_H_INCLUDED */
