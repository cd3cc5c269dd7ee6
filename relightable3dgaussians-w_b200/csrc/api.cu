// api.cu - C-ABI entry points (include/r3dgw_b200.h) and stage orchestration.
// Reference orchestration being replaced: CudaRasterizer::Rasterizer::forward / backward,
// diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:198-433.
#include <algorithm>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "common.cuh"
#include "geometry.cu"

namespace r3dgw {

std::atomic<uint64_t> g_launches{0};

// ---- per-kernel timing registry ---------------------------------------------------------------------
bool g_profile_on = false;
int g_sort_prof_kid = -1;
namespace {
struct ProfSpan { int id; cudaEvent_t e0, e1; };
std::mutex g_prof_mu;
std::vector<ProfSpan> g_prof_open;     // spans recorded since the last collect
std::vector<cudaEvent_t> g_prof_pool;
double g_prof_ms[KID_COUNT] = {0};
uint64_t g_prof_n[KID_COUNT] = {0};
const char* const g_prof_names[KID_COUNT] = {
    "preprocess_scan", "duplicate_keys", "sort_histogram", "sort_scan", "sort_onesweep_pass", "ranges_gather",
    "blend_forward", "blend_backward", "geometry_backward", "shade_forward", "shade_backward", "knn", "mark_visible",
    "ssim_l1_forward", "ssim_l1_backward", "masked_l1", "depth_to_normal", "blend_forward_cached",
    "blend_backward_geometry", "blend_backward_color_cached", "depth_sort",
    "envmap_sh_project", "sh_rotate_sweep", "gaussian_regularisers"};
cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
}  // namespace
void prof_mark(int id, cudaStream_t s, bool end) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!end) {
        ProfSpan sp{id, prof_event(), prof_event()};
        cudaEventRecord(sp.e0, s);
        g_prof_open.push_back(sp);
    } else {
        for (size_t i = g_prof_open.size(); i-- > 0;)
            if (g_prof_open[i].id == id) { cudaEventRecord(g_prof_open[i].e1, s); break; }
    }
}

int blend_forward_dispatch(int C, dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                           const float* features, const float* bg, float* final_T, uint32_t* n_contrib,
                           float* out_color, float* alpha_cache, int cache_mode, cudaStream_t stream);
int blend_backward_dispatch(int C, dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                            const float* bg, const float* features, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* dL_dmean2D,
                            float* dL_dconic, float* dL_dopacity, float* dL_dfeatures, cudaStream_t stream);
int blend_backward_color_cached_dispatch(int C, dim3 grid, const uint2* ranges, const float4* B, int W, int H,
                                         const uint32_t* n_contrib, const float* alpha_cache, const float* dL_dpix,
                                         float* dL_dfeatures, int grad_stride, cudaStream_t stream);
int blend_backward_geometry_cached_dispatch(int n_calls, const int* channels, dim3 grid, const uint2* ranges, const float4* A,
                                            const float4* B, int W, int H, const float* const* features,
                                            const uint32_t* n_contrib, const float* alpha_cache,
                                            const float* const* out_images, const float* const* bg,
                                            const float* const* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                                            float* dL_dopacity, cudaStream_t stream);
int sort_num_passes(int end_bit);

static inline int pre_blocks(int P) { return (P + PRE_THREADS - 1) / PRE_THREADS; }

// ---- early read-back of R ------------------------------------------------------------------------------
// R (num_rendered) is final when preprocess_scan_kernel ends, but the depth sort is enqueued behind it on the caller's
// stream, and a read-back ordered on that stream only returns once the sort has drained - the host then allocates
// the binning buffers and enqueues duplicate / sort / ranges / blend into an idle GPU (~30 us per geometry).  The
// 4-byte copy therefore runs on a private non-blocking stream behind an event recorded right after the preprocess
// launch, into pinned memory: the host learns R while the sort still runs and the next kernels queue up behind it.
// One side stream per device and a pool of (event, pinned word) slots, created on first use; a call owns its slot from
// the record to the read.  R3DGW_EARLY_R=0 restores the read-back on the caller's stream.
namespace {
struct EarlyRSlot { cudaEvent_t ev = nullptr; uint32_t* pinned = nullptr; };
struct EarlyRDev { cudaStream_t side = nullptr; std::vector<EarlyRSlot> free_slots; };
EarlyRDev g_early_r[64];
std::mutex g_early_r_mu;
}  // namespace

// takes a (event, pinned word) slot of the current device and records "preprocess done" on the caller's stream;
// call right after the preprocess launch.  Every call owns its slot until read_R_early returns it, so concurrent
// callers (other host threads / streams) never wait on each other's events.
static int mark_R_ready(cudaStream_t stream, EarlyRSlot* slot, int* dev_out) {
    int dev = -1;
    R3DGW_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return R3DGW_ERR_UNSUPPORTED;
    *dev_out = dev;
    {
        std::lock_guard<std::mutex> lk(g_early_r_mu);
        EarlyRDev& d = g_early_r[dev];
        if (!d.side) R3DGW_CUDA_TRY(cudaStreamCreateWithFlags(&d.side, cudaStreamNonBlocking));
        if (!d.free_slots.empty()) { *slot = d.free_slots.back(); d.free_slots.pop_back(); }
    }
    if (!slot->ev) {
        R3DGW_CUDA_TRY(cudaEventCreateWithFlags(&slot->ev, cudaEventDisableTiming));
        R3DGW_CUDA_TRY(cudaHostAlloc((void**)&slot->pinned, sizeof(uint32_t), cudaHostAllocDefault));
    }
    R3DGW_CUDA_TRY(cudaEventRecord(slot->ev, stream));
    return 0;
}

static int read_R_early(int dev, EarlyRSlot slot, const uint32_t* total_out, uint32_t* R) {
    cudaStream_t side;
    { std::lock_guard<std::mutex> lk(g_early_r_mu); side = g_early_r[dev].side; }
    cudaError_t err = cudaStreamWaitEvent(side, slot.ev, 0);
    if (err == cudaSuccess) err = cudaMemcpyAsync(slot.pinned, total_out, sizeof(uint32_t), cudaMemcpyDeviceToHost, side);
    // the copy is the last thing queued on the side stream by this thread; a later copy of another thread may also be
    // waited for here (harmless)
    if (err == cudaSuccess) err = cudaStreamSynchronize(side);
    if (err == cudaSuccess) *R = *slot.pinned;
    { std::lock_guard<std::mutex> lk(g_early_r_mu); g_early_r[dev].free_slots.push_back(slot); }
    return (int)err;
}

// depth-first binning scratch inside the geometry buffer: keys | keys_alt | ids | ids_alt | radix workspace |
// K2' look-back states (+ ticket)
struct DepthSortLayout { size_t keys, keys_alt, ids, ids_alt, ws, dup_state, total; };
static DepthSortLayout depth_sort_layout(int P) {
    DepthSortLayout D;
    const size_t Ps = (size_t)(P > 0 ? P : 0);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    D.keys = take(Ps * 4); D.keys_alt = take(Ps * 4); D.ids = take(Ps * 4); D.ids_alt = take(Ps * 4);
    D.ws = take(sort_workspace_bytes(P));
    D.dup_state = take(((Ps + 255) / 256) * 8 + 16);
    D.total = o;
    return D;
}
static size_t depth_sort_bytes(int P) { return depth_sort_layout(P).total; }

void geom_layout(int P, int has_sh, r3dgw_geom_layout* L) {
    size_t o = 0;
    const size_t Ps = (size_t)(P > 0 ? P : 0);
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L->depths = take(Ps * 4);
    L->means2D = take(Ps * 8);
    L->conic_opacity = take(Ps * 16);
    L->cov3D = take(Ps * 24);
    L->tiles_touched = take(Ps * 4);
    L->point_offsets = take(Ps * 4);
    L->rgb = take(has_sh ? Ps * 12 : 0);
    L->clamped = take(has_sh ? Ps * 3 : 0);
    L->scan_state = take((size_t)pre_blocks(P) * 8 + 16);   // +16: device scalar holding R
    L->depth_sort = take(depth_sort_bytes(P));
    L->total = o;
}

void binning_layout(int R, int W, int H, r3dgw_binning_layout* L) {
    size_t o = 0;
    const size_t Rs = (size_t)(R > 0 ? R : 0);
    const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L->point_list = take(Rs * 4);
    L->point_list_keys = take(Rs * 8);
    L->point_list_alt = take(Rs * 4);
    L->point_list_keys_alt = take(Rs * 8);
    L->ranges = take(tiles * 8);
    L->splats = take(Rs * 32);
    L->sort_hist = take(sort_workspace_bytes(R));
    L->total = o;
}

void image_layout(int W, int H, r3dgw_image_layout* L) {
    size_t o = 0;
    const size_t N = (size_t)W * H;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L->final_T = take(N * 4);
    L->n_contrib = take(N * 4);
    L->total = o;
}

static inline int bit_length(uint32_t n) {   // == getHigherMsb for the tile count (rasterizer_impl.cu:35-50)
    int b = 0;
    while (n) { ++b; n >>= 1; }
    return b;
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" {

const char* r3dgw_version(void) { return "r3dgw_b200 0.1 sm_100a"; }
uint64_t r3dgw_launch_count(void) { return g_launches.load(); }

void r3dgw_profile_enable(int on) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_profile_on = on != 0;
}
int r3dgw_profile_kernel_count(void) { return KID_COUNT; }
const char* r3dgw_profile_kernel_name(int id) { return (id >= 0 && id < KID_COUNT) ? g_prof_names[id] : ""; }
int r3dgw_profile_read(double* total_ms, uint64_t* launches, int reset) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& sp : g_prof_open) {
        if (cudaEventSynchronize(sp.e1) == cudaSuccess) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, sp.e0, sp.e1) == cudaSuccess) { g_prof_ms[sp.id] += ms; g_prof_n[sp.id] += 1; }
        }
        g_prof_pool.push_back(sp.e0);
        g_prof_pool.push_back(sp.e1);
    }
    g_prof_open.clear();
    (void)cudaGetLastError();
    for (int i = 0; i < KID_COUNT; ++i) {
        if (total_ms) total_ms[i] = g_prof_ms[i];
        if (launches) launches[i] = g_prof_n[i];
        if (reset) { g_prof_ms[i] = 0; g_prof_n[i] = 0; }
    }
    return 0;
}

int r3dgw_get_geom_layout(int P, int has_sh, r3dgw_geom_layout* out) {
    if (!out || P < 0) return R3DGW_ERR_BAD_ARG;
    geom_layout(P, has_sh, out);
    return 0;
}
int r3dgw_get_binning_layout(int R, int width, int height, r3dgw_binning_layout* out) {
    if (!out || R < 0 || width <= 0 || height <= 0) return R3DGW_ERR_BAD_ARG;
    binning_layout(R, width, height, out);
    return 0;
}
int r3dgw_get_image_layout(int width, int height, r3dgw_image_layout* out) {
    if (!out || width <= 0 || height <= 0) return R3DGW_ERR_BAD_ARG;
    image_layout(width, height, out);
    return 0;
}

int r3dgw_geometry_forward(r3dgw_alloc_fn geom_alloc, void* geom_ctx, r3dgw_alloc_fn binning_alloc,
                           void* binning_ctx, int P, int D, int M, const r3dgw_camera* cam,
                           const float* means3D, const float* shs, const float* opacities,
                           const float* scales, const float* rotations, const float* cov3D_precomp,
                           int* radii, int* num_rendered, void* stream_) {
    if (!geom_alloc || !binning_alloc || !cam || !num_rendered || P < 0) return R3DGW_ERR_BAD_ARG;
    if (cam->width <= 0 || cam->height <= 0) return R3DGW_ERR_BAD_ARG;
    *num_rendered = 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int W = cam->width, H = cam->height;
    const unsigned gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    if (P > 0) {
        if (!means3D || !opacities || !radii) return R3DGW_ERR_BAD_ARG;
        if (!cov3D_precomp && (!scales || !rotations)) return R3DGW_ERR_BAD_ARG;
        if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1))) return R3DGW_ERR_UNSUPPORTED;
    }

    r3dgw_geom_layout GL;
    geom_layout(P, shs != nullptr, &GL);
    char* g = geom_alloc(geom_ctx, GL.total);
    if (!g) return R3DGW_ERR_ALLOC;

    uint32_t R = 0;
    if (P > 0) {
        const int blocks = pre_blocks(P);
        R3DGW_CUDA_TRY(cudaMemsetAsync(g + GL.scan_state, 0, (size_t)blocks * 8 + 16, stream));
        PreprocessParams pp;
        pp.P = P; pp.D = D; pp.M = M; pp.W = W; pp.H = H;
        pp.tan_fovx = cam->tan_fovx; pp.tan_fovy = cam->tan_fovy;
        pp.focal_y = H / (2.0f * cam->tan_fovy);
        pp.focal_x = W / (2.0f * cam->tan_fovx);
        pp.scale_modifier = cam->scale_modifier;
        pp.prefiltered = cam->prefiltered;
        pp.grid_x = gx; pp.grid_y = gy;
        pp.means3D = means3D; pp.scales = scales; pp.rotations = rotations; pp.opacities = opacities;
        pp.shs = shs; pp.cov3D_precomp = cov3D_precomp;
        pp.viewmatrix = cam->viewmatrix; pp.projmatrix = cam->projmatrix; pp.cam_pos = cam->cam_pos;
        pp.radii = radii;
        pp.means2D = (float2*)(g + GL.means2D);
        pp.depths = (float*)(g + GL.depths);
        pp.cov3Ds = (float*)(g + GL.cov3D);
        pp.rgb = (float*)(g + GL.rgb);
        pp.clamped = (unsigned char*)(g + GL.clamped);
        pp.conic_opacity = (float4*)(g + GL.conic_opacity);
        pp.tiles_touched = (uint32_t*)(g + GL.tiles_touched);
        pp.point_offsets = (uint32_t*)(g + GL.point_offsets);
        pp.scan_state = (unsigned long long*)(g + GL.scan_state);
        pp.total_out = (uint32_t*)(g + GL.scan_state + (size_t)blocks * 8);
        {
            ProfScope ps(KID_PREPROCESS, stream);
            preprocess_scan_kernel<<<blocks, PRE_THREADS, 0, stream>>>(pp);
        }
        R3DGW_LAUNCH_CHECK();
        static const bool early_r = [] { const char* e = getenv("R3DGW_EARLY_R"); return !(e && e[0] == '0'); }();
        EarlyRSlot r_slot; int r_dev = -1;
        if (early_r) { const int rc = mark_R_ready(stream, &r_slot, &r_dev); if (rc) return rc; }
        // depth-first binning, step 1: stable sort of the P Gaussians by depth bits - enqueued BEFORE the read-back of
        // R so that it runs while the host waits
        {
            const DepthSortLayout DS = depth_sort_layout(P);
            char* ds = g + GL.depth_sort;
            R3DGW_CUDA_TRY(cudaMemsetAsync(ds + DS.dup_state, 0, ((size_t)(P + 255) / 256) * 8 + 16, stream));
            {
                {
                    ProfScope ps(KID_DEPTH_SORT, stream);
                    depth_keys_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, pp.depths, radii, (uint32_t*)(ds + DS.keys),
                                                                          (uint32_t*)(ds + DS.ids));
                }
                R3DGW_LAUNCH_CHECK();
                uint32_t* dks; uint32_t* dvs;
                g_sort_prof_kid = KID_DEPTH_SORT;
                int rc = sort_pairs_u32_u32((uint32_t*)(ds + DS.keys), (uint32_t*)(ds + DS.keys_alt), (uint32_t*)(ds + DS.ids),
                                            (uint32_t*)(ds + DS.ids_alt), P, 32, ds + DS.ws, stream, &dks, &dvs);
                g_sort_prof_kid = -1;
                if (rc) return rc;
                if (dvs != (uint32_t*)(ds + DS.ids)) return R3DGW_ERR_BAD_ARG;   // 4 passes: back in the main buffer
            }
        }
        // the one blocking read-back of the path (reference: rasterizer_impl.cu:281)
        if (early_r) {
            const int rc = read_R_early(r_dev, r_slot, pp.total_out, &R);
            if (rc) return rc;
        } else {
            R3DGW_CUDA_TRY(cudaMemcpyAsync(&R, pp.total_out, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
            R3DGW_CUDA_TRY(cudaStreamSynchronize(stream));
        }
        if (R > (uint32_t)INT_MAX) return R3DGW_ERR_OVERFLOW;
    }

    r3dgw_binning_layout BL;
    binning_layout((int)R, W, H, &BL);
    char* b = binning_alloc(binning_ctx, BL.total);
    if (!b) return R3DGW_ERR_ALLOC;
    R3DGW_CUDA_TRY(cudaMemsetAsync(b + BL.ranges, 0, (size_t)gx * gy * sizeof(uint2), stream));
    *num_rendered = (int)R;
    if (R == 0) return 0;

    // depth-first binning, steps 2-3: duplicates in depth order with tile-id keys, stable sort on the tile bits only
    const int end_bit = bit_length(gx * gy);
    const int passes = sort_num_passes(end_bit);
    uint32_t* tk_main = (uint32_t*)(b + BL.point_list_keys_alt);
    uint32_t* tk_alt = tk_main + R;
    uint32_t* vals_main = (uint32_t*)(b + BL.point_list);
    uint32_t* vals_alt = (uint32_t*)(b + BL.point_list_alt);
    // choose the starting buffer so that the sorted result lands in point_list / the first key half
    uint32_t* k0 = (passes & 1) ? tk_alt : tk_main;
    uint32_t* k1 = (passes & 1) ? tk_main : tk_alt;
    uint32_t* v0 = (passes & 1) ? vals_alt : vals_main;
    uint32_t* v1 = (passes & 1) ? vals_main : vals_alt;
    const DepthSortLayout DS = depth_sort_layout(P);
    char* ds = g + GL.depth_sort;
    {
        ProfScope ps(KID_DUPLICATE, stream);
        duplicate_sorted_kernel<<<(P + 255) / 256, 256, 0, stream>>>(
            P, (const uint32_t*)(ds + DS.ids), (const uint32_t*)(g + GL.tiles_touched), (const float2*)(g + GL.means2D),
            radii, (unsigned long long*)(ds + DS.dup_state),
            (uint32_t*)(ds + DS.dup_state + ((size_t)(P + 255) / 256) * 8), k0, v0, gx, gy);
    }
    R3DGW_LAUNCH_CHECK();

    uint32_t* ks; uint32_t* vs;
    int rc = sort_pairs_u32_u32(k0, k1, v0, v1, (int)R, end_bit, b + BL.sort_hist, stream, &ks, &vs);
    if (rc) return rc;
    if (ks != tk_main || vs != vals_main) return R3DGW_ERR_BAD_ARG;   // internal invariant

    {
        ProfScope ps(KID_RANGES, stream);
        // R3DGW_CULL_V2=0: the per-footprint edge minimisation (splat_may_touch) instead of the shared-line form
        static const bool cull_v2 = [] { const char* e = getenv("R3DGW_CULL_V2"); return !(e && e[0] == '0'); }();
        if (cull_v2)
            ranges_gather_kernel<true><<<(R + 255) / 256, 256, 0, stream>>>(
                (int)R, tk_main, vals_main, (const float2*)(g + GL.means2D), (const float4*)(g + GL.conic_opacity),
                (uint2*)(b + BL.ranges), (float4*)(b + BL.splats), (float4*)(b + BL.splats) + R, gx);
        else
            ranges_gather_kernel<false><<<(R + 255) / 256, 256, 0, stream>>>(
                (int)R, tk_main, vals_main, (const float2*)(g + GL.means2D), (const float4*)(g + GL.conic_opacity),
                (uint2*)(b + BL.ranges), (float4*)(b + BL.splats), (float4*)(b + BL.splats) + R, gx);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_materialize_sorted_keys(int P, int R, int width, int height, const char* geom_buffer, char* binning_buffer,
                                  void* stream_) {
    if (P < 0 || R < 0 || width <= 0 || height <= 0 || !geom_buffer || !binning_buffer) return R3DGW_ERR_BAD_ARG;
    if (R == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    r3dgw_geom_layout GL; geom_layout(P, 0, &GL);      // depths precede the SH-dependent fields
    r3dgw_binning_layout BL; binning_layout(R, width, height, &BL);
    materialize_keys_kernel<<<(R + 255) / 256, 256, 0, stream>>>(
        R, (const uint32_t*)(binning_buffer + BL.point_list_keys_alt), (const uint32_t*)(binning_buffer + BL.point_list),
        (const float*)(geom_buffer + GL.depths), (uint64_t*)(binning_buffer + BL.point_list_keys));
    R3DGW_LAUNCH_CHECK();
    return 0;
}

size_t r3dgw_alpha_cache_bytes(int R) { return R > 0 ? (size_t)R * TILE_PIX * sizeof(float) : 0; }

int r3dgw_blend_forward_cached(int P, int R, int C, const r3dgw_camera* cam, const char* geom_buffer,
                               const char* binning_buffer, char* image_buffer, const float* features,
                               const float* background, float* out_color, float* alpha_cache, int cache_mode,
                               void* stream_) {
    if (!cam || !binning_buffer || !image_buffer || !background || !out_color || R < 0 || P < 0) return R3DGW_ERR_BAD_ARG;
    if (cache_mode < 0 || cache_mode > 2 || (cache_mode != 0 && !alpha_cache && R > 0)) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int W = cam->width, H = cam->height;
    r3dgw_binning_layout BL; binning_layout(R, W, H, &BL);
    r3dgw_image_layout IL; image_layout(W, H, &IL);
    if (!features) {
        if (C != 3 || !geom_buffer) return R3DGW_ERR_BAD_ARG;
        r3dgw_geom_layout GL; geom_layout(P, 1, &GL);
        features = (const float*)(geom_buffer + GL.rgb);
    }
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    const float4* A = (const float4*)(binning_buffer + BL.splats);
    return blend_forward_dispatch(C, grid, (const uint2*)(binning_buffer + BL.ranges), A, A + R, W, H, features,
                                  background, (float*)(image_buffer + IL.final_T),
                                  (uint32_t*)(image_buffer + IL.n_contrib), out_color, alpha_cache, cache_mode, stream);
}

int r3dgw_blend_forward(int P, int R, int C, const r3dgw_camera* cam, const char* geom_buffer,
                        const char* binning_buffer, char* image_buffer, const float* features,
                        const float* background, float* out_color, void* stream_) {
    return r3dgw_blend_forward_cached(P, R, C, cam, geom_buffer, binning_buffer, image_buffer, features, background,
                                      out_color, nullptr, 0, stream_);
}

int r3dgw_blend_backward_color_cached(int P, int R, int C, const r3dgw_camera* cam, const char* binning_buffer,
                                      const char* image_buffer, const float* alpha_cache, const float* dL_dpixels,
                                      float* dL_dfeatures, int grad_stride, void* stream_) {
    if (!cam || !binning_buffer || !image_buffer || !dL_dpixels || !dL_dfeatures || R < 0 || P < 0) return R3DGW_ERR_BAD_ARG;
    if (R == 0) return 0;
    if (!alpha_cache) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int W = cam->width, H = cam->height;
    r3dgw_binning_layout BL; binning_layout(R, W, H, &BL);
    r3dgw_image_layout IL; image_layout(W, H, &IL);
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    const float4* A = (const float4*)(binning_buffer + BL.splats);
    return blend_backward_color_cached_dispatch(C, grid, (const uint2*)(binning_buffer + BL.ranges), A + R, W, H,
                                                (const uint32_t*)(image_buffer + IL.n_contrib), alpha_cache,
                                                dL_dpixels, dL_dfeatures, grad_stride, stream);
}

int r3dgw_blend_backward_geometry_cached_calls(int P, int R, int n_calls, const int* channels, const r3dgw_camera* cam,
                                               const char* binning_buffer, const char* image_buffer,
                                               const float* alpha_cache, const float* const* features,
                                               const float* const* out_images, const float* const* backgrounds,
                                               const float* const* dL_dpixels, float* dL_dmean2D, float* dL_dconic,
                                               float* dL_dopacity, void* stream_) {
    if (!cam || !binning_buffer || !image_buffer || n_calls <= 0 || !channels || !features || !out_images || !backgrounds ||
        !dL_dpixels || !dL_dmean2D || !dL_dconic || !dL_dopacity || R < 0 || P < 0)
        return R3DGW_ERR_BAD_ARG;
    if (R == 0) return 0;
    if (!alpha_cache) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int W = cam->width, H = cam->height;
    r3dgw_binning_layout BL; binning_layout(R, W, H, &BL);
    r3dgw_image_layout IL; image_layout(W, H, &IL);
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    const float4* A = (const float4*)(binning_buffer + BL.splats);
    return blend_backward_geometry_cached_dispatch(n_calls, channels, grid, (const uint2*)(binning_buffer + BL.ranges), A, A + R,
                                                   W, H, features, (const uint32_t*)(image_buffer + IL.n_contrib), alpha_cache,
                                                   out_images, backgrounds, dL_dpixels, dL_dmean2D, dL_dconic, dL_dopacity,
                                                   stream);
}

int r3dgw_blend_backward_geometry_cached(int P, int R, int C, const r3dgw_camera* cam, const char* binning_buffer,
                                         const char* image_buffer, const float* alpha_cache, const float* features,
                                         const float* out_images, const float* background, const float* dL_dpixels,
                                         float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, void* stream_) {
    return r3dgw_blend_backward_geometry_cached_calls(P, R, 1, &C, cam, binning_buffer, image_buffer, alpha_cache, &features,
                                                      &out_images, &background, &dL_dpixels, dL_dmean2D, dL_dconic,
                                                      dL_dopacity, stream_);
}

int r3dgw_blend_backward(int P, int R, int C, const r3dgw_camera* cam, const char* geom_buffer,
                         const char* binning_buffer, const char* image_buffer, const float* features,
                         const float* background, const float* dL_dpixels, float* dL_dmean2D,
                         float* dL_dconic, float* dL_dopacity, float* dL_dfeatures, void* stream_) {
    if (!cam || !binning_buffer || !image_buffer || !background || !dL_dpixels || !dL_dmean2D || !dL_dconic ||
        !dL_dopacity || R < 0 || P < 0)   // dL_dfeatures == NULL: geometry-only pass
        return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int W = cam->width, H = cam->height;
    r3dgw_binning_layout BL; binning_layout(R, W, H, &BL);
    r3dgw_image_layout IL; image_layout(W, H, &IL);
    if (!features) {
        if (C != 3 || !geom_buffer) return R3DGW_ERR_BAD_ARG;
        r3dgw_geom_layout GL; geom_layout(P, 1, &GL);
        features = (const float*)(geom_buffer + GL.rgb);
    }
    if (R == 0) return 0;
    dim3 grid((W + TILE_X - 1) / TILE_X, (H + TILE_Y - 1) / TILE_Y, 1);
    const float4* A = (const float4*)(binning_buffer + BL.splats);
    return blend_backward_dispatch(C, grid, (const uint2*)(binning_buffer + BL.ranges), A, A + R, W, H, background,
                                   features, (const float*)(image_buffer + IL.final_T),
                                   (const uint32_t*)(image_buffer + IL.n_contrib), dL_dpixels, dL_dmean2D,
                                   dL_dconic, dL_dopacity, dL_dfeatures, stream);
}

int r3dgw_geometry_backward(int P, int D, int M, const r3dgw_camera* cam, const float* means3D,
                            const float* shs, const float* scales, const float* rotations,
                            const float* cov3D_precomp, const int* radii, const char* geom_buffer,
                            const float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D,
                            float* dL_dcolor, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                            float* dL_drot, void* stream_) {
    if (!cam || P < 0) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!means3D || !radii || !geom_buffer || !dL_dmean2D || !dL_dconic || !dL_dmean3D || !dL_dcov3D) return R3DGW_ERR_BAD_ARG;
    if (!cov3D_precomp && (!scales || !rotations || !dL_dscale || !dL_drot)) return R3DGW_ERR_BAD_ARG;
    if (shs && (!dL_dcolor || !dL_dsh)) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    r3dgw_geom_layout GL; geom_layout(P, shs != nullptr, &GL);
    GeomBackwardParams gp;
    gp.P = P; gp.D = D; gp.M = M;
    gp.focal_y = cam->height / (2.0f * cam->tan_fovy);
    gp.focal_x = cam->width / (2.0f * cam->tan_fovx);
    gp.tan_fovx = cam->tan_fovx; gp.tan_fovy = cam->tan_fovy; gp.scale_modifier = cam->scale_modifier;
    gp.means3D = means3D; gp.radii = radii;
    gp.cov3Ds = cov3D_precomp ? cov3D_precomp : (const float*)(geom_buffer + GL.cov3D);
    gp.scales = cov3D_precomp ? nullptr : scales;
    gp.rotations = rotations; gp.shs = shs;
    gp.clamped = (const unsigned char*)(geom_buffer + GL.clamped);
    gp.viewmatrix = cam->viewmatrix; gp.projmatrix = cam->projmatrix; gp.cam_pos = cam->cam_pos;
    gp.dL_dmean2D = dL_dmean2D; gp.dL_dconic = dL_dconic; gp.dL_dmean3D = dL_dmean3D;
    gp.dL_dcolor = dL_dcolor; gp.dL_dcov3D = dL_dcov3D; gp.dL_dsh = dL_dsh;
    gp.dL_dscale = dL_dscale; gp.dL_drot = dL_drot;
    {
        ProfScope ps(KID_GEOM_BWD, stream);
        static const bool f32_chain = [] { const char* e = getenv("R3DGW_GEOM_BWD_F32"); return e && e[0] == '1'; }();
        if (f32_chain) geometry_backward_kernel<float><<<(P + 255) / 256, 256, 0, stream>>>(gp);
        else geometry_backward_kernel<double><<<(P + 255) / 256, 256, 0, stream>>>(gp);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_rasterize_forward(r3dgw_alloc_fn geom_alloc, void* geom_ctx, r3dgw_alloc_fn binning_alloc,
                            void* binning_ctx, r3dgw_alloc_fn image_alloc, void* image_ctx, int P, int D,
                            int M, const float* background, const r3dgw_camera* cam, const float* means3D,
                            const float* shs, const float* colors_precomp, const float* opacities,
                            const float* scales, const float* rotations, const float* cov3D_precomp,
                            float* out_color, int* radii, int* num_rendered, void* stream) {
    if (!image_alloc || !cam) return R3DGW_ERR_BAD_ARG;
    if (!colors_precomp && !shs) return R3DGW_ERR_BAD_ARG;
    struct Cap { r3dgw_alloc_fn fn; void* ctx; char* ptr; };
    Cap gc{geom_alloc, geom_ctx, nullptr}, bc{binning_alloc, binning_ctx, nullptr};
    auto tramp = [](void* c, size_t n) -> char* { Cap* k = (Cap*)c; k->ptr = k->fn(k->ctx, n); return k->ptr; };
    int rc = r3dgw_geometry_forward(tramp, &gc, tramp, &bc, P, D, M, cam, means3D, colors_precomp ? nullptr : shs,
                                    opacities, scales, rotations, cov3D_precomp, radii, num_rendered, stream);
    if (rc) return rc;
    r3dgw_image_layout IL; image_layout(cam->width, cam->height, &IL);
    char* img = image_alloc(image_ctx, IL.total);
    if (!img) return R3DGW_ERR_ALLOC;
    return r3dgw_blend_forward(P, *num_rendered, 3, cam, gc.ptr, bc.ptr, img, colors_precomp, background, out_color, stream);
}

int r3dgw_rasterize_backward(int P, int D, int M, int R, const float* background, const r3dgw_camera* cam,
                             const float* means3D, const float* shs, const float* colors_precomp,
                             const float* scales, const float* rotations, const float* cov3D_precomp,
                             const int* radii, const char* geom_buffer, const char* binning_buffer,
                             const char* image_buffer, const float* dL_dpix, float* dL_dmean2D,
                             float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D,
                             float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, void* stream) {
    int rc = r3dgw_blend_backward(P, R, 3, cam, geom_buffer, binning_buffer, image_buffer, colors_precomp,
                                  background, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, stream);
    if (rc) return rc;
    return r3dgw_geometry_backward(P, D, M, cam, means3D, colors_precomp ? nullptr : shs, scales, rotations,
                                   cov3D_precomp, radii, geom_buffer, dL_dmean2D, dL_dconic, dL_dmean3D, dL_dcolor,
                                   dL_dcov3D, dL_dsh, dL_dscale, dL_drot, stream);
}

int r3dgw_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, void* stream) {
    (void)projmatrix;
    if (P < 0) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!means3D || !viewmatrix || !present) return R3DGW_ERR_BAD_ARG;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, means3D, viewmatrix, present);
    R3DGW_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
