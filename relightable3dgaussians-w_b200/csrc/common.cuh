// common.cuh - shared device helpers, state-buffer layouts and launch bookkeeping.
// Part of the B200-native (sm_100a) splat rasterizer.  See include/r3dgw_b200.h for the C-ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <math.h>
#include <atomic>
#include "../../include/r3dgw_b200.h"

namespace r3dgw {

constexpr int TILE_X = 16;      // reference: BLOCK_X, cuda_rasterizer/config.h:16
constexpr int TILE_Y = 16;      // reference: BLOCK_Y, cuda_rasterizer/config.h:17
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int NUM_SMS = 148;    // B200

extern std::atomic<uint64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

#define R3DGW_CUDA_TRY(expr)                                \
    do {                                                    \
        cudaError_t _e = (expr);                            \
        if (_e != cudaSuccess) return (int)_e;              \
    } while (0)

#define R3DGW_LAUNCH_CHECK()                                \
    do {                                                    \
        r3dgw::count_launch();                              \
        cudaError_t _e = cudaGetLastError();                \
        if (_e != cudaSuccess) return (int)_e;              \
    } while (0)

// ---- optional per-kernel device timing (CUDA events on the launching stream) ---------------------
// Off by default; bench.py switches it on for a separate, untimed pass to attribute the step to kernels.
enum KernelId {
    KID_PREPROCESS = 0, KID_DUPLICATE, KID_SORT_HIST, KID_SORT_SCAN, KID_SORT_PASS, KID_RANGES, KID_BLEND_FWD,
    KID_BLEND_BWD, KID_GEOM_BWD, KID_SHADE_FWD, KID_SHADE_BWD, KID_KNN, KID_MARK_VISIBLE, KID_SSIM_FWD, KID_SSIM_BWD,
    KID_MASKED_L1, KID_DEPTH_NORMAL, KID_BLEND_FWD_CACHED, KID_BLEND_BWD_GEOM, KID_BLEND_BWD_COLOR, KID_DEPTH_SORT,
    KID_ENVMAP_SH, KID_SH_ROTATE, KID_GAUSS_REG, KID_COUNT
};
extern bool g_profile_on;
extern int g_sort_prof_kid;   // >= 0: the radix sort's launches are attributed to this kernel id (P-sized depth sort)
void prof_mark(int id, cudaStream_t s, bool end);
struct ProfScope {
    int id; cudaStream_t s;
    ProfScope(int id_, cudaStream_t s_) : id(id_), s(s_) { if (g_profile_on) prof_mark(id, s, false); }
    ~ProfScope() { if (g_profile_on) prof_mark(id, s, true); }
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- 3x3 helper with GLM's storage (c[col][row]) and GLM's accumulate order -------------------
// The reference builds its matrices with glm::mat3 (column constructor) and multiplies with
// glm's operator* (third_party/glm/glm/detail/type_mat3x3.inl:486-520): every element is
// a0*b0 + a1*b1 + a2*b2 evaluated left to right.  radii / tiles_touched are only bit-exact if the
// expression trees are the same, so this mirrors that order (not the library).
struct M3 {
    float c[3][3];
    __host__ __device__ M3() {}
    // column-wise fill, like glm::mat3(x0,y0,z0, x1,y1,z1, x2,y2,z2)
    __host__ __device__ M3(float x0, float y0, float z0, float x1, float y1, float z1,
                           float x2, float y2, float z2) {
        c[0][0] = x0; c[0][1] = y0; c[0][2] = z0;
        c[1][0] = x1; c[1][1] = y1; c[1][2] = z1;
        c[2][0] = x2; c[2][1] = y2; c[2][2] = z2;
    }
};

__host__ __device__ __forceinline__ M3 m3_mul(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int col = 0; col < 3; ++col)
#pragma unroll
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}

__host__ __device__ __forceinline__ M3 m3_transpose(const M3& m) {
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) r.c[i][j] = m.c[j][i];
    return r;
}

// reference: transformPoint4x3 / transformPoint4x4, cuda_rasterizer/auxiliary.h:58-77
__device__ __forceinline__ float3 xform_4x3(const float3& p, const float* m) {
    float3 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
    };
    return t;
}
__device__ __forceinline__ float4 xform_4x4(const float3& p, const float* m) {
    float4 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
        m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]
    };
    return t;
}

// reference: ndc2Pix, cuda_rasterizer/auxiliary.h:41-44 - the literals are double, so the
// expression is evaluated in FP64 and rounded once on return.
__device__ __forceinline__ float ndc_to_pix(float v, int S) {
    return ((v + 1.0) * S - 1.0) * 0.5;
}

// reference: getRect, cuda_rasterizer/auxiliary.h:46-56 ((int) casts truncate toward zero)
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rmin, uint2& rmax,
                                          unsigned gx, unsigned gy) {
    rmin.x = min(gx, (unsigned)max((int)0, (int)((p.x - max_radius) / TILE_X)));
    rmin.y = min(gy, (unsigned)max((int)0, (int)((p.y - max_radius) / TILE_Y)));
    rmax.x = min(gx, (unsigned)max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X)));
    rmax.y = min(gy, (unsigned)max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)));
}

// ---- state layouts (host side) ------------------------------------------------------------------
constexpr int PRE_THREADS = 256;          // K1 block size; also the scan tile size
constexpr int SORT_RADIX_BITS = 8;
constexpr int SORT_RADIX = 1 << SORT_RADIX_BITS;
constexpr int SORT_MAX_PASSES = 8;

struct SortPlan {
    int passes;
    int begin_bit[SORT_MAX_PASSES];
    int bits[SORT_MAX_PASSES];
};

void geom_layout(int P, int has_sh, r3dgw_geom_layout* L);
void binning_layout(int R, int W, int H, r3dgw_binning_layout* L);
void image_layout(int W, int H, r3dgw_image_layout* L);

// onesweep radix sort (radix_sort.cu)
size_t sort_workspace_bytes(int n);
int sort_pairs_u64_u32(uint64_t* keys_in, uint64_t* keys_alt, uint32_t* vals_in, uint32_t* vals_alt,
                       int n, int end_bit, char* workspace, cudaStream_t stream,
                       uint64_t** keys_out, uint32_t** vals_out);
int sort_pairs_u32_u32(uint32_t* keys_in, uint32_t* keys_alt, uint32_t* vals_in, uint32_t* vals_alt,
                       int n, int end_bit, char* workspace, cudaStream_t stream,
                       uint32_t** keys_out, uint32_t** vals_out);

}  // namespace r3dgw
