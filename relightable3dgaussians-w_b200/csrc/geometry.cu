// geometry.cu - per-Gaussian stages of the splat pipeline for sm_100a:
//   K1  preprocess_scan_kernel : cull, project, cov3D -> conic, radius, tile rect, FUSED with the
//                                prefix sum of tiles_touched (decoupled look-back), so the separate
//                                scan pass of the reference (rasterizer_impl.cu:276) disappears.
//   K2  duplicate_keys_kernel  : (tile|depth) keys, warp-cooperative for large rects.
//   K3  ranges_gather_kernel   : tile ranges + gather of the sorted-order splat stream.
//   K8  mark_visible_kernel
//   K6+K7 geometry_backward_kernel (fused).
// Reference behaviour: cuda_rasterizer/forward.cu:74-256, rasterizer_impl.cu:54-138,
// backward.cu:144-396.  Arithmetic order of the decision path mirrors the reference so that
// radii / tiles_touched / keys are bit-identical.
#include "common.cuh"
#include "sh.cuh"

namespace r3dgw {

// ---------------------------------------------------------------------------------------------
// cov3D from scale + (un-normalised) quaternion.  reference: computeCov3D, forward.cu:118-152
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 rot,
                                                     float* cov3D) {
    M3 S(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S.c[0][0] = mod * scale.x;
    S.c[1][1] = mod * scale.y;
    S.c[2][2] = mod * scale.z;
    float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    M3 R(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
         2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
         2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    M3 M = m3_mul(S, R);
    M3 Sigma = m3_mul(m3_transpose(M), M);
    cov3D[0] = Sigma.c[0][0];
    cov3D[1] = Sigma.c[0][1];
    cov3D[2] = Sigma.c[0][2];
    cov3D[3] = Sigma.c[1][1];
    cov3D[4] = Sigma.c[1][2];
    cov3D[5] = Sigma.c[2][2];
}

// EWA projection of the 3D covariance.  reference: computeCov2D, forward.cu:74-113
__device__ __forceinline__ float3 cov2d_project(const float3& mean, float focal_x, float focal_y,
                                                float tan_fovx, float tan_fovy, const float* cov3D,
                                                const float* viewmatrix) {
    float3 t = xform_4x3(mean, viewmatrix);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;

    M3 J(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
         0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
         0, 0, 0);
    M3 W(viewmatrix[0], viewmatrix[4], viewmatrix[8],
         viewmatrix[1], viewmatrix[5], viewmatrix[9],
         viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    M3 T = m3_mul(W, J);
    M3 Vrk(cov3D[0], cov3D[1], cov3D[2],
           cov3D[1], cov3D[3], cov3D[4],
           cov3D[2], cov3D[4], cov3D[5]);
    M3 cov = m3_mul(m3_mul(m3_transpose(T), m3_transpose(Vrk)), T);
    cov.c[0][0] += 0.3f;
    cov.c[1][1] += 0.3f;
    return {float(cov.c[0][0]), float(cov.c[0][1]), float(cov.c[1][1])};
}

// ---------------------------------------------------------------------------------------------
// Decoupled look-back tile state: high 32 bits flag (0 none, 1 aggregate, 2 inclusive), low 32 value.
__device__ __forceinline__ void scan_state_store(unsigned long long* p, unsigned flag, unsigned v) {
    unsigned long long w = ((unsigned long long)flag << 32) | v;
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long scan_state_load(const unsigned long long* p) {
    unsigned long long w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}

struct PreprocessParams {
    int P, D, M;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
    int prefiltered;
    unsigned grid_x, grid_y;
    const float* means3D;
    const float* scales;
    const float* rotations;
    const float* opacities;
    const float* shs;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    int* radii;
    float2* means2D;
    float* depths;
    float* cov3Ds;
    float* rgb;
    unsigned char* clamped;
    float4* conic_opacity;
    uint32_t* tiles_touched;
    uint32_t* point_offsets;
    unsigned long long* scan_state;
    uint32_t* total_out;   // device scalar: R
};

__global__ void __launch_bounds__(PRE_THREADS) preprocess_scan_kernel(const PreprocessParams p) {
    __shared__ float s_view[16], s_proj[16];
    __shared__ uint32_t s_warp_sum[PRE_THREADS / 32];
    __shared__ uint32_t s_block_prefix, s_tile;
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    // dynamic tile id: the look-back below waits on lower-numbered tiles, so ids are handed out in the order CTAs
    // actually start (forward progress does not depend on the hardware's block dispatch order)
    if (threadIdx.x == 0) s_tile = atomicAdd(p.total_out + 1, 1u);
    __syncthreads();
    const unsigned bid = s_tile;

    const int idx = bid * PRE_THREADS + threadIdx.x;
    uint32_t touched = 0;
    if (idx < p.P) {
        int my_radius_i = 0;
        // reference: in_frustum (auxiliary.h:139-164): near-plane cull on view-space z only
        float3 p_orig = {p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]};
        float3 p_view = xform_4x3(p_orig, s_view);
        if (p_view.z > 0.2f) {
            float4 p_hom = xform_4x4(p_orig, s_proj);
            float p_w = 1.0f / (p_hom.w + 0.0000001f);
            float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
            const float* cov3D;
            if (p.cov3D_precomp != nullptr) {
                cov3D = p.cov3D_precomp + idx * 6;
            } else {
                float3 sc = {p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]};
                float4 rq = reinterpret_cast<const float4*>(p.rotations)[idx];
                cov3d_from_scale_rot(sc, p.scale_modifier, rq, p.cov3Ds + idx * 6);
                cov3D = p.cov3Ds + idx * 6;
            }
            float3 cov = cov2d_project(p_orig, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, cov3D, s_view);
            float det = (cov.x * cov.z - cov.y * cov.y);
            if (det != 0.0f) {
                float det_inv = 1.f / det;
                float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
                float mid = 0.5f * (cov.x + cov.z);
                float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
                float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
                float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
                float2 point_image = {ndc_to_pix(p_proj.x, p.W), ndc_to_pix(p_proj.y, p.H)};
                uint2 rmin, rmax;
                tile_rect(point_image, my_radius, rmin, rmax, p.grid_x, p.grid_y);
                uint32_t n = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                if (n != 0) {
                    if (p.shs != nullptr) {
                        float3 c = sh_to_rgb_fwd(idx, p.D, p.M, p.means3D, p.cam_pos, p.shs, p.clamped);
                        p.rgb[idx * 3 + 0] = c.x;
                        p.rgb[idx * 3 + 1] = c.y;
                        p.rgb[idx * 3 + 2] = c.z;
                    }
                    p.depths[idx] = p_view.z;
                    my_radius_i = my_radius;
                    p.means2D[idx] = point_image;
                    p.conic_opacity[idx] = {conic.x, conic.y, conic.z, p.opacities[idx]};
                    touched = n;
                }
            }
        } else if (p.prefiltered) {
            // reference traps here (auxiliary.h:156-160)
            __trap();
        }
        p.radii[idx] = my_radius_i;
        p.tiles_touched[idx] = touched;
    }

    // ---- fused inclusive prefix sum over tiles_touched (replaces cub::DeviceScan::InclusiveSum)
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = touched;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t ws = lane < PRE_THREADS / 32 ? s_warp_sum[lane] : 0;
        uint32_t wi = ws;
#pragma unroll
        for (int o = 1; o < PRE_THREADS / 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        if (lane < PRE_THREADS / 32) s_warp_sum[lane] = wi - ws;   // exclusive warp offsets
        uint32_t block_total = __shfl_sync(0xffffffffu, wi, PRE_THREADS / 32 - 1);
        // publish aggregate, look back
        uint32_t excl = 0;
        if (bid == 0) {
            if (lane == 0) scan_state_store(p.scan_state, 2u, block_total);
        } else {
            if (lane == 0) scan_state_store(p.scan_state + bid, 1u, block_total);
            int base = (int)bid - 1;
            while (true) {
                int j = base - (int)lane;
                unsigned long long w = 0;
                unsigned flag = 2;  // virtual predecessor before block 0: inclusive 0
                if (j >= 0) {
                    do { w = scan_state_load(p.scan_state + j); flag = (unsigned)(w >> 32); } while (flag == 0);
                }
                unsigned incl_mask = __ballot_sync(0xffffffffu, flag == 2);
                // lanes up to and including the first inclusive one contribute
                int first = __ffs(incl_mask) - 1;   // incl_mask non-zero once j<0 lanes exist, else may be 0
                uint32_t v = (j >= 0 && (first < 0 || (int)lane <= first)) ? (uint32_t)w : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                excl += v;
                if (first >= 0) break;
                base -= 32;
            }
            if (lane == 0) scan_state_store(p.scan_state + bid, 2u, excl + block_total);
        }
        if (lane == 0) {
            s_block_prefix = excl;
            if (bid == gridDim.x - 1) *p.total_out = excl + block_total;
        }
    }
    __syncthreads();
    if (idx < p.P) p.point_offsets[idx] = s_block_prefix + s_warp_sum[warp] + incl;
}

// ---------------------------------------------------------------------------------------------
// K2.  reference: duplicateWithKeys, rasterizer_impl.cu:70-111.  Same keys, same emission order
// (row-major over the rect, ascending Gaussian index through point_offsets); rects larger than
// DUP_SMALL tiles are written by the whole warp so that stores coalesce and the loop is balanced.
constexpr int DUP_SMALL = 6;
__global__ void __launch_bounds__(256) duplicate_keys_kernel(
    int P, const float2* __restrict__ points_xy, const float* __restrict__ depths,
    const uint32_t* __restrict__ offsets, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
    const int* __restrict__ radii, unsigned gx, unsigned gy) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    uint32_t n = 0, off = 0, w = 0, depth_bits = 0;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    if (idx < P) {
        int r = radii[idx];
        if (r > 0) {
            off = (idx == 0) ? 0 : offsets[idx - 1];
            tile_rect(points_xy[idx], r, rmin, rmax, gx, gy);
            w = rmax.x - rmin.x;
            n = w * (rmax.y - rmin.y);
            depth_bits = __float_as_uint(depths[idx]);
        }
    }
    if (n <= DUP_SMALL) {
        uint32_t o = off;
        for (unsigned y = rmin.y; y < rmax.y; y++)
            for (unsigned x = rmin.x; x < rmax.x; x++) {
                uint64_t key = y * gx + x;
                key <<= 32;
                key |= depth_bits;
                keys[o] = key;
                vals[o] = idx;
                o++;
            }
    }
    unsigned big = __ballot_sync(0xffffffffu, n > DUP_SMALL);
    while (big) {
        int src = __ffs(big) - 1;
        big &= big - 1;
        uint32_t bn = __shfl_sync(0xffffffffu, n, src);
        uint32_t bo = __shfl_sync(0xffffffffu, off, src);
        uint32_t bw = __shfl_sync(0xffffffffu, w, src);
        uint32_t bx = __shfl_sync(0xffffffffu, rmin.x, src);
        uint32_t by = __shfl_sync(0xffffffffu, rmin.y, src);
        uint32_t bd = __shfl_sync(0xffffffffu, depth_bits, src);
        uint32_t bi = (uint32_t)(idx - (int)lane + src);
        for (uint32_t k = lane; k < bn; k += 32) {
            uint32_t y = by + k / bw, x = bx + k % bw;
            uint64_t key = y * gx + x;
            key <<= 32;
            key |= bd;
            keys[bo + k] = key;
            vals[bo + k] = bi;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Depth-first binning (B200-first replacement of the reference's one big sort).  The reference sorts R ~ 5-6 P
// (tile | depth) 64-bit keys in one 6-pass radix sort (rasterizer_impl.cu:303-308).  The same order results from
//   (1) a stable sort of the P Gaussians by depth bits (4 passes over P 32-bit keys - 5x fewer items),
//   (2) emitting the duplicates in that order (K2'), and
//   (3) a STABLE sort of the R duplicates by tile id only (bit_length(tiles) bits: 2 passes of 32-bit keys at 1080p).
// Stability makes ties come out exactly as in the reference: within a tile by depth, equal depths by ascending
// Gaussian index.  Traffic on the R-sized arrays drops from 152 B to 36 B per duplicate.

// (1) keys for the per-Gaussian depth sort; culled Gaussians (radius 0) go to the end
__global__ void __launch_bounds__(256) depth_keys_kernel(int P, const float* __restrict__ depths,
                                                         const int* __restrict__ radii, uint32_t* __restrict__ keys,
                                                         uint32_t* __restrict__ vals) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    keys[idx] = radii[idx] > 0 ? __float_as_uint(depths[idx]) : 0xFFFFFFFFu;
    vals[idx] = (uint32_t)idx;
}

// (2) K2': duplicates in depth order.  One thread per sorted position; the exclusive prefix sum of tiles_touched in
// that order is fused (block scan + decoupled look-back, dynamic tile ids), so no offsets array is materialised.
// Keys are the bare tile ids (32-bit); values the Gaussian ids.  Same rect expression as the reference's
// duplicateWithKeys (rasterizer_impl.cu:70-111); rects larger than DUP_SMALL tiles are written by the whole warp.
__global__ void __launch_bounds__(256) duplicate_sorted_kernel(
    int P, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ tiles_touched,
    const float2* __restrict__ points_xy, const int* __restrict__ radii, unsigned long long* __restrict__ scan_state,
    uint32_t* __restrict__ ticket, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, unsigned gx, unsigned gy) {
    __shared__ uint32_t s_warp_sum[8];
    __shared__ uint32_t s_block_prefix, s_tile;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const unsigned bid = s_tile;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int spos = (int)(bid * 256 + threadIdx.x);
    uint32_t n = 0, w = 0, gid = 0;
    uint2 rmin = {0, 0}, rmax = {0, 0};
    if (spos < P) {
        gid = perm[spos];
        n = tiles_touched[gid];
        if (n) {
            tile_rect(points_xy[gid], radii[gid], rmin, rmax, gx, gy);
            w = rmax.x - rmin.x;
        }
    }
    // block-wide inclusive scan of n
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t ws = lane < 8 ? s_warp_sum[lane] : 0;
        uint32_t wi = ws;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += v;
        }
        if (lane < 8) s_warp_sum[lane] = wi - ws;
        const uint32_t block_total = __shfl_sync(0xffffffffu, wi, 7);
        uint32_t excl = 0;
        if (bid == 0) {
            if (lane == 0) scan_state_store(scan_state, 2u, block_total);
        } else {
            if (lane == 0) scan_state_store(scan_state + bid, 1u, block_total);
            int base = (int)bid - 1;
            while (true) {
                const int j = base - (int)lane;
                unsigned long long sw = 0;
                unsigned flag = 2;
                if (j >= 0) {
                    do { sw = scan_state_load(scan_state + j); flag = (unsigned)(sw >> 32); } while (flag == 0);
                }
                const unsigned incl_mask = __ballot_sync(0xffffffffu, flag == 2);
                const int first = __ffs(incl_mask) - 1;
                uint32_t v = (j >= 0 && (first < 0 || (int)lane <= first)) ? (uint32_t)sw : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                excl += v;
                if (first >= 0) break;
                base -= 32;
            }
            if (lane == 0) scan_state_store(scan_state + bid, 2u, excl + block_total);
        }
        if (lane == 0) s_block_prefix = excl;
    }
    __syncthreads();
    const uint32_t off = s_block_prefix + s_warp_sum[warp] + incl - n;

    if (n && n <= DUP_SMALL) {
        uint32_t o = off;
        for (unsigned y = rmin.y; y < rmax.y; y++)
            for (unsigned x = rmin.x; x < rmax.x; x++) {
                keys[o] = y * gx + x;
                vals[o] = gid;
                o++;
            }
    }
    unsigned big = __ballot_sync(0xffffffffu, n > DUP_SMALL);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t bn = __shfl_sync(0xffffffffu, n, src);
        const uint32_t bo = __shfl_sync(0xffffffffu, off, src);
        const uint32_t bw = __shfl_sync(0xffffffffu, w, src);
        const uint32_t bx = __shfl_sync(0xffffffffu, rmin.x, src);
        const uint32_t by = __shfl_sync(0xffffffffu, rmin.y, src);
        const uint32_t bi = __shfl_sync(0xffffffffu, gid, src);
        for (uint32_t k = lane; k < bn; k += 32) {
            keys[bo + k] = (by + k / bw) * gx + (bx + k % bw);
            vals[bo + k] = bi;
        }
    }
}

// The reference's sorted 64-bit keys (tile << 32 | depth bits), rebuilt on demand for tests / tooling.
__global__ void __launch_bounds__(256) materialize_keys_kernel(int L, const uint32_t* __restrict__ tile_keys,
                                                               const uint32_t* __restrict__ point_list,
                                                               const float* __restrict__ depths,
                                                               uint64_t* __restrict__ keys64) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    keys64[idx] = ((uint64_t)tile_keys[idx] << 32) | (uint64_t)__float_as_uint(depths[point_list[idx]]);
}

// ---------------------------------------------------------------------------------------------
// Exact footprint test used to build the per-warp cull mask: q(d) = A dx^2 + 2 B dx dy + C dy^2 is convex, so its
// minimum over the box of offsets d = centre - pixel of an 8x4 pixel footprint is 0 when the centre is inside and
// otherwise lies on one of the four edges (a clamped 1-D minimisation each).  power = -q/2; if even the largest power
// in the footprint is safely below the splat's skip threshold no pixel of that warp can use the splat.
// Conservative: margins absorb rounding, and anything odd (A <= 0, NaN) keeps the splat.
__device__ __forceinline__ bool splat_may_touch(float cx, float cy, float A, float B, float C, float thr, float rx0,
                                                float rx1, float ry0, float ry1) {
    const float x0 = cx - rx1, x1 = cx - rx0, y0 = cy - ry1, y1 = cy - ry0;
    if (x0 <= 0.f && x1 >= 0.f && y0 <= 0.f && y1 >= 0.f) return true;
    if (!(A > 0.f) || !(C > 0.f)) return true;
    const float nBiC = -B / C, nBiA = -B / A;
    float t, q, qmin;
    t = fminf(fmaxf(nBiC * x0, y0), y1); qmin = A * x0 * x0 + (2.f * B * x0 + C * t) * t;
    t = fminf(fmaxf(nBiC * x1, y0), y1); q = A * x1 * x1 + (2.f * B * x1 + C * t) * t; qmin = fminf(qmin, q);
    t = fminf(fmaxf(nBiA * y0, x0), x1); q = C * y0 * y0 + (2.f * B * y0 + A * t) * t; qmin = fminf(qmin, q);
    t = fminf(fmaxf(nBiA * y1, x0), x1); q = C * y1 * y1 + (2.f * B * y1 + A * t) * t; qmin = fminf(qmin, q);
    return !(-0.5f * qmin < thr - (2e-3f + 1e-5f * qmin));
}

// K3.  reference: identifyTileRanges, rasterizer_impl.cu:116-138.  Fused with the gather that
// builds the sorted-order splat stream consumed by the blend kernels through bulk async copies:
//   A[i] = (x, y, conic.x, conic.y)   B[i] = (conic.z, opacity, skip_threshold | cull mask, id bits)
// skip_threshold = -log(255*opacity) - 0.01: power below it can never reach alpha >= 1/255
// (1 % safety margin), so the blend can skip the exp without changing any decision.  Its low 8 mantissa bits
// (a perturbation < 2e-4, inside the margin) carry the cull mask: bit w set <=> the splat can reach the 8x4
// footprint of warp w of its tile (warp w covers columns (w&1)*8.., rows (w>>1)*4.. - blend.cu tile_pixel).
// The same test for all eight footprints of a tile at once.  On a line x = X the quadratic is
// q(X, t) = (A - B^2/C) X^2 + C (t - y*)^2 with y* = -(B/C) X, so the clamped 1-D minimum over t in [y0, y1] is the
// line's base term plus C times the squared distance from y* to the interval; the eight boxes share 4 vertical lines
// (x offsets of columns 0, 7, 8, 15) with 4 row intervals each and 8 horizontal lines (rows 0, 3, 4, 7, ...) with 2
// column intervals each: 32 five-instruction evaluations + 12 line set-ups instead of 8 x 4 independent edge
// minimisations (K3 was FP32-issue bound on them: 711 -> ~300 instructions per duplicate).
// Conservative like splat_may_touch: absolute margin 2e-3 on the power, relative 2e-5 on q, plus 1e-6 of the largest
// cancelling terms (A X^2, C Y^2 at the far corner of the tile); A <= 0, C <= 0 or NaN keeps every footprint.
__device__ __forceinline__ uint32_t tile_cull_mask(float cx, float cy, float A, float B, float C, float thr, float tx,
                                                   float ty) {
    if (!(A > 0.f) || !(C > 0.f) || !(B == B)) return 0xFFu;
    float invC, invA;      // approximate reciprocals (1 ulp-ish): y*, x* and the base terms move by ~1e-7 relative, inside the guard
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(invC) : "f"(C));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(invA) : "f"(A));
    const float nBiC = -B * invC, nBiA = -B * invA;
    const float Dx = fmaf(B, nBiC, A), Dy = fmaf(B, nBiA, C);     // A - B^2/C, C - B^2/A
    // offsets d = centre - pixel of the 4 column lines / 8 row lines (descending pixel = ascending offset order)
    const float X0 = cx - tx, Y0 = cy - ty;
    const float ax = fmaxf(fabsf(X0), fabsf(X0 - 15.f)), ay = fmaxf(fabsf(Y0), fabsf(Y0 - 15.f));
    // cull iff q_min (1 - 2e-5) > 2 (2e-3 - thr) + guard   <=>   -q_min/2 < thr - (2e-3 + 1e-5 q_min) - guard/2
    const float qcut = fmaf(2.f, 2e-3f - thr, 1e-6f * fmaf(A * ax, ax, C * ay * ay));
    // column lines (pixels 0, 7, 8, 15): y* and base term, kept across the four footprint rows
    float ysl[4], basel[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const float X = X0 - (float)((l >> 1) * 8 + (l & 1) * 7);
        ysl[l] = nBiC * X;
        basel[l] = Dx * X * X;
    }
    const bool inx0 = (X0 - 7.f <= 0.f) && (X0 >= 0.f), inx1 = (X0 - 15.f <= 0.f) && (X0 - 8.f >= 0.f);
    uint32_t mask = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {          // footprint row k: pixels 4k .. 4k+3, both footprints of the row
        const float y1 = Y0 - (float)(4 * k), y0 = y1 - 3.f;
        float qv[4];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const float e = fminf(fmaxf(ysl[l], y0), y1) - ysl[l];
            qv[l] = fmaf(C * e, e, basel[l]);
        }
        float qh[2] = {3.0e38f, 3.0e38f};   // min over the row's two lines (pixels 4k and 4k+3), per column interval
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const float Y = r ? y0 : y1;
            const float xs = nBiA * Y, base = Dy * Y * Y;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const float x1 = X0 - (float)(8 * b), x0 = x1 - 7.f;
                const float e = fminf(fmaxf(xs, x0), x1) - xs;
                qh[b] = fminf(qh[b], fmaf(A * e, e, base));
            }
        }
        const bool iny = (y0 <= 0.f) && (y1 >= 0.f);
        const float q0 = fminf(fminf(qv[0], qv[1]), qh[0]), q1 = fminf(fminf(qv[2], qv[3]), qh[1]);
        if ((inx0 && iny) || !(q0 * (1.f - 2e-5f) > qcut)) mask |= 1u << (2 * k);
        if ((inx1 && iny) || !(q1 * (1.f - 2e-5f) > qcut)) mask |= 1u << (2 * k + 1);
    }
    return mask;
}

template <bool CULL_V2>
__global__ void __launch_bounds__(256) ranges_gather_kernel(
    int L, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
    uint2* __restrict__ ranges, float4* __restrict__ splatA, float4* __restrict__ splatB, unsigned gx) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t currtile = keys[idx];   // tile-id keys of the depth-first binning
    if (idx == 0)
        ranges[currtile].x = 0;
    else {
        uint32_t prevtile = keys[idx - 1];
        if (currtile != prevtile) {
            ranges[prevtile].y = idx;
            ranges[currtile].x = idx;
        }
    }
    if (idx == L - 1) ranges[currtile].y = L;

    uint32_t id = point_list[idx];
    float2 xy = means2D[id];
    float4 co = conic_opacity[id];
    float thr = (co.w > 0.0f) ? (-logf(255.0f * co.w) - 0.01f) : 1e30f;   // opacity <= 0: every power (<= 0) skips
    if (!(co.w == co.w)) thr = -1e30f;   // NaN opacity: never skip, let the exact path decide
    const float tx = (float)((currtile % gx) * TILE_X), ty = (float)((currtile / gx) * TILE_Y);
    uint32_t mask = 0;
    if constexpr (CULL_V2) {
        mask = tile_cull_mask(xy.x, xy.y, co.x, co.y, co.z, thr, tx, ty);
    } else {
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            const float rx0 = tx + (float)((w & 1) * 8), ry0 = ty + (float)((w >> 1) * 4);
            if (splat_may_touch(xy.x, xy.y, co.x, co.y, co.z, thr, rx0, rx0 + 7.f, ry0, ry0 + 3.f)) mask |= 1u << w;
        }
    }
    const float thr_packed = __uint_as_float((__float_as_uint(thr) & ~0xFFu) | mask);
    splatA[idx] = make_float4(xy.x, xy.y, co.x, co.y);
    splatB[idx] = make_float4(co.z, co.w, thr_packed, __uint_as_float(id));
}

// K8. reference: checkFrustum, rasterizer_impl.cu:54-66
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D,
                                    const float* __restrict__ viewmatrix, uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    float3 v = xform_4x3(p, viewmatrix);
    present[idx] = v.z > 0.2f;
}


// ---------------------------------------------------------------------------------------------
// K6 + K7 fused: one thread per visible Gaussian turns (dL_dconic, dL_dmean2D) into gradients of
// means3D, cov3D, scale and rotation.  reference: computeCov2DCUDA backward.cu:144-274,
// preprocessCUDA backward.cu:346-396, computeCov3D backward.cu:278-341.  The reference runs two
// kernels and round-trips dL_dmean3D / dL_dcov3D through HBM between them.
struct GeomBackwardParams {
    int P, D, M;
    float focal_x, focal_y, tan_fovx, tan_fovy, scale_modifier;
    const float* means3D;
    const int* radii;
    const float* cov3Ds;          // geometry-state cov3D or cov3D_precomp
    const float* scales;          // may be null (cov3D_precomp path)
    const float* rotations;
    const float* shs;             // may be null
    const unsigned char* clamped;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    const float* dL_dmean2D;      // [P,3]
    const float* dL_dconic;       // [P,4]
    float* dL_dmean3D;            // [P,3]
    float* dL_dcolor;             // [P,3] (SH path input)
    float* dL_dcov3D;             // [P,6]
    float* dL_dsh;
    float* dL_dscale;             // [P,3]
    float* dL_drot;               // [P,4]
};

// `real` is the type the chain is EVALUATED in.  The reference evaluates it in float (backward.cu:144-396), where the
// cov2D -> cov3D -> scale / rotation chain cancels heavily ((denom - a c) is -b^2 by subtraction, the quaternion
// gradient is a sum of differences of O(s^2 dL/dSigma) terms): against an FP64 evaluation of the same formulas
// (oracle/raster_f64.cu) a float chain is off by 2e-4 (scales) / 9e-4 (rotations) relative L2 at config B - the
// reference's own error is the same size - while the float screen-space gradients pushed through an FP64 chain are
// off by 6e-6 / 1.4e-5 (scripts/diag_f64_split.py, profiles/r02_grad_f64.txt).  The kernel is HBM-bound (160 B per
// visible Gaussian, ~600 flops), so real = double is the default: same formulas, same float decisions (clamp flags,
// radii > 0), float inputs and outputs, FP64 in between.  R3DGW_GEOM_BWD_F32=1 selects the float instantiation.
template <typename real>
__global__ void __launch_bounds__(256) geometry_backward_kernel(const GeomBackwardParams p) {
    __shared__ float s_view[16], s_proj[16];
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= p.P || !(p.radii[idx] > 0)) return;

    const float* view = s_view;
    const float* proj = s_proj;
    const float3 mean = {p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]};
    const float* cov3D = p.cov3Ds + 6 * (size_t)idx;
    const float4 gconic4 = reinterpret_cast<const float4*>(p.dL_dconic)[idx];
    const real gcx = gconic4.x, gcy = gconic4.y, gcz = gconic4.w;

    // ---- recompute the forward projection ----------------------------------------------------
    // clamp DECISIONS on the float values the forward pass saw (backward.cu:166-176)
    const float3 tf = xform_4x3(mean, view);
    const float limxf = 1.3f * p.tan_fovx, limyf = 1.3f * p.tan_fovy;
    const float txtzf = tf.x / tf.z, tytzf = tf.y / tf.z;
    const bool clamp_x = txtzf < -limxf || txtzf > limxf, clamp_y = tytzf < -limyf || tytzf > limyf;
    const real x_grad_mul = clamp_x ? real(0) : real(1), y_grad_mul = clamp_y ? real(0) : real(1);
    const real mx = mean.x, my = mean.y, mz = mean.z;
    real tx = real(view[0]) * mx + real(view[4]) * my + real(view[8]) * mz + real(view[12]);
    real ty = real(view[1]) * mx + real(view[5]) * my + real(view[9]) * mz + real(view[13]);
    const real tz_ = real(view[2]) * mx + real(view[6]) * my + real(view[10]) * mz + real(view[14]);
    if (sizeof(real) == sizeof(float)) {      // the reference's float expression, clamp included
        tx = real(fminf(limxf, fmaxf(-limxf, txtzf)) * tf.z);
        ty = real(fminf(limyf, fmaxf(-limyf, tytzf)) * tf.z);
    } else {
        if (clamp_x) tx = real(txtzf < -limxf ? -limxf : limxf) * tz_;
        if (clamp_y) ty = real(tytzf < -limyf ? -limyf : limyf) * tz_;
    }
    const real hx = p.focal_x, hy = p.focal_y;
    const real j00 = hx / tz_, j02 = -(hx * tx) / (tz_ * tz_);
    const real j11 = hy / tz_, j12 = -(hy * ty) / (tz_ * tz_);
    // u, v: first two columns of T = W * J;  W[k][r] = view[k + 4 r]
    real u[3], v[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        u[r] = real(view[0 + 4 * r]) * j00 + real(view[2 + 4 * r]) * j02;
        v[r] = real(view[1 + 4 * r]) * j11 + real(view[2 + 4 * r]) * j12;
    }
    const real V[3][3] = {{real(cov3D[0]), real(cov3D[1]), real(cov3D[2])}, {real(cov3D[1]), real(cov3D[3]), real(cov3D[4])},
                          {real(cov3D[2]), real(cov3D[4]), real(cov3D[5])}};
    real Vu[3], Vv[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        Vu[i] = V[i][0] * u[0] + V[i][1] * u[1] + V[i][2] * u[2];
        Vv[i] = V[i][0] * v[0] + V[i][1] * v[1] + V[i][2] * v[2];
    }
    const real a = (u[0] * Vu[0] + u[1] * Vu[1] + u[2] * Vu[2]) + real(0.3f);
    const real b = u[0] * Vv[0] + u[1] * Vv[1] + u[2] * Vv[2];
    const real c = (v[0] * Vv[0] + v[1] * Vv[1] + v[2] * Vv[2]) + real(0.3f);

    const real denom = a * c - b * b;
    real dL_da = 0, dL_db = 0, dL_dc = 0;
    const real denom2inv = real(1) / ((denom * denom) + real(0.0000001f));
    real gcov[6] = {0, 0, 0, 0, 0, 0};
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * gcx + 2 * b * c * gcy + (denom - a * c) * gcz);
        dL_dc = denom2inv * (-a * a * gcz + 2 * a * b * gcy + (denom - a * c) * gcx);
        dL_db = denom2inv * 2 * (b * c * gcx - (denom + 2 * b * b) * gcy + a * b * gcz);
        // a = u'Vu, b = u'Vv, c = v'Vv  (off-diagonal entries appear twice)
        gcov[0] = u[0] * u[0] * dL_da + u[0] * v[0] * dL_db + v[0] * v[0] * dL_dc;
        gcov[3] = u[1] * u[1] * dL_da + u[1] * v[1] * dL_db + v[1] * v[1] * dL_dc;
        gcov[5] = u[2] * u[2] * dL_da + u[2] * v[2] * dL_db + v[2] * v[2] * dL_dc;
        gcov[1] = 2 * u[0] * u[1] * dL_da + (u[0] * v[1] + u[1] * v[0]) * dL_db + 2 * v[0] * v[1] * dL_dc;
        gcov[2] = 2 * u[0] * u[2] * dL_da + (u[0] * v[2] + u[2] * v[0]) * dL_db + 2 * v[0] * v[2] * dL_dc;
        gcov[4] = 2 * u[2] * u[1] * dL_da + (u[1] * v[2] + u[2] * v[1]) * dL_db + 2 * v[1] * v[2] * dL_dc;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) p.dL_dcov3D[6 * (size_t)idx + i] = (float)gcov[i];

    real gu[3], gv[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        gu[i] = 2 * Vu[i] * dL_da + Vv[i] * dL_db;
        gv[i] = 2 * Vv[i] * dL_dc + Vu[i] * dL_db;
    }
    const real dL_dJ00 = real(view[0]) * gu[0] + real(view[4]) * gu[1] + real(view[8]) * gu[2];
    const real dL_dJ02 = real(view[2]) * gu[0] + real(view[6]) * gu[1] + real(view[10]) * gu[2];
    const real dL_dJ11 = real(view[1]) * gv[0] + real(view[5]) * gv[1] + real(view[9]) * gv[2];
    const real dL_dJ12 = real(view[2]) * gv[0] + real(view[6]) * gv[1] + real(view[10]) * gv[2];
    const real tz = real(1) / tz_, tz2 = tz * tz, tz3 = tz2 * tz;
    const real dL_dtx = x_grad_mul * -hx * tz2 * dL_dJ02;
    const real dL_dty = y_grad_mul * -hy * tz2 * dL_dJ12;
    const real dL_dtz = -hx * tz2 * dL_dJ00 - hy * tz2 * dL_dJ11 + (2 * hx * tx) * tz3 * dL_dJ02 + (2 * hy * ty) * tz3 * dL_dJ12;
    real gmx = real(view[0]) * dL_dtx + real(view[1]) * dL_dty + real(view[2]) * dL_dtz;
    real gmy = real(view[4]) * dL_dtx + real(view[5]) * dL_dty + real(view[6]) * dL_dtz;
    real gmz = real(view[8]) * dL_dtx + real(view[9]) * dL_dty + real(view[10]) * dL_dtz;

    // ---- screen-space mean gradient through the projection (backward.cu:366-387) --------------
    {
        const real hw = real(proj[3]) * mx + real(proj[7]) * my + real(proj[11]) * mz + real(proj[15]);
        const real m_w = real(1) / (hw + real(0.0000001f));
        const real mul1 = (real(proj[0]) * mx + real(proj[4]) * my + real(proj[8]) * mz + real(proj[12])) * m_w * m_w;
        const real mul2 = (real(proj[1]) * mx + real(proj[5]) * my + real(proj[9]) * mz + real(proj[13])) * m_w * m_w;
        const real gx = p.dL_dmean2D[3 * (size_t)idx], gy = p.dL_dmean2D[3 * (size_t)idx + 1];
        gmx += (real(proj[0]) * m_w - real(proj[3]) * mul1) * gx + (real(proj[1]) * m_w - real(proj[3]) * mul2) * gy;
        gmy += (real(proj[4]) * m_w - real(proj[7]) * mul1) * gx + (real(proj[5]) * m_w - real(proj[7]) * mul2) * gy;
        gmz += (real(proj[8]) * m_w - real(proj[11]) * mul1) * gx + (real(proj[9]) * m_w - real(proj[11]) * mul2) * gy;
    }
    float3 gmean = {(float)gmx, (float)gmy, (float)gmz};
    if (p.shs != nullptr) {
        float3 g = sh_to_rgb_bwd(idx, p.D, p.M, p.means3D, p.cam_pos, p.shs, p.clamped, p.dL_dcolor, p.dL_dsh);
        gmean.x += g.x; gmean.y += g.y; gmean.z += g.z;
    }
    p.dL_dmean3D[3 * (size_t)idx + 0] = gmean.x;
    p.dL_dmean3D[3 * (size_t)idx + 1] = gmean.y;
    p.dL_dmean3D[3 * (size_t)idx + 2] = gmean.z;

    // ---- cov3D -> scale / rotation (backward.cu:278-341) ---------------------------------------
    if (p.scales != nullptr) {
        const float4 q = reinterpret_cast<const float4*>(p.rotations)[idx];
        const real r = q.x, x = q.y, y = q.z, z = q.w;
        // R[c][r_]: GLM storage of the reference's column-constructed matrix
        const real R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)},
                              {2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)},
                              {2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)}};
        const real s[3] = {real(p.scale_modifier) * real(p.scales[3 * idx]), real(p.scale_modifier) * real(p.scales[3 * idx + 1]),
                           real(p.scale_modifier) * real(p.scales[3 * idx + 2])};
        real Mm[3][3];   // M = S * R  ->  M[c][r] = s_r * R[c][r]
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
#pragma unroll
            for (int rr = 0; rr < 3; ++rr) Mm[cc][rr] = s[rr] * R[cc][rr];
        const real gS[3][3] = {{gcov[0], real(0.5) * gcov[1], real(0.5) * gcov[2]},
                               {real(0.5) * gcov[1], gcov[3], real(0.5) * gcov[4]},
                               {real(0.5) * gcov[2], real(0.5) * gcov[4], gcov[5]}};
        real gM[3][3];   // 2 * M * dL_dSigma with glm's operator*: (A*B)[c][r] = sum_k A[k][r] B[c][k]
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
#pragma unroll
            for (int rr = 0; rr < 3; ++rr)
                gM[cc][rr] = 2 * (Mm[0][rr] * gS[cc][0] + Mm[1][rr] * gS[cc][1] + Mm[2][rr] * gS[cc][2]);
        // gMt[i][j] = gM[j][i];  dL_dscale_i = sum_j R[j][i] * gM[j][i]
#pragma unroll
        for (int i = 0; i < 3; ++i)
            p.dL_dscale[3 * (size_t)idx + i] = (float)(R[0][i] * gM[0][i] + R[1][i] * gM[1][i] + R[2][i] * gM[2][i]);
        // A[i][j] = dL_dMt[i][j] * s_i   with dL_dMt[i][j] = gM[j][i]
        real A[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) A[i][j] = gM[j][i] * s[i];
        float4 gq;
        gq.x = (float)(2 * z * (A[0][1] - A[1][0]) + 2 * y * (A[2][0] - A[0][2]) + 2 * x * (A[1][2] - A[2][1]));
        gq.y = (float)(2 * y * (A[1][0] + A[0][1]) + 2 * z * (A[2][0] + A[0][2]) + 2 * r * (A[1][2] - A[2][1]) - 4 * x * (A[2][2] + A[1][1]));
        gq.z = (float)(2 * x * (A[1][0] + A[0][1]) + 2 * r * (A[2][0] - A[0][2]) + 2 * z * (A[1][2] + A[2][1]) - 4 * y * (A[2][2] + A[0][0]));
        gq.w = (float)(2 * r * (A[0][1] - A[1][0]) + 2 * x * (A[2][0] + A[0][2]) + 2 * y * (A[1][2] + A[2][1]) - 4 * z * (A[1][1] + A[0][0]));
        reinterpret_cast<float4*>(p.dL_drot)[idx] = gq;
    }
}

}  // namespace r3dgw
