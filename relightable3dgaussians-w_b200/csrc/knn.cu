// knn.cu - simple-knn `distCUDA2`: mean squared distance to the 3 nearest neighbours of every point.
//
// Reference: SimpleKNN::knn, submodules/simple-knn/simple_knn.cu:185-221 (kernels :63-183).
// The reference's search is an EXACT 3-NN (Morton order only drives pruning), and each squared
// distance is d.x*d.x + d.y*d.y + d.z*d.z of (other - self); so any exact search that evaluates the
// same expression returns bit-identical means.  B200-first differences:
//   * no host round trips: the scene AABB stays on the device (the reference does two blocking
//     cudaMemcpy D2H, simple_knn.cu:197,200) and no cudaMalloc/cudaFree per call (caller workspace);
//   * Morton sort with the in-house onesweep sort instead of cub::DeviceRadixSort;
//   * points are re-laid out in sorted order as float4, leaf boxes hold 256 points with a second
//     level of 4096-point super boxes; a CTA of 256 consecutive (spatially coherent) queries votes on
//     each box, stages it in shared memory once and scans it with broadcast reads, instead of every
//     thread walking global memory on its own (boxMeanDist, simple_knn.cu:147-183).
#include <cfloat>
#include "common.cuh"

namespace r3dgw {

constexpr int KNN_BOX = 256;
constexpr int KNN_SUPER = 16;   // leaf boxes per super box

struct Aabb { float3 lo, hi; };

__device__ __forceinline__ Aabb aabb_empty() {
    Aabb a;
    a.lo = {FLT_MAX, FLT_MAX, FLT_MAX};
    a.hi = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    return a;
}
__device__ __forceinline__ void aabb_add(Aabb& a, const float3 p) {
    a.lo.x = fminf(a.lo.x, p.x); a.lo.y = fminf(a.lo.y, p.y); a.lo.z = fminf(a.lo.z, p.z);
    a.hi.x = fmaxf(a.hi.x, p.x); a.hi.y = fmaxf(a.hi.y, p.y); a.hi.z = fmaxf(a.hi.z, p.z);
}
__device__ __forceinline__ void aabb_merge(Aabb& a, const Aabb& b) {
    a.lo.x = fminf(a.lo.x, b.lo.x); a.lo.y = fminf(a.lo.y, b.lo.y); a.lo.z = fminf(a.lo.z, b.lo.z);
    a.hi.x = fmaxf(a.hi.x, b.hi.x); a.hi.y = fmaxf(a.hi.y, b.hi.y); a.hi.z = fmaxf(a.hi.z, b.hi.z);
}
__device__ __forceinline__ Aabb aabb_warp_reduce(Aabb a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Aabb b;
        b.lo.x = __shfl_xor_sync(0xffffffffu, a.lo.x, o); b.lo.y = __shfl_xor_sync(0xffffffffu, a.lo.y, o);
        b.lo.z = __shfl_xor_sync(0xffffffffu, a.lo.z, o); b.hi.x = __shfl_xor_sync(0xffffffffu, a.hi.x, o);
        b.hi.y = __shfl_xor_sync(0xffffffffu, a.hi.y, o); b.hi.z = __shfl_xor_sync(0xffffffffu, a.hi.z, o);
        aabb_merge(a, b);
    }
    return a;
}
// block-wide AABB (256 threads); result valid in thread 0
__device__ __forceinline__ Aabb aabb_block_reduce(Aabb a, Aabb* s_tmp) {
    a = aabb_warp_reduce(a);
    if ((threadIdx.x & 31) == 0) s_tmp[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x < 32) {
        Aabb b = threadIdx.x < (blockDim.x >> 5) ? s_tmp[threadIdx.x] : aabb_empty();
        a = aabb_warp_reduce(b);
    }
    return a;
}

__global__ void __launch_bounds__(256) knn_minmax_partial(int P, const float* __restrict__ pts, Aabb* __restrict__ part) {
    __shared__ Aabb s_tmp[8];
    Aabb a = aabb_empty();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
        aabb_add(a, {pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]});
    a = aabb_block_reduce(a, s_tmp);
    if (threadIdx.x == 0) part[blockIdx.x] = a;
}
__global__ void __launch_bounds__(256) knn_minmax_final(int n, const Aabb* __restrict__ part, Aabb* __restrict__ out) {
    __shared__ Aabb s_tmp[8];
    Aabb a = aabb_empty();
    for (int i = threadIdx.x; i < n; i += blockDim.x) aabb_merge(a, part[i]);
    a = aabb_block_reduce(a, s_tmp);
    if (threadIdx.x == 0) *out = a;
}

// reference: prepMorton / coord2Morton, simple_knn.cu:45-70
__device__ __forceinline__ uint32_t morton_spread(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}
__global__ void __launch_bounds__(256) knn_morton(int P, const float* __restrict__ pts, const Aabb* __restrict__ box,
                                                  uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float3 minn = box->lo, maxx = box->hi;
    const float3 c = {pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]};
    uint32_t x = morton_spread(((c.x - minn.x) / (maxx.x - minn.x)) * ((1 << 10) - 1));
    uint32_t y = morton_spread(((c.y - minn.y) / (maxx.y - minn.y)) * ((1 << 10) - 1));
    uint32_t z = morton_spread(((c.z - minn.z) / (maxx.z - minn.z)) * ((1 << 10) - 1));
    codes[i] = x | (y << 1) | (z << 2);
    idx[i] = i;
}

// sorted-order float4 copy of the points + leaf AABBs (one CTA == one leaf box)
__global__ void __launch_bounds__(KNN_BOX) knn_gather_boxes(int P, const float* __restrict__ pts,
                                                            const uint32_t* __restrict__ order,
                                                            float4* __restrict__ sorted, Aabb* __restrict__ boxes) {
    __shared__ Aabb s_tmp[8];
    const int i = blockIdx.x * KNN_BOX + threadIdx.x;
    Aabb a = aabb_empty();
    if (i < P) {
        const uint32_t src = order[i];
        const float3 p = {pts[3 * src], pts[3 * src + 1], pts[3 * src + 2]};
        sorted[i] = make_float4(p.x, p.y, p.z, __uint_as_float(src));
        aabb_add(a, p);
    }
    a = aabb_block_reduce(a, s_tmp);
    if (threadIdx.x == 0) boxes[blockIdx.x] = a;
}
__global__ void knn_super_boxes(int nboxes, const Aabb* __restrict__ boxes, Aabb* __restrict__ supers) {
    const int sb = blockIdx.x * blockDim.x + threadIdx.x;
    if (sb * KNN_SUPER >= nboxes) return;
    Aabb a = aabb_empty();
    for (int b = sb * KNN_SUPER; b < min(nboxes, (sb + 1) * KNN_SUPER); ++b) aabb_merge(a, boxes[b]);
    supers[sb] = a;
}

// reference: distBoxPoint, simple_knn.cu:119-129
__device__ __forceinline__ float box_dist2(const Aabb& box, const float3& p) {
    float3 diff = {0, 0, 0};
    if (p.x < box.lo.x || p.x > box.hi.x) diff.x = min(abs(p.x - box.lo.x), abs(p.x - box.hi.x));
    if (p.y < box.lo.y || p.y > box.hi.y) diff.y = min(abs(p.y - box.lo.y), abs(p.y - box.hi.y));
    if (p.z < box.lo.z || p.z > box.hi.z) diff.z = min(abs(p.z - box.lo.z), abs(p.z - box.hi.z));
    return diff.x * diff.x + diff.y * diff.y + diff.z * diff.z;
}
// reference: updateKBest<3>, simple_knn.cu:131-145
__device__ __forceinline__ void update_best3(const float3& ref, const float3& point, float* knn) {
    float3 d = {point.x - ref.x, point.y - ref.y, point.z - ref.z};
    float dist = d.x * d.x + d.y * d.y + d.z * d.z;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (knn[j] > dist) {
            float t = knn[j];
            knn[j] = dist;
            dist = t;
        }
    }
}

__global__ void __launch_bounds__(KNN_BOX) knn_search(int P, const float4* __restrict__ sorted,
                                                      const Aabb* __restrict__ boxes, const Aabb* __restrict__ supers,
                                                      int nboxes, int nsupers, float* __restrict__ dists) {
    __shared__ float4 s_pts[KNN_BOX];
    const int idx = blockIdx.x * KNN_BOX + threadIdx.x;
    const bool live = idx < P;
    float3 point = {0, 0, 0};
    uint32_t orig = 0;
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float reject = FLT_MAX;
    if (live) {
        const float4 me = sorted[idx];
        point = {me.x, me.y, me.z};
        orig = __float_as_uint(me.w);
        // seed bound from the +-3 Morton neighbours (simple_knn.cu:156-161)
        for (int i = max(0, idx - 3); i <= min(P - 1, idx + 3); i++) {
            if (i == idx) continue;
            const float4 q = sorted[i];
            update_best3(point, {q.x, q.y, q.z}, best);
        }
        reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
    }
    for (int sb = 0; sb < nsupers; ++sb) {
        bool need_sb = false;
        if (live) {
            const float ds = box_dist2(supers[sb], point);
            need_sb = !(ds > reject || ds > best[2]);
        }
        if (!__syncthreads_or(need_sb)) continue;
        const int b_end = min(nboxes, (sb + 1) * KNN_SUPER);
        for (int b = sb * KNN_SUPER; b < b_end; ++b) {
            bool need = false;
            if (need_sb) {
                const float db = box_dist2(boxes[b], point);
                need = !(db > reject || db > best[2]);
            }
            if (!__syncthreads_or(need)) continue;
            const int base = b * KNN_BOX;
            const int cnt = min(KNN_BOX, P - base);
            if ((int)threadIdx.x < cnt) s_pts[threadIdx.x] = sorted[base + threadIdx.x];
            __syncthreads();
            if (need) {
                for (int i = 0; i < cnt; ++i) {
                    if (base + i == idx) continue;
                    const float4 q = s_pts[i];
                    update_best3(point, {q.x, q.y, q.z}, best);
                }
            }
            __syncthreads();
        }
    }
    if (live) dists[orig] = (best[0] + best[1] + best[2]) / 3.0f;
}

struct KnnLayout {
    size_t codes, codes_alt, idx, idx_alt, sorted, boxes, supers, partial, scene, sortws, total;
};
static KnnLayout knn_layout(int P) {
    KnnLayout L;
    size_t o = 0;
    const size_t Ps = (size_t)(P > 0 ? P : 0);
    const size_t nb = (Ps + KNN_BOX - 1) / KNN_BOX, ns = (nb + KNN_SUPER - 1) / KNN_SUPER;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L.codes = take(Ps * 4); L.codes_alt = take(Ps * 4); L.idx = take(Ps * 4); L.idx_alt = take(Ps * 4);
    L.sorted = take(Ps * 16);
    L.boxes = take(nb * sizeof(Aabb)); L.supers = take(ns * sizeof(Aabb));
    L.partial = take(1024 * sizeof(Aabb)); L.scene = take(sizeof(Aabb));
    L.sortws = take(sort_workspace_bytes(P));
    L.total = o;
    return L;
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" size_t r3dgw_knn_workspace_bytes(int P) { return knn_layout(P).total; }

extern "C" int r3dgw_knn_dist2(int P, const float* points, float* mean_dist2, char* ws, void* stream_) {
    if (P < 0) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!points || !mean_dist2 || !ws) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    KnnLayout L = knn_layout(P);
    ProfScope ps(KID_KNN, stream);
    const int nb = (P + KNN_BOX - 1) / KNN_BOX, ns = (nb + KNN_SUPER - 1) / KNN_SUPER;
    const int pb = std::min(1024, (P + 255) / 256);
    Aabb* scene = (Aabb*)(ws + L.scene);
    knn_minmax_partial<<<pb, 256, 0, stream>>>(P, points, (Aabb*)(ws + L.partial));
    R3DGW_LAUNCH_CHECK();
    knn_minmax_final<<<1, 256, 0, stream>>>(pb, (const Aabb*)(ws + L.partial), scene);
    R3DGW_LAUNCH_CHECK();
    uint32_t* codes = (uint32_t*)(ws + L.codes);
    uint32_t* idx = (uint32_t*)(ws + L.idx);
    knn_morton<<<(P + 255) / 256, 256, 0, stream>>>(P, points, scene, codes, idx);
    R3DGW_LAUNCH_CHECK();
    uint32_t *ks, *vs;
    int rc = sort_pairs_u32_u32(codes, (uint32_t*)(ws + L.codes_alt), idx, (uint32_t*)(ws + L.idx_alt), P, 30,
                                ws + L.sortws, stream, &ks, &vs);
    if (rc) return rc;
    knn_gather_boxes<<<nb, KNN_BOX, 0, stream>>>(P, points, vs, (float4*)(ws + L.sorted), (Aabb*)(ws + L.boxes));
    R3DGW_LAUNCH_CHECK();
    knn_super_boxes<<<(ns + 127) / 128, 128, 0, stream>>>(nb, (const Aabb*)(ws + L.boxes), (Aabb*)(ws + L.supers));
    R3DGW_LAUNCH_CHECK();
    knn_search<<<nb, KNN_BOX, 0, stream>>>(P, (const float4*)(ws + L.sorted), (const Aabb*)(ws + L.boxes),
                                          (const Aabb*)(ws + L.supers), nb, ns, mean_dist2);
    R3DGW_LAUNCH_CHECK();
    return 0;
}
