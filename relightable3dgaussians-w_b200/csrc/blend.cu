// blend.cu - K4 / K5: tile-based front-to-back alpha blending of C-channel per-Gaussian features and
// its backward, for sm_100a.
//
// Reference behaviour: renderCUDA fwd  diff-gaussian-rasterization/cuda_rasterizer/forward.cu:261-374
//                      renderCUDA bwd  diff-gaussian-rasterization/cuda_rasterizer/backward.cu:399-557
// B200-first structure (not the reference's):
//  * each tile's Gaussians arrive as a CONTIGUOUS slice of the sorted-order splat stream built by K3
//    (geometry.cu), so a round of 256/128 splats is fetched with two 1-D bulk async copies
//    (cp.async.bulk -> SASS UBLKCP) completing on an mbarrier, double-buffered against the blend loop;
//    the reference does a dependent id -> xy/conic gather per thread per round;
//  * per-call features are staged in shared memory once per round (the reference re-reads them from
//    global per contributing pixel);
//  * a per-splat conservative threshold skips the exp for pairs that cannot reach alpha >= 1/255;
//  * warps cover 8x4 pixel footprints; backward pre-reduces the 6+C partial gradients over the warp
//    with a halving butterfly, parks them in warp-private shared slots, and issues ONE global
//    atomic per (tile, Gaussian, component) instead of one per (pixel, Gaussian, component);
//  * backward starts at the tile's deepest contributor instead of the end of the list.
// Every decision (power > 0, alpha < 1/255, T*(1-alpha) < 1e-4) is evaluated with the reference's
// expression trees so that n_contrib is bit-identical and colours agree to rounding.
#include <stdlib.h>
#include <cstdlib>
#include "common.cuh"
#include "warp_reduce.cuh"

namespace r3dgw {

// ---- mbarrier / bulk-copy PTX wrappers ---------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// streaming row load that the compiler may not sink below later volatile accesses
__device__ __forceinline__ float ldcs_ordered(const float* p) {
    float v;
    asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// lane -> pixel inside the 16x16 tile: warp w covers the 8x4 block (w%2, w/2)
__device__ __forceinline__ void tile_pixel(unsigned tid, unsigned& px, unsigned& py) {
    unsigned lane = tid & 31, warp = tid >> 5;
    px = (warp & 1) * 8 + (lane & 7);
    py = (warp >> 1) * 4 + (lane >> 3);
}

// Warp-level culling: K3 (geometry.cu) tested every splat of a tile list against the eight 8x4 warp footprints of
// its tile (exact minimum of the conic quadratic over the footprint) and left the result in the low 8 mantissa bits
// of the skip threshold B.z.  A lane reads one splat's bit, the warp ballots, and only survivors are walked.
__device__ __forceinline__ bool splat_touches_warp(const float4 b, unsigned warp) {
    return (__float_as_uint(b.z) >> warp) & 1u;
}

// =================================================================================================
// K4 forward
// =================================================================================================
constexpr int FWD_BATCH = 256;

// channels padded to a multiple of 4 in shared memory: one LDS.128 fetches 4 channels and the accumulation runs
// on packed FP32 pairs (fma.rn.f32x2 -> SASS FFMA2, new on sm_100: two FP32 FMAs per issue slot)
template <int C>
struct ChanPad { static constexpr int CP = (C + 3) / 4 * 4; };

template <int C>
struct FwdSmem {
    float4 A[2][FWD_BATCH];
    float4 B[2][FWD_BATCH];
    float4 feat[2][FWD_BATCH * ChanPad<C>::CP / 4];
    uint64_t bar[2];
};

// WRITE_CACHE: additionally stores, for every (list entry, warp) pair that survives the footprint cull, the 32 lanes'
// effective alphas (0 where the pixel did not take the splat) as one 128-byte row of `alpha_cache`
// (row = list position, 256 floats per entry indexed by thread id).  Later blends over the same geometry
// (blend_forward_cached_kernel) replay those rows instead of re-evaluating the conic / exp / threshold chain.
template <int C, bool WRITE_CACHE, bool FLAT = false>
__global__ void __launch_bounds__(TILE_PIX) blend_forward_kernel(
    const uint2* __restrict__ ranges, const float4* __restrict__ splatA, const float4* __restrict__ splatB,
    int W, int H, const float* __restrict__ features, const float* __restrict__ bg_color,
    float* __restrict__ final_T, uint32_t* __restrict__ n_contrib, float* __restrict__ out_color,
    float* __restrict__ alpha_cache) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    FwdSmem<C>& s = *reinterpret_cast<FwdSmem<C>*>(smem_raw);
    const unsigned tid = threadIdx.x;
    unsigned lx, ly;
    tile_pixel(tid, lx, ly);
    const uint2 pix = {blockIdx.x * TILE_X + lx, blockIdx.y * TILE_Y + ly};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    int done = inside ? 0 : 1;     // int, not bool: a byte-typed flag costs PRMT / SEL pairs per row
    const unsigned lane = tid & 31, warp = tid >> 5;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + FWD_BATCH - 1) / FWD_BATCH;

    if (tid == 0) {
        mbar_init(&s.bar[0], 1);
        mbar_init(&s.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0 && rounds > 0) {
        uint32_t cnt = (uint32_t)min(FWD_BATCH, total);
        mbar_expect_tx(&s.bar[0], cnt * 32u);
        bulk_g2s(s.A[0], splatA + range.x, cnt * 16u, &s.bar[0]);
        bulk_g2s(s.B[0], splatB + range.x, cnt * 16u, &s.bar[0]);
    }

    constexpr int CP = ChanPad<C>::CP;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    uint32_t last_row = 0xFFFFFFFFu;   // FLAT: this round's last taken list slot
    // alpha-cache layout: tile t owns rows [range.x * 8, range.y * 8) of 32 floats; warp w's rows are the `total`
    // rows from range.x * 8 + w * total, filled in the order the warp walks its survivors (row k = its k-th
    // surviving list entry): every replay streams one contiguous run per warp
    float* cache_rows = WRITE_CACHE ? alpha_cache + ((size_t)range.x * 8 + (size_t)warp * total) * 32 + lane : nullptr;
    uint32_t krow = 0;
    float2 Cacc[CP / 2];
#pragma unroll
    for (int q = 0; q < CP / 2; ++q) Cacc[q] = make_float2(0.f, 0.f);

    for (int i = 0; i < rounds; ++i) {
        const int buf = i & 1;
        const int cnt = min(FWD_BATCH, total - i * FWD_BATCH);
        mbar_wait(&s.bar[buf], (uint32_t)(i >> 1) & 1u);
        // stage this round's features (ids ride in B.w)
        if ((int)tid < cnt) {
            const uint32_t id = __float_as_uint(s.B[buf][tid].w);
            const float* f = features + (size_t)id * C;
            float fv[CP];
#pragma unroll
            for (int ch = 0; ch < CP; ++ch) fv[ch] = ch < C ? f[ch] : 0.f;
#pragma unroll
            for (int q = 0; q < CP / 4; ++q)
                s.feat[buf][tid * (CP / 4) + q] = make_float4(fv[4 * q], fv[4 * q + 1], fv[4 * q + 2], fv[4 * q + 3]);
        }
        // one barrier per round: publishes the features, retires the other buffer, and votes
        const int num_done = __syncthreads_count(done);
        if (num_done == TILE_PIX) break;
        if (tid == 0 && i + 1 < rounds) {
            const int nb = buf ^ 1;
            uint32_t ncnt = (uint32_t)min(FWD_BATCH, total - (i + 1) * FWD_BATCH);
            const size_t off = (size_t)range.x + (size_t)(i + 1) * FWD_BATCH;
            mbar_expect_tx(&s.bar[nb], ncnt * 32u);
            bulk_g2s(s.A[nb], splatA + off, ncnt * 16u, &s.bar[nb]);
            bulk_g2s(s.B[nb], splatB + off, ncnt * 16u, &s.bar[nb]);
        }
        if (__all_sync(0xffffffffu, done)) continue;   // whole warp finished: only keep the barriers company
        const uint32_t base = (uint32_t)i * FWD_BATCH;
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int jl = c0 + (int)lane;
            bool keep = false;
            if (jl < cnt) keep = splat_touches_warp(s.B[buf][jl], warp);
            unsigned m = __ballot_sync(0xffffffffu, keep);
            while (m) {
                const int j = c0 + __ffs(m) - 1;
                m &= m - 1;
                float a_eff = 0.f;   // alpha this pixel actually blended with (0: skipped / done)
                if constexpr (FLAT) {
                    // branch-free walk: same expressions and decisions, evaluated by every lane and applied under
                    // one predicate (the breaks of the branchy form cost a reconvergence scope, three branches and
                    // byte-packed flags per row; a lane that would have left early only skips work when the whole
                    // warp does, which the footprint cull has already made rare)
                    const float4 a = s.A[buf][j];
                    const float4 b = s.B[buf][j];
                    const float2 d = {a.x - pixf.x, a.y - pixf.y};
                    const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
                    const float alpha = min(0.99f, b.y * exp(power));
                    const float test_T = T * (1 - alpha);
                    // !(x > 0) && !(x < thr) keeps the branchy form's NaN behaviour (a NaN power goes on to alpha)
                    const bool cand = (done == 0) & !(power > 0.0f) & !(power < b.z) & !(alpha < 1.0f / 255.0f);
                    const bool stop = cand & (test_T < 0.0001f);
                    const bool take = cand & !stop;
                    if (stop) done = 1;
                    const float w = take ? alpha * T : 0.f;
                    const float2 w2 = make_float2(w, w);
                    if (take) {
#pragma unroll
                        for (int q = 0; q < CP / 4; ++q) {
                            const float4 f = s.feat[buf][j * (CP / 4) + q];
                            Cacc[2 * q] = __ffma2_rn(make_float2(f.x, f.y), w2, Cacc[2 * q]);
                            Cacc[2 * q + 1] = __ffma2_rn(make_float2(f.z, f.w), w2, Cacc[2 * q + 1]);
                        }
                        T = test_T;
                        last_row = (uint32_t)j;          // resolved to base + j + 1 once per round, below
                        a_eff = alpha;
                    }
                } else if (!done) {
                    do {
                        const float4 a = s.A[buf][j];
                        const float4 b = s.B[buf][j];
                        const float2 d = {a.x - pixf.x, a.y - pixf.y};
                        const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
                        if (power > 0.0f) break;
                        if (power < b.z) break;   // cannot reach alpha >= 1/255 (conservative, see K3)
                        const float alpha = min(0.99f, b.y * exp(power));
                        if (alpha < 1.0f / 255.0f) break;
                        const float test_T = T * (1 - alpha);
                        if (test_T < 0.0001f) {
                            done = 1;
                            break;
                        }
                        const float w = alpha * T;   // (f*alpha)*T in the reference: differs by <= 1 ulp per term
                        const float2 w2 = make_float2(w, w);
#pragma unroll
                        for (int q = 0; q < CP / 4; ++q) {
                            const float4 f = s.feat[buf][j * (CP / 4) + q];
                            Cacc[2 * q] = __ffma2_rn(make_float2(f.x, f.y), w2, Cacc[2 * q]);
                            Cacc[2 * q + 1] = __ffma2_rn(make_float2(f.z, f.w), w2, Cacc[2 * q + 1]);
                        }
                        T = test_T;
                        last_contributor = base + (uint32_t)j + 1u;
                        a_eff = alpha;
                    } while (false);
                }
                if constexpr (WRITE_CACHE) {   // one coalesced 128-B row per surviving (entry, warp), in walk order
                    __stcs(cache_rows + (size_t)krow * 32, a_eff);
                    ++krow;
                }
            }
            if (__all_sync(0xffffffffu, done)) break;
        }
        if constexpr (FLAT) {
            if (last_row != 0xFFFFFFFFu) last_contributor = base + last_row + 1u;
            last_row = 0xFFFFFFFFu;
        }
    }

    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
#pragma unroll
        for (int ch = 0; ch < C; ++ch) {
            const float c = (ch & 1) ? Cacc[ch >> 1].y : Cacc[ch >> 1].x;
            out_color[(size_t)ch * H * W + pix_id] = c + T * bg_color[ch];
        }
    }
}


// -------------------------------------------------------------------------------------------------
// K4 over cached alphas: every later blend on the same geometry (the reference's render() blends six
// feature sets over identical geometry; a relit-eval grid re-blends a view once per environment map).
// The per-(pixel, splat) decision chain - conic quadratic, exp, alpha threshold, transmittance test - does not
// depend on the features, so the first pass's effective alphas are replayed from HBM: one coalesced 128-byte
// row per surviving (list entry, warp) instead of ~50 issue slots.  T and the colour sums go through exactly
// the same operations as blend_forward_kernel, so the output is bit-identical to the uncached kernel.
// final_T / n_contrib are the producing pass's (identical for every feature set) and are only read.
// -------------------------------------------------------------------------------------------------
constexpr int CACHED_DEPTH = 8;
template <int C>
struct CachedSmem {
    float4 B[2][FWD_BATCH];
    float4 feat[2][FWD_BATCH * ChanPad<C>::CP / 4];
    // per warp: this round's entries that touch its 8x4 footprint; padded by the prefetch depth so the last group can
    // load its list slots unconditionally (a slot past nsurv is only ever paired with alpha = 0)
    uint16_t list[TILE_PIX / 32][FWD_BATCH + CACHED_DEPTH];
    uint32_t red[TILE_PIX / 32];
    uint64_t bar[2];
};

template <int C>
__global__ void __launch_bounds__(TILE_PIX) blend_forward_cached_kernel(
    const uint2* __restrict__ ranges, const float4* __restrict__ splatB, int W, int H,
    const float* __restrict__ features, const float* __restrict__ bg_color, const float* __restrict__ final_T,
    const uint32_t* __restrict__ n_contrib, const float* __restrict__ alpha_cache, float* __restrict__ out_color) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    CachedSmem<C>& s = *reinterpret_cast<CachedSmem<C>*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned lx, ly;
    tile_pixel(tid, lx, ly);
    const uint2 pix = {blockIdx.x * TILE_X + lx, blockIdx.y * TILE_Y + ly};
    const uint32_t pix_id = W * pix.y + pix.x;
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

    // nothing behind the warp's deepest contributor was blended by any of its pixels
    const uint32_t my_last = inside ? n_contrib[pix_id] : 0u;
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, my_last);
    if (lane == 0) s.red[warp] = warp_last;
    if (tid == 0) {
        mbar_init(&s.bar[0], 1);
        mbar_init(&s.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t tile_last = 0;
#pragma unroll
    for (int w = 0; w < TILE_PIX / 32; ++w) tile_last = max(tile_last, s.red[w]);
    const int total = (int)min(range.y - range.x, tile_last);
    const int rounds = (total + FWD_BATCH - 1) / FWD_BATCH;
    if (tid == 0 && rounds > 0) {
        const uint32_t cnt = (uint32_t)min(FWD_BATCH, total);
        mbar_expect_tx(&s.bar[0], cnt * 16u);
        bulk_g2s(s.B[0], splatB + range.x, cnt * 16u, &s.bar[0]);
    }

    constexpr int CP = ChanPad<C>::CP;
    float T = 1.0f;
    float2 Cacc[CP / 2];
#pragma unroll
    for (int q = 0; q < CP / 2; ++q) Cacc[q] = make_float2(0.f, 0.f);
    // this warp's contiguous run of cached rows (layout: blend_forward_kernel); row k = its k-th surviving entry
    const float* my_rows = alpha_cache + ((size_t)range.x * 8 + (size_t)warp * (range.y - range.x)) * 32 + lane;

    for (int i = 0; i < rounds; ++i) {
        const int buf = i & 1;
        const int cnt = min(FWD_BATCH, total - i * FWD_BATCH);
        mbar_wait(&s.bar[buf], (uint32_t)(i >> 1) & 1u);
        if ((int)tid < cnt) {
            const uint32_t id = __float_as_uint(s.B[buf][tid].w);
            const float* f = features + (size_t)id * C;
            float fv[CP];
#pragma unroll
            for (int ch = 0; ch < CP; ++ch) fv[ch] = ch < C ? f[ch] : 0.f;
#pragma unroll
            for (int q = 0; q < CP / 4; ++q)
                s.feat[buf][tid * (CP / 4) + q] = make_float4(fv[4 * q], fv[4 * q + 1], fv[4 * q + 2], fv[4 * q + 3]);
        }
        __syncthreads();   // features staged; the other buffer is free
        if (tid == 0 && i + 1 < rounds) {
            const int nb = buf ^ 1;
            const uint32_t ncnt = (uint32_t)min(FWD_BATCH, total - (i + 1) * FWD_BATCH);
            mbar_expect_tx(&s.bar[nb], ncnt * 16u);
            bulk_g2s(s.B[nb], splatB + (size_t)range.x + (size_t)(i + 1) * FWD_BATCH, ncnt * 16u, &s.bar[nb]);
        }
        const uint32_t base = (uint32_t)i * FWD_BATCH;
        if (base >= warp_last) continue;
        // compact this round's survivors of the footprint cull into a warp-private list, then stream their rows with
        // CACHED_DEPTH loads in flight per lane (the loads do not depend on the T chain)
        uint16_t* list = s.list[warp];
        int nsurv = 0;
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int jl = c0 + (int)lane;
            bool keep = false;
            if (jl < cnt && base + (uint32_t)jl < warp_last) keep = splat_touches_warp(s.B[buf][jl], warp);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) list[nsurv + __popc(m & ((1u << lane) - 1u))] = (uint16_t)jl;
            nsurv += __popc(m);
        }
        __syncwarp();
        const float* rows = my_rows;      // next unread row of the warp's run
        my_rows += (size_t)nsurv * 32;
        // (measured: full groups without the bounds predicate on the eight row loads - the blocked walk that pays in the
        // colour-only backward - make this kernel SLOWER, 0.333 vs 0.234 ms: ptxas then interleaves the loads with the
        // T chain and the lane no longer has eight rows in flight)
        for (int g = 0; g < nsurv; g += CACHED_DEPTH) {
            int js[CACHED_DEPTH];
            float al[CACHED_DEPTH];
#pragma unroll
            for (int u = 0; u < CACHED_DEPTH; ++u) {
                js[u] = (int)list[g + u];
                al[u] = (g + u < nsurv) ? __ldcs(rows + (g + u) * 32) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < CACHED_DEPTH; ++u) {
                if (al[u] > 0.f) {
                    const float alpha = al[u];
                    const float w = alpha * T;
                    const float2 w2 = make_float2(w, w);
#pragma unroll
                    for (int q = 0; q < CP / 4; ++q) {
                        const float4 f = s.feat[buf][js[u] * (CP / 4) + q];
                        Cacc[2 * q] = __ffma2_rn(make_float2(f.x, f.y), w2, Cacc[2 * q]);
                        Cacc[2 * q + 1] = __ffma2_rn(make_float2(f.z, f.w), w2, Cacc[2 * q + 1]);
                    }
                    T = T * (1 - alpha);
                }
            }
        }
        __syncwarp();   // the list is rewritten next round
    }

    if (inside) {
        const float Tf = final_T[pix_id];
#pragma unroll
        for (int ch = 0; ch < C; ++ch) {
            const float c = (ch & 1) ? Cacc[ch >> 1].y : Cacc[ch >> 1].x;
            out_color[(size_t)ch * H * W + pix_id] = c + Tf * bg_color[ch];
        }
    }
}


// =================================================================================================
// K5 backward
// =================================================================================================
constexpr int BWD_WARPS = TILE_PIX / 32;
// true : the warp-reduced partials leave the warp as one RED per (warp, splat, component) straight to L2
// false: they are parked in warp-private smem and flushed as one RED per (tile, splat, component)
#ifndef R3DGW_BWD_DIRECT_RED
#define R3DGW_BWD_DIRECT_RED 0
#endif
constexpr bool BWD_DIRECT = R3DGW_BWD_DIRECT_RED != 0;

// Backward-only shortcuts.  1 - alpha lies in [0.01, 1], so the bare MUFU.RCP needs none of __fdividef's range scaling
// (2 instructions instead of 6; no decision depends on it): on.  exp(power) as ex2.approx(power * log2 e) (2
// instructions instead of expf's 10, ~5 ulp) was measured too - geometry-only pass 1.18 -> 1.12 ms - but is OFF:
// a pair whose alpha then lands on the other side of the 1/255 threshold than in the forward pass (a few per frame)
// perturbs that pixel's transmittance by up to 0.4 % for the rest of its walk, and the colour / opacity gradients of
// some scenes went from the reference's own run-to-run noise (relative L2 5e-7) to 1e-5 - inside the 1e-3 tolerance,
// but not worth 0.6 % of a step.  The forward pass always uses the reference's expf (bit-exact decisions, n_contrib).
#ifndef R3DGW_BWD_FAST_EXP
#define R3DGW_BWD_FAST_EXP 0
#endif
#ifndef R3DGW_BWD_FAST_RCP
#define R3DGW_BWD_FAST_RCP 1
#endif
__device__ __forceinline__ float bwd_exp(float x) {
#if R3DGW_BWD_FAST_EXP
    return __expf(x);
#else
    return exp(x);
#endif
}
__device__ __forceinline__ float bwd_rcp(float x) {
#if R3DGW_BWD_FAST_RCP
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return __fdividef(1.f, x);
#endif
}

// COLOR = false: the geometry-only pass of the deferred drop-in backward (dL/dfeatures already produced per call by
// blend_backward_color_cached_kernel): 6 reduced values per (warp, splat) instead of 6 + C
// QROWS > 0: the six geometry sums of a (warp, splat) pair are not reduced with a shuffle butterfly per splat.  All six
// are sums over the warp's pixels of ONE per-pixel scalar q = G * dL/dalpha times a polynomial of the pixel offset
// d = mean - pixel, and with d = (mean - box origin) - (u, v) they are combinations of six moments of q over the
// box-local integer pixel coordinates: S0, Su, Sv, Suu, Suv, Svv.  Each lane therefore parks only q (one STS);
// every QROWS splats the warp computes the moments as a small matrix product Q[QROWS x 32] . Phi[32 x 6] with
// compile-time Phi (32 / QROWS lanes share a row), and the row's first lane turns them into the six gradients.
// ~10 issue slots per (warp, splat) instead of ~50 (six products + an 11-shuffle halving butterfly).
template <int C, int BATCH, bool COLOR = true, int QROWS = 0>
struct BwdSmem {
    static constexpr int NV = 6 + (COLOR ? C : 0);
    float4 A[2][BATCH];
    float4 B[2][BATCH];
    float4 feat[2][BATCH * ChanPad<C>::CP / 4];
    float acc[BWD_DIRECT ? 1 : BWD_WARPS][BWD_DIRECT ? 1 : BATCH * NV];
    float Q[QROWS > 0 ? BWD_WARPS : 1][QROWS > 0 ? QROWS : 1][33];
    uint16_t Qj[QROWS > 0 ? BWD_WARPS : 1][QROWS > 0 ? QROWS : 1];
    uint32_t mask[BWD_WARPS][(BATCH + 31) / 32];
    uint32_t red[BWD_WARPS];
    uint64_t bar[2];
};

template <int C, int BATCH, bool COLOR, int QROWS, int MINB = 1>
__global__ void __launch_bounds__(TILE_PIX, (MINB > 1 ? MINB : 0)) blend_backward_kernel(
    const uint2* __restrict__ ranges, const float4* __restrict__ splatA, const float4* __restrict__ splatB,
    int W, int H, const float* __restrict__ bg_color, const float* __restrict__ features,
    const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpixels, float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic2D,
    float* __restrict__ dL_dopacity, float* __restrict__ dL_dfeatures) {
    constexpr int NV = 6 + (COLOR ? C : 0);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr bool MOM = QROWS > 0;
    static_assert(!MOM || !BWD_DIRECT, "the moment reduction parks its results per tile");
    BwdSmem<C, BATCH, COLOR, QROWS>& s = *reinterpret_cast<BwdSmem<C, BATCH, COLOR, QROWS>*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned lx, ly;
    tile_pixel(tid, lx, ly);
    const uint2 pix = {blockIdx.x * TILE_X + lx, blockIdx.y * TILE_Y + ly};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    // pixel coordinates of the warp's 8x4 box origin (lane 0's pixel)
    const float box_x = (float)(blockIdx.x * TILE_X + (warp & 1) * 8), box_y = (float)(blockIdx.y * TILE_Y + (warp >> 1) * 4);

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const float T_final = inside ? final_Ts[pix_id] : 0;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0;

    // deepest contributor of the warp / tile: nothing behind it receives gradient
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) s.red[warp] = warp_last;
    if (tid == 0) {
        mbar_init(&s.bar[0], 1);
        mbar_init(&s.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t tile_last = 0;
#pragma unroll
    for (int w = 0; w < BWD_WARPS; ++w) tile_last = max(tile_last, s.red[w]);
    const int total = (int)min(range.y - range.x, tile_last);
    const int rounds = (total + BATCH - 1) / BATCH;
    if (rounds == 0) return;

    if (tid == 0) {
        const int cnt = min(BATCH, total);
        const size_t off = (size_t)range.x + (size_t)(total - cnt);
        mbar_expect_tx(&s.bar[0], (uint32_t)cnt * 32u);
        bulk_g2s(s.A[0], splatA + off, (uint32_t)cnt * 16u, &s.bar[0]);
        bulk_g2s(s.B[0], splatB + off, (uint32_t)cnt * 16u, &s.bar[0]);
    }

    constexpr int CP = ChanPad<C>::CP;
    // The reference carries accum_rec[ch] / last_color[ch] per channel (backward.cu:492-503); both only ever enter
    // dL/dalpha through their dot product with this pixel's dL/dpixel, and the accum_rec recurrence is linear, so the
    // whole channel dimension collapses into two scalars: s = <colour_k, dL/dpixel> and A = <accum_rec, dL/dpixel>,
    // A <- last_alpha * s_last + (1 - last_alpha) * A.  2C fewer live registers and C-independent recurrence cost.
    // (s - A subtracts two rounded dot products: an error of ~1e-7 |s|, three orders of magnitude below the
    // atomic-order noise of the reference's own gradients, profiles/r01_grad_l2.txt.  The cancellation-free form
    // E_k = <colour_k - colour_last, dL/dpixel> + (1 - last_alpha) E_last was measured: same errors, 18 % slower.)
    float2 dL_dpixel[CP / 2];   // channel pairs (packed FP32 math below)
    float rec_A = 0.f, rec_s_last = 0.f;
    float bg_dot_dpixel = 0;
#pragma unroll
    for (int q = 0; q < CP / 2; ++q) {
        const int c0 = 2 * q, c1 = 2 * q + 1;
        dL_dpixel[q].x = (inside && c0 < C) ? dL_dpixels[(size_t)c0 * H * W + pix_id] : 0.f;
        dL_dpixel[q].y = (inside && c1 < C) ? dL_dpixels[(size_t)c1 * H * W + pix_id] : 0.f;
        if (c0 < C) bg_dot_dpixel += bg_color[c0] * dL_dpixel[q].x;
        if (c1 < C) bg_dot_dpixel += bg_color[c1] * dL_dpixel[q].y;
    }
    float last_alpha = 0;
    bool bg_nonzero = false;   // kernel-uniform: the background term vanishes for the usual black background
#pragma unroll
    for (int ch = 0; ch < C; ++ch) bg_nonzero |= (bg_color[ch] != 0.f);
    const float ddelx_dx = 0.5 * W;
    const float ddely_dy = 0.5 * H;

    for (int i = 0; i < rounds; ++i) {
        const int buf = i & 1;
        const int hi = total - i * BATCH;
        const int cnt = min(BATCH, hi);
        const int lo = hi - cnt;
        mbar_wait(&s.bar[buf], (uint32_t)(i >> 1) & 1u);
        if ((int)tid < cnt) {
            const uint32_t id = __float_as_uint(s.B[buf][tid].w);
            const float* f = features + (size_t)id * C;
            float fv[CP];
#pragma unroll
            for (int ch = 0; ch < CP; ++ch) fv[ch] = ch < C ? f[ch] : 0.f;
#pragma unroll
            for (int q = 0; q < CP / 4; ++q)
                s.feat[buf][tid * (CP / 4) + q] = make_float4(fv[4 * q], fv[4 * q + 1], fv[4 * q + 2], fv[4 * q + 3]);
        }
        __syncthreads();   // features staged; previous round (and its flush) finished -> other buffer / acc / mask free
        if constexpr (!BWD_DIRECT) {
            if (lane < (BATCH + 31) / 32) s.mask[warp][lane] = 0;
            __syncwarp();
        }
        if (tid == 0 && i + 1 < rounds) {
            const int nb = buf ^ 1;
            const int nhi = total - (i + 1) * BATCH;
            const int ncnt = min(BATCH, nhi);
            const size_t off = (size_t)range.x + (size_t)(nhi - ncnt);
            mbar_expect_tx(&s.bar[nb], (uint32_t)ncnt * 32u);
            bulk_g2s(s.A[nb], splatA + off, (uint32_t)ncnt * 16u, &s.bar[nb]);
            bulk_g2s(s.B[nb], splatB + off, (uint32_t)ncnt * 16u, &s.bar[nb]);
        }

        int qpos = 0;   // pending rows of Q (warp-uniform)
        auto reduce_q = [&](int n) {
            if constexpr (MOM) {
                __syncwarp();
                constexpr int PARTS = 32 / QROWS, PP = QROWS;      // PP pixels of the box per lane
                const int row = (int)lane % QROWS, part = (int)lane / QROWS;
                float S0 = 0.f, Su = 0.f, Sv = 0.f, Suu = 0.f, Suv = 0.f, Svv = 0.f;
                const float* qrow = &s.Q[warp][row][part * PP];
#pragma unroll
                for (int k = 0; k < PP; ++k) {                     // pixel p = part*PP + k: u = p & 7, v = p >> 3
                    const float q = qrow[k];
                    const float u = (float)(k & 7), v = (float)(k >> 3);
                    S0 += q;
                    if ((k & 7) != 0) { Su = fmaf(q, u, Su); Suu = fmaf(q, u * u, Suu); }
                    if ((k >> 3) != 0) { Sv = fmaf(q, v, Sv); Svv = fmaf(q, v * v, Svv); }
                    if ((k & 7) != 0 && (k >> 3) != 0) Suv = fmaf(q, u * v, Suv);
                }
                if (PARTS > 1) {
                    const float vb = (float)(part * (PP / 8));     // rows of the box before this lane's slice
                    Svv = fmaf(vb * vb, S0, fmaf(2.f * vb, Sv, Svv));
                    Suv = fmaf(vb, Su, Suv);
                    Sv = fmaf(vb, S0, Sv);
#pragma unroll
                    for (int off = QROWS; off < 32; off <<= 1) {
                        S0 += __shfl_xor_sync(0xffffffffu, S0, off);
                        Su += __shfl_xor_sync(0xffffffffu, Su, off);
                        Sv += __shfl_xor_sync(0xffffffffu, Sv, off);
                        Suu += __shfl_xor_sync(0xffffffffu, Suu, off);
                        Suv += __shfl_xor_sync(0xffffffffu, Suv, off);
                        Svv += __shfl_xor_sync(0xffffffffu, Svv, off);
                    }
                }
                if (part == 0 && row < n) {
                    const int jj = (int)s.Qj[warp][row];
                    const float4 a = s.A[buf][jj];
                    const float4 b = s.B[buf][jj];
                    const float mx = a.x - box_x, my = a.y - box_y;          // d = (mx - u, my - v)
                    const float X1 = fmaf(mx, S0, -Su), Y1 = fmaf(my, S0, -Sv);
                    const float XX = fmaf(mx, fmaf(mx, S0, -2.f * Su), Suu);
                    const float YY = fmaf(my, fmaf(my, S0, -2.f * Sv), Svv);
                    const float XY = fmaf(mx, fmaf(my, S0, -Sv), fmaf(-my, Su, Suv));
                    const float o = b.y;                                     // q = G dL/dalpha; dL/dG G = o q
                    float* slot = &s.acc[warp][jj * NV];
                    slot[0] = -o * (a.z * X1 + a.w * Y1) * ddelx_dx;
                    slot[1] = -o * (b.x * Y1 + a.w * X1) * ddely_dy;
                    const float ho = -0.5f * o;
                    slot[2] = ho * XX;
                    slot[3] = ho * XY;
                    slot[4] = ho * YY;
                    slot[5] = S0;
                    atomicOr(&s.mask[warp][jj >> 5], 1u << (jj & 31));
                }
                __syncwarp();
            }
        };
        if ((uint32_t)lo < warp_last) {
            // 32 splats at a time, deepest first: lane-parallel footprint cull, then the survivors back to front
            for (int c0 = ((cnt - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
                const int jl = c0 + (int)lane;
                bool keep = false;
                if (jl < cnt && (uint32_t)(lo + jl) < warp_last)
                    keep = splat_touches_warp(s.B[buf][jl], warp);
                unsigned m = __ballot_sync(0xffffffffu, keep);
                uint32_t my_mask = 0;   // bit b: splat c0+b received something from this warp
                while (m) {
                    const int bit = 31 - __clz(m);
                    m &= ~(1u << bit);
                    const int jj = c0 + bit;
                    const uint32_t k = (uint32_t)(lo + jj);
                    const float4 a = s.A[buf][jj];
                    const float4 b = s.B[buf][jj];
                    const float2 d = {a.x - pixf.x, a.y - pixf.y};
                    const float power = -0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y;
                    bool valid = (k < last_contributor) && !(power > 0.0f) && !(power < b.z);
                    float G = 0.f, alpha = 0.f;
                    if (valid) {
                        G = bwd_exp(power);
                        alpha = min(0.99f, b.y * G);
                        valid = !(alpha < 1.0f / 255.0f);
                    }
                    if (__any_sync(0xffffffffu, valid)) {
                        float g[NV];
                        float qgeo = 0.f;   // MOM: the one geometry scalar of this pixel, G * dL/dalpha
#pragma unroll
                        for (int q = 0; q < NV; ++q) g[q] = 0.f;
                        if (valid) {
                            // 1/(1-alpha) once (alpha <= 0.99): the reference divides twice; the 2-ulp reciprocal is far
                            // inside the gradient tolerance and the reference's own atomic-order noise
                            const float inv_1ma = bwd_rcp(1.f - alpha);
                            T = T * inv_1ma;
                            const float dchannel_dcolor = alpha * T;
                            // per-channel recurrences on packed FP32 pairs (FFMA2 / FMUL2)
                            const float2 dcd2 = make_float2(dchannel_dcolor, dchannel_dcolor);
                            float2 dot2 = make_float2(0.f, 0.f);
#pragma unroll
                            for (int q = 0; q < CP / 4; ++q) {
                                const float4 f = s.feat[buf][jj * (CP / 4) + q];
#pragma unroll
                                for (int h = 0; h < 2; ++h) {
                                    const int p2 = 2 * q + h;
                                    const float2 c2 = h ? make_float2(f.z, f.w) : make_float2(f.x, f.y);
                                    dot2 = __ffma2_rn(c2, dL_dpixel[p2], dot2);
                                    if constexpr (COLOR) {
                                        const float2 gc = __fmul2_rn(dcd2, dL_dpixel[p2]);
                                        if (2 * p2 < C) g[6 + 2 * p2] = gc.x;
                                        if (2 * p2 + 1 < C) g[6 + 2 * p2 + 1] = gc.y;
                                    }
                                }
                            }
                            const float sdot = dot2.x + dot2.y;
                            rec_A = last_alpha * rec_s_last + (1.f - last_alpha) * rec_A;
                            rec_s_last = sdot;
                            float dL_dalpha = (sdot - rec_A) * T;
                            last_alpha = alpha;
                            if (bg_nonzero) dL_dalpha += (-T_final * inv_1ma) * bg_dot_dpixel;
                            qgeo = G * dL_dalpha;
                            if constexpr (!MOM) {
                            const float dL_dG = b.y * dL_dalpha;
                            const float gdx = G * d.x;
                            const float gdy = G * d.y;
                            const float dG_ddelx = -gdx * a.z - gdy * a.w;
                            const float dG_ddely = -gdy * b.x - gdx * a.w;
                            g[0] = dL_dG * dG_ddelx * ddelx_dx;
                            g[1] = dL_dG * dG_ddely * ddely_dy;
                            const float hG = -0.5f * dL_dG;
                            g[2] = hG * gdx * d.x;
                            g[3] = hG * gdx * d.y;
                            g[4] = hG * gdy * d.y;
                            g[5] = G * dL_dalpha;
                            }
                        }
                        if constexpr (MOM) {
                            s.Q[warp][qpos][lane] = qgeo;
                            if (lane == 0) s.Qj[warp][qpos] = (uint16_t)jj;
                            if constexpr (COLOR) {
                                float* slot = &s.acc[warp][jj * NV + 6];
                                warp_reduce_multi<C>(g + 6, lane, [&](int vi, float v) { slot[vi] = v; });
                                my_mask |= 1u << bit;
                            }
                            if (++qpos == QROWS) {
                                reduce_q(QROWS);
                                qpos = 0;
                            }
                        } else if constexpr (BWD_DIRECT) {
                            const uint32_t id = __float_as_uint(b.w);
                            warp_reduce_multi<NV>(g, lane, [&](int vi, float v) {
                                float* dst;
                                if (vi < 2) dst = dL_dmean2D + (size_t)id * 3 + vi;
                                else if (vi < 5) dst = dL_dconic2D + (size_t)id * 4 + (vi == 4 ? 3 : vi - 2);
                                else if (vi == 5) dst = dL_dopacity + id;
                                else dst = dL_dfeatures + (size_t)id * C + (vi - 6);
                                atomicAdd(dst, v);
                            });
                        } else {
                            float* slot = &s.acc[warp][jj * NV];
                            warp_reduce_multi<NV>(g, lane, [&](int vi, float v) { slot[vi] = v; });
                            my_mask |= 1u << bit;
                        }
                    }
                }
                if constexpr (MOM) {
                    if (lane == 0 && my_mask) atomicOr(&s.mask[warp][c0 >> 5], my_mask);
                } else if constexpr (!BWD_DIRECT) {
                    if (lane == 0) s.mask[warp][c0 >> 5] = my_mask;
                }
            }
            if (qpos > 0) reduce_q(qpos);   // before the flush reads acc and the next round overwrites A / B
        }
        if constexpr (BWD_DIRECT) continue;
        __syncthreads();   // all warps' partial sums parked

        // flush: one atomic per (splat, component) that received anything in this tile
        for (int e = (int)tid; e < cnt * NV; e += TILE_PIX) {
            const int jj = e / NV;
            const int vi = e - jj * NV;
            float sum = 0.f;
            bool any = false;
#pragma unroll
            for (int w = 0; w < BWD_WARPS; ++w) {
                if ((s.mask[w][jj >> 5] >> (jj & 31)) & 1u) {
                    sum += s.acc[w][e];
                    any = true;
                }
            }
            if (any) {
                const uint32_t id = __float_as_uint(s.B[buf][jj].w);
                float* dst;
                if (vi < 2) dst = dL_dmean2D + (size_t)id * 3 + vi;
                else if (vi < 5) dst = dL_dconic2D + (size_t)id * 4 + (vi == 4 ? 3 : vi - 2);
                else if (vi == 5) dst = dL_dopacity + id;
                else dst = dL_dfeatures + (size_t)id * C + (vi - 6);
                atomicAdd(dst, sum);
            }
        }
    }
}

// -------------------------------------------------------------------------------------------------
// Colour-only backward over cached alphas (deferred drop-in backward).  render() blends six feature sets over one
// geometry and autograd runs their backwards one by one; the colour gradient of call i only needs
//     dL/dfeature[id][c] = sum over pixels of  alpha * T_before * dL/dpixel[c]
// which is the forward recurrence again, so it is replayed FRONT TO BACK from the alpha cache (no conic, no exp, no
// T division), reduced over the warp, parked in warp-private shared slots and flushed as one atomic per
// (tile, splat, channel).  The geometry gradients of all calls are then produced ONCE by blend_backward_kernel
// <sum C, BATCH, COLOR=false> over the concatenated features / pixel gradients (api: r3dgw_blend_backward with
// dL_dfeatures == NULL), instead of once per call.
// -------------------------------------------------------------------------------------------------
constexpr int BWDC_BATCH = 256;
template <int C, int WROWS>
struct BwdColorSmem {
    float4 B[2][BWDC_BATCH];
    uint16_t list[TILE_PIX / 32][BWDC_BATCH];
    // per warp: the blend weights alpha*T of up to WROWS pending splats x 32 pixels (row padded to 33: both the
    // row-wise stores and the column-wise reads are bank-conflict free), their list positions, and the warp's
    // 32 pixel gradients.  The per-splat reduction over pixels is then W[WROWS x 32] . D[32 x C]: 32 / WROWS lanes
    // share a row (each sums its slice of the pixels, one shuffle step per halving joins them), ~5 issue slots per
    // splat instead of a shuffle butterfly per splat (~25); the row's first lane then issues the reduction of its
    // splat (on average 1.6 warps of a tile touch a splat, so parking the partial sums per tile first would save
    // few atomics and cost a second block barrier per round).  WROWS = 16 keeps the CTA at 33.6 KB (6 per SM).
    float W[TILE_PIX / 32][WROWS][33];
    uint16_t Wj[TILE_PIX / 32][WROWS];   // LEAN: index into `list` of the row's entry, else the entry itself
    float4 D[TILE_PIX / 32][32 * ((C + 3) / 4)];   // the warp's 32 pixel gradients, (C + 3) / 4 channel quads each
    uint32_t red[TILE_PIX / 32];
    uint64_t bar[2];
};

// one 16-byte reduction per (warp, splat) when the gradient rows are padded to 4 floats (sm_90+: red.v4.f32)
__device__ __forceinline__ void red_add_v4(float* dst, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void sts_u16(uint32_t addr, uint16_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}

// LEAN: the per-row bookkeeping of the walk trimmed (42 SASS instructions per cached row before, ~29 after).  The W
// row is stored through a running 32-bit shared address (the compiler rebuilt the address of s.W[warp][pos][lane]
// from the CTA's shared window on every row: 8 of the 20 instructions of the store), and the row's list entry is not
// prefetched next to its alpha (8 LDS.U16 + 24 PRMT/MOV per group of 8 for the -1 default): lane 0 records the
// row's position in the survivor list (a register value) and the reducing lane resolves list -> entry -> id once per
// row.  Same operations on the same values in the same order: bit-identical partial sums.
template <int C, int DEPTH, int WROWS, bool VEC, int LEAN>
__global__ void __launch_bounds__(TILE_PIX, C > 4 ? 4 : 6) blend_backward_color_cached_kernel(
    const uint2* __restrict__ ranges, const float4* __restrict__ splatB, int W, int H,
    const uint32_t* __restrict__ n_contrib, const float* __restrict__ alpha_cache,
    const float* __restrict__ dL_dpixels, float* __restrict__ dL_dfeatures) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BwdColorSmem<C, WROWS>& s = *reinterpret_cast<BwdColorSmem<C, WROWS>*>(smem_raw);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned lx, ly;
    tile_pixel(tid, lx, ly);
    const uint2 pix = {blockIdx.x * TILE_X + lx, blockIdx.y * TILE_Y + ly};
    const uint32_t pix_id = W * pix.y + pix.x;
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const uint32_t my_last = inside ? n_contrib[pix_id] : 0u;
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, my_last);
    if (lane == 0) s.red[warp] = warp_last;
    if (tid == 0) {
        mbar_init(&s.bar[0], 1);
        mbar_init(&s.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t tile_last = 0;
#pragma unroll
    for (int w = 0; w < TILE_PIX / 32; ++w) tile_last = max(tile_last, s.red[w]);
    const int total = (int)min(range.y - range.x, tile_last);
    const int rounds = (total + BWDC_BATCH - 1) / BWDC_BATCH;
    if (rounds == 0) return;
    if (tid == 0) {
        const uint32_t cnt = (uint32_t)min(BWDC_BATCH, total);
        mbar_expect_tx(&s.bar[0], cnt * 16u);
        bulk_g2s(s.B[0], splatB + range.x, cnt * 16u, &s.bar[0]);
    }
    constexpr int CQ = (C + 3) / 4;   // channel quads (C <= 4: one; the fused single-pass call: 16 channels = 4)
    {
        float dp[4 * CQ];
#pragma unroll
        for (int ch = 0; ch < 4 * CQ; ++ch) dp[ch] = (ch < C && inside) ? dL_dpixels[(size_t)ch * H * W + pix_id] : 0.f;
#pragma unroll
        for (int q = 0; q < CQ; ++q) s.D[warp][lane * CQ + q] = make_float4(dp[4 * q], dp[4 * q + 1], dp[4 * q + 2], dp[4 * q + 3]);
    }
    __syncwarp();
    float T = 1.0f;
    const float* my_rows = alpha_cache + ((size_t)range.x * 8 + (size_t)warp * (range.y - range.x)) * 32 + lane;

    for (int i = 0; i < rounds; ++i) {
        const int buf = i & 1;
        const int cnt = min(BWDC_BATCH, total - i * BWDC_BATCH);
        mbar_wait(&s.bar[buf], (uint32_t)(i >> 1) & 1u);
        __syncthreads();   // every warp is done with the other B buffer
        if (tid == 0 && i + 1 < rounds) {
            const int nb = buf ^ 1;
            const uint32_t ncnt = (uint32_t)min(BWDC_BATCH, total - (i + 1) * BWDC_BATCH);
            mbar_expect_tx(&s.bar[nb], ncnt * 16u);
            bulk_g2s(s.B[nb], splatB + (size_t)range.x + (size_t)(i + 1) * BWDC_BATCH, ncnt * 16u, &s.bar[nb]);
        }
        const uint32_t base = (uint32_t)i * BWDC_BATCH;
        if (base >= warp_last) continue;
        uint16_t* list = s.list[warp];
        int nsurv = 0;
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int jl = c0 + (int)lane;
            bool keep = false;
            if (jl < cnt && base + (uint32_t)jl < warp_last) keep = splat_touches_warp(s.B[buf][jl], warp);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) list[nsurv + __popc(m & ((1u << lane) - 1u))] = (uint16_t)jl;
            nsurv += __popc(m);
        }
        __syncwarp();
        const float* rows = my_rows;      // next unread row of the warp's contiguous run (layout: blend_forward_kernel)
        my_rows += (size_t)nsurv * 32;
        int pos = 0;   // pending rows of W (warp-uniform)
        // reduce the pending rows over the 32 pixels: lane l owns row l and issues its splat's C atomics
        int blk_base = 0;   // LEAN == 2: survivor-list index of W row 0
        auto reduce_rows = [&](int n) {
            __syncwarp();
            constexpr int PARTS = 32 / WROWS;          // lanes per row: each sums 32 / PARTS pixels
            const int row = (int)lane % WROWS, part = (int)lane / WROWS;
            float o[4 * CQ];
#pragma unroll
            for (int ch = 0; ch < 4 * CQ; ++ch) o[ch] = 0.f;
            if (PARTS > 1 || row < n) {
                const float* wrow = s.W[warp][row];
#pragma unroll 8
                for (int k = 0; k < 32 / PARTS; ++k) {
                    const int p = part * (32 / PARTS) + k;
                    const float wv = wrow[p];
#pragma unroll
                    for (int q = 0; q < CQ; ++q) {
                        const float4 d = s.D[warp][p * CQ + q];
                        o[4 * q] = fmaf(wv, d.x, o[4 * q]);
                        if (4 * q + 1 < C) o[4 * q + 1] = fmaf(wv, d.y, o[4 * q + 1]);
                        if (4 * q + 2 < C) o[4 * q + 2] = fmaf(wv, d.z, o[4 * q + 2]);
                        if (4 * q + 3 < C) o[4 * q + 3] = fmaf(wv, d.w, o[4 * q + 3]);
                    }
                }
            }
            if (PARTS > 1) {
#pragma unroll
                for (int ch = 0; ch < C; ++ch)
#pragma unroll
                    for (int off = WROWS; off < 32; off <<= 1) o[ch] += __shfl_xor_sync(0xffffffffu, o[ch], off);
            }
            bool nz = LEAN != 2;   // LEAN == 2 also parks rows no pixel blended (1.3 % at config B): nothing to add
            if (LEAN == 2) {
#pragma unroll
                for (int ch = 0; ch < C; ++ch) nz |= (o[ch] != 0.f);
            }
            if (part == 0 && row < n && nz) {
                const uint32_t j = LEAN == 2 ? (uint32_t)list[blk_base + row]
                                 : LEAN == 1 ? (uint32_t)list[s.Wj[warp][row]] : (uint32_t)s.Wj[warp][row];
                const uint32_t id = __float_as_uint(s.B[buf][j].w);
                if (VEC) {   // rows of 4 CQ floats (C <= 4: padded to 4; C = 8, 12, 16: dense)
#pragma unroll
                    for (int q = 0; q < CQ; ++q)
                        red_add_v4(dL_dfeatures + (size_t)id * (4 * CQ) + 4 * q, o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
                } else {
                    float* dst = dL_dfeatures + (size_t)id * C;
#pragma unroll
                    for (int ch = 0; ch < C; ++ch) atomicAdd(dst + ch, o[ch]);
                }
            }
            __syncwarp();
        };
        const uint32_t w_base = smem_u32(&s.W[warp][0][lane]), wj_base = smem_u32(&s.Wj[warp][0]);
        uint32_t w_addr = w_base;   // LEAN: &s.W[warp][pos][lane] as a running shared address
        if constexpr (LEAN == 2) {
            // blocks of WROWS rows at fixed W slots: no per-row vote / counter / list-position store - a row no pixel
            // blended is parked as zeros like any other - and no bounds predicate on the loads of a full block
            static_assert(LEAN != 2 || WROWS % DEPTH == 0, "a block is a whole number of load groups");
            int g0 = 0;
            for (; g0 + WROWS <= nsurv; g0 += WROWS) {
                const float* rp = rows + (size_t)g0 * 32;
#pragma unroll
                for (int h = 0; h < WROWS / DEPTH; ++h) {
                    float al[DEPTH];
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) al[u] = ldcs_ordered(rp + (h * DEPTH + u) * 32);
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) {
                        sts_f32(w_base + (h * DEPTH + u) * 33 * 4, al[u] * T);
                        T = T * (1 - al[u]);
                    }
                }
                blk_base = g0;
                reduce_rows(WROWS);
            }
            const int rem = nsurv - g0;
            if (rem > 0) {
                const float* rp = rows + (size_t)g0 * 32;
#pragma unroll
                for (int h = 0; h < WROWS / DEPTH; ++h) {
                    if (h * DEPTH >= rem) break;
                    float al[DEPTH];
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) al[u] = (h * DEPTH + u < rem) ? __ldcs(rp + (h * DEPTH + u) * 32) : 0.f;
#pragma unroll
                    for (int u = 0; u < DEPTH; ++u) {
                        if (h * DEPTH + u < rem) {
                            sts_f32(w_base + (h * DEPTH + u) * 33 * 4, al[u] * T);
                            T = T * (1 - al[u]);
                        }
                    }
                }
                blk_base = g0;
                reduce_rows(rem);
            }
            continue;   // next round
        }
        for (int g0 = 0; g0 < nsurv; g0 += DEPTH) {
            int js[DEPTH];
            float al[DEPTH];
#pragma unroll
            for (int u = 0; u < DEPTH; ++u) {
                if (!LEAN) js[u] = (g0 + u < nsurv) ? (int)list[g0 + u] : -1;
                al[u] = (g0 + u < nsurv) ? __ldcs(rows + (g0 + u) * 32) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < DEPTH; ++u) {
                const float alpha = al[u];
                if (!__any_sync(0xffffffffu, alpha > 0.f)) continue;
                if (LEAN) {
                    sts_f32(w_addr, alpha * T);       // 0 for pixels that did not take the splat
                    w_addr += 33 * 4;
                    T = T * (1 - alpha);
                    if (lane == 0) sts_u16(wj_base + 2 * pos, (uint16_t)(g0 + u));
                } else {
                    s.W[warp][pos][lane] = alpha * T;
                    T = T * (1 - alpha);
                    if (lane == 0) s.Wj[warp][pos] = (uint16_t)js[u];
                }
                if (++pos == WROWS) {
                    reduce_rows(WROWS);
                    pos = 0;
                    w_addr = w_base;
                }
            }
        }
        if (pos > 0) reduce_rows(pos);   // before the next round's barrier lets B[buf] be overwritten
    }
}

template <int C, int DEPTH, int WROWS, bool VEC, int LEAN>
static int launch_backward_color_cached(dim3 grid, const uint2* ranges, const float4* B, int W, int H,
                                        const uint32_t* n_contrib, const float* alpha_cache, const float* dL_dpix,
                                        float* dL_dfeatures, cudaStream_t stream) {
    const int smem = (int)sizeof(BwdColorSmem<C, WROWS>);
    R3DGW_CUDA_TRY(cudaFuncSetAttribute((blend_backward_color_cached_kernel<C, DEPTH, WROWS, VEC, LEAN>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    {
        ProfScope ps(KID_BLEND_BWD_COLOR, stream);
        blend_backward_color_cached_kernel<C, DEPTH, WROWS, VEC, LEAN><<<grid, TILE_PIX, smem, stream>>>(ranges, B, W, H, n_contrib, alpha_cache, dL_dpix,
                                                                              dL_dfeatures);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

template <int C>
static int launch_backward_color_cached_depth(dim3 grid, const uint2* ranges, const float4* B, int W, int H,
                                              const uint32_t* n_contrib, const float* alpha_cache, const float* dL_dpix,
                                              float* dL_dfeatures, bool vec, cudaStream_t stream) {
    // measured on B200 (config B, 5 calls per step): rows-per-reduce 32 / 8 loads in flight 2.27 ms, 16 / 8 2.05 ms
    // (33.6 KB of smem instead of 50.7 KB: 6 CTAs per SM), deeper load queues (12, 16) no faster; with 16-byte
    // reductions into padded gradient rows 1.98 ms
    // C > 4 (the fused 16-channel call): 4 CTAs per SM by launch bounds (64 registers) 0.75 -> 0.64 ms at C = 16; 4 loads
    // in flight 0.67 ms, 8 rows per reduce 0.88 ms
    // R3DGW_BWDC_LEAN=0 selects the pre-diet walk (A/B timing only: 0.369 -> 0.333 ms per call at config B)
    static const int lean = [] { const char* e = getenv("R3DGW_BWDC_LEAN"); return e ? atoi(e) : 2; }();
#define R3DGW_BWDC_GO(L) (vec ? launch_backward_color_cached<C, 8, 16, true, L>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, stream) \
                              : launch_backward_color_cached<C, 8, 16, false, L>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, stream))
    if (lean == 0) return R3DGW_BWDC_GO(0);
    if (lean == 1) return R3DGW_BWDC_GO(1);
    return R3DGW_BWDC_GO(2);
#undef R3DGW_BWDC_GO
}

int blend_backward_color_cached_dispatch(int C, dim3 grid, const uint2* ranges, const float4* B, int W, int H,
                                         const uint32_t* n_contrib, const float* alpha_cache, const float* dL_dpix,
                                         float* dL_dfeatures, int grad_stride, cudaStream_t stream) {
    const int padded = (C + 3) / 4 * 4;
    if (grad_stride != C && grad_stride != padded) return R3DGW_ERR_BAD_ARG;
    const bool vec = grad_stride == padded && (reinterpret_cast<uintptr_t>(dL_dfeatures) & 15) == 0;
    if (grad_stride != C && !vec) return R3DGW_ERR_BAD_ARG;
    switch (C) {
        case 1: return launch_backward_color_cached_depth<1>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 2: return launch_backward_color_cached_depth<2>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 3: return launch_backward_color_cached_depth<3>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 4: return launch_backward_color_cached_depth<4>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 8: return launch_backward_color_cached_depth<8>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 12: return launch_backward_color_cached_depth<12>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 15: return launch_backward_color_cached_depth<15>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        case 16: return launch_backward_color_cached_depth<16>(grid, ranges, B, W, H, n_contrib, alpha_cache, dL_dpix, dL_dfeatures, vec, stream);
        default: return R3DGW_ERR_UNSUPPORTED;
    }
}

// =================================================================================================
// host-side dispatch
// =================================================================================================
template <int C>
static int launch_forward(dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                          const float* features, const float* bg, float* final_T, uint32_t* n_contrib,
                          float* out_color, float* alpha_cache, int cache_mode, cudaStream_t stream) {
    if (cache_mode == 2) {   // replay cached alphas
        const int smem = (int)sizeof(CachedSmem<C>);
        R3DGW_CUDA_TRY(cudaFuncSetAttribute(blend_forward_cached_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        {
            ProfScope ps(KID_BLEND_FWD_CACHED, stream);
            blend_forward_cached_kernel<C><<<grid, TILE_PIX, smem, stream>>>(ranges, B, W, H, features, bg, final_T, n_contrib,
                                                                            alpha_cache, out_color);
        }
        R3DGW_LAUNCH_CHECK();
        return 0;
    }
    const int smem = (int)sizeof(FwdSmem<C>);
    // R3DGW_FWD_FLAT=0: the branchy per-row walk (A/B)
    static const bool flat = [] { const char* e = getenv("R3DGW_FWD_FLAT"); return !(e && e[0] == '0'); }();
#define R3DGW_FWD_LAUNCH(WC, FL, CACHE)                                                                                  \
    do {                                                                                                                \
        R3DGW_CUDA_TRY(cudaFuncSetAttribute((blend_forward_kernel<C, WC, FL>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
        ProfScope ps(KID_BLEND_FWD, stream);                                                                            \
        blend_forward_kernel<C, WC, FL><<<grid, TILE_PIX, smem, stream>>>(ranges, A, B, W, H, features, bg, final_T,    \
                                                                         n_contrib, out_color, CACHE);                  \
    } while (0)
    if (cache_mode == 1) {   // compute and record the alphas
        if (flat) R3DGW_FWD_LAUNCH(true, true, alpha_cache); else R3DGW_FWD_LAUNCH(true, false, alpha_cache);
    } else {
        if (flat) R3DGW_FWD_LAUNCH(false, true, nullptr); else R3DGW_FWD_LAUNCH(false, false, nullptr);
    }
#undef R3DGW_FWD_LAUNCH
    R3DGW_LAUNCH_CHECK();
    return 0;
}

template <int C, int BATCH, bool COLOR, int QROWS = 0, int MINB = 1>
static int launch_backward(dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                           const float* bg, const float* features, const float* final_T,
                           const uint32_t* n_contrib, const float* dL_dpix, float* dL_dmean2D,
                           float* dL_dconic, float* dL_dopacity, float* dL_dfeatures, cudaStream_t stream) {
    const int smem = (int)sizeof(BwdSmem<C, BATCH, COLOR, QROWS>);
    R3DGW_CUDA_TRY(cudaFuncSetAttribute((blend_backward_kernel<C, BATCH, COLOR, QROWS, MINB>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    {
        ProfScope ps(COLOR ? KID_BLEND_BWD : KID_BLEND_BWD_GEOM, stream);
        blend_backward_kernel<C, BATCH, COLOR, QROWS, MINB><<<grid, TILE_PIX, smem, stream>>>(ranges, A, B, W, H, bg, features, final_T, n_contrib,
                                                                               dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dfeatures);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

// -------------------------------------------------------------------------------------------------
// Geometry-only backward over cached alphas, FRONT TO BACK (round 2).
// blend_backward_kernel<.., COLOR=false> walks every tile list back to front and re-derives each pair's alpha from the
// conic (quadratic, exp, thresholds) because the reference's recurrence needs the suffix sums
//     dL/dalpha_k = T_k s_k - ( sum_{j>k} w_j s_j + T_final <bg, g> ) / (1 - alpha_k),   w_j = alpha_j T_j, s_j = <c_j, g>
// (backward.cu:464-520: accum_rec / last_color, collapsed to the scalar s).  But the bracket is the pixel's whole
// dot product minus a PREFIX:  sum_{j>k} w_j s_j + T_final <bg, g>  =  <out_pixel, g> - sum_{j<=k} w_j s_j,  and
// <out_pixel, g> is known before the walk starts from the forward image.  So the pass runs in the direction the alpha
// cache was written - one coalesced 128-byte row per surviving (warp, splat) pair, the same rows in the same order as
// the forward replay and the colour-only backward - with T and the prefix as running products / sums: no conic
// quadratic, no exp (G = alpha / opacity; only a pixel whose alpha sits on the 0.99 clamp re-evaluates exp), no
// division chain T / (1 - alpha).  The six geometry sums per (warp, splat) come from the same moment product
// Q[16 x 32] . Phi[32 x 6] as the back-to-front kernel and leave as 4 atomics.
// Rounding: <out, g> - prefix cancels for the deepest contributors of a pixel, whose true bracket is T_final <bg, g>
// (0 for a black background); the absolute error is ~1e-7 |<out, g>| per pair, the size of the reference's own
// error in T = T_final / prod (1 - alpha) - measured against the FP64 oracle in tests/test_gpu_gradient_f64.py.
// -------------------------------------------------------------------------------------------------
constexpr int BWDG_BATCH = 64;
constexpr int BWDG_QROWS = 16;
constexpr int BWDG_DEPTH = 8;
constexpr int BWDG_MAX_C = 16;
template <int C>
struct BwdGeoCachedSmem {
    float4 A[2][BWDG_BATCH];
    float4 B[2][BWDG_BATCH];
    float4 feat[2][BWDG_BATCH * ChanPad<C>::CP / 4];
    uint16_t list[TILE_PIX / 32][BWDG_BATCH + BWDG_QROWS];
    float Q[TILE_PIX / 32][BWDG_QROWS][33];
    uint32_t red[TILE_PIX / 32];
    uint64_t bar[2];
};

// Where each of the pass's C channels lives: the deferred pass covers several blend calls, whose features / forward
// images / pixel gradients / backgrounds are separate tensors - the table addresses them in place (no concatenation
// copies: 0.17 ms of torch.cat per step at config B).  feat[ch] already points at the channel's column of its call's
// [P, c_call] matrix (row stride feat_stride[ch] floats); dpix / out point at the channel's [H, W] plane.
struct BwdGeoChannels {
    const float* feat[BWDG_MAX_C];
    const float* dpix[BWDG_MAX_C];
    const float* out[BWDG_MAX_C];
    const float* bg[BWDG_MAX_C];
    int feat_stride[BWDG_MAX_C];
};

template <int C>
__global__ void __launch_bounds__(TILE_PIX, 4) blend_backward_geometry_cached_kernel(
    const uint2* __restrict__ ranges, const float4* __restrict__ splatA, const float4* __restrict__ splatB, int W, int H,
    const __grid_constant__ BwdGeoChannels chn, const uint32_t* __restrict__ n_contrib, const float* __restrict__ alpha_cache,
    float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic2D, float* __restrict__ dL_dopacity) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BwdGeoCachedSmem<C>& s = *reinterpret_cast<BwdGeoCachedSmem<C>*>(smem_raw);
    constexpr int CP = ChanPad<C>::CP, QROWS = BWDG_QROWS, DEPTH = BWDG_DEPTH, BATCH = BWDG_BATCH;
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned lx, ly;
    tile_pixel(tid, lx, ly);
    const uint2 pix = {blockIdx.x * TILE_X + lx, blockIdx.y * TILE_Y + ly};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)W && pix.y < (unsigned)H;
    const float box_x = (float)(blockIdx.x * TILE_X + (warp & 1) * 8), box_y = (float)(blockIdx.y * TILE_Y + (warp >> 1) * 4);
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const uint32_t my_last = inside ? n_contrib[pix_id] : 0u;
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, my_last);
    if (lane == 0) s.red[warp] = warp_last;
    if (tid == 0) {
        mbar_init(&s.bar[0], 1);
        mbar_init(&s.bar[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t tile_last = 0;
#pragma unroll
    for (int w = 0; w < TILE_PIX / 32; ++w) tile_last = max(tile_last, s.red[w]);
    const int total = (int)min(range.y - range.x, tile_last);
    const int rounds = (total + BATCH - 1) / BATCH;
    if (rounds == 0) return;
    if (tid == 0) {
        const uint32_t cnt = (uint32_t)min(BATCH, total);
        mbar_expect_tx(&s.bar[0], cnt * 32u);
        bulk_g2s(s.A[0], splatA + range.x, cnt * 16u, &s.bar[0]);
        bulk_g2s(s.B[0], splatB + range.x, cnt * 16u, &s.bar[0]);
    }
    float2 dpix[CP / 2];
    float Sb = 0.f;          // <forward pixel, dL/dpixel> over all channels (background term included)
    float bg_dot = 0.f;      // <background, dL/dpixel>
#pragma unroll
    for (int q = 0; q < CP / 2; ++q) {
        const int c0 = 2 * q, c1 = 2 * q + 1;
        dpix[q].x = (inside && c0 < C) ? chn.dpix[c0 < C ? c0 : 0][pix_id] : 0.f;
        dpix[q].y = (inside && c1 < C) ? chn.dpix[c1 < C ? c1 : 0][pix_id] : 0.f;
        if (inside && c0 < C) Sb = fmaf(chn.out[c0 < C ? c0 : 0][pix_id], dpix[q].x, Sb);
        if (inside && c1 < C) Sb = fmaf(chn.out[c1 < C ? c1 : 0][pix_id], dpix[q].y, Sb);
        if (c0 < C) bg_dot = fmaf(*chn.bg[c0 < C ? c0 : 0], dpix[q].x, bg_dot);
        if (c1 < C) bg_dot = fmaf(*chn.bg[c1 < C ? c1 : 0], dpix[q].y, bg_dot);
    }
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    float T = 1.0f, prefix = 0.f;
    const float* my_rows = alpha_cache + ((size_t)range.x * 8 + (size_t)warp * (range.y - range.x)) * 32 + lane;

    for (int i = 0; i < rounds; ++i) {
        const int buf = i & 1;
        const int cnt = min(BATCH, total - i * BATCH);
        mbar_wait(&s.bar[buf], (uint32_t)(i >> 1) & 1u);
        if ((int)tid < cnt) {
            const uint32_t id = __float_as_uint(s.B[buf][tid].w);
            float fv[CP];
#pragma unroll
            for (int ch = 0; ch < CP; ++ch)
                fv[ch] = ch < C ? chn.feat[ch < C ? ch : 0][(size_t)id * chn.feat_stride[ch < C ? ch : 0]] : 0.f;
#pragma unroll
            for (int q = 0; q < CP / 4; ++q)
                s.feat[buf][tid * (CP / 4) + q] = make_float4(fv[4 * q], fv[4 * q + 1], fv[4 * q + 2], fv[4 * q + 3]);
        }
        __syncthreads();   // features staged; every warp is done with the other A / B / feat buffers
        if (tid == 0 && i + 1 < rounds) {
            const int nb = buf ^ 1;
            const uint32_t ncnt = (uint32_t)min(BATCH, total - (i + 1) * BATCH);
            const size_t off = (size_t)range.x + (size_t)(i + 1) * BATCH;
            mbar_expect_tx(&s.bar[nb], ncnt * 32u);
            bulk_g2s(s.A[nb], splatA + off, ncnt * 16u, &s.bar[nb]);
            bulk_g2s(s.B[nb], splatB + off, ncnt * 16u, &s.bar[nb]);
        }
        const uint32_t base = (uint32_t)i * BATCH;
        if (base >= warp_last) continue;
        uint16_t* list = s.list[warp];
        int nsurv = 0;
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int jl = c0 + (int)lane;
            bool keep = false;
            if (jl < cnt && base + (uint32_t)jl < warp_last) keep = splat_touches_warp(s.B[buf][jl], warp);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) list[nsurv + __popc(m & ((1u << lane) - 1u))] = (uint16_t)jl;
            nsurv += __popc(m);
        }
        __syncwarp();
        const float* rows = my_rows;
        my_rows += (size_t)nsurv * 32;

        for (int g0 = 0; g0 < nsurv; g0 += QROWS) {
            const int n = min(QROWS, nsurv - g0);
            const float* rp = rows + (size_t)g0 * 32;
#pragma unroll
            for (int h = 0; h < QROWS / DEPTH; ++h) {
                if (h * DEPTH >= n) break;
                float al[DEPTH];
#pragma unroll
                for (int u = 0; u < DEPTH; ++u) al[u] = (h * DEPTH + u < n) ? __ldcs(rp + (h * DEPTH + u) * 32) : 0.f;
#pragma unroll
                for (int u = 0; u < DEPTH; ++u) {
                    const int r = h * DEPTH + u;
                    if (r >= n) break;
                    const int jj = (int)list[g0 + r];
                    const float alpha = al[u];
                    float2 dot2 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int q = 0; q < CP / 4; ++q) {
                        const float4 f = s.feat[buf][jj * (CP / 4) + q];
                        dot2 = __ffma2_rn(make_float2(f.x, f.y), dpix[2 * q], dot2);
                        dot2 = __ffma2_rn(make_float2(f.z, f.w), dpix[2 * q + 1], dot2);
                    }
                    const float sdot = dot2.x + dot2.y;
                    prefix = fmaf(alpha * T, sdot, prefix);
                    const float one_m = 1.f - alpha;
                    const float T_next = T * one_m;
                    // behind the pixel's LAST contributor only the background is left: exactly T_final <bg, g>, no
                    // cancellation there (that pair carries the largest 1 / (1 - alpha) of an opaque surface)
                    const float bracket = (base + (uint32_t)jj + 1u == my_last) ? T_next * bg_dot : Sb - prefix;
                    const float dL_dalpha = fmaf(T, sdot, -bracket * bwd_rcp(one_m));
                    T = T_next;
                    // parked: q' = opacity * G * dL/dalpha.  Off the clamp opacity * G IS alpha, so the row walk needs
                    // neither the splat's B entry nor a reciprocal; the common factor 1 / opacity is applied once per
                    // (warp, splat) where the six sums are turned into gradients (it cancels in five of them)
                    float oG = alpha;
                    if (alpha == 0.99f) {      // on the clamp: alpha no longer determines G
                        const float4 a = s.A[buf][jj];
                        const float4 b = s.B[buf][jj];
                        const float2 d = {a.x - pixf.x, a.y - pixf.y};
                        oG = b.y * bwd_exp(-0.5f * (a.z * d.x * d.x + b.x * d.y * d.y) - a.w * d.x * d.y);
                    }
                    s.Q[warp][r][lane] = oG * dL_dalpha;     // alpha == 0 (pixel did not take the splat): exactly 0
                }
            }
            // ---- six moments of q over the box-local pixel coordinates, 32 / QROWS lanes per row ----
            __syncwarp();
            {
                constexpr int PARTS = 32 / QROWS, PP = QROWS;
                const int row = (int)lane % QROWS, part = (int)lane / QROWS;
                float S0 = 0.f, Su = 0.f, Sv = 0.f, Suu = 0.f, Suv = 0.f, Svv = 0.f;
                const float* qrow = &s.Q[warp][row][part * PP];
#pragma unroll
                for (int k = 0; k < PP; ++k) {                     // pixel p = part*PP + k: u = p & 7, v = p >> 3
                    const float q = qrow[k];
                    const float u = (float)(k & 7), v = (float)(k >> 3);
                    S0 += q;
                    if ((k & 7) != 0) { Su = fmaf(q, u, Su); Suu = fmaf(q, u * u, Suu); }
                    if ((k >> 3) != 0) { Sv = fmaf(q, v, Sv); Svv = fmaf(q, v * v, Svv); }
                    if ((k & 7) != 0 && (k >> 3) != 0) Suv = fmaf(q, u * v, Suv);
                }
                if (PARTS > 1) {
                    const float vb = (float)(part * (PP / 8));     // rows of the box before this lane's slice
                    Svv = fmaf(vb * vb, S0, fmaf(2.f * vb, Sv, Svv));
                    Suv = fmaf(vb, Su, Suv);
                    Sv = fmaf(vb, S0, Sv);
#pragma unroll
                    for (int off = QROWS; off < 32; off <<= 1) {
                        S0 += __shfl_xor_sync(0xffffffffu, S0, off);
                        Su += __shfl_xor_sync(0xffffffffu, Su, off);
                        Sv += __shfl_xor_sync(0xffffffffu, Sv, off);
                        Suu += __shfl_xor_sync(0xffffffffu, Suu, off);
                        Suv += __shfl_xor_sync(0xffffffffu, Suv, off);
                        Svv += __shfl_xor_sync(0xffffffffu, Svv, off);
                    }
                }
                const bool nz = (S0 != 0.f) | (Su != 0.f) | (Sv != 0.f) | (Suu != 0.f) | (Suv != 0.f) | (Svv != 0.f);
                if (part == 0 && row < n && nz) {
                    const int jj = (int)list[g0 + row];
                    const float4 a = s.A[buf][jj];
                    const float4 b = s.B[buf][jj];
                    const float mx = a.x - box_x, my = a.y - box_y;          // d = (mx - u, my - v)
                    const float X1 = fmaf(mx, S0, -Su), Y1 = fmaf(my, S0, -Sv);
                    const float XX = fmaf(mx, fmaf(mx, S0, -2.f * Su), Suu);
                    const float YY = fmaf(my, fmaf(my, S0, -2.f * Sv), Svv);
                    const float XY = fmaf(mx, fmaf(my, S0, -Sv), fmaf(-my, Su, Suv));
                    // the sums carry opacity * G: dL/dG * G = opacity * q cancels the factor, dL/dopacity = sum(q) keeps it
                    const uint32_t id = __float_as_uint(b.w);
                    atomicAdd(dL_dmean2D + (size_t)id * 3, -(a.z * X1 + a.w * Y1) * ddelx_dx);
                    atomicAdd(dL_dmean2D + (size_t)id * 3 + 1, -(b.x * Y1 + a.w * X1) * ddely_dy);
                    red_add_v4(dL_dconic2D + (size_t)id * 4, -0.5f * XX, -0.5f * XY, 0.f, -0.5f * YY);
                    atomicAdd(dL_dopacity + id, S0 * bwd_rcp(b.y));
                }
            }
            __syncwarp();   // Q is rewritten by the next block
        }
    }
}

template <int C>
static int launch_backward_geometry_cached(dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                                           const BwdGeoChannels& chn, const uint32_t* n_contrib, const float* alpha_cache,
                                           float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, cudaStream_t stream) {
    const int smem = (int)sizeof(BwdGeoCachedSmem<C>);
    R3DGW_CUDA_TRY(cudaFuncSetAttribute((blend_backward_geometry_cached_kernel<C>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    {
        ProfScope ps(KID_BLEND_BWD_GEOM, stream);
        blend_backward_geometry_cached_kernel<C><<<grid, TILE_PIX, smem, stream>>>(ranges, A, B, W, H, chn, n_contrib, alpha_cache,
                                                                                  dL_dmean2D, dL_dconic, dL_dopacity);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

// n_calls blend calls of channels[i] channels each (sum = an instantiated total); features[i] f32[P, channels[i]],
// out_images[i] / dL_dpix[i] f32[channels[i], H, W], bg[i] f32[channels[i]] - all device pointers, in host arrays
int blend_backward_geometry_cached_dispatch(int n_calls, const int* channels, dim3 grid, const uint2* ranges, const float4* A,
                                            const float4* B, int W, int H, const float* const* features,
                                            const uint32_t* n_contrib, const float* alpha_cache,
                                            const float* const* out_images, const float* const* bg,
                                            const float* const* dL_dpix, float* dL_dmean2D, float* dL_dconic,
                                            float* dL_dopacity, cudaStream_t stream) {
    BwdGeoChannels chn;
    memset(&chn, 0, sizeof(chn));
    int C = 0;
    const size_t plane = (size_t)H * W;
    for (int i = 0; i < n_calls; ++i) {
        const int c = channels[i];
        if (c <= 0 || C + c > BWDG_MAX_C || !features[i] || !out_images[i] || !bg[i] || !dL_dpix[i]) return R3DGW_ERR_BAD_ARG;
        for (int k = 0; k < c; ++k, ++C) {
            chn.feat[C] = features[i] + k;
            chn.feat_stride[C] = c;
            chn.dpix[C] = dL_dpix[i] + (size_t)k * plane;
            chn.out[C] = out_images[i] + (size_t)k * plane;
            chn.bg[C] = bg[i] + k;
        }
    }
    switch (C) {
#define R3DGW_BWDG_CASE(N) case N: return launch_backward_geometry_cached<N>(grid, ranges, A, B, W, H, chn, n_contrib, alpha_cache, dL_dmean2D, dL_dconic, dL_dopacity, stream);
        R3DGW_BWDG_CASE(3) R3DGW_BWDG_CASE(4) R3DGW_BWDG_CASE(6) R3DGW_BWDG_CASE(8) R3DGW_BWDG_CASE(9) R3DGW_BWDG_CASE(12)
        R3DGW_BWDG_CASE(15) R3DGW_BWDG_CASE(16)
#undef R3DGW_BWDG_CASE
        default: return R3DGW_ERR_UNSUPPORTED;
    }
}

int blend_forward_dispatch(int C, dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                           const float* features, const float* bg, float* final_T, uint32_t* n_contrib,
                           float* out_color, float* alpha_cache, int cache_mode, cudaStream_t stream) {
    switch (C) {
#define R3DGW_FWD_CASE(N) case N: return launch_forward<N>(grid, ranges, A, B, W, H, features, bg, final_T, n_contrib, out_color, alpha_cache, cache_mode, stream);
        R3DGW_FWD_CASE(1) R3DGW_FWD_CASE(2) R3DGW_FWD_CASE(3) R3DGW_FWD_CASE(4) R3DGW_FWD_CASE(6)
        R3DGW_FWD_CASE(8) R3DGW_FWD_CASE(9) R3DGW_FWD_CASE(12) R3DGW_FWD_CASE(15) R3DGW_FWD_CASE(16)
        R3DGW_FWD_CASE(18) R3DGW_FWD_CASE(24) R3DGW_FWD_CASE(27) R3DGW_FWD_CASE(32)
#undef R3DGW_FWD_CASE
        default: return R3DGW_ERR_UNSUPPORTED;
    }
}

template <int C>
static int launch_backward_geom(dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                                const float* bg, const float* features, const float* final_T,
                                const uint32_t* n_contrib, const float* dL_dpix, float* dL_dmean2D,
                                float* dL_dconic, float* dL_dopacity, cudaStream_t stream) {
    // measured on B200 (config B, C = 15): butterfly, BATCH 128: 1.42 ms; moments QROWS 16 / BATCH 128: 1.29 ms
    // (66 KB of smem, 3 CTAs per SM); QROWS 16 / BATCH 64: 1.18 ms (41 KB, 4 CTAs); QROWS 8 / BATCH 64: 1.19 ms.
    // With COLOR = true the C colour sums still need their butterfly and the extra smem costs occupancy: no gain
    // (C = 3: 2.49 -> 2.59 ms fwd+bwd; C = 14: 2.46 -> 2.42 ms), so those kernels keep QROWS = 0.
    if constexpr (C == 15 || C == 16) {      // occupancy experiment: R3DGW_BWDG_MINB=5 caps the kernel at 51 registers
        static const int minb = [] { const char* e = getenv("R3DGW_BWDG_MINB"); return e ? atoi(e) : 1; }();
        if (minb == 5)
            return launch_backward<C, 64, false, 16, 5>(grid, ranges, A, B, W, H, bg, features, final_T, n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, nullptr, stream);
    }
    return launch_backward<C, 64, false, 16>(grid, ranges, A, B, W, H, bg, features, final_T, n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, nullptr, stream);
}

int blend_backward_dispatch(int C, dim3 grid, const uint2* ranges, const float4* A, const float4* B, int W, int H,
                            const float* bg, const float* features, const float* final_T,
                            const uint32_t* n_contrib, const float* dL_dpix, float* dL_dmean2D,
                            float* dL_dconic, float* dL_dopacity, float* dL_dfeatures, cudaStream_t stream) {
    const bool color = dL_dfeatures != nullptr;   // NULL: geometry-only pass (6 reduced values per splat)
    switch (C) {
#define R3DGW_BWD_CASE(N, BT) case N: return color \
        ? launch_backward<N, BT, true>(grid, ranges, A, B, W, H, bg, features, final_T, n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dfeatures, stream) \
        : launch_backward_geom<N>(grid, ranges, A, B, W, H, bg, features, final_T, n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, stream);
#define BT_(direct, parked) (BWD_DIRECT ? direct : parked)
        R3DGW_BWD_CASE(1, BT_(256, 128)) R3DGW_BWD_CASE(2, BT_(256, 128)) R3DGW_BWD_CASE(3, BT_(256, 128))
        R3DGW_BWD_CASE(4, BT_(256, 128)) R3DGW_BWD_CASE(6, BT_(256, 128)) R3DGW_BWD_CASE(8, BT_(128, 64))
        R3DGW_BWD_CASE(9, BT_(128, 64)) R3DGW_BWD_CASE(12, BT_(128, 64)) R3DGW_BWD_CASE(15, BT_(128, 64))
        R3DGW_BWD_CASE(16, BT_(128, 64)) R3DGW_BWD_CASE(18, BT_(64, 32)) R3DGW_BWD_CASE(24, BT_(64, 32))
        R3DGW_BWD_CASE(27, BT_(64, 32)) R3DGW_BWD_CASE(32, BT_(64, 32))
#undef BT_
#undef R3DGW_BWD_CASE
        default: return R3DGW_ERR_UNSUPPORTED;
    }
}

}  // namespace r3dgw
