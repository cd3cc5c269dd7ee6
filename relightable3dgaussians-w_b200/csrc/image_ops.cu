// image_ops.cu - image-space post-ops of the training step for sm_100a:
//   * fused masked-L1 + 11x11 Gaussian SSIM, forward and backward
//       reference: utils/loss_utils.py:27-36 (l1_loss), :41-99 (gaussian / create_window / ssim / _ssim) - a torch
//       graph of five grouped 11x11 conv2d + ~25 elementwise kernels forward and as many backward;
//   * depth_to_normal forward and backward
//       reference: utils/graphics_utils.py:141-169 (depths_to_points, depth_to_normal) - meshgrid / matmul / cat /
//       cross / normalize torch graph;
//   * masked L1 against zero or a target (sky-BRDF regulariser, train.py:86-87).
// All three are HBM-bound stencils on [C,H,W] images: every input is read once from HBM (halo re-reads hit L2),
// the separable window runs out of shared memory, per-block partial sums are reduced in a FIXED order by a
// one-block finalize kernel (no float atomics: the loss is bit-reproducible run to run).
#include "common.cuh"

namespace r3dgw {

constexpr int IMG_T = 32;                  // output tile edge
constexpr int WIN = 11, HALO = 5;          // window_size = 11, padding = 5 (loss_utils.py:53, 70)
constexpr int IMG_IN = IMG_T + 2 * HALO;   // 42
constexpr int IMG_THREADS = 256;
constexpr int IMG_WARPS = IMG_THREADS / 32;
constexpr int IMG_RB = 4;                  // outputs per thread per separable pass (register blocking)

// gaussian(11, 1.5) / sum, exactly the float32 values torch produces for loss_utils.py:41-44
// (exp evaluated in double by Python, rounded to f32 by torch.Tensor, divided by the f32 sum)
struct Window { float w[WIN]; };
static Window make_window() {
    Window g = {{0x1.0d956cp-10f, 0x1.f1fe02p-8f, 0x1.26eb18p-5f, 0x1.bff0fep-4f, 0x1.b43c3ep-3f, 0x1.10656p-2f,
                 0x1.b43c3ep-3f, 0x1.bff0fep-4f, 0x1.26eb18p-5f, 0x1.f1fe02p-8f, 0x1.0d956cp-10f}};
    return g;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// per-block partial sums -> partials[block*3 + {0,1,2}] (fixed order inside the block)
__device__ __forceinline__ void block_partials3(float a, float b, float c, float (*red)[IMG_WARPS], float* dst) {
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    if (lane == 0) { red[0][warp] = a; red[1][warp] = b; red[2][warp] = c; }
    __syncthreads();
    if (threadIdx.x < 3) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < IMG_WARPS; ++w) s += red[threadIdx.x][w];
        dst[threadIdx.x] = s;
    }
}

// ---- SSIM + masked L1 forward ------------------------------------------------------------------------
// out per pixel (only when dmap != NULL): mask * dS/dmu1, mask * dS/dE[x^2], mask * dS/dE[xy]  (S = ssim_map)
__global__ void __launch_bounds__(IMG_THREADS) ssim_l1_forward_kernel(
    int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
    const float* __restrict__ mask, size_t mask_cstride, Window win, float* __restrict__ dmap,
    float* __restrict__ partials) {
    // the two images interleaved (x, y) and the five horizontal moments as (mu1, mu2, E[xx], E[yy]) + E[xy]: one 8-byte /
    // 16-byte shared load feeds packed FP32 pairs (FMUL2 / FADD2 / FFMA2: IEEE per lane, the same operations in the same
    // order per output as the scalar form - the kernel is issue-bound, ncu 76 % issue-active)
    __shared__ __align__(16) float2 s12[IMG_IN][IMG_IN + 1];
    __shared__ __align__(16) float4 hs4[IMG_IN][IMG_T];
    __shared__ float hs1[IMG_IN][IMG_T];
    __shared__ float red[3][IMG_WARPS];
    const int tid = threadIdx.x, c = blockIdx.z;
    const int x0 = blockIdx.x * IMG_T, y0 = blockIdx.y * IMG_T;
    const size_t plane = (size_t)H * W;
    const float* p1 = img1 + (size_t)c * plane;
    const float* p2 = img2 + (size_t)c * plane;
    for (int i = tid; i < IMG_IN * IMG_IN; i += IMG_THREADS) {
        const int r = i / IMG_IN, q = i - r * IMG_IN;
        const int y = y0 + r - HALO, x = x0 + q - HALO;
        const bool in = (x >= 0) && (x < W) && (y >= 0) && (y < H);   // zero padding (conv2d padding=5)
        s12[r][q] = make_float2(in ? p1[(size_t)y * W + x] : 0.f, in ? p2[(size_t)y * W + x] : 0.f);
    }
    __syncthreads();
    // horizontal pass, register-blocked: one item = IMG_RB consecutive outputs of one row, fed from WIN + IMG_RB - 1
    // loads instead of WIN per output; a warp takes 8 column groups x 4 rows
    for (int i = tid; i < IMG_IN * (IMG_T / IMG_RB); i += IMG_THREADS) {
        const int r = i / (IMG_T / IMG_RB), q0 = (i - r * (IMG_T / IMG_RB)) * IMG_RB;
        float2 xy[WIN + IMG_RB - 1];
#pragma unroll
        for (int t = 0; t < WIN + IMG_RB - 1; ++t) xy[t] = s12[r][q0 + t];
#pragma unroll
        for (int o = 0; o < IMG_RB; ++o) {
            float2 a01 = make_float2(0.f, 0.f), a23 = make_float2(0.f, 0.f);
            float a4 = 0.f;
#pragma unroll
            for (int k = 0; k < WIN; ++k) {
                const float2 v = xy[o + k];
                const float2 wv = __fmul2_rn(make_float2(win.w[k], win.w[k]), v);      // (w x, w y)
                a01 = __fadd2_rn(a01, wv);                                            // mu1 += w x; mu2 += w y
                a23 = __ffma2_rn(wv, v, a23);                                         // E[xx] += (w x) x; E[yy] += (w y) y
                a4 = fmaf(wv.x, v.y, a4);                                             // E[xy] += (w x) y
            }
            hs4[r][q0 + o] = make_float4(a01.x, a01.y, a23.x, a23.y);
            hs1[r][q0 + o] = a4;
        }
    }
    __syncthreads();
    // vertical pass: a thread owns column tx and the IMG_RB consecutive rows from IMG_RB * ty; the WIN + IMG_RB - 1 rows
    // it needs stream through once, each feeding every output whose window covers it (taps in ascending order per
    // output, as before)
    const int tx = tid & 31, ty = tid >> 5;
    static_assert(IMG_T / IMG_WARPS == IMG_RB, "one thread per column and IMG_RB consecutive rows");
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    float accS = 0.f, accL = 0.f, accN = 0.f;
    float mom[5][IMG_RB];
    {
        float2 m01[IMG_RB], m23[IMG_RB];
        float m4[IMG_RB];
#pragma unroll
        for (int j = 0; j < IMG_RB; ++j) { m01[j] = make_float2(0.f, 0.f); m23[j] = make_float2(0.f, 0.f); m4[j] = 0.f; }
#pragma unroll
        for (int t = 0; t < WIN + IMG_RB - 1; ++t) {
            const float4 h = hs4[IMG_RB * ty + t][tx];
            const float h4 = hs1[IMG_RB * ty + t][tx];
#pragma unroll
            for (int j = 0; j < IMG_RB; ++j) {
                const int k = t - j;
                if (k >= 0 && k < WIN) {
                    const float2 w2 = make_float2(win.w[k], win.w[k]);
                    m01[j] = __ffma2_rn(w2, make_float2(h.x, h.y), m01[j]);
                    m23[j] = __ffma2_rn(w2, make_float2(h.z, h.w), m23[j]);
                    m4[j] = fmaf(win.w[k], h4, m4[j]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < IMG_RB; ++j) {
            mom[0][j] = m01[j].x; mom[1][j] = m01[j].y; mom[2][j] = m23[j].x; mom[3][j] = m23[j].y; mom[4][j] = m4[j];
        }
    }
#pragma unroll
    for (int j = 0; j < IMG_RB; ++j) {
        const int r = IMG_RB * ty + j;
        const int y = y0 + r, x = x0 + tx;
        const float mu1 = mom[0][j], mu2 = mom[1][j], e11 = mom[2][j], e22 = mom[3][j], e12 = mom[4][j];
        if (x < W && y < H) {
            const size_t idx = (size_t)y * W + x;
            const float m = mask ? mask[(size_t)c * mask_cstride + idx] : 1.f;
            // loss_utils.py:73-87
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
            const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1_mu2;
            const float A1 = 2.f * mu1_mu2 + C1, A2 = 2.f * sigma12 + C2;
            const float B1 = mu1_sq + mu2_sq + C1, B2 = sigma1_sq + sigma2_sq + C2;
            const float S = (A1 * A2) / (B1 * B2);
            accS += S * m;
            const float2 ab = s12[r + HALO][tx + HALO];
            const float a = ab.x, b = ab.y;
            accL += fabsf(a * m - b * m);                      // loss_utils.py:34
            accN += (m == 1.f) ? 1.f : 0.f;                     // torch.sum(mask == 1)
            if (dmap) {
                const float inv = 1.f / (B1 * B2);
                const float dS_dmu1 = (2.f * mu2 * (A2 - A1) - S * 2.f * mu1 * (B2 - B1)) * inv;
                const float dS_de11 = -S / B2;
                const float dS_de12 = 2.f * A1 * inv;
                const size_t o = (size_t)c * plane + idx;
                dmap[o] = m * dS_dmu1;
                dmap[(size_t)C * plane + o] = m * dS_de11;
                dmap[2 * (size_t)C * plane + o] = m * dS_de12;
            }
        }
    }
    const size_t block = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    block_partials3(accS, accL, accN, red, partials + block * 3);
}

// one block: fixed-order reduction of the per-block partials in double
// out[0] = ssim mean (1 if count == 0, loss_utils.py:48-49), out[1] = l1 mean (0 if count == 0, :29-30),
// out[2] = count, out[3] = 1/count (0 if count == 0)
__global__ void __launch_bounds__(256) loss_finalize_kernel(const float* __restrict__ partials, int nblocks,
                                                            float* __restrict__ out) {
    __shared__ double sh[3][256];
    double a = 0, b = 0, c = 0;
    for (int i = threadIdx.x; i < nblocks; i += 256) {
        a += partials[3 * i];
        b += partials[3 * i + 1];
        c += partials[3 * i + 2];
    }
    sh[0][threadIdx.x] = a; sh[1][threadIdx.x] = b; sh[2][threadIdx.x] = c;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) {
            sh[0][threadIdx.x] += sh[0][threadIdx.x + o];
            sh[1][threadIdx.x] += sh[1][threadIdx.x + o];
            sh[2][threadIdx.x] += sh[2][threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double n = sh[2][0];
        out[0] = n > 0 ? (float)(sh[0][0] / n) : 1.f;
        out[1] = n > 0 ? (float)(sh[1][0] / n) : 0.f;
        out[2] = (float)n;
        out[3] = n > 0 ? (float)(1.0 / n) : 0.f;
    }
}

// ---- SSIM + masked L1 backward ------------------------------------------------------------------------
// dL/dimg1 = kS * (conv(D0) + 2*img1*conv(D1) + img2*conv(D2)) + kL * sign(img1*m - img2*m) * m
// with kS = coef[0]/count, kL = coef[1]/count (coef = upstream gradients of the two means, device scalars)
__global__ void __launch_bounds__(IMG_THREADS) ssim_l1_backward_kernel(
    int C, int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
    const float* __restrict__ mask, size_t mask_cstride, Window win, const float* __restrict__ dmap,
    const float* __restrict__ sums, const float* __restrict__ coef, float* __restrict__ dL_dimg1) {
    __shared__ float sd[3][IMG_IN][IMG_IN + 1];
    __shared__ __align__(16) float hs[3][IMG_IN][IMG_T];
    const int tid = threadIdx.x, c = blockIdx.z;
    const int x0 = blockIdx.x * IMG_T, y0 = blockIdx.y * IMG_T;
    const size_t plane = (size_t)H * W;
    for (int i = tid; i < IMG_IN * IMG_IN; i += IMG_THREADS) {
        const int r = i / IMG_IN, q = i - r * IMG_IN;
        const int y = y0 + r - HALO, x = x0 + q - HALO;
        const bool in = (x >= 0) && (x < W) && (y >= 0) && (y < H);
        const size_t o = (size_t)c * plane + (size_t)(in ? y : 0) * W + (in ? x : 0);
#pragma unroll
        for (int t = 0; t < 3; ++t) sd[t][r][q] = in ? dmap[(size_t)t * C * plane + o] : 0.f;
    }
    __syncthreads();
    // both passes register-blocked like the forward kernel's (IMG_RB outputs from WIN + IMG_RB - 1 loads)
    for (int i = tid; i < IMG_IN * (IMG_T / IMG_RB); i += IMG_THREADS) {
        const int r = i / (IMG_T / IMG_RB), q0 = (i - r * (IMG_T / IMG_RB)) * IMG_RB;
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            float v[WIN + IMG_RB - 1];
#pragma unroll
            for (int u = 0; u < WIN + IMG_RB - 1; ++u) v[u] = sd[t][r][q0 + u];
            float o[IMG_RB];
#pragma unroll
            for (int j = 0; j < IMG_RB; ++j) {
                float a = 0.f;
#pragma unroll
                for (int k = 0; k < WIN; ++k) a = fmaf(win.w[k], v[j + k], a);
                o[j] = a;
            }
            *reinterpret_cast<float4*>(&hs[t][r][q0]) = make_float4(o[0], o[1], o[2], o[3]);
        }
    }
    __syncthreads();
    const int tx = tid & 31, ty = tid >> 5;
    const float inv_n = sums[3];
    const float kS = coef[0] * inv_n, kL = coef[1] * inv_n;
    float cv[3][IMG_RB];
#pragma unroll
    for (int t = 0; t < 3; ++t) {
        float col[WIN + IMG_RB - 1];
#pragma unroll
        for (int u = 0; u < WIN + IMG_RB - 1; ++u) col[u] = hs[t][IMG_RB * ty + u][tx];
#pragma unroll
        for (int j = 0; j < IMG_RB; ++j) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < WIN; ++k) a = fmaf(win.w[k], col[j + k], a);
            cv[t][j] = a;
        }
    }
#pragma unroll
    for (int j = 0; j < IMG_RB; ++j) {
        const int r = IMG_RB * ty + j;
        const int y = y0 + r, x = x0 + tx;
        const float c0 = cv[0][j], c1 = cv[1][j], c2 = cv[2][j];
        if (x < W && y < H) {
            const size_t idx = (size_t)y * W + x;
            const size_t o = (size_t)c * plane + idx;
            const float a = img1[o], b = img2[o];
            const float m = mask ? mask[(size_t)c * mask_cstride + idx] : 1.f;
            const float d = a * m - b * m;
            const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
            dL_dimg1[o] = kS * (c0 + 2.f * a * c1 + b * c2) + kL * sgn * m;
        }
    }
}

// ---- masked L1 (no SSIM): sum |x*m - t*m| and count(m == 1); target may be NULL (zeros) ---------------
// complement != 0 uses m = 1 - mask (train.py:86: mask = 1 - sky_mask)
// grid = (blocks per plane, C): no per-element division; four pixels per thread per step through 16-byte loads when
// the plane is a multiple of 4 and the pointers are aligned (VEC), scalar otherwise.
template <bool VEC>
__global__ void __launch_bounds__(IMG_THREADS) masked_l1_forward_kernel(
    size_t plane, const float* __restrict__ x, const float* __restrict__ target,
    const float* __restrict__ mask, size_t mask_cstride, int complement, float* __restrict__ partials) {
    __shared__ float red[3][IMG_WARPS];
    const size_t c = blockIdx.y;
    const float* xc = x + c * plane;
    const float* tc = target ? target + c * plane : nullptr;
    const float* mc = mask ? mask + c * mask_cstride : nullptr;
    float accL = 0.f, accN = 0.f;
    auto one = [&](float xv, float tv, float m) {
        if (complement) m = 1.f - m;
        accL += fabsf(xv * m - tv * m);
        accN += (m == 1.f) ? 1.f : 0.f;
    };
    if (VEC) {
        const size_t n4 = plane / 4;
        for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < n4; i += (size_t)gridDim.x * IMG_THREADS) {
            const float4 xv = reinterpret_cast<const float4*>(xc)[i];
            const float4 tv = tc ? reinterpret_cast<const float4*>(tc)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 mv = mc ? reinterpret_cast<const float4*>(mc)[i] : make_float4(1.f, 1.f, 1.f, 1.f);
            one(xv.x, tv.x, mv.x); one(xv.y, tv.y, mv.y); one(xv.z, tv.z, mv.z); one(xv.w, tv.w, mv.w);
        }
    } else {
        for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < plane; i += (size_t)gridDim.x * IMG_THREADS)
            one(xc[i], tc ? tc[i] : 0.f, mc ? mc[i] : 1.f);
    }
    block_partials3(0.f, accL, accN, red, partials + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 3);
}

template <bool VEC>
__global__ void __launch_bounds__(IMG_THREADS) masked_l1_backward_kernel(
    size_t plane, const float* __restrict__ x, const float* __restrict__ target,
    const float* __restrict__ mask, size_t mask_cstride, int complement, const float* __restrict__ sums,
    const float* __restrict__ coef, float* __restrict__ dL_dx) {
    const float k = coef[0] * sums[3];
    const size_t c = blockIdx.y;
    const float* xc = x + c * plane;
    const float* tc = target ? target + c * plane : nullptr;
    const float* mc = mask ? mask + c * mask_cstride : nullptr;
    float* gc = dL_dx + c * plane;
    auto one = [&](float xv, float tv, float m) {
        if (complement) m = 1.f - m;
        const float d = xv * m - tv * m;
        const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);
        return k * sgn * m;
    };
    if (VEC) {
        const size_t n4 = plane / 4;
        for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < n4; i += (size_t)gridDim.x * IMG_THREADS) {
            const float4 xv = reinterpret_cast<const float4*>(xc)[i];
            const float4 tv = tc ? reinterpret_cast<const float4*>(tc)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 mv = mc ? reinterpret_cast<const float4*>(mc)[i] : make_float4(1.f, 1.f, 1.f, 1.f);
            reinterpret_cast<float4*>(gc)[i] = make_float4(one(xv.x, tv.x, mv.x), one(xv.y, tv.y, mv.y),
                                                           one(xv.z, tv.z, mv.z), one(xv.w, tv.w, mv.w));
        }
    } else {
        for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < plane; i += (size_t)gridDim.x * IMG_THREADS)
            gc[i] = one(xc[i], tc ? tc[i] : 0.f, mc ? mc[i] : 1.f);
    }
}

// ---- depth_to_normal -------------------------------------------------------------------------------------
// rays: device f32[21] = A (3x3 row-major, intrins.inverse().T), B (3x3 row-major, c2w[:3,:3].T), rays_o[3]:
// rays_d = ([x, y, 1] @ A) @ B, points = depth * rays_d + rays_o (graphics_utils.py:146-155), evaluated in the
// reference's two-step order so that the rounding of the points stays close to the torch graph's.
constexpr int RAYS_N = 21;

__device__ __forceinline__ float3 ray_dir(const float* M, int x, int y) {
    const float fx = (float)x, fy = (float)y;
    const float c0 = fmaf(fx, M[0], fmaf(fy, M[3], M[6]));
    const float c1 = fmaf(fx, M[1], fmaf(fy, M[4], M[7]));
    const float c2 = fmaf(fx, M[2], fmaf(fy, M[5], M[8]));
    const float* B = M + 9;
    return make_float3(fmaf(c0, B[0], fmaf(c1, B[3], c2 * B[6])), fmaf(c0, B[1], fmaf(c1, B[4], c2 * B[7])),
                       fmaf(c0, B[2], fmaf(c1, B[5], c2 * B[8])));
}
__device__ __forceinline__ float3 cross3(const float3 a, const float3 b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float dot3(const float3 a, const float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 world_point(const float* __restrict__ depth, const float* M, int W, int x, int y) {
    const float d = depth[(size_t)y * W + x];
    const float3 r = ray_dir(M, x, y);
    const float* o = M + 18;
    return make_float3(__fadd_rn(__fmul_rn(d, r.x), o[0]), __fadd_rn(__fmul_rn(d, r.y), o[1]),
                       __fadd_rn(__fmul_rn(d, r.z), o[2]));
}

// dx / dy of graphics_utils.py:164-165 at an interior pixel: dx is the ROW difference, dy the COLUMN difference
__device__ __forceinline__ void normal_diffs(const float* __restrict__ depth, const float* M, int W, int x, int y,
                                             float3& dx, float3& dy) {
    const float3 pu = world_point(depth, M, W, x, y + 1), pd = world_point(depth, M, W, x, y - 1);
    const float3 pr = world_point(depth, M, W, x + 1, y), pl = world_point(depth, M, W, x - 1, y);
    dx = make_float3(pu.x - pd.x, pu.y - pd.y, pu.z - pd.z);
    dy = make_float3(pr.x - pl.x, pr.y - pl.y, pr.z - pl.z);
}

__global__ void __launch_bounds__(256) depth_to_normal_forward_kernel(int H, int W, const float* __restrict__ depth,
                                                                      const float* __restrict__ rays,
                                                                      float* __restrict__ normal /* [H,W,3] */) {
    __shared__ float M[RAYS_N];
    if (threadIdx.x < RAYS_N) M[threadIdx.x] = rays[threadIdx.x];
    __syncthreads();
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    float3 n = make_float3(0.f, 0.f, 0.f);
    if (x >= 1 && x < W - 1 && y >= 1 && y < H - 1) {
        float3 dx, dy;
        normal_diffs(depth, M, W, x, y, dx, dy);
        const float3 v = cross3(dx, dy);
        const float len = fmaxf(sqrtf(dot3(v, v)), 1e-12f);   // F.normalize eps
        n = make_float3(v.x / len, v.y / len, v.z / len);
    }
    float* o = normal + ((size_t)y * W + x) * 3;
    o[0] = n.x; o[1] = n.y; o[2] = n.z;
}

// gradient of the normal at interior pixel (cx, cy) w.r.t. its dx / dy vectors, given g = dL/dn there
__device__ __forceinline__ void normal_backprop(const float* __restrict__ depth, const float* M,
                                                const float* __restrict__ gn, int W, int cx, int cy, float3& g_dx,
                                                float3& g_dy) {
    float3 dx, dy;
    normal_diffs(depth, M, W, cx, cy, dx, dy);
    const float3 v = cross3(dx, dy);
    const float len = sqrtf(dot3(v, v));
    const float* gp = gn + ((size_t)cy * W + cx) * 3;
    const float3 g = make_float3(gp[0], gp[1], gp[2]);
    float3 u;
    if (len > 1e-12f) {
        const float inv = 1.f / len;
        const float3 n = make_float3(v.x * inv, v.y * inv, v.z * inv);
        const float ng = dot3(n, g);
        u = make_float3((g.x - n.x * ng) * inv, (g.y - n.y * ng) * inv, (g.z - n.z * ng) * inv);
    } else {
        u = make_float3(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f);   // clamp_min(eps) branch: v / eps
    }
    g_dx = cross3(dy, u);   // d(v.u)/d(dx), v = dx x dy
    g_dy = cross3(u, dx);
}

__global__ void __launch_bounds__(256) depth_to_normal_backward_kernel(int H, int W, const float* __restrict__ depth,
                                                                       const float* __restrict__ rays,
                                                                       const float* __restrict__ dL_dnormal,
                                                                       float* __restrict__ dL_ddepth) {
    __shared__ float M[RAYS_N];
    if (threadIdx.x < RAYS_N) M[threadIdx.x] = rays[threadIdx.x];
    __syncthreads();
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const float3 r = ray_dir(M, x, y);
    float acc = 0.f;
    float3 gdx, gdy;
    auto interior = [&](int cx, int cy) { return cx >= 1 && cx < W - 1 && cy >= 1 && cy < H - 1; };
    if (interior(x, y - 1)) { normal_backprop(depth, M, dL_dnormal, W, x, y - 1, gdx, gdy); acc += dot3(r, gdx); }
    if (interior(x, y + 1)) { normal_backprop(depth, M, dL_dnormal, W, x, y + 1, gdx, gdy); acc -= dot3(r, gdx); }
    if (interior(x - 1, y)) { normal_backprop(depth, M, dL_dnormal, W, x - 1, y, gdx, gdy); acc += dot3(r, gdy); }
    if (interior(x + 1, y)) { normal_backprop(depth, M, dL_dnormal, W, x + 1, y, gdx, gdy); acc -= dot3(r, gdy); }
    dL_ddepth[(size_t)y * W + x] = acc;
}

// ---- image-space tail of render() + the normal-consistency term, fused ----------------------------------
// gaussian_renderer/__init__.py:231-237, 266-271 on the rasterized normal / depth / alpha images:
//     normal     = ((n01 - 0.5) * 2) * sky + 1 * (1 - sky)
//     normal_ref = depth_to_normal(view, depth[0] * sky).permute(2,0,1) * alpha.detach() + 1 * (1 - sky)
// one thread per pixel, same expression order as the torch ops (the depth fed to depth_to_normal is the MASKED depth of
// every neighbour); a dozen elementwise launches + the permute each way become one kernel each way.
__device__ __forceinline__ float3 world_point_m(const float* __restrict__ depth, const float* __restrict__ sky, const float* M,
                                                int W, int x, int y) {
    const size_t i = (size_t)y * W + x;
    const float d = __fmul_rn(depth[i], sky[i]);
    const float3 r = ray_dir(M, x, y);
    const float* o = M + 18;
    return make_float3(__fadd_rn(__fmul_rn(d, r.x), o[0]), __fadd_rn(__fmul_rn(d, r.y), o[1]),
                       __fadd_rn(__fmul_rn(d, r.z), o[2]));
}
__device__ __forceinline__ void normal_diffs_m(const float* __restrict__ depth, const float* __restrict__ sky, const float* M,
                                               int W, int x, int y, float3& dx, float3& dy) {
    const float3 pu = world_point_m(depth, sky, M, W, x, y + 1), pd = world_point_m(depth, sky, M, W, x, y - 1);
    const float3 pr = world_point_m(depth, sky, M, W, x + 1, y), pl = world_point_m(depth, sky, M, W, x - 1, y);
    dx = make_float3(pu.x - pd.x, pu.y - pd.y, pu.z - pd.z);
    dy = make_float3(pr.x - pl.x, pr.y - pl.y, pr.z - pl.z);
}

__global__ void __launch_bounds__(256) normal_tail_forward_kernel(
    int H, int W, const float* __restrict__ n01, const float* __restrict__ depth, const float* __restrict__ alpha,
    size_t alpha_cstride, const float* __restrict__ sky, const float* __restrict__ rays, float* __restrict__ normal,
    float* __restrict__ normal_ref) {
    __shared__ float M[RAYS_N];
    if (threadIdx.x < RAYS_N) M[threadIdx.x] = rays[threadIdx.x];
    __syncthreads();
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const size_t plane = (size_t)H * W, i = (size_t)y * W + x;
    const float m = sky[i], one_m = __fsub_rn(1.f, m);
    float3 n = make_float3(0.f, 0.f, 0.f);
    if (x >= 1 && x < W - 1 && y >= 1 && y < H - 1) {
        float3 dx, dy;
        normal_diffs_m(depth, sky, M, W, x, y, dx, dy);
        const float3 v = cross3(dx, dy);
        const float len = fmaxf(sqrtf(dot3(v, v)), 1e-12f);
        n = make_float3(v.x / len, v.y / len, v.z / len);
    }
    const float nr[3] = {n.x, n.y, n.z};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v = __fmul_rn(__fsub_rn(n01[c * plane + i], 0.5f), 2.0f);
        normal[c * plane + i] = __fadd_rn(__fmul_rn(v, m), one_m);
        normal_ref[c * plane + i] = __fadd_rn(__fmul_rn(nr[c], alpha[c * alpha_cstride + i]), one_m);
    }
}

// gradient of the (alpha-weighted) normal at interior pixel (cx, cy) w.r.t. its dx / dy vectors; g planar [3,H,W]
__device__ __forceinline__ void normal_backprop_m(const float* __restrict__ depth, const float* __restrict__ sky,
                                                  const float* M, const float* __restrict__ g_ref,
                                                  const float* __restrict__ alpha, size_t acs, int W, size_t plane,
                                                  int cx, int cy, float3& g_dx, float3& g_dy) {
    float3 dx, dy;
    normal_diffs_m(depth, sky, M, W, cx, cy, dx, dy);
    const float3 v = cross3(dx, dy);
    const float len = sqrtf(dot3(v, v));
    const size_t j = (size_t)cy * W + cx;
    const float3 g = make_float3(g_ref[j] * alpha[j], g_ref[plane + j] * alpha[acs + j], g_ref[2 * plane + j] * alpha[2 * acs + j]);
    float3 u;
    if (len > 1e-12f) {
        const float inv = 1.f / len;
        const float3 n = make_float3(v.x * inv, v.y * inv, v.z * inv);
        const float ng = dot3(n, g);
        u = make_float3((g.x - n.x * ng) * inv, (g.y - n.y * ng) * inv, (g.z - n.z * ng) * inv);
    } else {
        u = make_float3(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f);
    }
    g_dx = cross3(dy, u);
    g_dy = cross3(u, dx);
}

__global__ void __launch_bounds__(256) normal_tail_backward_kernel(
    int H, int W, const float* __restrict__ depth, const float* __restrict__ alpha, size_t acs,
    const float* __restrict__ sky, const float* __restrict__ rays, const float* __restrict__ g_normal,
    const float* __restrict__ g_ref, float* __restrict__ g_n01, float* __restrict__ g_depth) {
    __shared__ float M[RAYS_N];
    if (threadIdx.x < RAYS_N) M[threadIdx.x] = rays[threadIdx.x];
    __syncthreads();
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const size_t plane = (size_t)H * W, i = (size_t)y * W + x;
    const float m = sky[i];
    if (g_n01) {
#pragma unroll
        for (int c = 0; c < 3; ++c) g_n01[c * plane + i] = g_normal ? (g_normal[c * plane + i] * m) * 2.0f : 0.f;
    }
    if (g_depth) {
        float acc = 0.f;
        if (g_ref) {
            const float3 r = ray_dir(M, x, y);
            float3 gdx, gdy;
            auto interior = [&](int cx, int cy) { return cx >= 1 && cx < W - 1 && cy >= 1 && cy < H - 1; };
            if (interior(x, y - 1)) { normal_backprop_m(depth, sky, M, g_ref, alpha, acs, W, plane, x, y - 1, gdx, gdy); acc += dot3(r, gdx); }
            if (interior(x, y + 1)) { normal_backprop_m(depth, sky, M, g_ref, alpha, acs, W, plane, x, y + 1, gdx, gdy); acc -= dot3(r, gdx); }
            if (interior(x - 1, y)) { normal_backprop_m(depth, sky, M, g_ref, alpha, acs, W, plane, x - 1, y, gdx, gdy); acc += dot3(r, gdy); }
            if (interior(x + 1, y)) { normal_backprop_m(depth, sky, M, g_ref, alpha, acs, W, plane, x + 1, y, gdx, gdy); acc -= dot3(r, gdy); }
        }
        g_depth[i] = acc * m;          // depth enters as depth * sky
    }
}

// normal-consistency term of train.py:92-96: mean over pixels of 1 - sum_c (normal_c w)(normal_ref_c w), w = occl * sky
__global__ void __launch_bounds__(IMG_THREADS) normal_loss_forward_kernel(
    size_t plane, const float* __restrict__ normal, const float* __restrict__ normal_ref,
    const float* __restrict__ occl, const float* __restrict__ sky, float* __restrict__ partials) {
    __shared__ float red[3][IMG_WARPS];
    float acc = 0.f, cnt = 0.f;
    for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < plane; i += (size_t)gridDim.x * IMG_THREADS) {
        const float o = occl ? occl[i] : 1.f, m = sky ? sky[i] : 1.f;
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float rn = (normal[c * plane + i] * o) * m, rs = (normal_ref[c * plane + i] * o) * m;
            d += rn * rs;
        }
        acc += 1.f - d;
        cnt += 1.f;
    }
    block_partials3(acc, 0.f, cnt, red, partials + (size_t)blockIdx.x * 3);
}

__global__ void __launch_bounds__(IMG_THREADS) normal_loss_backward_kernel(
    size_t plane, const float* __restrict__ normal, const float* __restrict__ normal_ref,
    const float* __restrict__ occl, const float* __restrict__ sky, const float* __restrict__ coef,
    float* __restrict__ g_normal, float* __restrict__ g_ref) {
    const float k = -coef[0] / (float)plane;
    for (size_t i = (size_t)blockIdx.x * IMG_THREADS + threadIdx.x; i < plane; i += (size_t)gridDim.x * IMG_THREADS) {
        const float o = occl ? occl[i] : 1.f, m = sky ? sky[i] : 1.f;
        const float w2 = k * (o * m) * (o * m);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float a = normal[c * plane + i], b = normal_ref[c * plane + i];
            if (g_normal) g_normal[c * plane + i] = w2 * b;
            if (g_ref) g_ref[c * plane + i] = w2 * a;
        }
    }
}

// blocks per channel plane: 16 pixels per thread, at most NUM_SMS * 8 blocks over all channels (the workspace bound)
static inline int l1_blocks(size_t plane, int C) {
    size_t b = (plane + IMG_THREADS * 16 - 1) / (IMG_THREADS * 16);
    const size_t cap = (size_t)NUM_SMS * 8 / (size_t)C;
    return (int)(b < 1 ? 1 : (b > cap ? (cap < 1 ? 1 : cap) : b));
}

static inline bool l1_vec_ok(size_t plane, size_t mask_cstride, const void* x, const void* t, const void* m, const void* g) {
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return plane % 4 == 0 && mask_cstride % 4 == 0 && al(x) && al(t) && al(m) && al(g);
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" {

size_t r3dgw_image_loss_workspace_bytes(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return 256;
    const size_t blocks = (size_t)((W + IMG_T - 1) / IMG_T) * ((H + IMG_T - 1) / IMG_T) * C;
    const size_t l1b = (size_t)NUM_SMS * 8;
    return align_up((blocks > l1b ? blocks : l1b) * 3 * sizeof(float), 256);
}

int r3dgw_ssim_l1_forward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                          size_t mask_channel_stride, float* dmap, char* workspace, float* out, void* stream_) {
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !workspace || !out) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + IMG_T - 1) / IMG_T, (H + IMG_T - 1) / IMG_T, C);
    float* partials = reinterpret_cast<float*>(workspace);
    {
        ProfScope ps(KID_SSIM_FWD, stream);
        ssim_l1_forward_kernel<<<grid, IMG_THREADS, 0, stream>>>(C, H, W, img1, img2, mask, mask_channel_stride,
                                                                 make_window(), dmap, partials);
    }
    R3DGW_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, stream>>>(partials, (int)(grid.x * grid.y * grid.z), out);
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_ssim_l1_backward(int C, int H, int W, const float* img1, const float* img2, const float* mask,
                           size_t mask_channel_stride, const float* dmap, const float* sums, const float* coef,
                           float* dL_dimg1, void* stream_) {
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !dmap || !sums || !coef || !dL_dimg1) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + IMG_T - 1) / IMG_T, (H + IMG_T - 1) / IMG_T, C);
    {
        ProfScope ps(KID_SSIM_BWD, stream);
        ssim_l1_backward_kernel<<<grid, IMG_THREADS, 0, stream>>>(C, H, W, img1, img2, mask, mask_channel_stride,
                                                                  make_window(), dmap, sums, coef, dL_dimg1);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_masked_l1_forward(int C, int H, int W, const float* x, const float* target, const float* mask,
                            size_t mask_channel_stride, int complement, char* workspace, float* out, void* stream_) {
    if (C <= 0 || H <= 0 || W <= 0 || !x || !workspace || !out) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t plane = (size_t)H * W;
    if (C > NUM_SMS * 8) return R3DGW_ERR_UNSUPPORTED;
    const dim3 grid(l1_blocks(plane, C), C);
    float* partials = reinterpret_cast<float*>(workspace);
    {
        ProfScope ps(KID_MASKED_L1, stream);
        if (l1_vec_ok(plane, mask_channel_stride, x, target, mask, nullptr))
            masked_l1_forward_kernel<true><<<grid, IMG_THREADS, 0, stream>>>(plane, x, target, mask, mask_channel_stride,
                                                                             complement, partials);
        else
            masked_l1_forward_kernel<false><<<grid, IMG_THREADS, 0, stream>>>(plane, x, target, mask, mask_channel_stride,
                                                                              complement, partials);
    }
    R3DGW_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, stream>>>(partials, (int)(grid.x * grid.y), out);
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_masked_l1_backward(int C, int H, int W, const float* x, const float* target, const float* mask,
                             size_t mask_channel_stride, int complement, const float* sums, const float* coef,
                             float* dL_dx, void* stream_) {
    if (C <= 0 || H <= 0 || W <= 0 || !x || !sums || !coef || !dL_dx) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t plane = (size_t)H * W;
    if (C > NUM_SMS * 8) return R3DGW_ERR_UNSUPPORTED;
    const dim3 grid(l1_blocks(plane, C), C);
    {
        ProfScope ps(KID_MASKED_L1, stream);
        if (l1_vec_ok(plane, mask_channel_stride, x, target, mask, dL_dx))
            masked_l1_backward_kernel<true><<<grid, IMG_THREADS, 0, stream>>>(plane, x, target, mask, mask_channel_stride,
                                                                              complement, sums, coef, dL_dx);
        else
            masked_l1_backward_kernel<false><<<grid, IMG_THREADS, 0, stream>>>(plane, x, target, mask, mask_channel_stride,
                                                                               complement, sums, coef, dL_dx);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_normal_tail_forward(int H, int W, const float* n01, const float* depth, const float* alpha,
                              size_t alpha_channel_stride, const float* sky, const float* rays, float* normal,
                              float* normal_ref, void* stream_) {
    if (H <= 0 || W <= 0 || !n01 || !depth || !alpha || !sky || !rays || !normal || !normal_ref) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + 31) / 32, (H + 7) / 8);
    {
        ProfScope ps(KID_DEPTH_NORMAL, stream);
        normal_tail_forward_kernel<<<grid, 256, 0, stream>>>(H, W, n01, depth, alpha, alpha_channel_stride, sky, rays, normal,
                                                             normal_ref);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_normal_tail_backward(int H, int W, const float* depth, const float* alpha, size_t alpha_channel_stride,
                               const float* sky, const float* rays, const float* dL_dnormal, const float* dL_dnormal_ref,
                               float* dL_dn01, float* dL_ddepth, void* stream_) {
    if (H <= 0 || W <= 0 || !depth || !alpha || !sky || !rays) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + 31) / 32, (H + 7) / 8);
    {
        ProfScope ps(KID_DEPTH_NORMAL, stream);
        normal_tail_backward_kernel<<<grid, 256, 0, stream>>>(H, W, depth, alpha, alpha_channel_stride, sky, rays, dL_dnormal,
                                                              dL_dnormal_ref, dL_dn01, dL_ddepth);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_normal_loss_forward(int H, int W, const float* normal, const float* normal_ref, const float* occl,
                              const float* sky, char* workspace, float* out, void* stream_) {
    if (H <= 0 || W <= 0 || !normal || !normal_ref || !workspace || !out) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t plane = (size_t)H * W;
    const int blocks = l1_blocks(plane, 1);
    float* partials = reinterpret_cast<float*>(workspace);
    {
        ProfScope ps(KID_MASKED_L1, stream);
        normal_loss_forward_kernel<<<blocks, IMG_THREADS, 0, stream>>>(plane, normal, normal_ref, occl, sky, partials);
    }
    R3DGW_LAUNCH_CHECK();
    loss_finalize_kernel<<<1, 256, 0, stream>>>(partials, blocks, out);       // out[0] = mean over the H*W pixels
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_normal_loss_backward(int H, int W, const float* normal, const float* normal_ref, const float* occl,
                               const float* sky, const float* coef, float* dL_dnormal, float* dL_dnormal_ref,
                               void* stream_) {
    if (H <= 0 || W <= 0 || !normal || !normal_ref || !coef) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t plane = (size_t)H * W;
    {
        ProfScope ps(KID_MASKED_L1, stream);
        normal_loss_backward_kernel<<<l1_blocks(plane, 1), IMG_THREADS, 0, stream>>>(plane, normal, normal_ref, occl, sky,
                                                                                      coef, dL_dnormal, dL_dnormal_ref);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_depth_to_normal_forward(int H, int W, const float* depth, const float* rays, float* normal, void* stream_) {
    if (H <= 0 || W <= 0 || !depth || !rays || !normal) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + 31) / 32, (H + 7) / 8);
    {
        ProfScope ps(KID_DEPTH_NORMAL, stream);
        depth_to_normal_forward_kernel<<<grid, 256, 0, stream>>>(H, W, depth, rays, normal);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_depth_to_normal_backward(int H, int W, const float* depth, const float* rays, const float* dL_dnormal,
                                   float* dL_ddepth, void* stream_) {
    if (H <= 0 || W <= 0 || !depth || !rays || !dL_dnormal || !dL_ddepth) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    dim3 grid((W + 31) / 32, (H + 7) / 8);
    {
        ProfScope ps(KID_DEPTH_NORMAL, stream);
        depth_to_normal_backward_kernel<<<grid, 256, 0, stream>>>(H, W, depth, rays, dL_dnormal, dL_ddepth);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
