// gauss_reg.cu - the per-Gaussian regularisers of the training step as ONE masked fixed-order reduction.
//
// Replaces two torch graphs of the reference (boolean-mask gathers, a [P_vis,3] torch.sort, two host syncs):
//   min_scale_loss        utils/loss_utils.py:210-220   mean over visible foreground Gaussians of their smallest scale
//   depth_loss_gaussians  utils/loss_utils.py:140-148   exp(-gamma (mean depth of visible sky Gaussians -
//                                                       mean depth of visible foreground Gaussians [detached]))
//   (depth = GaussianModel.get_depth, scene/gaussian_model.py:125-130: z of the view-space position)
// visible = radii > 0 (train.py:77 visibility_filter).  One streaming pass over radii / is_sky / scales / means3D
// (29 B per Gaussian); per-CTA partial sums are reduced in a fixed order by a one-block finalize kernel, so the two
// losses are bit-reproducible and nothing returns to the host.
#include "common.cuh"
#include <algorithm>
#include <cstdio>

namespace r3dgw {

constexpr int REG_THREADS = 256;
constexpr int REG_MAX_BLOCKS = NUM_SMS * 4;
constexpr int REG_VALS = 6;       // sum min-scale (fg), count fg, sum depth sky, count sky, sum depth fg, count visible

__device__ __forceinline__ int min_scale_axis(float s0, float s1, float s2) {
    return (s1 < s0) ? ((s2 < s1) ? 2 : 1) : ((s2 < s0) ? 2 : 0);
}

__global__ void __launch_bounds__(REG_THREADS)
gaussian_reg_partial_kernel(int P, const int* __restrict__ radii, const unsigned char* __restrict__ is_sky,
                            const float* __restrict__ scales, const float* __restrict__ means3D,
                            const float* __restrict__ vm, double* __restrict__ partial) {
    float acc[REG_VALS] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const float m2 = vm[2], m6 = vm[6], m10 = vm[10], m14 = vm[14];
    for (int i = blockIdx.x * REG_THREADS + threadIdx.x; i < P; i += gridDim.x * REG_THREADS) {
        if (radii[i] <= 0) continue;
        acc[5] += 1.f;
        const float depth = m2 * means3D[3 * i] + m6 * means3D[3 * i + 1] + m10 * means3D[3 * i + 2] + m14;
        if (is_sky && is_sky[i]) {
            acc[2] += depth; acc[3] += 1.f;
        } else {
            acc[0] += fminf(fminf(scales[3 * i], scales[3 * i + 1]), scales[3 * i + 2]);
            acc[1] += 1.f;
            acc[4] += depth;
        }
    }
    __shared__ double red[REG_THREADS / 32][REG_VALS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < REG_VALS; ++k) {
        double a = (double)acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) red[warp][k] = a;
    }
    __syncthreads();
    if (threadIdx.x < REG_VALS) {
        double a = 0.0;
#pragma unroll
        for (int w = 0; w < REG_THREADS / 32; ++w) a += red[w][threadIdx.x];
        partial[(size_t)blockIdx.x * REG_VALS + threadIdx.x] = a;
    }
}

// out[8] = { scale_loss, depth_loss, count fg, count sky, mean depth sky, mean depth fg, count visible, 0 }
__global__ void gaussian_reg_finalize_kernel(int blocks, const double* __restrict__ partial, float gamma,
                                             float* __restrict__ out) {
    // warp k sums value k: lane l takes rows l, l + 32, ... (independent loads in flight instead of one dependent
    // chain of `blocks` L2 round trips: 39 -> ~5 us), then a fixed-order butterfly - still bit-reproducible
    __shared__ double tot[REG_VALS];
    const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k < REG_VALS) {
        double a = 0.0;
        for (int b = lane; b < blocks; b += 32) a += partial[(size_t)b * REG_VALS + k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) tot[k] = a;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // torch.mean of an empty selection is NaN (0 / 0); no visible Gaussian at all: the reference returns None
        // (and train.py then fails) - reported as 0 here
        const float scale_loss = tot[5] > 0.0 ? (float)(tot[0] / tot[1]) : 0.f;
        const float avg_sky = (float)(tot[2] / tot[3]);
        const float avg_fg = (float)(tot[4] / tot[1]);
        out[0] = scale_loss;
        out[1] = expf(-gamma * (avg_sky - avg_fg));
        out[2] = (float)tot[1]; out[3] = (float)tot[3];
        out[4] = avg_sky; out[5] = avg_fg; out[6] = (float)tot[5]; out[7] = 0.f;
    }
}

__global__ void __launch_bounds__(REG_THREADS)
gaussian_reg_backward_kernel(int P, const int* __restrict__ radii, const unsigned char* __restrict__ is_sky,
                             const float* __restrict__ scales, const float* __restrict__ vm, float gamma,
                             const float* __restrict__ out, const float* __restrict__ g,
                             float* __restrict__ d_scales, float* __restrict__ d_means3D) {
    const int i = blockIdx.x * REG_THREADS + threadIdx.x;
    if (i >= P) return;
    float ds[3] = {0.f, 0.f, 0.f}, dm[3] = {0.f, 0.f, 0.f};
    if (radii[i] > 0) {
        if (is_sky && is_sky[i]) {
            // d exp(-gamma (avg_sky - avg_fg)) / d depth_i = -gamma loss / count_sky  (avg_fg is detached)
            const float c = g[1] * (-gamma) * out[1] / out[3];
            dm[0] = c * vm[2]; dm[1] = c * vm[6]; dm[2] = c * vm[10];
        } else {
            ds[min_scale_axis(scales[3 * i], scales[3 * i + 1], scales[3 * i + 2])] = g[0] / out[2];
        }
    }
    if (d_scales) { d_scales[3 * i] = ds[0]; d_scales[3 * i + 1] = ds[1]; d_scales[3 * i + 2] = ds[2]; }
    if (d_means3D) { d_means3D[3 * i] = dm[0]; d_means3D[3 * i + 1] = dm[1]; d_means3D[3 * i + 2] = dm[2]; }
}

static inline int reg_blocks(int P) { return std::max(1, std::min((P + REG_THREADS - 1) / REG_THREADS, REG_MAX_BLOCKS)); }

}  // namespace r3dgw

using namespace r3dgw;

extern "C" size_t r3dgw_gaussian_reg_workspace_bytes(int P) {
    return align_up((size_t)reg_blocks(P > 0 ? P : 1) * REG_VALS * sizeof(double), 256);
}

extern "C" int r3dgw_gaussian_regularisers_forward(int P, const int* radii, const uint8_t* is_sky, const float* scales,
                                                   const float* means3D, const float* viewmatrix, float gamma,
                                                   char* workspace, float* out, void* stream_) {
    if (P < 0 || !workspace || !out || !viewmatrix) return R3DGW_ERR_BAD_ARG;
    if (P > 0 && (!radii || !scales || !means3D)) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int blocks = reg_blocks(P);
    {
        ProfScope ps(KID_GAUSS_REG, stream);
        gaussian_reg_partial_kernel<<<blocks, REG_THREADS, 0, stream>>>(P, radii, is_sky, scales, means3D, viewmatrix,
                                                                        reinterpret_cast<double*>(workspace));
        R3DGW_LAUNCH_CHECK();
        gaussian_reg_finalize_kernel<<<1, REG_VALS * 32, 0, stream>>>(blocks, reinterpret_cast<const double*>(workspace), gamma, out);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

extern "C" int r3dgw_gaussian_regularisers_backward(int P, const int* radii, const uint8_t* is_sky, const float* scales,
                                                    const float* viewmatrix, float gamma, const float* out,
                                                    const float* grads, float* dL_dscales, float* dL_dmeans3D,
                                                    void* stream_) {
    if (P < 0 || !out || !grads || !viewmatrix) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!radii || !scales) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    {
        ProfScope ps(KID_GAUSS_REG, stream);
        gaussian_reg_backward_kernel<<<(P + REG_THREADS - 1) / REG_THREADS, REG_THREADS, 0, stream>>>(
            P, radii, is_sky, scales, viewmatrix, gamma, out, grads, dL_dscales, dL_dmeans3D);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

// ---- allocation-free variant (no counterpart in the reference) -----
// Callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it)
// get the forward pass with its per-CTA partial sums in a file-scope device array instead of a caller workspace.
// Every GPU that loads the library holds the array (REG_MAX_BLOCKS * REG_VALS doubles = 28 416 B) until exit.
// g_reg_scratch is a HOST-SIDE STAND-IN in host code, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, resolved once per device with that device current.
namespace r3dgw {

constexpr int REG_NOALLOC_MAX_P = 1 << 24;   // the f32 counts in out[2], out[3], out[6] are exact up to 2^24
constexpr int REG_NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) double g_reg_scratch[REG_MAX_BLOCKS * REG_VALS];
static std::atomic<double*> g_reg_scratch_addr[REG_NOALLOC_MAX_DEVICES];   // zero-initialised: not resolved yet

// -> 0 and *ptr = device address of g_reg_scratch on the CURRENT device, or a cudaError_t value
static int reg_scratch_for_current_device(double** ptr) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= REG_NOALLOC_MAX_DEVICES) return R3DGW_ERR_UNSUPPORTED;
    double* p = g_reg_scratch_addr[dev].load(std::memory_order_acquire);
    if (!p) {
        // set-up, once per device (a concurrent first call resolves the same address twice: harmless)
        e = cudaGetSymbolAddress((void**)&p, g_reg_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!p) return (int)cudaErrorInvalidSymbol;
        g_reg_scratch_addr[dev].store(p, std::memory_order_release);
    }
    *ptr = p;
    return 0;
}

}  // namespace r3dgw

extern "C" int r3dgw_gaussian_regularisers_noalloc_max_gaussians(void) { return REG_NOALLOC_MAX_P; }

extern "C" int r3dgw_gaussian_regularisers_forward_noalloc(int P, const int* radii, const uint8_t* is_sky,
                                                           const float* scales, const float* means3D,
                                                           const float* viewmatrix, float gamma, float* out,
                                                           void* stream_) {
    if (P < 0 || !out || !viewmatrix) return R3DGW_ERR_BAD_ARG;
    if (P > 0 && (!radii || !scales || !means3D)) return R3DGW_ERR_BAD_ARG;
    if (P > REG_NOALLOC_MAX_P) {
        fprintf(stderr, "r3dgw_gaussian_regularisers_forward_noalloc: P = %d exceeds the maximum of %d Gaussians of the "
                        "allocation-free entry point; use r3dgw_gaussian_regularisers_forward with a caller workspace\n",
                P, REG_NOALLOC_MAX_P);
        return R3DGW_ERR_TOO_LARGE;
    }
    double* scratch = nullptr;
    const int rc = reg_scratch_for_current_device(&scratch);
    if (rc != 0) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int blocks = reg_blocks(P);   // <= REG_MAX_BLOCKS for every P: the array always holds the partial sums
    // nothing in the scratch has to start at zero: every CTA overwrites its REG_VALS slots unconditionally and the
    // finalize kernel reads exactly the `blocks` rows this call wrote (whatever the last call left behind the row
    // `blocks` is never read).  No ProfScope here: the profiler creates events.
    gaussian_reg_partial_kernel<<<blocks, REG_THREADS, 0, stream>>>(P, radii, is_sky, scales, means3D, viewmatrix, scratch);
    R3DGW_LAUNCH_CHECK();
    gaussian_reg_finalize_kernel<<<1, REG_VALS * 32, 0, stream>>>(blocks, scratch, gamma, out);
    R3DGW_LAUNCH_CHECK();
    return 0;
}
