// warp_reduce.cuh - multi-value warp reductions (halving butterfly) shared by blend.cu and shade.cu.
#pragma once
#include <cuda_runtime.h>

namespace r3dgw {

// Sum NV per-lane values over the 32 lanes.  Values are processed in power-of-two chunks; a chunk of
// S = 2^k values is reduced with a halving butterfly (S-1 shuffles) plus 5-k plain steps, after
// which lane l (low 5-k bits zero) holds component (l >> (5-k)) of the chunk.  `sink(index, value)`
// is called by exactly one lane per component.
template <int S, typename Sink>
__device__ __forceinline__ void warp_reduce_chunk(const float* v, int off, unsigned lane, Sink&& sink) {
    constexpr int K = (S == 1) ? 0 : (S == 2) ? 1 : (S == 4) ? 2 : (S == 8) ? 3 : (S == 16) ? 4 : 5;
    float t[S];
#pragma unroll
    for (int q = 0; q < S; ++q) t[q] = v[off + q];
    int cur = S;
#pragma unroll
    for (int step = 0; step < K; ++step) {
        const int half = cur >> 1;
        const unsigned bit = 16u >> step;
        const bool upper = (lane & bit) != 0;
#pragma unroll
        for (int q = 0; q < half; ++q) {
            const float send = upper ? t[q] : t[q + half];
            const float keep = upper ? t[q + half] : t[q];
            t[q] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
        }
        cur = half;
    }
#pragma unroll
    for (unsigned bit = (16u >> K); bit >= 1; bit >>= 1) t[0] += __shfl_xor_sync(0xffffffffu, t[0], bit);
    constexpr unsigned low = (1u << (5 - K)) - 1u;
    if ((lane & low) == 0) sink(off + (int)(lane >> (5 - K)), t[0]);
}

template <int NV, typename Sink>
__device__ __forceinline__ void warp_reduce_multi(const float* v, unsigned lane, Sink&& sink) {
    int off = 0;
    if constexpr ((NV & 32) != 0) { warp_reduce_chunk<32>(v, off, lane, sink); off += 32; }
    if constexpr ((NV & 16) != 0) { warp_reduce_chunk<16>(v, off, lane, sink); off += 16; }
    if constexpr ((NV & 8) != 0) { warp_reduce_chunk<8>(v, off, lane, sink); off += 8; }
    if constexpr ((NV & 4) != 0) { warp_reduce_chunk<4>(v, off, lane, sink); off += 4; }
    if constexpr ((NV & 2) != 0) { warp_reduce_chunk<2>(v, off, lane, sink); off += 2; }
    if constexpr ((NV & 1) != 0) { warp_reduce_chunk<1>(v, off, lane, sink); off += 1; }
}


}  // namespace r3dgw
