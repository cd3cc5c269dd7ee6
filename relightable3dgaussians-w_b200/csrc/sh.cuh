// sh.cuh - real spherical-harmonics bases (degree <= 4) and their direction gradients.
// Used by (i) the in-rasterizer SH->RGB option (reference: computeColorFromSH,
// cuda_rasterizer/forward.cu:20-71, backward.cu:20-139; API completeness - the relighting repo
// always passes colors_precomp) and (ii) the fused relit shading kernel (reference: eval_sh,
// utils/sh_utils.py:81-151).  Organised as "basis vector, then dot with coefficients".
#pragma once
#include <cuda_runtime.h>

namespace r3dgw {

// Polynomial basis in the Inria/PlenOctree sign convention. n = (deg+1)^2 values written.
template <int DEG>
__host__ __device__ __forceinline__ void sh_basis(float x, float y, float z, float* b) {
    b[0] = 0.28209479177387814f;
    if (DEG > 0) {
        b[1] = -0.4886025119029199f * y;
        b[2] = 0.4886025119029199f * z;
        b[3] = -0.4886025119029199f * x;
    }
    if (DEG > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        b[4] = 1.0925484305920792f * xy;
        b[5] = -1.0925484305920792f * yz;
        b[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
        b[7] = -1.0925484305920792f * xz;
        b[8] = 0.5462742152960396f * (xx - yy);
        if (DEG > 2) {
            b[9] = -0.5900435899266435f * y * (3.0f * xx - yy);
            b[10] = 2.890611442640554f * xy * z;
            b[11] = -0.4570457994644658f * y * (4.0f * zz - xx - yy);
            b[12] = 0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
            b[13] = -0.4570457994644658f * x * (4.0f * zz - xx - yy);
            b[14] = 1.445305721320277f * z * (xx - yy);
            b[15] = -0.5900435899266435f * x * (xx - 3.0f * yy);
        }
        if (DEG > 3) {
            b[16] = 2.5033429417967046f * xy * (xx - yy);
            b[17] = -1.7701307697799304f * yz * (3.0f * xx - yy);
            b[18] = 0.9461746957575601f * xy * (7.0f * zz - 1.0f);
            b[19] = -0.6690465435572892f * yz * (7.0f * zz - 3.0f);
            b[20] = 0.10578554691520431f * (zz * (35.0f * zz - 30.0f) + 3.0f);
            b[21] = -0.6690465435572892f * xz * (7.0f * zz - 3.0f);
            b[22] = 0.47308734787878004f * (xx - yy) * (7.0f * zz - 1.0f);
            b[23] = -1.7701307697799304f * xz * (xx - 3.0f * yy);
            b[24] = 0.6258357354491761f * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
        }
    }
}

// Partial derivatives of the basis polynomials w.r.t. x, y, z (treating them as independent).
template <int DEG>
__host__ __device__ __forceinline__ void sh_basis_grad(float x, float y, float z, float* dx, float* dy, float* dz) {
    dx[0] = dy[0] = dz[0] = 0.f;
    if (DEG > 0) {
        const float c1 = 0.4886025119029199f;
        dx[1] = 0.f; dy[1] = -c1; dz[1] = 0.f;
        dx[2] = 0.f; dy[2] = 0.f; dz[2] = c1;
        dx[3] = -c1; dy[3] = 0.f; dz[3] = 0.f;
    }
    if (DEG > 1) {
        const float a = 1.0925484305920792f, c6 = 0.31539156525252005f, c8 = 0.5462742152960396f;
        dx[4] = a * y;  dy[4] = a * x;  dz[4] = 0.f;
        dx[5] = 0.f;    dy[5] = -a * z; dz[5] = -a * y;
        dx[6] = -2.f * c6 * x; dy[6] = -2.f * c6 * y; dz[6] = 4.f * c6 * z;
        dx[7] = -a * z; dy[7] = 0.f;    dz[7] = -a * x;
        dx[8] = 2.f * c8 * x; dy[8] = -2.f * c8 * y; dz[8] = 0.f;
    }
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    if (DEG > 2) {
        const float k9 = -0.5900435899266435f, k10 = 2.890611442640554f, k11 = -0.4570457994644658f,
                    k12 = 0.3731763325901154f, k14 = 1.445305721320277f;
        dx[9] = k9 * 6.f * xy;              dy[9] = k9 * 3.f * (xx - yy);        dz[9] = 0.f;
        dx[10] = k10 * yz;                  dy[10] = k10 * xz;                   dz[10] = k10 * xy;
        dx[11] = k11 * -2.f * xy;           dy[11] = k11 * (4.f * zz - xx - 3.f * yy); dz[11] = k11 * 8.f * yz;
        dx[12] = k12 * -6.f * xz;           dy[12] = k12 * -6.f * yz;            dz[12] = k12 * 3.f * (2.f * zz - xx - yy);
        dx[13] = k11 * (4.f * zz - 3.f * xx - yy); dy[13] = k11 * -2.f * xy;     dz[13] = k11 * 8.f * xz;
        dx[14] = k14 * 2.f * xz;            dy[14] = k14 * -2.f * yz;            dz[14] = k14 * (xx - yy);
        dx[15] = k9 * 3.f * (xx - yy);      dy[15] = k9 * -6.f * xy;             dz[15] = 0.f;
    }
    if (DEG > 3) {
        const float k16 = 2.5033429417967046f, k17 = -1.7701307697799304f, k18 = 0.9461746957575601f,
                    k19 = -0.6690465435572892f, k20 = 0.10578554691520431f, k22 = 0.47308734787878004f,
                    k24 = 0.6258357354491761f;
        // b16 = k16 (x^3 y - x y^3)
        dx[16] = k16 * (3.f * xx * y - yy * y); dy[16] = k16 * (xx * x - 3.f * x * yy); dz[16] = 0.f;
        // b17 = k17 (3 x^2 y z - y^3 z)
        dx[17] = k17 * 6.f * xy * z; dy[17] = k17 * (3.f * xx - 3.f * yy) * z; dz[17] = k17 * y * (3.f * xx - yy);
        // b18 = k18 xy (7 z^2 - 1)
        dx[18] = k18 * y * (7.f * zz - 1.f); dy[18] = k18 * x * (7.f * zz - 1.f); dz[18] = k18 * 14.f * xy * z;
        // b19 = k19 yz (7 z^2 - 3)
        dx[19] = 0.f; dy[19] = k19 * z * (7.f * zz - 3.f); dz[19] = k19 * y * (21.f * zz - 3.f);
        // b20 = k20 (35 z^4 - 30 z^2 + 3)
        dx[20] = 0.f; dy[20] = 0.f; dz[20] = k20 * (140.f * zz * z - 60.f * z);
        // b21 = k19 xz (7 z^2 - 3)
        dx[21] = k19 * z * (7.f * zz - 3.f); dy[21] = 0.f; dz[21] = k19 * x * (21.f * zz - 3.f);
        // b22 = k22 (x^2 - y^2)(7 z^2 - 1)
        dx[22] = k22 * 2.f * x * (7.f * zz - 1.f); dy[22] = k22 * -2.f * y * (7.f * zz - 1.f); dz[22] = k22 * (xx - yy) * 14.f * z;
        // b23 = k17 xz (x^2 - 3 y^2)
        dx[23] = k17 * z * (3.f * xx - 3.f * yy); dy[23] = k17 * -6.f * xy * z; dz[23] = k17 * x * (xx - 3.f * yy);
        // b24 = k24 (x^4 - 6 x^2 y^2 + y^4)
        dx[24] = k24 * (4.f * xx * x - 12.f * x * yy); dy[24] = k24 * (4.f * yy * y - 12.f * xx * y); dz[24] = 0.f;
    }
}

// The same derivatives handed to a callback f(k, dYk/dx, dYk/dy, dYk/dz) one coefficient at a time, in index order:
// a consumer that contracts them on the spot (sum_k c_k grad Y_k) keeps 3 accumulators instead of 3 (DEG+1)^2 values live.
template <int DEG, class F>
__host__ __device__ __forceinline__ void sh_basis_grad_each(float x, float y, float z, F&& f) {
    f(0, 0.f, 0.f, 0.f);
    if (DEG > 0) {
        const float c1 = 0.4886025119029199f;
        f(1, 0.f, -c1, 0.f);
        f(2, 0.f, 0.f, c1);
        f(3, -c1, 0.f, 0.f);
    }
    if (DEG > 1) {
        const float a = 1.0925484305920792f, c6 = 0.31539156525252005f, c8 = 0.5462742152960396f;
        f(4, a * y, a * x, 0.f);
        f(5, 0.f, -a * z, -a * y);
        f(6, -2.f * c6 * x, -2.f * c6 * y, 4.f * c6 * z);
        f(7, -a * z, 0.f, -a * x);
        f(8, 2.f * c8 * x, -2.f * c8 * y, 0.f);
    }
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    if (DEG > 2) {
        const float k9 = -0.5900435899266435f, k10 = 2.890611442640554f, k11 = -0.4570457994644658f,
                    k12 = 0.3731763325901154f, k14 = 1.445305721320277f;
        f(9, k9 * 6.f * xy, k9 * 3.f * (xx - yy), 0.f);
        f(10, k10 * yz, k10 * xz, k10 * xy);
        f(11, k11 * -2.f * xy, k11 * (4.f * zz - xx - 3.f * yy), k11 * 8.f * yz);
        f(12, k12 * -6.f * xz, k12 * -6.f * yz, k12 * 3.f * (2.f * zz - xx - yy));
        f(13, k11 * (4.f * zz - 3.f * xx - yy), k11 * -2.f * xy, k11 * 8.f * xz);
        f(14, k14 * 2.f * xz, k14 * -2.f * yz, k14 * (xx - yy));
        f(15, k9 * 3.f * (xx - yy), k9 * -6.f * xy, 0.f);
    }
    if (DEG > 3) {
        const float k16 = 2.5033429417967046f, k17 = -1.7701307697799304f, k18 = 0.9461746957575601f,
                    k19 = -0.6690465435572892f, k20 = 0.10578554691520431f, k22 = 0.47308734787878004f,
                    k24 = 0.6258357354491761f;
        // b16 = k16 (x^3 y - x y^3)
        f(16, k16 * (3.f * xx * y - yy * y), k16 * (xx * x - 3.f * x * yy), 0.f);
        // b17 = k17 (3 x^2 y z - y^3 z)
        f(17, k17 * 6.f * xy * z, k17 * (3.f * xx - 3.f * yy) * z, k17 * y * (3.f * xx - yy));
        // b18 = k18 xy (7 z^2 - 1)
        f(18, k18 * y * (7.f * zz - 1.f), k18 * x * (7.f * zz - 1.f), k18 * 14.f * xy * z);
        // b19 = k19 yz (7 z^2 - 3)
        f(19, 0.f, k19 * z * (7.f * zz - 3.f), k19 * y * (21.f * zz - 3.f));
        // b20 = k20 (35 z^4 - 30 z^2 + 3)
        f(20, 0.f, 0.f, k20 * (140.f * zz * z - 60.f * z));
        // b21 = k19 xz (7 z^2 - 3)
        f(21, k19 * z * (7.f * zz - 3.f), 0.f, k19 * x * (21.f * zz - 3.f));
        // b22 = k22 (x^2 - y^2)(7 z^2 - 1)
        f(22, k22 * 2.f * x * (7.f * zz - 1.f), k22 * -2.f * y * (7.f * zz - 1.f), k22 * (xx - yy) * 14.f * z);
        // b23 = k17 xz (x^2 - 3 y^2)
        f(23, k17 * z * (3.f * xx - 3.f * yy), k17 * -6.f * xy * z, k17 * x * (xx - 3.f * yy));
        // b24 = k24 (x^4 - 6 x^2 y^2 + y^4)
        f(24, k24 * (4.f * xx * x - 12.f * x * yy), k24 * (4.f * yy * y - 12.f * xx * y), 0.f);
    }
}

// d(v/|v|)/dv applied to a cotangent.  (same math as dnormvdv, cuda_rasterizer/auxiliary.h:107-118)
__device__ __forceinline__ float3 normalize_bwd(float3 v, float3 g) {
    float s2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float inv32 = 1.0f / sqrtf(s2 * s2 * s2);
    float dot = v.x * g.x + v.y * g.y + v.z * g.z;
    return {(s2 * g.x - v.x * dot) * inv32, (s2 * g.y - v.y * dot) * inv32, (s2 * g.z - v.z * dot) * inv32};
}

// In-rasterizer SH->RGB (degree <= 3). Returns max(rgb + 0.5, 0) and records the clamp flags.
__device__ inline float3 sh_to_rgb_fwd(int idx, int deg, int max_coeffs, const float* means,
                                       const float* campos, const float* shs, unsigned char* clamped) {
    float3 d = {means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1], means[3 * idx + 2] - campos[2]};
    float inv = 1.0f / sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
    d.x *= inv; d.y *= inv; d.z *= inv;
    float b[16];
    sh_basis<3>(d.x, d.y, d.z, b);
    const float* sh = shs + (size_t)idx * max_coeffs * 3;
    int n = (deg + 1) * (deg + 1);
    float3 r = {0.f, 0.f, 0.f};
    for (int k = 0; k < n; ++k) {
        r.x += b[k] * sh[3 * k + 0];
        r.y += b[k] * sh[3 * k + 1];
        r.z += b[k] * sh[3 * k + 2];
    }
    r.x += 0.5f; r.y += 0.5f; r.z += 0.5f;
    clamped[3 * idx + 0] = (r.x < 0);
    clamped[3 * idx + 1] = (r.y < 0);
    clamped[3 * idx + 2] = (r.z < 0);
    return {fmaxf(r.x, 0.f), fmaxf(r.y, 0.f), fmaxf(r.z, 0.f)};
}

// Backward of the above: writes dL_dsh rows, returns the position gradient contribution.
__device__ inline float3 sh_to_rgb_bwd(int idx, int deg, int max_coeffs, const float* means,
                                       const float* campos, const float* shs, const unsigned char* clamped,
                                       const float* dL_dcolor, float* dL_dshs) {
    float3 d0 = {means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1], means[3 * idx + 2] - campos[2]};
    float inv = 1.0f / sqrtf(d0.x * d0.x + d0.y * d0.y + d0.z * d0.z);
    float3 d = {d0.x * inv, d0.y * inv, d0.z * inv};
    float g[3] = {clamped[3 * idx + 0] ? 0.f : dL_dcolor[3 * idx + 0],
                  clamped[3 * idx + 1] ? 0.f : dL_dcolor[3 * idx + 1],
                  clamped[3 * idx + 2] ? 0.f : dL_dcolor[3 * idx + 2]};
    float b[16], bx[16], by[16], bz[16];
    sh_basis<3>(d.x, d.y, d.z, b);
    sh_basis_grad<3>(d.x, d.y, d.z, bx, by, bz);
    const float* sh = shs + (size_t)idx * max_coeffs * 3;
    float* dsh = dL_dshs + (size_t)idx * max_coeffs * 3;
    int n = (deg + 1) * (deg + 1);
    float3 gd = {0.f, 0.f, 0.f};
    for (int k = 0; k < n; ++k) {
        float s = sh[3 * k] * g[0] + sh[3 * k + 1] * g[1] + sh[3 * k + 2] * g[2];
        gd.x += bx[k] * s; gd.y += by[k] * s; gd.z += bz[k] * s;
        dsh[3 * k + 0] = b[k] * g[0];
        dsh[3 * k + 1] = b[k] * g[1];
        dsh[3 * k + 2] = b[k] * g[2];
    }
    return normalize_bwd(d0, gd);
}

}  // namespace r3dgw
