// raster_f64.cu - TEST INFRASTRUCTURE: double-precision evaluation of the reference rasterizer's BACKWARD pass.
//
// Only tests/ may load this (through oracle/raster_f64.py).  It is an oracle, not product code: a restatement of
//   renderCUDA (backward)     diff-gaussian-rasterization/cuda_rasterizer/backward.cu:399-557
//   computeCov2DCUDA          backward.cu:144-274
//   preprocessCUDA (backward) backward.cu:346-396   (+ computeCov3D backward, :278-341)
// with every DECISION taken exactly as the reference takes it in float (power > 0, alpha < 1/255 through float expf,
// contributor >= last_contributor from the forward pass' n_contrib, the 1.3 tan(fov) clamp flags, radii > 0) and
// only the ARITHMETIC widened to FP64 (sums accumulated with FP64 atomics: order-insensitive at the 1e-16 level).
// Inputs are the forward intermediates both implementations hold bit-identically (means2D, conic_opacity, cov3D,
// the sorted point list, tile ranges, n_contrib); T is rebuilt as the FP64 product of (1 - alpha) over the pixel's
// contributors.  The reference's float atomics make ITS gradients differ run to run; this evaluation is the fixed
// point both the reference and the new kernels are measured against
// (tests/test_gpu_gradient_f64.py: ||new - f64|| <= 1.5 ||ref - f64|| and <= 1e-3 ||f64|| per tensor).
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

constexpr int BX = 16, BY = 16;

// ---- renderCUDA backward: one thread per pixel, whole tile list walked by every thread -----------------------------
__global__ void render_backward_f64(int W, int H, int C, const uint2* __restrict__ ranges,
                                    const uint32_t* __restrict__ point_list, const float* __restrict__ bg,
                                    const float2* __restrict__ xy, const float4* __restrict__ conic_opacity,
                                    const float* __restrict__ colors, const uint32_t* __restrict__ n_contrib,
                                    const float* __restrict__ dL_dpixels, double* __restrict__ dL_dmean2D,
                                    double* __restrict__ dL_dconic, double* __restrict__ dL_dopacity,
                                    double* __restrict__ dL_dcolors, double* __restrict__ image) {
    const int px = blockIdx.x * BX + threadIdx.x, py = blockIdx.y * BY + threadIdx.y;
    if (px >= W || py >= H) return;
    const int pix_id = W * py + px;
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int last = (int)n_contrib[pix_id];
    const float pfx = (float)px, pfy = (float)py;
    constexpr int MAXC = 16;
    double dpix[MAXC], accum_rec[MAXC], last_color[MAXC], Cacc[MAXC];
    for (int c = 0; c < C; ++c) {
        dpix[c] = (double)dL_dpixels[(size_t)c * H * W + pix_id];
        accum_rec[c] = 0.0; last_color[c] = 0.0; Cacc[c] = 0.0;
    }
    // forward in FP64 with the float decisions: T_final = prod (1 - alpha) over the contributors (forward.cu:325-361)
    double T = 1.0;
    for (int k = 0; k < last; ++k) {
        const uint32_t id = point_list[range.x + k];
        const float2 m = xy[id];
        const float4 co = conic_opacity[id];
        const float dxf = m.x - pfx, dyf = m.y - pfy;
        const float power = -0.5f * (co.x * dxf * dxf + co.z * dyf * dyf) - co.y * dxf * dyf;
        if (power > 0.0f) continue;
        const float alpha_f = fminf(0.99f, co.w * expf(power));
        if (alpha_f < 1.0f / 255.0f) continue;
        const double dx = (double)m.x - (double)pfx, dy = (double)m.y - (double)pfy;
        const double pw = -0.5 * ((double)co.x * dx * dx + (double)co.z * dy * dy) - (double)co.y * dx * dy;
        const double alpha = fmin(0.99, (double)co.w * exp(pw));
        for (int c = 0; c < C; ++c) Cacc[c] += (double)colors[(size_t)id * C + c] * alpha * T;
        T *= (1.0 - alpha);
    }
    const double T_final = T;
    if (image)
        for (int c = 0; c < C; ++c) image[(size_t)c * H * W + pix_id] = Cacc[c] + T_final * (double)bg[c];
    double bg_dot = 0.0;
    for (int c = 0; c < C; ++c) bg_dot += (double)bg[c] * dpix[c];
    const double ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
    double last_alpha = 0.0;
    for (int k = last - 1; k >= 0; --k) {
        const uint32_t id = point_list[range.x + k];
        const float2 m = xy[id];
        const float4 co = conic_opacity[id];
        const float dxf = m.x - pfx, dyf = m.y - pfy;
        const float power = -0.5f * (co.x * dxf * dxf + co.z * dyf * dyf) - co.y * dxf * dyf;
        if (power > 0.0f) continue;
        const float alpha_f = fminf(0.99f, co.w * expf(power));
        if (alpha_f < 1.0f / 255.0f) continue;
        const double dx = (double)m.x - (double)pfx, dy = (double)m.y - (double)pfy;
        const double pw = -0.5 * ((double)co.x * dx * dx + (double)co.z * dy * dy) - (double)co.y * dx * dy;
        const double G = exp(pw);
        const double alpha = fmin(0.99, (double)co.w * G);
        T = T / (1.0 - alpha);
        const double dchannel_dcolor = alpha * T;
        double dL_dalpha = 0.0;
        for (int c = 0; c < C; ++c) {
            const double col = (double)colors[(size_t)id * C + c];
            accum_rec[c] = last_alpha * last_color[c] + (1.0 - last_alpha) * accum_rec[c];
            last_color[c] = col;
            dL_dalpha += (col - accum_rec[c]) * dpix[c];
            atomicAdd(&dL_dcolors[(size_t)id * C + c], dchannel_dcolor * dpix[c]);
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final / (1.0 - alpha)) * bg_dot;
        const double dL_dG = (double)co.w * dL_dalpha;
        const double gdx = G * dx, gdy = G * dy;
        const double dG_ddelx = -gdx * (double)co.x - gdy * (double)co.y;
        const double dG_ddely = -gdy * (double)co.z - gdx * (double)co.y;
        atomicAdd(&dL_dmean2D[2 * (size_t)id], dL_dG * dG_ddelx * ddelx_dx);
        atomicAdd(&dL_dmean2D[2 * (size_t)id + 1], dL_dG * dG_ddely * ddely_dy);
        atomicAdd(&dL_dconic[3 * (size_t)id], -0.5 * gdx * dx * dL_dG);
        atomicAdd(&dL_dconic[3 * (size_t)id + 1], -0.5 * gdx * dy * dL_dG);
        atomicAdd(&dL_dconic[3 * (size_t)id + 2], -0.5 * gdy * dy * dL_dG);
        atomicAdd(&dL_dopacity[id], G * dL_dalpha);
    }
}

// 3x3 with GLM's storage convention m[col][row] and column-wise constructor, in double
struct D3 {
    double m[3][3];
    __device__ D3() {}
    __device__ D3(double a, double b, double c, double d, double e, double f, double g, double h, double i) {
        m[0][0] = a; m[0][1] = b; m[0][2] = c; m[1][0] = d; m[1][1] = e; m[1][2] = f; m[2][0] = g; m[2][1] = h; m[2][2] = i;
    }
};
__device__ D3 mul(const D3& A, const D3& B) {      // glm operator*: (A*B)[c][r] = sum_k A[k][r] B[c][k]
    D3 R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}
__device__ D3 transpose(const D3& A) {
    D3 R;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) R.m[c][r] = A.m[r][c];
    return R;
}

// ---- computeCov2DCUDA + preprocessCUDA backward + computeCov3D backward, one thread per Gaussian ------------------
__global__ void geometry_backward_f64(int P, const float* __restrict__ means, const int* __restrict__ radii,
                                      const float* __restrict__ cov3Ds, const float* __restrict__ scales,
                                      const float* __restrict__ rotations, float scale_modifier, float h_xf, float h_yf,
                                      float tan_fovx, float tan_fovy, const float* __restrict__ view,
                                      const float* __restrict__ proj, const double* __restrict__ dL_dmean2D,
                                      const double* __restrict__ dL_dconic3, double* __restrict__ dL_dmeans,
                                      double* __restrict__ dL_dscales, double* __restrict__ dL_drots) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P || !(radii[idx] > 0)) return;
    const double h_x = (double)h_xf, h_y = (double)h_yf;
    const float mxf = means[3 * idx], myf = means[3 * idx + 1], mzf = means[3 * idx + 2];
    const double mx = mxf, my = myf, mz = mzf;
    const double cxx = dL_dconic3[3 * (size_t)idx], cxy = dL_dconic3[3 * (size_t)idx + 1], cyy = dL_dconic3[3 * (size_t)idx + 2];
    // float copy of t for the clamp DECISIONS (backward.cu:166-176), double for the values
    const float tfx = view[0] * mxf + view[4] * myf + view[8] * mzf + view[12];
    const float tfy = view[1] * mxf + view[5] * myf + view[9] * mzf + view[13];
    const float tfz = view[2] * mxf + view[6] * myf + view[10] * mzf + view[14];
    const float limxf = 1.3f * tan_fovx, limyf = 1.3f * tan_fovy;
    const float txtzf = tfx / tfz, tytzf = tfy / tfz;
    const double x_grad_mul = (txtzf < -limxf || txtzf > limxf) ? 0.0 : 1.0;
    const double y_grad_mul = (tytzf < -limyf || tytzf > limyf) ? 0.0 : 1.0;
    double tx = (double)view[0] * mx + (double)view[4] * my + (double)view[8] * mz + (double)view[12];
    double ty = (double)view[1] * mx + (double)view[5] * my + (double)view[9] * mz + (double)view[13];
    const double tz = (double)view[2] * mx + (double)view[6] * my + (double)view[10] * mz + (double)view[14];
    const double limx = (double)limxf, limy = (double)limyf;
    if (x_grad_mul == 0.0) tx = (txtzf < -limxf ? -limx : limx) * tz;
    if (y_grad_mul == 0.0) ty = (tytzf < -limyf ? -limy : limy) * tz;

    const D3 J(h_x / tz, 0.0, -(h_x * tx) / (tz * tz), 0.0, h_y / tz, -(h_y * ty) / (tz * tz), 0.0, 0.0, 0.0);
    const D3 Wm(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    const float* c3 = cov3Ds + 6 * (size_t)idx;
    const D3 Vrk(c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]);
    const D3 T = mul(Wm, J);
    D3 cov2D = mul(mul(transpose(T), transpose(Vrk)), T);
    // the low-pass 0.3f is the float constant (backward.cu:196-198)
    const double a_ = cov2D.m[0][0] + (double)0.3f, b = cov2D.m[0][1], c_ = cov2D.m[1][1] + (double)0.3f;
    const double denom = a_ * c_ - b * b;
    const double denom2inv = 1.0 / ((denom * denom) + (double)0.0000001f);
    double dL_da = 0, dL_db = 0, dL_dc = 0;
    double dcov[6] = {0, 0, 0, 0, 0, 0};
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c_ * c_ * cxx + 2 * b * c_ * cxy + (denom - a_ * c_) * cyy);
        dL_dc = denom2inv * (-a_ * a_ * cyy + 2 * a_ * b * cxy + (denom - a_ * c_) * cxx);
        dL_db = denom2inv * 2 * (b * c_ * cxx - (denom + 2 * b * b) * cxy + a_ * b * cyy);
        const double (*Tm)[3] = T.m;
        dcov[0] = Tm[0][0] * Tm[0][0] * dL_da + Tm[0][0] * Tm[1][0] * dL_db + Tm[1][0] * Tm[1][0] * dL_dc;
        dcov[3] = Tm[0][1] * Tm[0][1] * dL_da + Tm[0][1] * Tm[1][1] * dL_db + Tm[1][1] * Tm[1][1] * dL_dc;
        dcov[5] = Tm[0][2] * Tm[0][2] * dL_da + Tm[0][2] * Tm[1][2] * dL_db + Tm[1][2] * Tm[1][2] * dL_dc;
        dcov[1] = 2 * Tm[0][0] * Tm[0][1] * dL_da + (Tm[0][0] * Tm[1][1] + Tm[0][1] * Tm[1][0]) * dL_db + 2 * Tm[1][0] * Tm[1][1] * dL_dc;
        dcov[2] = 2 * Tm[0][0] * Tm[0][2] * dL_da + (Tm[0][0] * Tm[1][2] + Tm[0][2] * Tm[1][0]) * dL_db + 2 * Tm[1][0] * Tm[1][2] * dL_dc;
        dcov[4] = 2 * Tm[0][2] * Tm[0][1] * dL_da + (Tm[0][1] * Tm[1][2] + Tm[0][2] * Tm[1][1]) * dL_db + 2 * Tm[1][1] * Tm[1][2] * dL_dc;
    }
    const double (*Tm)[3] = T.m;
    const double (*V)[3] = Vrk.m;
    const double dT00 = 2 * (Tm[0][0] * V[0][0] + Tm[0][1] * V[0][1] + Tm[0][2] * V[0][2]) * dL_da + (Tm[1][0] * V[0][0] + Tm[1][1] * V[0][1] + Tm[1][2] * V[0][2]) * dL_db;
    const double dT01 = 2 * (Tm[0][0] * V[1][0] + Tm[0][1] * V[1][1] + Tm[0][2] * V[1][2]) * dL_da + (Tm[1][0] * V[1][0] + Tm[1][1] * V[1][1] + Tm[1][2] * V[1][2]) * dL_db;
    const double dT02 = 2 * (Tm[0][0] * V[2][0] + Tm[0][1] * V[2][1] + Tm[0][2] * V[2][2]) * dL_da + (Tm[1][0] * V[2][0] + Tm[1][1] * V[2][1] + Tm[1][2] * V[2][2]) * dL_db;
    const double dT10 = 2 * (Tm[1][0] * V[0][0] + Tm[1][1] * V[0][1] + Tm[1][2] * V[0][2]) * dL_dc + (Tm[0][0] * V[0][0] + Tm[0][1] * V[0][1] + Tm[0][2] * V[0][2]) * dL_db;
    const double dT11 = 2 * (Tm[1][0] * V[1][0] + Tm[1][1] * V[1][1] + Tm[1][2] * V[1][2]) * dL_dc + (Tm[0][0] * V[1][0] + Tm[0][1] * V[1][1] + Tm[0][2] * V[1][2]) * dL_db;
    const double dT12 = 2 * (Tm[1][0] * V[2][0] + Tm[1][1] * V[2][1] + Tm[1][2] * V[2][2]) * dL_dc + (Tm[0][0] * V[2][0] + Tm[0][1] * V[2][1] + Tm[0][2] * V[2][2]) * dL_db;
    const double (*Wd)[3] = Wm.m;
    const double dJ00 = Wd[0][0] * dT00 + Wd[0][1] * dT01 + Wd[0][2] * dT02;
    const double dJ02 = Wd[2][0] * dT00 + Wd[2][1] * dT01 + Wd[2][2] * dT02;
    const double dJ11 = Wd[1][0] * dT10 + Wd[1][1] * dT11 + Wd[1][2] * dT12;
    const double dJ12 = Wd[2][0] * dT10 + Wd[2][1] * dT11 + Wd[2][2] * dT12;
    const double itz = 1.0 / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const double dtx = x_grad_mul * -h_x * itz2 * dJ02;
    const double dty = y_grad_mul * -h_y * itz2 * dJ12;
    const double dtz = -h_x * itz2 * dJ00 - h_y * itz2 * dJ11 + (2 * h_x * tx) * itz3 * dJ02 + (2 * h_y * ty) * itz3 * dJ12;
    // transformVec4x3Transpose (auxiliary.h:97-105)
    double gmx = (double)view[0] * dtx + (double)view[1] * dty + (double)view[2] * dtz;
    double gmy = (double)view[4] * dtx + (double)view[5] * dty + (double)view[6] * dtz;
    double gmz = (double)view[8] * dtx + (double)view[9] * dty + (double)view[10] * dtz;

    // preprocessCUDA backward (:346-396): the projection path of the mean gradient
    const double p0 = proj[0], p1 = proj[1], p3 = proj[3], p4 = proj[4], p5 = proj[5], p7 = proj[7], p8 = proj[8],
                 p9 = proj[9], p11 = proj[11], p12 = proj[12], p13 = proj[13], p15 = proj[15];
    const double hw = p3 * mx + p7 * my + p11 * mz + p15;
    const double m_w = 1.0 / (hw + (double)0.0000001f);
    const double mul1 = (p0 * mx + p4 * my + p8 * mz + p12) * m_w * m_w;
    const double mul2 = (p1 * mx + p5 * my + p9 * mz + p13) * m_w * m_w;
    const double g2x = dL_dmean2D[2 * (size_t)idx], g2y = dL_dmean2D[2 * (size_t)idx + 1];
    gmx += (p0 * m_w - p3 * mul1) * g2x + (p1 * m_w - p3 * mul2) * g2y;
    gmy += (p4 * m_w - p7 * mul1) * g2x + (p5 * m_w - p7 * mul2) * g2y;
    gmz += (p8 * m_w - p11 * mul1) * g2x + (p9 * m_w - p11 * mul2) * g2y;
    dL_dmeans[3 * (size_t)idx] = gmx; dL_dmeans[3 * (size_t)idx + 1] = gmy; dL_dmeans[3 * (size_t)idx + 2] = gmz;

    // computeCov3D backward (:278-341)
    if (!scales) return;
    const double r = rotations[4 * idx], x = rotations[4 * idx + 1], y = rotations[4 * idx + 2], z = rotations[4 * idx + 3];
    const D3 R(1.0 - 2.0 * (y * y + z * z), 2.0 * (x * y - r * z), 2.0 * (x * z + r * y),
               2.0 * (x * y + r * z), 1.0 - 2.0 * (x * x + z * z), 2.0 * (y * z - r * x),
               2.0 * (x * z - r * y), 2.0 * (y * z + r * x), 1.0 - 2.0 * (x * x + y * y));
    const double sx = (double)scale_modifier * scales[3 * idx], sy = (double)scale_modifier * scales[3 * idx + 1],
                 sz = (double)scale_modifier * scales[3 * idx + 2];
    D3 S(sx, 0, 0, 0, sy, 0, 0, 0, sz);
    const D3 M = mul(S, R);
    const D3 dSigma(dcov[0], 0.5 * dcov[1], 0.5 * dcov[2], 0.5 * dcov[1], dcov[3], 0.5 * dcov[4], 0.5 * dcov[2],
                    0.5 * dcov[4], dcov[5]);
    D3 dM = mul(M, dSigma);
    for (int cc = 0; cc < 3; ++cc)
        for (int rr = 0; rr < 3; ++rr) dM.m[cc][rr] *= 2.0;
    const D3 Rt = transpose(R);
    D3 dMt = transpose(dM);
    for (int k = 0; k < 3; ++k)
        dL_dscales[3 * (size_t)idx + k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
    const double s3[3] = {sx, sy, sz};
    for (int k = 0; k < 3; ++k)
        for (int rr = 0; rr < 3; ++rr) dMt.m[k][rr] *= s3[k];
    const double (*D)[3] = dMt.m;
    dL_drots[4 * (size_t)idx + 0] = 2 * z * (D[0][1] - D[1][0]) + 2 * y * (D[2][0] - D[0][2]) + 2 * x * (D[1][2] - D[2][1]);
    dL_drots[4 * (size_t)idx + 1] = 2 * y * (D[1][0] + D[0][1]) + 2 * z * (D[2][0] + D[0][2]) + 2 * r * (D[1][2] - D[2][1]) - 4 * x * (D[2][2] + D[1][1]);
    dL_drots[4 * (size_t)idx + 2] = 2 * x * (D[1][0] + D[0][1]) + 2 * r * (D[2][0] - D[0][2]) + 2 * z * (D[1][2] + D[2][1]) - 4 * y * (D[2][2] + D[0][0]);
    dL_drots[4 * (size_t)idx + 3] = 2 * r * (D[0][1] - D[1][0]) + 2 * x * (D[2][0] + D[0][2]) + 2 * y * (D[1][2] + D[2][1]) - 4 * z * (D[1][1] + D[0][0]);
}

}  // namespace

extern "C" int raster_f64_render_backward(int W, int H, int C, const void* ranges, const void* point_list,
                                          const float* bg, const void* xy, const void* conic_opacity,
                                          const float* colors, const void* n_contrib, const float* dL_dpixels,
                                          double* dL_dmean2D, double* dL_dconic, double* dL_dopacity,
                                          double* dL_dcolors, double* image, void* stream) {
    if (C < 1 || C > 16) return -2;
    dim3 grid((W + BX - 1) / BX, (H + BY - 1) / BY), block(BX, BY);
    render_backward_f64<<<grid, block, 0, (cudaStream_t)stream>>>(
        W, H, C, (const uint2*)ranges, (const uint32_t*)point_list, bg, (const float2*)xy,
        (const float4*)conic_opacity, colors, (const uint32_t*)n_contrib, dL_dpixels, dL_dmean2D, dL_dconic,
        dL_dopacity, dL_dcolors, image);
    return (int)cudaGetLastError();
}

extern "C" int raster_f64_geometry_backward(int P, const float* means, const int* radii, const float* cov3Ds,
                                            const float* scales, const float* rotations, float scale_modifier,
                                            float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                                            const float* view, const float* proj, const double* dL_dmean2D,
                                            const double* dL_dconic3, double* dL_dmeans, double* dL_dscales,
                                            double* dL_drots, void* stream) {
    geometry_backward_f64<<<(P + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        P, means, radii, cov3Ds, scales, rotations, scale_modifier, focal_x, focal_y, tan_fovx, tan_fovy, view, proj,
        dL_dmean2D, dL_dconic3, dL_dmeans, dL_dscales, dL_drots);
    return (int)cudaGetLastError();
}
