// shade.cu - fused per-Gaussian relit shading (forward + backward) for sm_100a.
//
// Replaces the reference's torch graph (dozens of elementwise kernels and an [N,25,3] temporary):
//   EnvironmentLight.get_diffuse_irradiance   scene/NVDIFFREC/light.py:65-94
//   EnvironmentLight.get_specular_light_sh    scene/NVDIFFREC/light.py:97-119  (gauss_kernel utils/sh_utils.py:162-181)
//   EnvironmentLight.shade                    scene/NVDIFFREC/light.py:131-193
//   eval_sh                                   utils/sh_utils.py:81-151
//   gamma_correction / safe_normalize / reflect / dot   scene/NVDIFFREC/util.py:21-31, 523-526
//   sky colour + masked scatters              gaussian_renderer/__init__.py:132-156
//   dr.texture(FG_LUT, uv, 'linear', 'clamp') (nvdiffrast, un-vendored): bilinear, texel centres at
//   (i+0.5)/256, clamp to edge - restated here; parity for it is unpinned by the reference.
// One pass: 44 B in + 36 B out per Gaussian; the 25x3 environment SH and 4x3 sky SH live in shared
// memory; the 512 KB FG LUT is served from L2.  HBM-bound streaming kernel.
#include "common.cuh"
#include "sh.cuh"
#include "warp_reduce.cuh"

namespace r3dgw {

constexpr float IRR_C1 = 0.429043f, IRR_C2 = 0.511664f, IRR_C3 = 0.743125f, IRR_C4 = 0.886227f, IRR_C5 = 0.247708f;
constexpr float INV_GAMMA = 1.0f / 2.2f;
constexpr int LUT_RES = 256;

struct ShadeParams {
    int P, specular, fix_sky, sh_degree;
    const float* positions; const float* normals; const float* albedo; const float* roughness;
    const float* metalness; const unsigned char* is_sky; const int* fg_index; const float* cam_pos;
    const float* env_sh; const float* sky_sh; const float* fg_lut;
};

__device__ __forceinline__ float gamma_fwd(float x) {
    return powf(fminf(fmaxf(x, 0.0f), 1.0f) + 1e-4f, INV_GAMMA);
}
__device__ __forceinline__ float gamma_bwd(float x) {   // d gamma / dx (torch.clamp passes gradient on [0,1])
    if (!(x >= 0.0f && x <= 1.0f)) return 0.0f;
    return INV_GAMMA * powf(x + 1e-4f, INV_GAMMA - 1.0f);
}
__device__ __forceinline__ float3 safe_normalize3(float3 v, float& len) {
    len = sqrtf(fmaxf(v.x * v.x + v.y * v.y + v.z * v.z, 1e-20f));
    return {v.x / len, v.y / len, v.z / len};
}
// cotangent of v -> v / sqrt(max(|v|^2, eps))
__device__ __forceinline__ float3 safe_normalize3_bwd(float3 v, float len, float3 g) {
    const float s2 = v.x * v.x + v.y * v.y + v.z * v.z;
    const float inv = 1.0f / len;
    if (s2 <= 1e-20f) return {g.x * inv, g.y * inv, g.z * inv};
    const float d = (v.x * g.x + v.y * g.y + v.z * g.z) * inv * inv * inv;
    return {g.x * inv - v.x * d, g.y * inv - v.y * d, g.z * inv - v.z * d};
}

// order-2 irradiance polynomial coefficients per SH index (light.py:79-92)
__device__ __forceinline__ void irr_coeffs(float x, float y, float z, float* w) {
    w[0] = IRR_C4;
    w[1] = 2.f * IRR_C2 * y;
    w[2] = 2.f * IRR_C2 * z;
    w[3] = 2.f * IRR_C2 * x;
    w[4] = 2.f * IRR_C1 * x * y;
    w[5] = 2.f * IRR_C1 * y * z;
    w[6] = IRR_C3 * (z * z) - IRR_C5;
    w[7] = 2.f * IRR_C1 * x * z;
    w[8] = IRR_C1 * (x * x - y * y);
}

struct LutTap { int i00, i01, i10, i11; float fx, fy; };
__device__ __forceinline__ LutTap lut_tap(float u, float v) {
    // nvdiffrast linear/clamp: clamp uv to [0,1], texel centres at (i + 0.5) / res
    u = fminf(fmaxf(u, 0.f), 1.f);
    v = fminf(fmaxf(v, 0.f), 1.f);
    const float x = u * LUT_RES - 0.5f, y = v * LUT_RES - 0.5f;
    const float x0f = floorf(x), y0f = floorf(y);
    LutTap t;
    t.fx = x - x0f; t.fy = y - y0f;
    const int x0 = min(max((int)x0f, 0), LUT_RES - 1), x1 = min(max((int)x0f + 1, 0), LUT_RES - 1);
    const int y0 = min(max((int)y0f, 0), LUT_RES - 1), y1 = min(max((int)y0f + 1, 0), LUT_RES - 1);
    t.i00 = y0 * LUT_RES + x0; t.i01 = y0 * LUT_RES + x1; t.i10 = y1 * LUT_RES + x0; t.i11 = y1 * LUT_RES + x1;
    return t;
}

// band index of SH coefficient k (k < 25)
__device__ __forceinline__ int sh_band(int k) { return k == 0 ? 0 : k < 4 ? 1 : k < 9 ? 2 : k < 16 ? 3 : 4; }

template <int DEG>
__global__ void __launch_bounds__(256) shade_forward_kernel(const ShadeParams p, float* __restrict__ rgb,
                                                            float* __restrict__ diffuse, float* __restrict__ specular) {
    constexpr int NSH = (DEG + 1) * (DEG + 1);
    __shared__ float s_env[NSH * 3];
    __shared__ float s_sky[12];
    __shared__ float s_cam[3];
    for (int i = threadIdx.x; i < NSH * 3; i += blockDim.x) s_env[i] = p.env_sh[i];
    if (threadIdx.x < 12) s_sky[threadIdx.x] = p.sky_sh ? p.sky_sh[threadIdx.x] : 0.f;
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.cam_pos[threadIdx.x];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.P) return;
    const float3 pos = {p.positions[3 * i], p.positions[3 * i + 1], p.positions[3 * i + 2]};
    float o_rgb[3], o_dif[3] = {0.f, 0.f, 0.f}, o_spec[3] = {0.f, 0.f, 0.f};
    const bool sky = p.is_sky && p.is_sky[i];
    if (sky) {
        if (p.fix_sky) {
            o_rgb[0] = o_rgb[1] = o_rgb[2] = 1.0f;
        } else {
            float len;
            const float3 d = safe_normalize3({pos.x - s_cam[0], pos.y - s_cam[1], pos.z - s_cam[2]}, len);
            float b[4];
            sh_basis<1>(d.x, d.y, d.z, b);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float r = b[0] * s_sky[c] + b[1] * s_sky[3 + c] + b[2] * s_sky[6 + c] + b[3] * s_sky[9 + c];
                o_rgb[c] = fmaxf(r + 0.5f, 0.0f);
            }
        }
    } else {
        const int fi = p.fg_index ? p.fg_index[i] : i;
        const float3 n = {p.normals[3 * i], p.normals[3 * i + 1], p.normals[3 * i + 2]};
        const float alb[3] = {p.albedo[3 * fi], p.albedo[3 * fi + 1], p.albedo[3 * fi + 2]};
        float w[9];
        irr_coeffs(n.x, n.y, n.z, w);
        float D[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float irr = 0.f;
#pragma unroll
            for (int k = 0; k < 9; ++k) irr += w[k] * s_env[3 * k + c];
            D[c] = alb[c] * fmaxf(irr, 1e-4f);
            o_dif[c] = gamma_fwd(D[c]);
        }
        if (!p.specular) {
#pragma unroll
            for (int c = 0; c < 3; ++c) o_rgb[c] = o_dif[c];
        } else {
            const float kr = p.roughness[fi];
            const bool has_km = p.metalness != nullptr;
            const float km = has_km ? p.metalness[fi] : 0.f;
            float lwo, lr;
            const float3 wo = safe_normalize3({s_cam[0] - pos.x, s_cam[1] - pos.y, s_cam[2] - pos.z}, lwo);
            const float dwn = wo.x * n.x + wo.y * n.y + wo.z * n.z;
            const float3 r = safe_normalize3({2.f * dwn * n.x - wo.x, 2.f * dwn * n.y - wo.y, 2.f * dwn * n.z - wo.z}, lr);
            const float NdotV = fmaxf(dwn, 1e-4f);
            const LutTap t = lut_tap(NdotV, kr);
            const float2* lut = reinterpret_cast<const float2*>(p.fg_lut);
            const float2 t00 = __ldg(lut + t.i00), t01 = __ldg(lut + t.i01), t10 = __ldg(lut + t.i10), t11 = __ldg(lut + t.i11);
            const float fgx = (1.f - t.fy) * ((1.f - t.fx) * t00.x + t.fx * t01.x) + t.fy * ((1.f - t.fx) * t10.x + t.fx * t11.x);
            const float fgy = (1.f - t.fy) * ((1.f - t.fx) * t00.y + t.fx * t01.y) + t.fy * ((1.f - t.fx) * t10.y + t.fx * t11.y);
            float b[NSH];
            sh_basis<DEG>(r.x, r.y, r.z, b);
            float gl[DEG + 1];
#pragma unroll
            for (int l = 0; l <= DEG; ++l) gl[l] = expf(-(float)(l * (l + 1)) * (0.3f * kr));
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float si = 0.f;
#pragma unroll
                for (int k = 0; k < NSH; ++k) si += b[k] * (gl[sh_band(k)] * s_env[3 * k + c]);
                si = fmaxf(si, 1e-4f);
                const float F0 = has_km ? ((1.0f - km) * 0.04f + alb[c] * km) : 0.04f;
                const float S = si * (F0 * fgx + fgy);
                const float shaded = has_km ? ((1.f - km) * D[c] + S) : (D[c] + S);
                o_rgb[c] = gamma_fwd(shaded);
                o_spec[c] = gamma_fwd(S);
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        rgb[3 * i + c] = o_rgb[c];
        if (diffuse) diffuse[3 * i + c] = o_dif[c];
        if (specular) specular[3 * i + c] = o_spec[c];
    }
}


// -------------------------------------------------------------------------------------------------
// Backward.  Grid-stride over Gaussians, one per lane.  d(env_sh) / d(sky_sh) are outer products
//   d env[k][c] += u_i[k] * g_i[c]   (specular: u = Y_k(r) * lobe(l_k), g = dL/d(spec irradiance);
//                                     diffuse:  u = irradiance weight w_k(n), g = dL/d(irradiance))
// so each lane parks its u column and g vector in warp-private shared memory and lane k of the warp sums
// coefficient k over the warp's 32 Gaussians (a [K x 32] . [32 x 3] product): three live accumulators per lane for
// the whole loop instead of 87 per thread, no shuffles.  (The first version kept all 87 sums in registers: 255
// registers + spills, one 256-thread CTA per SM, 0.24 ms at config B; this one: 168 registers, 3 CTAs of 128, 0.22 ms -
// ncu: 104 M warp instructions, 43 % issue-active at 12 warps/SM, stalled on the loads at the head of the body.
// Capping at 128 registers for 4 CTAs spills 544 B and is slower, 0.30 ms.)  Warps -> CTA through shared memory, one
// atomicAdd per coefficient per CTA.
constexpr int SHADE_BWD_THREADS = 128;
template <int DEG>
__global__ void __launch_bounds__(SHADE_BWD_THREADS, 3) shade_backward_kernel(
    const ShadeParams p, const float* __restrict__ g_rgb, const float* __restrict__ g_dif,
    const float* __restrict__ g_spc, float* __restrict__ d_pos, float* __restrict__ d_nrm,
    float* __restrict__ d_alb, float* __restrict__ d_kr, float* __restrict__ d_km,
    float* __restrict__ d_env, float* __restrict__ d_sky) {
    constexpr int NSH = (DEG + 1) * (DEG + 1);
    constexpr int NE = NSH * 3;
    constexpr int NW = SHADE_BWD_THREADS / 32;
    static_assert(NSH + 4 <= 32, "one lane per env / sky coefficient");
    __shared__ float s_env[NE];
    __shared__ float s_sky[12];
    __shared__ float s_cam[3];
    __shared__ float s_red[NW][NE + 12];
    __shared__ float s_U[NW][NSH][33];     // specular columns: Y_k(r_i) * lobe(l_k)
    __shared__ float s_W[NW][9][33];       // diffuse columns: irradiance weights w_k(n_i)
    __shared__ float s_S[NW][4][33];       // sky columns: Y_k(view dir_i), k < 4
    __shared__ float4 s_G[NW][3][32];      // per Gaussian: g_spec, g_irr, g_sky (xyz)
    for (int i = threadIdx.x; i < NE; i += blockDim.x) s_env[i] = p.env_sh[i];
    if (threadIdx.x < 12) s_sky[threadIdx.x] = p.sky_sh ? p.sky_sh[threadIdx.x] : 0.f;
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.cam_pos[threadIdx.x];
    __syncthreads();

    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float accE[3] = {0.f, 0.f, 0.f};   // lane k < NSH: d env[k][:]; lane NSH + k < NSH + 4: d sky[k][:]

    for (int i0 = blockIdx.x * blockDim.x; i0 < p.P; i0 += gridDim.x * blockDim.x) {
        const int i = i0 + (int)threadIdx.x;
        const bool active = i < p.P;
        bool any_spec = false, any_sky = false;   // what this lane parked (the warp votes below)
        float gs[3] = {0.f, 0.f, 0.f}, gi[3] = {0.f, 0.f, 0.f}, gk[3] = {0.f, 0.f, 0.f};
        if (active) {
        const float3 pos = {p.positions[3 * i], p.positions[3 * i + 1], p.positions[3 * i + 2]};
        const float gr[3] = {g_rgb[3 * i], g_rgb[3 * i + 1], g_rgb[3 * i + 2]};
        float3 gp = {0.f, 0.f, 0.f}, gn = {0.f, 0.f, 0.f};
        const bool sky = p.is_sky && p.is_sky[i];
        if (sky) {
            if (!p.fix_sky) {
                float len;
                const float3 v = {pos.x - s_cam[0], pos.y - s_cam[1], pos.z - s_cam[2]};
                const float3 d = safe_normalize3(v, len);
                float b[4], bx[4], by[4], bz[4];
                sh_basis<1>(d.x, d.y, d.z, b);
                sh_basis_grad<1>(d.x, d.y, d.z, bx, by, bz);
                float3 gd = {0.f, 0.f, 0.f};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float raw = b[0] * s_sky[c] + b[1] * s_sky[3 + c] + b[2] * s_sky[6 + c] + b[3] * s_sky[9 + c] + 0.5f;
                    const float g = raw >= 0.0f ? gr[c] : 0.f;
                    gk[c] = g;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        gd.x += g * s_sky[3 * k + c] * bx[k];
                        gd.y += g * s_sky[3 * k + c] * by[k];
                        gd.z += g * s_sky[3 * k + c] * bz[k];
                    }
                }
                gp = safe_normalize3_bwd(v, len, gd);
#pragma unroll
                for (int k = 0; k < 4; ++k) s_S[warp][k][lane] = b[k];
                any_sky = true;
            }
        } else {
            const int fi = p.fg_index ? p.fg_index[i] : i;
            const float3 n = {p.normals[3 * i], p.normals[3 * i + 1], p.normals[3 * i + 2]};
            const float alb[3] = {p.albedo[3 * fi], p.albedo[3 * fi + 1], p.albedo[3 * fi + 2]};
            const float gdv[3] = {g_dif ? g_dif[3 * i] : 0.f, g_dif ? g_dif[3 * i + 1] : 0.f, g_dif ? g_dif[3 * i + 2] : 0.f};
            float w[9];
            irr_coeffs(n.x, n.y, n.z, w);
            float irr_raw[3], irr[3], D[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float a = 0.f;
#pragma unroll
                for (int k = 0; k < 9; ++k) a += w[k] * s_env[3 * k + c];
                irr_raw[c] = a;
                irr[c] = fmaxf(a, 1e-4f);
                D[c] = alb[c] * irr[c];
            }
            float gD[3], galb[3] = {0.f, 0.f, 0.f};
            float gkr = 0.f, gkm = 0.f;
            if (!p.specular) {
#pragma unroll
                for (int c = 0; c < 3; ++c) gD[c] = (gr[c] + gdv[c]) * gamma_bwd(D[c]);
            } else {
                const float gsv[3] = {g_spc ? g_spc[3 * i] : 0.f, g_spc ? g_spc[3 * i + 1] : 0.f, g_spc ? g_spc[3 * i + 2] : 0.f};
                const float kr = p.roughness[fi];
                const bool has_km = p.metalness != nullptr;
                const float km = has_km ? p.metalness[fi] : 0.f;
                const float one_m_km = has_km ? (1.f - km) : 1.f;
                float lwo, lr;
                const float3 vwo = {s_cam[0] - pos.x, s_cam[1] - pos.y, s_cam[2] - pos.z};
                const float3 wo = safe_normalize3(vwo, lwo);
                const float dwn = wo.x * n.x + wo.y * n.y + wo.z * n.z;
                const float3 rraw = {2.f * dwn * n.x - wo.x, 2.f * dwn * n.y - wo.y, 2.f * dwn * n.z - wo.z};
                const float3 r = safe_normalize3(rraw, lr);
                const float NdotV = fmaxf(dwn, 1e-4f);
                const LutTap t = lut_tap(NdotV, kr);
                const float2* lut = reinterpret_cast<const float2*>(p.fg_lut);
                const float2 t00 = __ldg(lut + t.i00), t01 = __ldg(lut + t.i01), t10 = __ldg(lut + t.i10), t11 = __ldg(lut + t.i11);
                const float fgx = (1.f - t.fy) * ((1.f - t.fx) * t00.x + t.fx * t01.x) + t.fy * ((1.f - t.fx) * t10.x + t.fx * t11.x);
                const float fgy = (1.f - t.fy) * ((1.f - t.fx) * t00.y + t.fx * t01.y) + t.fy * ((1.f - t.fx) * t10.y + t.fx * t11.y);
                // d(fg)/du, d(fg)/dv of the bilinear tap (zero where uv was clamped)
                const bool u_in = NdotV >= 0.f && NdotV <= 1.f, v_in = kr >= 0.f && kr <= 1.f;
                const float dfgx_du = u_in ? LUT_RES * ((1.f - t.fy) * (t01.x - t00.x) + t.fy * (t11.x - t10.x)) : 0.f;
                const float dfgy_du = u_in ? LUT_RES * ((1.f - t.fy) * (t01.y - t00.y) + t.fy * (t11.y - t10.y)) : 0.f;
                const float dfgx_dv = v_in ? LUT_RES * ((1.f - t.fx) * (t10.x - t00.x) + t.fx * (t11.x - t01.x)) : 0.f;
                const float dfgy_dv = v_in ? LUT_RES * ((1.f - t.fx) * (t10.y - t00.y) + t.fx * (t11.y - t01.y)) : 0.f;
                float b[NSH], bx[NSH], by[NSH], bz[NSH];
                sh_basis<DEG>(r.x, r.y, r.z, b);
                sh_basis_grad<DEG>(r.x, r.y, r.z, bx, by, bz);
                float gl[DEG + 1], ggl[DEG + 1];
#pragma unroll
                for (int l = 0; l <= DEG; ++l) { gl[l] = expf(-(float)(l * (l + 1)) * (0.3f * kr)); ggl[l] = 0.f; }
                float g_fgx = 0.f, g_fgy = 0.f;
                float3 g_r = {0.f, 0.f, 0.f};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float si_raw = 0.f;
#pragma unroll
                    for (int k = 0; k < NSH; ++k) si_raw += b[k] * (gl[sh_band(k)] * s_env[3 * k + c]);
                    const float si = fmaxf(si_raw, 1e-4f);
                    const float F0 = has_km ? ((1.0f - km) * 0.04f + alb[c] * km) : 0.04f;
                    const float refl = F0 * fgx + fgy;
                    const float S = si * refl;
                    const float shaded = one_m_km * D[c] + S;
                    const float g_sh = gr[c] * gamma_bwd(shaded);
                    gD[c] = gdv[c] * gamma_bwd(D[c]) + one_m_km * g_sh;
                    const float gS = gsv[c] * gamma_bwd(S) + g_sh;
                    if (has_km) gkm += -D[c] * g_sh;
                    const float g_si = (si_raw >= 1e-4f) ? gS * refl : 0.f;
                    gs[c] = g_si;
                    const float g_refl = gS * si;
                    const float g_F0 = g_refl * fgx;
                    g_fgx += g_refl * F0;
                    g_fgy += g_refl;
                    if (has_km) {
                        galb[c] += g_F0 * km;
                        gkm += g_F0 * (alb[c] - 0.04f);
                    }
#pragma unroll
                    for (int k = 0; k < NSH; ++k) {
                        const int l = sh_band(k);
                        const float e = s_env[3 * k + c];
                        ggl[l] += g_si * b[k] * e;
                        const float ce = g_si * gl[l] * e;
                        g_r.x += ce * bx[k]; g_r.y += ce * by[k]; g_r.z += ce * bz[k];
                    }
                }
#pragma unroll
                for (int k = 0; k < NSH; ++k) s_U[warp][k][lane] = b[k] * gl[sh_band(k)];
                any_spec = true;
#pragma unroll
                for (int l = 1; l <= DEG; ++l) gkr += ggl[l] * (-0.3f * (float)(l * (l + 1))) * gl[l];
                gkr += g_fgx * dfgx_dv + g_fgy * dfgy_dv;
                const float g_NdotV = g_fgx * dfgx_du + g_fgy * dfgy_du;
                float g_dwn = (dwn >= 1e-4f) ? g_NdotV : 0.f;
                const float3 g_rraw = safe_normalize3_bwd(rraw, lr, g_r);
                g_dwn += 2.f * (n.x * g_rraw.x + n.y * g_rraw.y + n.z * g_rraw.z);
                gn.x += 2.f * dwn * g_rraw.x; gn.y += 2.f * dwn * g_rraw.y; gn.z += 2.f * dwn * g_rraw.z;
                float3 g_wo = {-g_rraw.x + g_dwn * n.x, -g_rraw.y + g_dwn * n.y, -g_rraw.z + g_dwn * n.z};
                gn.x += g_dwn * wo.x; gn.y += g_dwn * wo.y; gn.z += g_dwn * wo.z;
                const float3 g_v = safe_normalize3_bwd(vwo, lwo, g_wo);
                gp = {-g_v.x, -g_v.y, -g_v.z};
            }
            // diffuse branch: D = albedo * max(irr, 1e-4)
            float wx[9] = {0.f, 0.f, 0.f, 2.f * IRR_C2, 2.f * IRR_C1 * n.y, 0.f, 0.f, 2.f * IRR_C1 * n.z, 2.f * IRR_C1 * n.x};
            float wy[9] = {0.f, 2.f * IRR_C2, 0.f, 0.f, 2.f * IRR_C1 * n.x, 2.f * IRR_C1 * n.z, 0.f, 0.f, -2.f * IRR_C1 * n.y};
            float wz[9] = {0.f, 0.f, 2.f * IRR_C2, 0.f, 0.f, 2.f * IRR_C1 * n.y, 2.f * IRR_C3 * n.z, 2.f * IRR_C1 * n.x, 0.f};
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                galb[c] += gD[c] * irr[c];
                const float g_irr = (irr_raw[c] >= 1e-4f) ? gD[c] * alb[c] : 0.f;
                gi[c] = g_irr;
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const float e = s_env[3 * k + c];
                    gn.x += g_irr * e * wx[k]; gn.y += g_irr * e * wy[k]; gn.z += g_irr * e * wz[k];
                }
            }
#pragma unroll
            for (int k = 0; k < 9; ++k) s_W[warp][k][lane] = w[k];
            if (d_alb) { d_alb[3 * fi] = galb[0]; d_alb[3 * fi + 1] = galb[1]; d_alb[3 * fi + 2] = galb[2]; }
            if (d_kr) d_kr[fi] = gkr;
            if (d_km && p.metalness) d_km[fi] = gkm;
        }
        if (d_pos) { d_pos[3 * i] = gp.x; d_pos[3 * i + 1] = gp.y; d_pos[3 * i + 2] = gp.z; }
        if (d_nrm) { d_nrm[3 * i] = gn.x; d_nrm[3 * i + 1] = gn.y; d_nrm[3 * i + 2] = gn.z; }
        }   // active
        // ---- SH-coefficient gradients of this warp's 32 Gaussians: lane k sums coefficient k ----
        s_G[warp][0][lane] = make_float4(gs[0], gs[1], gs[2], 0.f);
        s_G[warp][1][lane] = make_float4(gi[0], gi[1], gi[2], 0.f);
        s_G[warp][2][lane] = make_float4(gk[0], gk[1], gk[2], 0.f);
        const unsigned m_spec = __ballot_sync(0xffffffffu, any_spec);
        const unsigned m_sky = __ballot_sync(0xffffffffu, any_sky);
        const unsigned m_fg = __ballot_sync(0xffffffffu, gi[0] != 0.f || gi[1] != 0.f || gi[2] != 0.f);
        // columns a lane did not write this iteration hold stale values: their g vector is zero, but 0 * NaN/Inf
        // must not leak in, so such lanes are skipped through the ballots
        if (lane < NSH) {
            for (unsigned m = m_spec; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const float u = s_U[warp][lane][j];
                const float4 g = s_G[warp][0][j];
                accE[0] = fmaf(u, g.x, accE[0]); accE[1] = fmaf(u, g.y, accE[1]); accE[2] = fmaf(u, g.z, accE[2]);
            }
            if (lane < 9) {
                for (unsigned m = m_fg; m; m &= m - 1) {
                    const int j = __ffs(m) - 1;
                    const float u = s_W[warp][lane][j];
                    const float4 g = s_G[warp][1][j];
                    accE[0] = fmaf(u, g.x, accE[0]); accE[1] = fmaf(u, g.y, accE[1]); accE[2] = fmaf(u, g.z, accE[2]);
                }
            }
        } else if (lane < NSH + 4) {
            for (unsigned m = m_sky; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const float u = s_S[warp][lane - NSH][j];
                const float4 g = s_G[warp][2][j];
                accE[0] = fmaf(u, g.x, accE[0]); accE[1] = fmaf(u, g.y, accE[1]); accE[2] = fmaf(u, g.z, accE[2]);
            }
        }
        __syncwarp();   // the columns are rewritten by the next iteration
    }

    // ---- warps -> CTA through shared memory -> one atomic per coefficient
    constexpr int NT = NE + 12;
    for (int k = lane; k < NT; k += 32) s_red[warp][k] = 0.f;
    __syncwarp();
    if (lane < NSH + 4) {
#pragma unroll
        for (int c = 0; c < 3; ++c) s_red[warp][3 * lane + c] = accE[c];   // sky coefficient k sits at NE + 3k + c = 3 (NSH + k) + c
    }
    __syncthreads();
    for (int k = threadIdx.x; k < NT; k += blockDim.x) {
        float v = 0.f;
#pragma unroll
        for (int wv = 0; wv < NW; ++wv) v += s_red[wv][k];
        if (v != 0.f) {
            if (k < NE) { if (d_env) atomicAdd(d_env + k, v); }
            else if (d_sky) atomicAdd(d_sky + (k - NE), v);
        }
    }
}

static ShadeParams to_params(const r3dgw_shade_args* a) {
    ShadeParams p;
    p.P = a->P; p.specular = a->specular; p.fix_sky = a->fix_sky; p.sh_degree = a->sh_degree;
    p.positions = a->positions; p.normals = a->normals; p.albedo = a->albedo; p.roughness = a->roughness;
    p.metalness = a->metalness; p.is_sky = a->is_sky; p.fg_index = a->fg_index; p.cam_pos = a->cam_pos;
    p.env_sh = a->env_sh; p.sky_sh = a->sky_sh; p.fg_lut = a->fg_lut;
    return p;
}
static int check_args(const r3dgw_shade_args* a) {
    if (!a || a->P < 0) return R3DGW_ERR_BAD_ARG;
    if (a->P == 0) return 0;
    if (!a->positions || !a->normals || !a->albedo || !a->cam_pos || !a->env_sh) return R3DGW_ERR_BAD_ARG;
    if (a->specular && (!a->roughness || !a->fg_lut)) return R3DGW_ERR_BAD_ARG;
    if (a->is_sky && !a->fix_sky && !a->sky_sh) return R3DGW_ERR_BAD_ARG;
    if (a->sh_degree < 2 || a->sh_degree > 4) return R3DGW_ERR_UNSUPPORTED;
    return 0;
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" int r3dgw_shade_forward(const r3dgw_shade_args* a, float* rgb, float* diffuse, float* specular, void* stream_) {
    int rc = check_args(a);
    if (rc) return rc;
    if (a->P == 0) return 0;
    if (!rgb) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    ShadeParams p = to_params(a);
    const int blocks = (a->P + 255) / 256;
    ProfScope ps(KID_SHADE_FWD, stream);
    switch (a->sh_degree) {
        case 2: shade_forward_kernel<2><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular); break;
        case 3: shade_forward_kernel<3><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular); break;
        default: shade_forward_kernel<4><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular); break;
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

extern "C" int r3dgw_shade_backward(const r3dgw_shade_args* a, const float* dL_drgb, const float* dL_ddiffuse,
                                    const float* dL_dspecular, float* dL_dpositions, float* dL_dnormals,
                                    float* dL_dalbedo, float* dL_droughness, float* dL_dmetalness,
                                    float* dL_denv_sh, float* dL_dsky_sh, void* stream_) {
    int rc = check_args(a);
    if (rc) return rc;
    if (a->P == 0) return 0;
    if (!dL_drgb) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    ShadeParams p = to_params(a);
    const int blocks = std::min((a->P + SHADE_BWD_THREADS - 1) / SHADE_BWD_THREADS, NUM_SMS * 12);
    ProfScope ps(KID_SHADE_BWD, stream);
    switch (a->sh_degree) {
        case 2: shade_backward_kernel<2><<<blocks, SHADE_BWD_THREADS, 0, stream>>>(p, dL_drgb, dL_ddiffuse, dL_dspecular, dL_dpositions, dL_dnormals, dL_dalbedo, dL_droughness, dL_dmetalness, dL_denv_sh, dL_dsky_sh); break;
        case 3: shade_backward_kernel<3><<<blocks, SHADE_BWD_THREADS, 0, stream>>>(p, dL_drgb, dL_ddiffuse, dL_dspecular, dL_dpositions, dL_dnormals, dL_dalbedo, dL_droughness, dL_dmetalness, dL_denv_sh, dL_dsky_sh); break;
        default: shade_backward_kernel<4><<<blocks, SHADE_BWD_THREADS, 0, stream>>>(p, dL_drgb, dL_ddiffuse, dL_dspecular, dL_dpositions, dL_dnormals, dL_dalbedo, dL_droughness, dL_dmetalness, dL_denv_sh, dL_dsky_sh); break;
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}
