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
#include <cstdlib>
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
    // pre-shading fusion (GEOM kernels): normals derived in-kernel from scales / rotations, the depth and normal
    // "colours" of render() produced beside the shaded ones.  Outputs / upstream gradients are addressed with a row
    // stride so that they are either five dense [P,3] arrays (stride 3) or columns of one [P,16] feature matrix.
    const float* scales; const float* rotations; const float* viewmatrix;
    int stride;
};

// ---- GaussianModel.get_normal (scene/gaussian_model.py:115-122) ---------------------------------------------------
// build_rotation(q) (utils/general_utils.py:98-119: q / |q|, then the 3x3 from (r,x,y,z)), get_minimum_axis (:158-162:
// the column of R at argmin(scale)), flip_align_view (:165-170: n *= (n . -viewdir >= 0 ? 1 : -1)).
struct MinAxisNormal { float3 n; float r, x, y, z, inv_len, sgn; int axis; };
__device__ __forceinline__ MinAxisNormal min_axis_normal(const float* __restrict__ s3, const float* __restrict__ q4,
                                                         bool flip, float3 viewdir) {
    MinAxisNormal m;
    const float qr = q4[0], qx = q4[1], qy = q4[2], qz = q4[3];
    const float len = sqrtf(qr * qr + qx * qx + qy * qy + qz * qz);
    m.inv_len = 1.0f / len;
    m.r = qr / len; m.x = qx / len; m.y = qy / len; m.z = qz / len;
    const float s0 = s3[0], s1 = s3[1], s2 = s3[2];
    m.axis = (s1 < s0) ? ((s2 < s1) ? 2 : 1) : ((s2 < s0) ? 2 : 0);     // first index of the minimum
    const float r = m.r, x = m.x, y = m.y, z = m.z;
    if (m.axis == 0) m.n = {1.f - 2.f * (y * y + z * z), 2.f * (x * y + r * z), 2.f * (x * z - r * y)};
    else if (m.axis == 1) m.n = {2.f * (x * y - r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + r * x)};
    else m.n = {2.f * (x * z + r * y), 2.f * (y * z - r * x), 1.f - 2.f * (x * x + y * y)};
    m.sgn = 1.f;
    if (flip) {
        const float d = m.n.x * -viewdir.x + m.n.y * -viewdir.y + m.n.z * -viewdir.z;
        m.sgn = d >= 0.f ? 1.f : -1.f;
        m.n = {m.n.x * m.sgn, m.n.y * m.sgn, m.n.z * m.sgn};
    }
    return m;
}
// cotangent of the (flipped) normal -> cotangent of the un-normalised quaternion the kernel was given
__device__ __forceinline__ void min_axis_normal_bwd(const MinAxisNormal& m, float3 g, float* gq) {
    const float g0 = g.x * m.sgn, g1 = g.y * m.sgn, g2 = g.z * m.sgn;
    const float r = m.r, x = m.x, y = m.y, z = m.z;
    float hr, hx, hy, hz;     // d / d(normalised quaternion)
    if (m.axis == 0) {
        hr = 2.f * (z * g1 - y * g2); hx = 2.f * (y * g1 + z * g2);
        hy = -4.f * y * g0 + 2.f * (x * g1 - r * g2); hz = -4.f * z * g0 + 2.f * (r * g1 + x * g2);
    } else if (m.axis == 1) {
        hr = 2.f * (x * g2 - z * g0); hx = 2.f * y * g0 - 4.f * x * g1 + 2.f * r * g2;
        hy = 2.f * (x * g0 + z * g2); hz = -2.f * r * g0 - 4.f * z * g1 + 2.f * y * g2;
    } else {
        hr = 2.f * (y * g0 - x * g1); hx = 2.f * (z * g0 - r * g1) - 4.f * x * g2;
        hy = 2.f * (r * g0 + z * g1) - 4.f * y * g2; hz = 2.f * (x * g0 + y * g1);
    }
    const float dt = r * hr + x * hx + y * hy + z * hz;
    gq[0] = (hr - r * dt) * m.inv_len; gq[1] = (hx - x * dt) * m.inv_len;
    gq[2] = (hy - y * dt) * m.inv_len; gq[3] = (hz - z * dt) * m.inv_len;
}

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

template <int DEG, bool GEOM>
__global__ void __launch_bounds__(256) shade_forward_kernel(const ShadeParams p, float* __restrict__ rgb,
                                                            float* __restrict__ diffuse, float* __restrict__ specular,
                                                            float* __restrict__ depth3, float* __restrict__ normal01,
                                                            float* __restrict__ ones, float* __restrict__ normals_out) {
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
    float3 n_geom = {0.f, 0.f, 0.f};
    if (GEOM) {
        // dir_pp_normalized = safe_normalize(xyz - campos) (gaussian_renderer/__init__.py:127-129), get_normal, and the
        // depth / normal "colours" (:182-189; get_depth scene/gaussian_model.py:125-130) for EVERY Gaussian, sky included
        float len;
        const float3 vd = safe_normalize3({pos.x - s_cam[0], pos.y - s_cam[1], pos.z - s_cam[2]}, len);
        n_geom = min_axis_normal(p.scales + 3 * i, p.rotations + 4 * i, true, vd).n;
        const float* vm = p.viewmatrix;
        const float depth = vm[2] * pos.x + vm[6] * pos.y + vm[10] * pos.z + vm[14];
        const size_t o = (size_t)p.stride * i;
        if (depth3) { depth3[o] = depth; depth3[o + 1] = depth; depth3[o + 2] = depth; }
        if (normal01) {
            normal01[o] = 0.5f * n_geom.x + 0.5f; normal01[o + 1] = 0.5f * n_geom.y + 0.5f; normal01[o + 2] = 0.5f * n_geom.z + 0.5f;
        }
        if (ones) ones[o] = 1.0f;
        if (normals_out) { normals_out[3 * i] = n_geom.x; normals_out[3 * i + 1] = n_geom.y; normals_out[3 * i + 2] = n_geom.z; }
    }
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
        float3 n;
        if (GEOM) n = n_geom;
        else n = {p.normals[3 * i], p.normals[3 * i + 1], p.normals[3 * i + 2]};
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
    const size_t so = (size_t)(GEOM ? p.stride : 3) * i;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        rgb[so + c] = o_rgb[c];
        if (diffuse) diffuse[so + c] = o_dif[c];
        if (specular) specular[so + c] = o_spec[c];
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
template <int DEG, bool GEOM, int MINB = 3>
__global__ void __launch_bounds__(SHADE_BWD_THREADS, MINB) shade_backward_kernel(
    const ShadeParams p, const float* __restrict__ g_rgb, const float* __restrict__ g_dif,
    const float* __restrict__ g_spc, const float* __restrict__ g_depth3, const float* __restrict__ g_n01,
    float* __restrict__ d_pos, float* __restrict__ d_nrm, float* __restrict__ d_rot,
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
        const size_t go = (size_t)(GEOM ? p.stride : 3) * i;
        const float gr[3] = {g_rgb[go], g_rgb[go + 1], g_rgb[go + 2]};
        float3 gp = {0.f, 0.f, 0.f}, gn = {0.f, 0.f, 0.f};
        const bool sky = p.is_sky && p.is_sky[i];
        MinAxisNormal mn;
        if (GEOM) {
            float len;
            const float3 vd = safe_normalize3({pos.x - s_cam[0], pos.y - s_cam[1], pos.z - s_cam[2]}, len);
            mn = min_axis_normal(p.scales + 3 * i, p.rotations + 4 * i, true, vd);
        }
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
            float3 n;
            if (GEOM) n = mn.n;
            else n = {p.normals[3 * i], p.normals[3 * i + 1], p.normals[3 * i + 2]};
            const float alb[3] = {p.albedo[3 * fi], p.albedo[3 * fi + 1], p.albedo[3 * fi + 2]};
            const float gdv[3] = {g_dif ? g_dif[go] : 0.f, g_dif ? g_dif[go + 1] : 0.f, g_dif ? g_dif[go + 2] : 0.f};
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
                const float gsv[3] = {g_spc ? g_spc[go] : 0.f, g_spc ? g_spc[go + 1] : 0.f, g_spc ? g_spc[go + 2] : 0.f};
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
                // b[k] stays live; the 3 NSH basis derivatives are contracted one coefficient at a time by
                // sh_basis_grad_each (the first version held b, bx, by, bz = 4 NSH values: 168 registers)
                float b[NSH];
                sh_basis<DEG>(r.x, r.y, r.z, b);
                float gl[DEG + 1], ggl[DEG + 1];
#pragma unroll
                for (int l = 0; l <= DEG; ++l) { gl[l] = expf(-(float)(l * (l + 1)) * (0.3f * kr)); ggl[l] = 0.f; }
                float g_fgx = 0.f, g_fgy = 0.f;
                float si_raw[3] = {0.f, 0.f, 0.f};
#pragma unroll
                for (int k = 0; k < NSH; ++k) {
                    const float glk = gl[sh_band(k)];
                    s_U[warp][k][lane] = b[k] * glk;
#pragma unroll
                    for (int c = 0; c < 3; ++c) si_raw[c] += b[k] * (glk * s_env[3 * k + c]);
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float si = fmaxf(si_raw[c], 1e-4f);
                    const float F0 = has_km ? ((1.0f - km) * 0.04f + alb[c] * km) : 0.04f;
                    const float refl = F0 * fgx + fgy;
                    const float S = si * refl;
                    const float shaded = one_m_km * D[c] + S;
                    const float g_sh = gr[c] * gamma_bwd(shaded);
                    gD[c] = gdv[c] * gamma_bwd(D[c]) + one_m_km * g_sh;
                    const float gS = gsv[c] * gamma_bwd(S) + g_sh;
                    if (has_km) gkm += -D[c] * g_sh;
                    gs[c] = (si_raw[c] >= 1e-4f) ? gS * refl : 0.f;
                    const float g_refl = gS * si;
                    const float g_F0 = g_refl * fgx;
                    g_fgx += g_refl * F0;
                    g_fgy += g_refl;
                    if (has_km) {
                        galb[c] += g_F0 * km;
                        gkm += g_F0 * (alb[c] - 0.04f);
                    }
                }
                // w_k = sum_c g_si[c] env[k][c]:  d/d lobe_l += w_k b_k,  d/d r += w_k lobe_l grad Y_k
                float3 g_r = {0.f, 0.f, 0.f};
                sh_basis_grad_each<DEG>(r.x, r.y, r.z, [&](int k, float dbx, float dby, float dbz) {
                    const float wk = gs[0] * s_env[3 * k] + gs[1] * s_env[3 * k + 1] + gs[2] * s_env[3 * k + 2];
                    const int l = sh_band(k);
                    ggl[l] += wk * b[k];
                    const float cw = wk * gl[l];
                    g_r.x += cw * dbx; g_r.y += cw * dby; g_r.z += cw * dbz;
                });
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
        if (GEOM) {
            // depth "colour" (3 copies of the view-space z) and normal "colour" (0.5 n + 0.5) of every Gaussian
            if (g_depth3) {
                const float gd = g_depth3[go] + g_depth3[go + 1] + g_depth3[go + 2];
                const float* vm = p.viewmatrix;
                gp.x += gd * vm[2]; gp.y += gd * vm[6]; gp.z += gd * vm[10];
            }
            if (g_n01) { gn.x += 0.5f * g_n01[go]; gn.y += 0.5f * g_n01[go + 1]; gn.z += 0.5f * g_n01[go + 2]; }
            if (d_rot) {
                float gq[4];
                min_axis_normal_bwd(mn, gn, gq);
                *reinterpret_cast<float4*>(d_rot + 4 * (size_t)i) = make_float4(gq[0], gq[1], gq[2], gq[3]);
            }
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
    p.scales = nullptr; p.rotations = nullptr; p.viewmatrix = nullptr; p.stride = 3;
    return p;
}
static int check_args(const r3dgw_shade_args* a, bool need_normals = true) {
    if (!a || a->P < 0) return R3DGW_ERR_BAD_ARG;
    if (a->P == 0) return 0;
    if (!a->positions || (need_normals && !a->normals) || !a->albedo || !a->cam_pos || !a->env_sh) return R3DGW_ERR_BAD_ARG;
    if (a->specular && (!a->roughness || !a->fg_lut)) return R3DGW_ERR_BAD_ARG;
    if (a->is_sky && !a->fix_sky && !a->sky_sh) return R3DGW_ERR_BAD_ARG;
    if (a->sh_degree < 2 || a->sh_degree > 4) return R3DGW_ERR_UNSUPPORTED;
    return 0;
}

template <bool GEOM>
static int launch_shade_forward(const ShadeParams& p, float* rgb, float* diffuse, float* specular, float* depth3,
                                float* normal01, float* ones, float* normals_out, cudaStream_t stream) {
    const int blocks = (p.P + 255) / 256;
    ProfScope ps(KID_SHADE_FWD, stream);
    switch (p.sh_degree) {
        case 2: shade_forward_kernel<2, GEOM><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular, depth3, normal01, ones, normals_out); break;
        case 3: shade_forward_kernel<3, GEOM><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular, depth3, normal01, ones, normals_out); break;
        default: shade_forward_kernel<4, GEOM><<<blocks, 256, 0, stream>>>(p, rgb, diffuse, specular, depth3, normal01, ones, normals_out); break;
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

template <bool GEOM>
static int launch_shade_backward(const ShadeParams& p, const float* g_rgb, const float* g_dif, const float* g_spc,
                                 const float* g_depth3, const float* g_n01, float* d_pos, float* d_nrm, float* d_rot,
                                 float* d_alb, float* d_kr, float* d_km, float* d_env, float* d_sky, cudaStream_t stream) {
    const int blocks = std::min((p.P + SHADE_BWD_THREADS - 1) / SHADE_BWD_THREADS, NUM_SMS * 12);
    ProfScope ps(KID_SHADE_BWD, stream);
#define R3DGW_SHADE_BWD(D) shade_backward_kernel<D, GEOM><<<blocks, SHADE_BWD_THREADS, 0, stream>>>( \
        p, g_rgb, g_dif, g_spc, g_depth3, g_n01, d_pos, d_nrm, d_rot, d_alb, d_kr, d_km, d_env, d_sky)
    static const int minb = [] { const char* e = getenv("R3DGW_SHADE_BWD_MINB"); return e ? atoi(e) : 3; }();
    switch (p.sh_degree) {
        case 2: R3DGW_SHADE_BWD(2); break;
        case 3: R3DGW_SHADE_BWD(3); break;
        default:
            if (minb == 4)      // occupancy experiment: 128 registers / 4 CTAs per SM instead of 168 / 3
                shade_backward_kernel<4, GEOM, 4><<<blocks, SHADE_BWD_THREADS, 0, stream>>>(
                    p, g_rgb, g_dif, g_spc, g_depth3, g_n01, d_pos, d_nrm, d_rot, d_alb, d_kr, d_km, d_env, d_sky);
            else R3DGW_SHADE_BWD(4);
            break;
    }
#undef R3DGW_SHADE_BWD
    R3DGW_LAUNCH_CHECK();
    return 0;
}

__global__ void gaussian_normal_forward_kernel(int P, const float* __restrict__ scales, const float* __restrict__ rotations,
                                               const float* __restrict__ positions, const float* __restrict__ cam_pos,
                                               float* __restrict__ normals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float3 vd = {0.f, 0.f, 0.f};
    if (cam_pos) {
        float len;
        vd = safe_normalize3({positions[3 * i] - cam_pos[0], positions[3 * i + 1] - cam_pos[1], positions[3 * i + 2] - cam_pos[2]}, len);
    }
    const MinAxisNormal m = min_axis_normal(scales + 3 * i, rotations + 4 * i, cam_pos != nullptr, vd);
    normals[3 * i] = m.n.x; normals[3 * i + 1] = m.n.y; normals[3 * i + 2] = m.n.z;
}

__global__ void gaussian_normal_backward_kernel(int P, const float* __restrict__ scales, const float* __restrict__ rotations,
                                                const float* __restrict__ positions, const float* __restrict__ cam_pos,
                                                const float* __restrict__ g_n, float* __restrict__ d_rot) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    float3 vd = {0.f, 0.f, 0.f};
    if (cam_pos) {
        float len;
        vd = safe_normalize3({positions[3 * i] - cam_pos[0], positions[3 * i + 1] - cam_pos[1], positions[3 * i + 2] - cam_pos[2]}, len);
    }
    const MinAxisNormal m = min_axis_normal(scales + 3 * i, rotations + 4 * i, cam_pos != nullptr, vd);
    float gq[4];
    min_axis_normal_bwd(m, {g_n[3 * i], g_n[3 * i + 1], g_n[3 * i + 2]}, gq);
    *reinterpret_cast<float4*>(d_rot + 4 * (size_t)i) = make_float4(gq[0], gq[1], gq[2], gq[3]);
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" int r3dgw_shade_forward(const r3dgw_shade_args* a, float* rgb, float* diffuse, float* specular, void* stream_) {
    int rc = check_args(a);
    if (rc) return rc;
    if (a->P == 0) return 0;
    if (!rgb) return R3DGW_ERR_BAD_ARG;
    return launch_shade_forward<false>(to_params(a), rgb, diffuse, specular, nullptr, nullptr, nullptr, nullptr,
                                       (cudaStream_t)stream_);
}

extern "C" int r3dgw_shade_backward(const r3dgw_shade_args* a, const float* dL_drgb, const float* dL_ddiffuse,
                                    const float* dL_dspecular, float* dL_dpositions, float* dL_dnormals,
                                    float* dL_dalbedo, float* dL_droughness, float* dL_dmetalness,
                                    float* dL_denv_sh, float* dL_dsky_sh, void* stream_) {
    int rc = check_args(a);
    if (rc) return rc;
    if (a->P == 0) return 0;
    if (!dL_drgb) return R3DGW_ERR_BAD_ARG;
    return launch_shade_backward<false>(to_params(a), dL_drgb, dL_ddiffuse, dL_dspecular, nullptr, nullptr, dL_dpositions,
                                        dL_dnormals, nullptr, dL_dalbedo, dL_droughness, dL_dmetalness, dL_denv_sh,
                                        dL_dsky_sh, (cudaStream_t)stream_);
}

static int preshade_params(const r3dgw_preshade_args* a, int stride, ShadeParams* p) {
    if (!a) return R3DGW_ERR_BAD_ARG;
    int rc = check_args(&a->shade, false);
    if (rc) return rc;
    if (a->shade.P > 0 && (!a->scales || !a->rotations || !a->viewmatrix || stride < 3)) return R3DGW_ERR_BAD_ARG;
    *p = to_params(&a->shade);
    p->scales = a->scales; p->rotations = a->rotations; p->viewmatrix = a->viewmatrix; p->stride = stride;
    return 0;
}

extern "C" int r3dgw_preshade_forward(const r3dgw_preshade_args* a, int out_stride, float* rgb, float* diffuse,
                                      float* specular, float* depth3, float* normal01, float* ones, float* normals,
                                      void* stream_) {
    ShadeParams p;
    int rc = preshade_params(a, out_stride, &p);
    if (rc) return rc;
    if (p.P == 0) return 0;
    if (!rgb) return R3DGW_ERR_BAD_ARG;
    return launch_shade_forward<true>(p, rgb, diffuse, specular, depth3, normal01, ones, normals, (cudaStream_t)stream_);
}

extern "C" int r3dgw_preshade_backward(const r3dgw_preshade_args* a, int g_stride, const float* dL_drgb,
                                       const float* dL_ddiffuse, const float* dL_dspecular, const float* dL_ddepth3,
                                       const float* dL_dnormal01, float* dL_dpositions, float* dL_drotations,
                                       float* dL_dalbedo, float* dL_droughness, float* dL_dmetalness,
                                       float* dL_denv_sh, float* dL_dsky_sh, void* stream_) {
    ShadeParams p;
    int rc = preshade_params(a, g_stride, &p);
    if (rc) return rc;
    if (p.P == 0) return 0;
    if (!dL_drgb) return R3DGW_ERR_BAD_ARG;
    return launch_shade_backward<true>(p, dL_drgb, dL_ddiffuse, dL_dspecular, dL_ddepth3, dL_dnormal01, dL_dpositions,
                                       nullptr, dL_drotations, dL_dalbedo, dL_droughness, dL_dmetalness, dL_denv_sh,
                                       dL_dsky_sh, (cudaStream_t)stream_);
}

extern "C" int r3dgw_gaussian_normal_forward(int P, const float* scales, const float* rotations, const float* positions,
                                             const float* cam_pos, float* normals, void* stream_) {
    if (P < 0) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!scales || !rotations || !normals || (cam_pos && !positions)) return R3DGW_ERR_BAD_ARG;
    gaussian_normal_forward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, scales, rotations, positions,
                                                                                       cam_pos, normals);
    R3DGW_LAUNCH_CHECK();
    return 0;
}

extern "C" int r3dgw_gaussian_normal_backward(int P, const float* scales, const float* rotations, const float* positions,
                                              const float* cam_pos, const float* dL_dnormals, float* dL_drotations,
                                              void* stream_) {
    if (P < 0) return R3DGW_ERR_BAD_ARG;
    if (P == 0) return 0;
    if (!scales || !rotations || !dL_dnormals || !dL_drotations || (cam_pos && !positions)) return R3DGW_ERR_BAD_ARG;
    gaussian_normal_backward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream_>>>(P, scales, rotations, positions,
                                                                                        cam_pos, dL_dnormals, dL_drotations);
    R3DGW_LAUNCH_CHECK();
    return 0;
}
