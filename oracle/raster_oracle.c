/*
 * raster_oracle.c - CPU restatement of the reference rasterizer + simple-knn.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this; the product path
 * (relightable3dgaussians-w_b200/) never does.  Scalar, single-threaded, written for readability.
 *
 * Parity status: the reference ships no tests/golden vectors for this path, so this
 * oracle is pinned against outputs of the reference CUDA extension itself, generated on a B200 by
 * tests/golden/make_golden.py and committed under tests/golden/.  CPU vs GPU float rounding differs
 * (FMA contraction, expf ulp), so integer outputs are compared with a small mismatch budget; the
 * bit-exact gate is new-CUDA vs reference-CUDA on the GPU box.
 *
 * Each function cites the reference source it follows (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/ unless stated otherwise).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16

/* Threads used by the loops below (1 = the scalar, deterministic order the parity tests use; bench.py's
 * cpu_baseline / reference legs raise it to the host core count). */
static int g_threads = 1;
void oracle_set_threads(int n) { g_threads = n > 0 ? n : 1; }

/* column-major 3x3 with glm semantics: m[col][row] */
typedef struct { float c[3][3]; } m3;

static m3 m3_cols(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
    m3 m; m.c[0][0]=a; m.c[0][1]=b; m.c[0][2]=c; m.c[1][0]=d; m.c[1][1]=e; m.c[1][2]=f; m.c[2][0]=g; m.c[2][1]=h; m.c[2][2]=i; return m;
}
/* glm operator* (third_party/glm/glm/detail/type_mat3x3.inl:486-520) */
static m3 m3_mul(m3 a, m3 b) {
    m3 r;
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            r.c[col][row] = a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1] + a.c[2][row] * b.c[col][2];
    return r;
}
static m3 m3_t(m3 a) { m3 r; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.c[i][j] = a.c[j][i]; return r; }

/* auxiliary.h:58-77 */
static void xform43(const float* p, const float* m, float* o) {
    o[0] = m[0]*p[0] + m[4]*p[1] + m[8]*p[2] + m[12];
    o[1] = m[1]*p[0] + m[5]*p[1] + m[9]*p[2] + m[13];
    o[2] = m[2]*p[0] + m[6]*p[1] + m[10]*p[2] + m[14];
}
static void xform44(const float* p, const float* m, float* o) {
    xform43(p, m, o);
    o[3] = m[3]*p[0] + m[7]*p[1] + m[11]*p[2] + m[15];
}
/* auxiliary.h:41-44 (double literals) */
static float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }
/* auxiliary.h:46-56 */
static void get_rect(const float* p, int r, int gx, int gy, int* mn, int* mx) {
    mn[0] = imin(gx, imax(0, (int)((p[0] - r) / TILE)));
    mn[1] = imin(gy, imax(0, (int)((p[1] - r) / TILE)));
    mx[0] = imin(gx, imax(0, (int)((p[0] + r + TILE - 1) / TILE)));
    mx[1] = imin(gy, imax(0, (int)((p[1] + r + TILE - 1) / TILE)));
}

/* forward.cu:118-152 */
static void cov3d(const float* s, float mod, const float* q, float* out) {
    m3 S = m3_cols(1,0,0, 0,1,0, 0,0,1);
    S.c[0][0] = mod * s[0]; S.c[1][1] = mod * s[1]; S.c[2][2] = mod * s[2];
    float r = q[0], x = q[1], y = q[2], z = q[3];
    m3 R = m3_cols(1.f - 2.f*(y*y + z*z), 2.f*(x*y - r*z), 2.f*(x*z + r*y),
                   2.f*(x*y + r*z), 1.f - 2.f*(x*x + z*z), 2.f*(y*z - r*x),
                   2.f*(x*z - r*y), 2.f*(y*z + r*x), 1.f - 2.f*(x*x + y*y));
    m3 M = m3_mul(S, R);
    m3 Sg = m3_mul(m3_t(M), M);
    out[0] = Sg.c[0][0]; out[1] = Sg.c[0][1]; out[2] = Sg.c[0][2];
    out[3] = Sg.c[1][1]; out[4] = Sg.c[1][2]; out[5] = Sg.c[2][2];
}

typedef struct { m3 T; float t[3]; float xmul, ymul; } proj_ctx;

/* forward.cu:74-113 */
static void cov2d(const float* mean, float fx, float fy, float tfx, float tfy, const float* c3, const float* view,
                  float* cov, proj_ctx* ctx) {
    float t[3];
    xform43(mean, view, t);
    const float limx = 1.3f * tfx, limy = 1.3f * tfy;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    m3 J = m3_cols(fx / t[2], 0.0f, -(fx * t[0]) / (t[2] * t[2]),
                   0.0f, fy / t[2], -(fy * t[1]) / (t[2] * t[2]),
                   0, 0, 0);
    m3 W = m3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
    m3 T = m3_mul(W, J);
    m3 V = m3_cols(c3[0], c3[1], c3[2], c3[1], c3[3], c3[4], c3[2], c3[4], c3[5]);
    m3 c = m3_mul(m3_mul(m3_t(T), m3_t(V)), T);
    c.c[0][0] += 0.3f; c.c[1][1] += 0.3f;
    cov[0] = c.c[0][0]; cov[1] = c.c[0][1]; cov[2] = c.c[1][1];
    if (ctx) {
        ctx->T = T; ctx->t[0] = t[0]; ctx->t[1] = t[1]; ctx->t[2] = t[2];
        ctx->xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        ctx->ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    }
}

/* K1: forward.cu:155-256 (+ in_frustum auxiliary.h:139-164) */
void oracle_preprocess(int P, const float* means, const float* scales, float mod, const float* rots,
                       const float* opac, const float* cov3D_precomp, const float* view, const float* proj,
                       int W, int H, float tfx, float tfy,
                       int* radii, float* means2D, float* depths, float* cov3Ds, float* conic_opacity,
                       uint32_t* tiles_touched) {
    const float fy = H / (2.0f * tfy), fx = W / (2.0f * tfx);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static) num_threads(g_threads)
    for (int i = 0; i < P; ++i) {
        radii[i] = 0; tiles_touched[i] = 0;
        const float* p = means + 3 * i;
        float pv[3], ph[4];
        xform43(p, view, pv);
        if (pv[2] <= 0.2f) continue;
        xform44(p, proj, ph);
        float pw = 1.0f / (ph[3] + 0.0000001f);
        float pp[3] = { ph[0] * pw, ph[1] * pw, ph[2] * pw };
        const float* c3;
        if (cov3D_precomp) c3 = cov3D_precomp + 6 * i;
        else { cov3d(scales + 3 * i, mod, rots + 4 * i, cov3Ds + 6 * i); c3 = cov3Ds + 6 * i; }
        float cov[3];
        cov2d(p, fx, fy, tfx, tfy, c3, view, cov, NULL);
        float det = cov[0] * cov[2] - cov[1] * cov[1];
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float conic[3] = { cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv };
        float mid = 0.5f * (cov[0] + cov[2]);
        float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        float pim[2] = { ndc2pix(pp[0], W), ndc2pix(pp[1], H) };
        int mn[2], mx[2];
        get_rect(pim, (int)my_radius, gx, gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        depths[i] = pv[2];
        radii[i] = (int)my_radius;
        means2D[2 * i] = pim[0]; means2D[2 * i + 1] = pim[1];
        conic_opacity[4 * i] = conic[0]; conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2]; conic_opacity[4 * i + 3] = opac[i];
        tiles_touched[i] = (uint32_t)((mx[1] - mn[1]) * (mx[0] - mn[0]));
    }
}

typedef struct { uint64_t key; uint32_t val; uint32_t seq; } kv;
static uint64_t g_mask;
static int kv_cmp(const void* a, const void* b) {
    const kv* x = (const kv*)a; const kv* y = (const kv*)b;
    uint64_t kx = x->key & g_mask, ky = y->key & g_mask;
    if (kx != ky) return kx < ky ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);   /* stability: emission order */
}

/* K2 + stable sort + K3: rasterizer_impl.cu:70-138, 276-318 (cub InclusiveSum / SortPairs semantics) */
int oracle_bin(int P, int W, int H, const int* radii, const float* means2D, const float* depths,
               const uint32_t* tiles_touched, uint32_t* point_offsets, int R,
               uint64_t* keys_sorted, uint32_t* list_sorted, uint32_t* ranges /* [tiles][2] */) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    uint32_t run = 0;
    for (int i = 0; i < P; ++i) { run += tiles_touched[i]; point_offsets[i] = run; }
    if ((int)run != R) return -1;
    kv* a = (kv*)malloc(sizeof(kv) * (size_t)(R > 0 ? R : 1));
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0 : point_offsets[i - 1];
        int mn[2], mx[2];
        get_rect(means2D + 2 * i, radii[i], gx, gy, mn, mx);
        uint32_t dbits; memcpy(&dbits, depths + i, 4);
        for (int y = mn[1]; y < mx[1]; ++y)
            for (int x = mn[0]; x < mx[0]; ++x) {
                uint64_t key = (uint64_t)(y * gx + x);
                key <<= 32; key |= dbits;
                a[off].key = key; a[off].val = (uint32_t)i; a[off].seq = off; off++;
            }
    }
    /* getHigherMsb (rasterizer_impl.cu:35-50): number of bits needed for the tile count */
    int bit = 0; { uint32_t n = (uint32_t)(gx * gy); while (n) { ++bit; n >>= 1; } }
    g_mask = (32 + bit) >= 64 ? ~0ull : ((1ull << (32 + bit)) - 1ull);
    qsort(a, (size_t)R, sizeof(kv), kv_cmp);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)(gx * gy));
    for (int i = 0; i < R; ++i) {
        keys_sorted[i] = a[i].key; list_sorted[i] = a[i].val;
        uint32_t cur = (uint32_t)(a[i].key >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(a[i - 1].key >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
    free(a);
    return 0;
}

/* K4: forward.cu:261-374, C channels */
void oracle_blend_forward(int C, int W, int H, const uint32_t* ranges, const uint32_t* list, const float* means2D,
                          const float* conic_opacity, const float* feat, const float* bg,
                          float* final_T, uint32_t* n_contrib, float* out) {
    const int gx = (W + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 4) num_threads(g_threads)
    for (int py = 0; py < H; ++py) for (int px = 0; px < W; ++px) {
        const int tile = (py / TILE) * gx + (px / TILE);
        const uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
        float T = 1.0f; uint32_t contributor = 0, last = 0;
        float acc[32]; for (int c = 0; c < C; ++c) acc[c] = 0.f;
        const float fx = (float)px, fy = (float)py;
        for (uint32_t k = lo; k < hi; ++k) {
            contributor++;
            const uint32_t id = list[k];
            const float dx = means2D[2 * id] - fx, dy = means2D[2 * id + 1] - fy;
            const float* co = conic_opacity + 4 * id;
            const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, co[3] * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break;
            for (int c = 0; c < C; ++c) acc[c] += feat[id * C + c] * alpha * T;
            T = test_T; last = contributor;
        }
        const int pid = py * W + px;
        final_T[pid] = T; n_contrib[pid] = last;
        for (int c = 0; c < C; ++c) out[(size_t)c * H * W + pid] = acc[c] + T * bg[c];
    }
}

#define ACC(dst, v) do { float _v = (v); _Pragma("omp atomic") (dst) += _v; } while (0)
/* K5: backward.cu:399-557 (sequential accumulation when single-threaded) */
void oracle_blend_backward(int C, int W, int H, const uint32_t* ranges, const uint32_t* list, const float* means2D,
                           const float* conic_opacity, const float* feat, const float* bg, const float* final_T,
                           const uint32_t* n_contrib, const float* dL_dpix,
                           float* dL_dmean2D /*[P,3]*/, float* dL_dconic /*[P,4]*/, float* dL_dopacity, float* dL_dfeat) {
    const int gx = (W + TILE - 1) / TILE;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
    /* with g_threads > 1 the += below become atomic adds in thread order (like the GPU's atomicAdd) */
#pragma omp parallel for schedule(dynamic, 4) num_threads(g_threads)
    for (int py = 0; py < H; ++py) for (int px = 0; px < W; ++px) {
        const int tile = (py / TILE) * gx + (px / TILE);
        const uint32_t lo = ranges[2 * tile], hi = ranges[2 * tile + 1];
        const int pid = py * W + px;
        const float T_final = final_T[pid];
        float T = T_final;
        const uint32_t last_contributor = n_contrib[pid];
        float accum_rec[32], last_color[32], dpx[32];
        float bg_dot = 0.f;
        for (int c = 0; c < C; ++c) { accum_rec[c] = 0; last_color[c] = 0; dpx[c] = dL_dpix[(size_t)c * H * W + pid]; }
        float last_alpha = 0;
        const float fx = (float)px, fy = (float)py;
        for (uint32_t k = hi; k-- > lo;) {
            const uint32_t contributor = k - lo;
            if (contributor >= last_contributor) continue;
            const uint32_t id = list[k];
            const float dx = means2D[2 * id] - fx, dy = means2D[2 * id + 1] - fy;
            const float* co = conic_opacity + 4 * id;
            const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
            if (power > 0.0f) continue;
            const float G = expf(power);
            const float alpha = fminf(0.99f, co[3] * G);
            if (alpha < 1.0f / 255.0f) continue;
            T = T / (1.f - alpha);
            const float dchannel_dcolor = alpha * T;
            float dL_dalpha = 0.0f;
            for (int c = 0; c < C; ++c) {
                const float col = feat[id * C + c];
                accum_rec[c] = last_alpha * last_color[c] + (1.f - last_alpha) * accum_rec[c];
                last_color[c] = col;
                dL_dalpha += (col - accum_rec[c]) * dpx[c];
                ACC(dL_dfeat[id * C + c], dchannel_dcolor * dpx[c]);
            }
            dL_dalpha *= T;
            last_alpha = alpha;
            bg_dot = 0.f;
            for (int c = 0; c < C; ++c) bg_dot += bg[c] * dpx[c];
            dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
            const float dL_dG = co[3] * dL_dalpha;
            const float gdx = G * dx, gdy = G * dy;
            const float dG_ddelx = -gdx * co[0] - gdy * co[1];
            const float dG_ddely = -gdy * co[2] - gdx * co[1];
            ACC(dL_dmean2D[3 * id], dL_dG * dG_ddelx * ddelx_dx);
            ACC(dL_dmean2D[3 * id + 1], dL_dG * dG_ddely * ddely_dy);
            ACC(dL_dconic[4 * id], -0.5f * gdx * dx * dL_dG);
            ACC(dL_dconic[4 * id + 1], -0.5f * gdx * dy * dL_dG);
            ACC(dL_dconic[4 * id + 3], -0.5f * gdy * dy * dL_dG);
            ACC(dL_dopacity[id], G * dL_dalpha);
        }
    }
}

/* K6 + K7: backward.cu:144-274 (computeCov2DCUDA), :346-396 (preprocessCUDA), :278-341 (computeCov3D) */
void oracle_geometry_backward(int P, const float* means, const int* radii, const float* cov3Ds, const float* scales,
                              float mod, const float* rots, const float* view, const float* proj, int W, int H,
                              float tfx, float tfy, const float* dL_dmean2D, const float* dL_dconic,
                              float* dL_dmean3D, float* dL_dcov3D, float* dL_dscale, float* dL_drot) {
    const float hy = H / (2.0f * tfy), hx = W / (2.0f * tfx);
#pragma omp parallel for schedule(static) num_threads(g_threads)
    for (int i = 0; i < P; ++i) {
        if (!(radii[i] > 0)) continue;
        const float* mean = means + 3 * i;
        const float* c3 = cov3Ds + 6 * i;
        float cov[3]; proj_ctx cx;
        cov2d(mean, hx, hy, tfx, tfy, c3, view, cov, &cx);
        const float gcx = dL_dconic[4 * i], gcy = dL_dconic[4 * i + 1], gcz = dL_dconic[4 * i + 3];
        const float a = cov[0], b = cov[1], c = cov[2];
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        m3 T = cx.T;
        float* gc = dL_dcov3D + 6 * i;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * gcx + 2 * b * c * gcy + (denom - a * c) * gcz);
            dL_dc = denom2inv * (-a * a * gcz + 2 * a * b * gcy + (denom - a * c) * gcx);
            dL_db = denom2inv * 2 * (b * c * gcx - (denom + 2 * b * b) * gcy + a * b * gcz);
            gc[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
            gc[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
            gc[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
            gc[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
            gc[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
            gc[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
        } else {
            for (int k = 0; k < 6; ++k) gc[k] = 0;
        }
        const float V[3][3] = { { c3[0], c3[1], c3[2] }, { c3[1], c3[3], c3[4] }, { c3[2], c3[4], c3[5] } };
        float dT0[3], dT1[3];
        for (int k = 0; k < 3; ++k) {
            dT0[k] = 2 * (T.c[0][0] * V[k][0] + T.c[0][1] * V[k][1] + T.c[0][2] * V[k][2]) * dL_da +
                     (T.c[1][0] * V[k][0] + T.c[1][1] * V[k][1] + T.c[1][2] * V[k][2]) * dL_db;
            dT1[k] = 2 * (T.c[1][0] * V[k][0] + T.c[1][1] * V[k][1] + T.c[1][2] * V[k][2]) * dL_dc +
                     (T.c[0][0] * V[k][0] + T.c[0][1] * V[k][1] + T.c[0][2] * V[k][2]) * dL_db;
        }
        /* W[col][row] = view[col + 4*row] */
        const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
        const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
        const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
        const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
        const float tz = 1.f / cx.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = cx.xmul * -hx * tz2 * dJ02;
        const float dty = cx.ymul * -hy * tz2 * dJ12;
        const float dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * cx.t[0]) * tz3 * dJ02 + (2 * hy * cx.t[1]) * tz3 * dJ12;
        float gm[3] = { view[0] * dtx + view[1] * dty + view[2] * dtz,
                        view[4] * dtx + view[5] * dty + view[6] * dtz,
                        view[8] * dtx + view[9] * dty + view[10] * dtz };
        /* projection path, backward.cu:366-387 */
        float mh[4]; xform44(mean, proj, mh);
        const float mw = 1.0f / (mh[3] + 0.0000001f);
        const float mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * mw * mw;
        const float mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * mw * mw;
        const float g2x = dL_dmean2D[3 * i], g2y = dL_dmean2D[3 * i + 1];
        gm[0] += (proj[0] * mw - proj[3] * mul1) * g2x + (proj[1] * mw - proj[3] * mul2) * g2y;
        gm[1] += (proj[4] * mw - proj[7] * mul1) * g2x + (proj[5] * mw - proj[7] * mul2) * g2y;
        gm[2] += (proj[8] * mw - proj[11] * mul1) * g2x + (proj[9] * mw - proj[11] * mul2) * g2y;
        dL_dmean3D[3 * i] = gm[0]; dL_dmean3D[3 * i + 1] = gm[1]; dL_dmean3D[3 * i + 2] = gm[2];
        if (!scales) continue;
        /* computeCov3D backward, backward.cu:278-341 */
        const float* q = rots + 4 * i;
        const float r = q[0], x = q[1], y = q[2], z = q[3];
        m3 R = m3_cols(1.f - 2.f*(y*y + z*z), 2.f*(x*y - r*z), 2.f*(x*z + r*y),
                       2.f*(x*y + r*z), 1.f - 2.f*(x*x + z*z), 2.f*(y*z - r*x),
                       2.f*(x*z - r*y), 2.f*(y*z + r*x), 1.f - 2.f*(x*x + y*y));
        m3 S = m3_cols(1,0,0, 0,1,0, 0,0,1);
        const float s[3] = { mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2] };
        S.c[0][0] = s[0]; S.c[1][1] = s[1]; S.c[2][2] = s[2];
        m3 M = m3_mul(S, R);
        m3 dSig = m3_cols(gc[0], 0.5f * gc[1], 0.5f * gc[2], 0.5f * gc[1], gc[3], 0.5f * gc[4], 0.5f * gc[2], 0.5f * gc[4], gc[5]);
        m3 dM = m3_mul(M, dSig);
        for (int u = 0; u < 3; ++u) for (int v = 0; v < 3; ++v) dM.c[u][v] *= 2.0f;
        m3 Rt = m3_t(R), dMt = m3_t(dM);
        for (int k = 0; k < 3; ++k)
            dL_dscale[3 * i + k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
        for (int k = 0; k < 3; ++k) for (int v = 0; v < 3; ++v) dMt.c[k][v] *= s[k];
        float* gq = dL_drot + 4 * i;
        gq[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
        gq[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
        gq[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
        gq[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    }
}

/* checkFrustum, rasterizer_impl.cu:54-66 */
void oracle_mark_visible(int P, const float* means, const float* view, uint8_t* present) {
    for (int i = 0; i < P; ++i) { float pv[3]; xform43(means + 3 * i, view, pv); present[i] = pv[2] > 0.2f; }
}

/* simple-knn: exact 3-NN mean squared distance (simple_knn.cu:131-183).  The reference's Morton/box scan is
 * an exact search, so a brute-force scan with the same distance expression and the same insertion rule
 * (updateKBest<3>) yields the same three values. */
void oracle_knn_dist2(int P, const float* pts, float* out) {
#pragma omp parallel for schedule(static) num_threads(g_threads)
    for (int i = 0; i < P; ++i) {
        float best[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
        for (int j = 0; j < P; ++j) {
            if (j == i) continue;
            float dx = pts[3 * j] - pts[3 * i], dy = pts[3 * j + 1] - pts[3 * i + 1], dz = pts[3 * j + 2] - pts[3 * i + 2];
            float dist = dx * dx + dy * dy + dz * dz;
            for (int k = 0; k < 3; ++k) if (best[k] > dist) { float t = best[k]; best[k] = dist; dist = t; }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
