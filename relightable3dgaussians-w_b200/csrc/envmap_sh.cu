// envmap_sh.cu - host side of the eval grid moved to the device:
//   * cubic down-scale of a lat-long environment map (cv2.resize INTER_CUBIC, as the reference's resize_image does,
//     utils/sh_additional_utils.py:48-70, for maps wider than 1000 px, :181-184),
//   * projection onto the real SH basis up to l_max = 4 in FP64 (utils/sh_additional_utils.py:176-206 with
//     utils/spherical_harmonics.py:43-216), with the u8 -> f32/255 + bright-pixel boost of
//     process_environment_map_image (eval_with_gt_envmaps.py:42-49) fused into the load,
//   * the two chained real-SH rotations of the sun-angle sweep for every angle in one launch
//     (spaudiopy.sph.rotate_sh, eval_with_gt_envmaps.py:95-100; Ivanic-Ruedenberg recursion).
// All three are tiny next to a frame; the point is that the sweep's 51 environment bases are produced on the device
// without a host round trip per angle.  FP64 throughout (the reference computes in numpy float64).
#include "common.cuh"

namespace r3dgw {

constexpr int SH_MAX_L = 4;
constexpr int SH_MAX_N = (SH_MAX_L + 1) * (SH_MAX_L + 1);
constexpr int PROJ_THREADS = 128;
constexpr int PROJ_ACC = 2 * SH_MAX_L + 1;     // 1, cos(m phi), sin(m phi) for m = 1..4

struct ShNorms { double k[SH_MAX_N]; };        // K(l,|m|) * (m != 0 ? sqrt(2) : 1), index l*l+l+m

// u8 source: f = u8 / 255 in f32, entries above the threshold multiplied by scale (both f32, like the torch ops)
struct PixelSource {
    const float* f32; const uint8_t* u8; int H, W, stride;   // stride = channels per pixel in memory (>= 3)
    float threshold, scale; int boost;
    __device__ __forceinline__ float at(int y, int x, int c) const {
        const size_t i = ((size_t)y * W + x) * stride + c;
        if (f32) return f32[i];
        float v = __fdiv_rn((float)u8[i], 255.0f);
        if (boost && v > threshold) v = __fmul_rn(v, scale);
        return v;
    }
};

// ---- cv2.resize(..., INTER_CUBIC) for f32 images --------------------------------------------------------------
// OpenCV's generic resize: per destination column fx = (float)((dx + 0.5) * (Ws / Wd) - 0.5), sx = floor(fx), the four
// taps sx-1 .. sx+2 clamped to the image, weights from the A = -0.75 cubic in f32; rows are filtered horizontally
// into f32 and then combined vertically with the same construction.
__device__ __forceinline__ void cubic_weights(float x, float w[4]) {
    const float A = -0.75f;
    w[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
    w[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
    w[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
    w[3] = 1.f - w[0] - w[1] - w[2];
}

__global__ void resize_cubic_kernel(PixelSource src, int Hd, int Wd, float* __restrict__ dst) {
    const int dx = blockIdx.x * blockDim.x + threadIdx.x;
    const int dy = blockIdx.y;
    if (dx >= Wd) return;
    const double scale_x = (double)src.W / Wd, scale_y = (double)src.H / Hd;
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = (int)floorf(fx); fx -= sx;
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = (int)floorf(fy); fy -= sy;
    float wx[4], wy[4];
    cubic_weights(fx, wx);
    cubic_weights(fy, wy);
    int xs[4], ys[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        xs[k] = min(max(sx - 1 + k, 0), src.W - 1);
        ys[k] = min(max(sy - 1 + k, 0), src.H - 1);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float rows[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float a = __fmul_rn(src.at(ys[r], xs[0], c), wx[0]);
            a = __fadd_rn(a, __fmul_rn(src.at(ys[r], xs[1], c), wx[1]));
            a = __fadd_rn(a, __fmul_rn(src.at(ys[r], xs[2], c), wx[2]));
            rows[r] = __fadd_rn(a, __fmul_rn(src.at(ys[r], xs[3], c), wx[3]));
        }
        float v = __fmul_rn(rows[0], wy[0]);
        v = __fadd_rn(v, __fmul_rn(rows[1], wy[1]));
        v = __fadd_rn(v, __fmul_rn(rows[2], wy[2]));
        v = __fadd_rn(v, __fmul_rn(rows[3], wy[3]));
        dst[((size_t)dy * Wd + dx) * 3 + c] = v;
    }
}

// ---- projection -----------------------------------------------------------------------------------------------
__device__ double legendre(int l, int m, double x) {       // spherical_harmonics.py:43-78
    double pmm = 1.0;
    if (m > 0) {
        const double s = sqrt((1.0 - x) * (1.0 + x));
        double f = 1.0;
        for (int i = 0; i < m; ++i) { pmm *= (-f) * s; f += 2.0; }
    }
    if (l == m) return pmm;
    double pm1 = x * (2.0 * m + 1.0) * pmm;
    if (l == m + 1) return pm1;
    double out = pm1;
    for (int ll = m + 2; ll <= l; ++ll) {
        out = ((2.0 * ll - 1.0) * x * pm1 - (ll + m - 1.0) * pmm) / (ll - m);
        pmm = pm1; pm1 = out;
    }
    return out;
}

// One block per image row: the basis separates into a per-row factor K P_l^m(cos theta) dOmega and a per-column
// cos/sin(m phi), so each pixel costs l_max sincos and 3 (2 l_max + 1) FMAs; the row factor is applied once.
__global__ void __launch_bounds__(PROJ_THREADS)
envmap_project_rows_kernel(PixelSource src, int l_max, ShNorms norms, double* __restrict__ partial) {
    const int y = blockIdx.x, H = src.H, W = src.W;
    double acc[3][PROJ_ACC];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < PROJ_ACC; ++k) acc[c][k] = 0.0;
    const double phi_div = (double)W / (M_PI * 2);          // xLocToLon, sh_additional_utils.py:422-423
    for (int x = threadIdx.x; x < W; x += PROJ_THREADS) {
        const double phi = (double)x / phi_div;
        double v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) v[c] = (double)src.at(y, x, c);
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[c][0] += v[c];
#pragma unroll
        for (int m = 1; m <= SH_MAX_L; ++m) {
            if (m <= l_max) {
                double sn, cs;
                sincos((double)m * phi, &sn, &cs);
#pragma unroll
                for (int c = 0; c < 3; ++c) { acc[c][2 * m - 1] += v[c] * cs; acc[c][2 * m] += v[c] * sn; }
            }
        }
    }
    __shared__ double red[PROJ_THREADS / 32][3 * PROJ_ACC];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int k = 0; k < PROJ_ACC; ++k) {
            double a = acc[c][k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0) red[warp][c * PROJ_ACC + k] = a;
        }
    __syncthreads();
    const int n = (l_max + 1) * (l_max + 1);
    if ((int)threadIdx.x < n * 3) {
        const int i = threadIdx.x / 3, c = threadIdx.x % 3;
        int l = 0;
        while ((l + 1) * (l + 1) <= i) ++l;
        const int m = i - l * l - l, am = m < 0 ? -m : m;
        const int k = m == 0 ? 0 : (m > 0 ? 2 * m - 1 : 2 * am);
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < PROJ_THREADS / 32; ++w) s += red[w][c * PROJ_ACC + k];
        const double theta = (double)y / ((double)H / M_PI);                     // yLocToLat, :420-421
        const double row = norms.k[i] * legendre(l, am, cos(theta));
        const double t2 = (1.0 - (((double)y + 0.5) / H)) * M_PI;                // get_solid_angle, :95-105
        const double half = (M_PI / H) / 2.0;
        const double omega = ((M_PI * 2) / W) * (cos(t2 - half) - cos(t2 + half));
        partial[(size_t)y * (SH_MAX_N * 3) + i * 3 + c] = s * row * omega;
    }
}

// fixed-order reduction over the rows: one block per (coefficient, channel)
__global__ void __launch_bounds__(PROJ_THREADS)
envmap_project_finalize_kernel(const double* __restrict__ partial, int H, double* __restrict__ coeffs) {
    const int j = blockIdx.x;                                  // i*3 + c
    double a = 0.0;
    for (int y = threadIdx.x; y < H; y += PROJ_THREADS) a += partial[(size_t)y * (SH_MAX_N * 3) + j];
    __shared__ double red[PROJ_THREADS];
    red[threadIdx.x] = a;
    __syncthreads();
    for (int o = PROJ_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) coeffs[j] = red[0];
}

// ---- EnvironmentLight.render_sh: sh_render (utils/sh_additional_utils.py:288-299) ------------------------------------
// image[y][x][c] = sum_idx Y_idx(theta_y, phi_x) * diffuse_coeff[l(idx)] * coeffs[idx][c], accumulated in index order in
// FP64 like the numpy loop; theta = y pi / H, phi = x 2 pi / W (get_coefficients_matrix, :159-170).
struct ShBandScale { double a[SH_MAX_L + 1]; };
__global__ void __launch_bounds__(128)
envmap_sh_render_kernel(int l_max, ShNorms norms, ShBandScale band, const double* __restrict__ coeffs, int H, int W,
                        double* __restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const double theta = (double)y / ((double)H / M_PI);
    const double phi = (double)x / ((double)W / (M_PI * 2));
    const double ct = cos(theta);
    double acc[3] = {0.0, 0.0, 0.0};
    for (int l = 0; l <= l_max; ++l)
        for (int m = -l; m <= l; ++m) {
            const int i = l * l + l + m, am = m < 0 ? -m : m;
            double v = norms.k[i] * legendre(l, am, ct);
            if (m > 0) v = norms.k[i] * cos((double)m * phi) * legendre(l, am, ct);
            else if (m < 0) v = norms.k[i] * sin((double)am * phi) * legendre(l, am, ct);
#pragma unroll
            for (int c = 0; c < 3; ++c) acc[c] += v * (band.a[l] * coeffs[i * 3 + c]);
        }
    double* o = out + ((size_t)y * W + x) * 3;
    o[0] = acc[0]; o[1] = acc[1]; o[2] = acc[2];
}

// ---- real-SH rotation (Ivanic & Ruedenberg 1996 / errata 1998; what spaudiopy.sph.sh_rotation_matrix computes) ----
constexpr int ROT_THREADS = 96;                 // >= 81 entries of the l = 4 block
constexpr int ROT_DOUBLES = 165;                // sum over l = 0..4 of (2l+1)^2
__device__ __forceinline__ int band_offset(int l) { return l * (4 * l * l - 1) / 3; }   // 0, 1, 10, 35, 84

struct RotBlocks {
    double d[ROT_DOUBLES];
    __device__ double& at(int l, int m, int n) { return d[band_offset(l) + (m + l) * (2 * l + 1) + (n + l)]; }
};

__device__ double ir_p(RotBlocks& R, int i, int l, int a, int b) {
    const double ri1 = R.at(1, i, 1), rim1 = R.at(1, i, -1), ri0 = R.at(1, i, 0);
    if (b == -l) return ri1 * R.at(l - 1, a, -(l - 1)) + rim1 * R.at(l - 1, a, l - 1);
    if (b == l) return ri1 * R.at(l - 1, a, l - 1) - rim1 * R.at(l - 1, a, -(l - 1));
    return ri0 * R.at(l - 1, a, b);
}

// fills R for the z-y-x Euler rotation (yaw, pitch, roll); all threads of the block take part
__device__ void build_rotation(RotBlocks& R, int l_max, double yaw, double pitch, double roll) {
    const int t = threadIdx.x;
    if (t == 0) {
        double sz, cz, sy, cy, sx, cx;
        sincos(yaw, &sz, &cz); sincos(pitch, &sy, &cy); sincos(roll, &sx, &cx);
        // Rz Ry Rx
        const double m[3][3] = {{cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx},
                                {sz * cy, sz * sy * sx + cz * cx, sz * sy * cx - cz * sx},
                                {-sy, cy * sx, cy * cx}};
        const int perm[3] = {1, 2, 0};          // (x, y, z) -> (y, z, x) = m = -1, 0, 1
        R.d[0] = 1.0;
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) R.at(1, a - 1, b - 1) = m[perm[a]][perm[b]];
    }
    __syncthreads();
    for (int l = 2; l <= l_max; ++l) {
        const int side = 2 * l + 1;
        if (t < side * side) {
            const int m = t / side - l, n = t % side - l;
            const int am = m < 0 ? -m : m, an = n < 0 ? -n : n;
            const double d = m == 0 ? 1.0 : 0.0;
            const double denom = an == l ? (double)((2 * l) * (2 * l - 1)) : (double)(l * l - n * n);
            const double u = sqrt((double)(l * l - m * m) / denom);
            const double v = sqrt((1 + d) * (l + am - 1) * (l + am) / denom) * (1 - 2 * d) * 0.5;
            const double w = sqrt((double)((l - am - 1) * (l - am)) / denom) * (1 - d) * (-0.5);
            double acc = 0.0;
            if (u != 0) acc += u * ir_p(R, 0, l, m, n);
            if (v != 0) {
                double V;
                if (m == 0) V = ir_p(R, 1, l, 1, n) + ir_p(R, -1, l, -1, n);
                else if (m > 0) {
                    const double e = m == 1 ? 1.0 : 0.0;
                    V = ir_p(R, 1, l, m - 1, n) * sqrt(1 + e) - ir_p(R, -1, l, -m + 1, n) * (1 - e);
                } else {
                    const double e = m == -1 ? 1.0 : 0.0;
                    V = ir_p(R, 1, l, m + 1, n) * (1 - e) + ir_p(R, -1, l, -m - 1, n) * sqrt(1 + e);
                }
                acc += v * V;
            }
            if (w != 0) {
                const double Wv = m > 0 ? ir_p(R, 1, l, m + 1, n) + ir_p(R, -1, l, -m - 1, n)
                                        : ir_p(R, 1, l, m - 1, n) - ir_p(R, -1, l, -m + 1, n);
                acc += w * Wv;
            }
            R.at(l, m, n) = acc;
        }
        __syncthreads();
    }
}

// f' = D f per band and channel (rows of coefficients times D^T in the reference's row layout)
__device__ void apply_rotation(RotBlocks& R, int l_max, int C, const double* in, double* out) {
    const int n = (l_max + 1) * (l_max + 1);
    for (int j = threadIdx.x; j < n * C; j += blockDim.x) {
        const int i = j / C, c = j % C;
        int l = 0;
        while ((l + 1) * (l + 1) <= i) ++l;
        const int m = i - l * l - l;
        double a = 0.0;
        for (int k = -l; k <= l; ++k) a += in[(l * l + l + k) * C + c] * R.at(l, m, k);
        out[j] = a;
    }
    __syncthreads();
}

// one block per sweep entry: coeffs [n, C] -> rotate by `first` (optional) -> rotate by angles[b] -> out[b]
__global__ void __launch_bounds__(ROT_THREADS)
sh_rotate_sweep_kernel(int l_max, int C, const double* __restrict__ coeffs, const double* __restrict__ first,
                       const double* __restrict__ angles, float* __restrict__ out32, double* __restrict__ out64) {
    __shared__ RotBlocks R;
    __shared__ double f0[SH_MAX_N * 4], f1[SH_MAX_N * 4];
    const int b = blockIdx.x, n = (l_max + 1) * (l_max + 1);
    for (int j = threadIdx.x; j < n * C; j += blockDim.x) f0[j] = coeffs[j];
    __syncthreads();
    double* cur = f0;
    double* nxt = f1;
    if (first) {
        build_rotation(R, l_max, first[0], first[1], first[2]);
        apply_rotation(R, l_max, C, cur, nxt);
        double* t = cur; cur = nxt; nxt = t;
    }
    build_rotation(R, l_max, angles[b * 3 + 0], angles[b * 3 + 1], angles[b * 3 + 2]);
    apply_rotation(R, l_max, C, cur, nxt);
    for (int j = threadIdx.x; j < n * C; j += blockDim.x) {
        if (out32) out32[(size_t)b * n * C + j] = (float)nxt[j];
        if (out64) out64[(size_t)b * n * C + j] = nxt[j];
    }
}

static ShNorms make_norms() {
    ShNorms s;
    for (int l = 0; l <= SH_MAX_L; ++l)
        for (int m = -l; m <= l; ++m) {
            const int a = m < 0 ? -m : m;
            double num = 1.0, den = 1.0;
            for (int k = 2; k <= l - a; ++k) num *= k;
            for (int k = 2; k <= l + a; ++k) den *= k;
            const double K = sqrt(((2 * l + 1) * num) / (4 * M_PI * den));      // spherical_harmonics.py:114-126
            s.k[l * l + l + m] = m == 0 ? K : sqrt(2.0) * K;
        }
    return s;
}

static int check_source(int H, int W, int stride, const float* f32, const uint8_t* u8) {
    if (H <= 0 || W <= 0 || stride < 3 || (!f32 && !u8) || (f32 && u8)) return R3DGW_ERR_BAD_ARG;
    return 0;
}

}  // namespace r3dgw

using namespace r3dgw;

extern "C" {

size_t r3dgw_envmap_sh_workspace_bytes(int H) {
    return align_up((size_t)(H > 0 ? H : 1) * SH_MAX_N * 3 * sizeof(double), 256);
}

int r3dgw_envmap_resize_cubic(int H, int W, int stride, const float* f32, const uint8_t* u8, int boost, float threshold,
                              float scale, int Hd, int Wd, float* dst, void* stream_) {
    if (int rc = check_source(H, W, stride, f32, u8)) return rc;
    if (Hd <= 0 || Wd <= 0 || !dst) return R3DGW_ERR_BAD_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    PixelSource src{f32, u8, H, W, stride, threshold, scale, boost};
    dim3 grid((Wd + 127) / 128, Hd);
    {
        ProfScope ps(KID_ENVMAP_SH, stream);
        resize_cubic_kernel<<<grid, 128, 0, stream>>>(src, Hd, Wd, dst);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_envmap_sh_project(int H, int W, int stride, const float* f32, const uint8_t* u8, int boost, float threshold,
                            float scale, int l_max, char* workspace, double* coeffs, void* stream_) {
    if (int rc = check_source(H, W, stride, f32, u8)) return rc;
    if (!workspace || !coeffs || l_max < 0 || H != W / 2) return R3DGW_ERR_BAD_ARG;   // yres = int(xres/2), :160
    if (l_max > SH_MAX_L) return R3DGW_ERR_UNSUPPORTED;
    cudaStream_t stream = (cudaStream_t)stream_;
    PixelSource src{f32, u8, H, W, stride, threshold, scale, boost};
    double* partial = reinterpret_cast<double*>(workspace);
    const int n = (l_max + 1) * (l_max + 1);
    {
        ProfScope ps(KID_ENVMAP_SH, stream);
        envmap_project_rows_kernel<<<H, PROJ_THREADS, 0, stream>>>(src, l_max, make_norms(), partial);
    }
    R3DGW_LAUNCH_CHECK();
    envmap_project_finalize_kernel<<<n * 3, PROJ_THREADS, 0, stream>>>(partial, H, coeffs);
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_envmap_sh_render(int l_max, const double* coeffs, int W, double* out, void* stream_) {
    if (l_max < 0 || !coeffs || !out || W < 2) return R3DGW_ERR_BAD_ARG;
    if (l_max > SH_MAX_L) return R3DGW_ERR_UNSUPPORTED;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int H = W / 2;
    // get_diffuse_coefficients (utils/sh_additional_utils.py:267-280): pi, 2 pi / 3, then 2 pi (a / b) c for even l, 0 for
    // odd l, all divided by pi
    ShBandScale band;
    for (int l = 0; l <= SH_MAX_L; ++l) {
        double v;
        if (l == 0) v = M_PI;
        else if (l == 1) v = (2 * M_PI) / 3;
        else if (l % 2 == 0) {
            const double a = pow(-1.0, l / 2.0 - 1.0), b = (l + 2.0) * (l - 1.0);
            double fl = 1.0, fh = 1.0;
            for (int k = 2; k <= l; ++k) fl *= k;
            for (int k = 2; k <= l / 2; ++k) fh *= k;
            const double c = fl / (pow(2.0, l) * fh * fh);
            v = 2 * M_PI * (a / b) * c;
        } else v = 0.0;
        band.a[l] = v / M_PI;
    }
    dim3 grid((W + 127) / 128, H);
    {
        ProfScope ps(KID_ENVMAP_SH, stream);
        envmap_sh_render_kernel<<<grid, 128, 0, stream>>>(l_max, make_norms(), band, coeffs, H, W, out);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

int r3dgw_sh_rotate_sweep(int l_max, int C, const double* coeffs, const double* first_ypr, int B,
                          const double* angles_ypr, float* out32, double* out64, void* stream_) {
    if (l_max < 0 || C <= 0 || B < 0 || !coeffs || (B > 0 && !angles_ypr) || (!out32 && !out64)) return R3DGW_ERR_BAD_ARG;
    if (l_max > SH_MAX_L || C > 4) return R3DGW_ERR_UNSUPPORTED;
    if (B == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_;
    {
        ProfScope ps(KID_SH_ROTATE, stream);
        sh_rotate_sweep_kernel<<<B, ROT_THREADS, 0, stream>>>(l_max, C, coeffs, first_ypr, angles_ypr, out32, out64);
    }
    R3DGW_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
