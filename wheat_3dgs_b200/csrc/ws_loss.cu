// ws_loss.cu -- the training loss either side of the rasterizer:
//   L = (1 - lambda) * mean|x - y| + lambda * (1 - mean(SSIM(x, y)))
// (train_vanilla_3dgs.py:77-80 with utils/loss_utils.py:17-62: l1_loss, ssim with an 11x11
// Gaussian window, sigma 1.5, zero padding, groups = channels), forward and backward, so that
// dL/dcolor goes straight into the blend backward.
//
// The reference runs five grouped 11x11 convolutions (mu1, mu2, E[xx], E[yy], E[xy]) through
// cuDNN plus ~20 elementwise kernels, and autograd runs them all again backwards.  Here:
//
//   ssim_fwd   one pass: a CTA loads a 32x32 tile of both images with its 5-pixel halo into
//              shared memory, convolves the five moments separably (11 + 11 taps instead of
//              121), evaluates the SSIM map and the three partial derivatives the backward needs
//              (dm/dmu1, dm/dE[xx], dm/dE[xy]) and stores those; per-CTA sums of the SSIM map
//              and of |x - y| go to a partials array (summed in a fixed order by loss_finish:
//              deterministic, no float atomics).
//   ssim_bwd   one pass: convolves the three partial maps with the same (symmetric) window and
//              combines  dL/dx = a * (conv(dm/dmu1) + 2x conv(dm/dE[xx]) + y conv(dm/dE[xy]))
//              + b * sign(x - y).
//
// Algorithmic HBM bytes per pixel-channel: fwd 8 read + 12 written, bwd 12 + 8 read + 4 written
// (44 B; 415 MB at 3x1536x2048).  HBM-bound.
#include "ws_common.cuh"

namespace ws {

constexpr int LS_TILE = 32;            // output tile edge
constexpr int LS_HALO = 5;             // window 11
constexpr int LS_IN = LS_TILE + 2 * LS_HALO; // 42
constexpr int LS_THREADS = 256;

__constant__ float c_gauss[11];
static bool g_gauss_ready[16] = {};

// utils/loss_utils.py:23-25: exp in double, stored as float32, normalised in float32
static int upload_window()
{
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 15;
    if (g_gauss_ready[dev]) return WS_OK;
    float w[11], sum = 0.0f;
    for (int x = 0; x < 11; x++) {
        w[x] = (float)exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5));
        sum += w[x];
    }
    for (int x = 0; x < 11; x++) w[x] = w[x] / sum;
    WS_CUDA_CHECK(cudaMemcpyToSymbol(c_gauss, w, sizeof(w)));
    g_gauss_ready[dev] = true;
    return WS_OK;
}

struct LossParams {
    const float *x, *y;   // [planes][H][W]
    int planes, H, W, tiles_x, tiles_y;
    float *dm_dmu1, *dm_dexx, *dm_dexy; // [planes][H][W]
    float *partials;      // [n_ctas][2]: sum of ssim map, sum of |x - y|
};

__global__ void __launch_bounds__(LS_THREADS) ssim_fwd(const LossParams p)
{
    __shared__ float s_x[LS_IN][LS_IN + 1];
    __shared__ float s_y[LS_IN][LS_IN + 1];
    __shared__ float s_h[5][LS_IN][LS_TILE + 1]; // horizontally convolved moments
    __shared__ float s_red[2][LS_THREADS / 32];

    const int tid = threadIdx.x;
    const int plane = blockIdx.z;
    const int x0 = blockIdx.x * LS_TILE, y0 = blockIdx.y * LS_TILE;
    const size_t poff = (size_t)plane * p.H * p.W;
    const float *X = p.x + poff, *Y = p.y + poff;

    for (int i = tid; i < LS_IN * LS_IN; i += LS_THREADS) {
        const int r = i / LS_IN, c = i - r * LS_IN;
        const int gy = y0 + r - LS_HALO, gx = x0 + c - LS_HALO;
        float vx = 0.0f, vy = 0.0f; // zero padding (F.conv2d padding = 5)
        if (gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) {
            vx = __ldg(X + (size_t)gy * p.W + gx);
            vy = __ldg(Y + (size_t)gy * p.W + gx);
        }
        s_x[r][c] = vx;
        s_y[r][c] = vy;
    }
    __syncthreads();

    // horizontal pass: 42 rows x 32 columns, FOUR adjacent columns per thread: 14 loads of x and y feed 4 x 11 taps
    // (the passes are shared-memory bound: one value per tap cost 22 LDS per output, this costs 7)
    for (int i = tid; i < LS_IN * (LS_TILE / 4); i += LS_THREADS) {
        const int r = i / (LS_TILE / 4), c0 = (i - r * (LS_TILE / 4)) * 4;
        float m1[4] = {0.f, 0.f, 0.f, 0.f}, m2[4] = {0.f, 0.f, 0.f, 0.f}, xx[4] = {0.f, 0.f, 0.f, 0.f},
              yy[4] = {0.f, 0.f, 0.f, 0.f}, xy[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kk = 0; kk < 14; kk++) {
            const float a = s_x[r][c0 + kk], b = s_y[r][c0 + kk];
            const float aa = a * a, bb = b * b, ab = a * b;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int k = kk - j;
                if (k >= 0 && k < 11) {
                    const float w = c_gauss[k];
                    m1[j] += w * a; m2[j] += w * b; xx[j] += w * aa; yy[j] += w * bb; xy[j] += w * ab;
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            s_h[0][r][c0 + j] = m1[j]; s_h[1][r][c0 + j] = m2[j]; s_h[2][r][c0 + j] = xx[j];
            s_h[3][r][c0 + j] = yy[j]; s_h[4][r][c0 + j] = xy[j];
        }
    }
    __syncthreads();

    // vertical pass + SSIM map and its partials (utils/loss_utils.py:43-57)
    float sum_m = 0.0f, sum_l1 = 0.0f;
    // FOUR adjacent rows per thread: 14 loads per moment feed 4 x 11 taps
    for (int i = tid; i < (LS_TILE / 4) * LS_TILE; i += LS_THREADS) {
        const int rg = i / LS_TILE, c = i - rg * LS_TILE, r0 = rg * 4;
        float acc[5][4];
#pragma unroll
        for (int m = 0; m < 5; m++) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[m][j] = 0.f;
#pragma unroll
            for (int kk = 0; kk < 14; kk++) {
                const float v = s_h[m][r0 + kk][c];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = kk - j;
                    if (k >= 0 && k < 11) acc[m][j] += c_gauss[k] * v;
                }
            }
        }
        const int gx = x0 + c;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int r = r0 + j, gy = y0 + r;
            if (gy >= p.H || gx >= p.W) continue;
            const float mu1 = acc[0][j], mu2 = acc[1][j], exx = acc[2][j], eyy = acc[3][j], exy = acc[4][j];
            const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = exx - mu1_sq, s2 = eyy - mu2_sq, s12 = exy - mu12;
            const float A1 = 2.0f * mu12 + C1, A2 = 2.0f * s12 + C2;
            const float B1 = mu1_sq + mu2_sq + C1, B2 = s1 + s2 + C2;
            const float inv = 1.0f / (B1 * B2);
            const float m = A1 * A2 * inv;
            // m = A1 A2 / (B1 B2) with sigma terms expanded in (mu1, E[xx], E[xy])
            const float dmu1 = 2.0f * inv * (mu2 * (A2 - A1) - mu1 * m * (B2 - B1));
            const float dexx = -m / B2;
            const float dexy = 2.0f * A1 * inv;
            const size_t o = poff + (size_t)gy * p.W + gx;
            p.dm_dmu1[o] = dmu1; p.dm_dexx[o] = dexx; p.dm_dexy[o] = dexy;
            sum_m += m;
            sum_l1 += fabsf(s_x[r + LS_HALO][c + LS_HALO] - s_y[r + LS_HALO][c + LS_HALO]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum_m += __shfl_xor_sync(0xffffffffu, sum_m, o);
        sum_l1 += __shfl_xor_sync(0xffffffffu, sum_l1, o);
    }
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = sum_m; s_red[1][tid >> 5] = sum_l1; }
    __syncthreads();
    if (tid == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int k = 0; k < LS_THREADS / 32; k++) { a += s_red[0][k]; b += s_red[1][k]; }
        const size_t cta = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.partials[2 * cta] = a;
        p.partials[2 * cta + 1] = b;
    }
}

// per-plane sums of the CTA partials, fixed order (deterministic)
__global__ void __launch_bounds__(256) loss_finish(const float *__restrict__ partials, int ctas_per_plane,
                                                   float *__restrict__ sums /*[planes][2]*/)
{
    __shared__ double s_a[256], s_b[256];
    const int plane = blockIdx.x, tid = threadIdx.x;
    double a = 0.0, b = 0.0;
    for (int i = tid; i < ctas_per_plane; i += 256) {
        a += (double)partials[2 * ((size_t)plane * ctas_per_plane + i)];
        b += (double)partials[2 * ((size_t)plane * ctas_per_plane + i) + 1];
    }
    s_a[tid] = a; s_b[tid] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (tid < o) { s_a[tid] += s_a[tid + o]; s_b[tid] += s_b[tid + o]; }
        __syncthreads();
    }
    if (tid == 0) { sums[2 * plane] = (float)s_a[0]; sums[2 * plane + 1] = (float)s_b[0]; }
}

struct LossBwdParams {
    const float *x, *y;
    int planes, H, W;
    const float *dm_dmu1, *dm_dexx, *dm_dexy;
    const float *coef;    // device: {a, b} (scalars) or per-plane pairs if per_plane
    int per_plane;
    float *dL_dx;
};

// dL/dx(q) = a * sum_p w(p - q) [dm_dmu1(p) + 2 x(q) dm_dexx(p) + y(q) dm_dexy(p)] + b * sign(x(q) - y(q))
__global__ void __launch_bounds__(LS_THREADS) ssim_bwd(const LossBwdParams p)
{
    __shared__ float s_in[3][LS_IN][LS_IN + 1];
    __shared__ float s_h[3][LS_IN][LS_TILE + 1];
    const int tid = threadIdx.x;
    const int plane = blockIdx.z;
    const int x0 = blockIdx.x * LS_TILE, y0 = blockIdx.y * LS_TILE;
    const size_t poff = (size_t)plane * p.H * p.W;
    const float *M0 = p.dm_dmu1 + poff, *M1 = p.dm_dexx + poff, *M2 = p.dm_dexy + poff;

    for (int i = tid; i < LS_IN * LS_IN; i += LS_THREADS) {
        const int r = i / LS_IN, c = i - r * LS_IN;
        const int gy = y0 + r - LS_HALO, gx = x0 + c - LS_HALO;
        float a = 0.f, b = 0.f, d = 0.f; // windows centred outside the image do not exist
        if (gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) {
            const size_t o = (size_t)gy * p.W + gx;
            a = __ldg(M0 + o); b = __ldg(M1 + o); d = __ldg(M2 + o);
        }
        s_in[0][r][c] = a; s_in[1][r][c] = b; s_in[2][r][c] = d;
    }
    __syncthreads();
    for (int i = tid; i < LS_IN * (LS_TILE / 4); i += LS_THREADS) {   // four adjacent columns per thread
        const int r = i / (LS_TILE / 4), c0 = (i - r * (LS_TILE / 4)) * 4;
        float acc[3][4];
#pragma unroll
        for (int m = 0; m < 3; m++) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[m][j] = 0.f;
#pragma unroll
            for (int kk = 0; kk < 14; kk++) {
                const float v = s_in[m][r][c0 + kk];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = kk - j;
                    if (k >= 0 && k < 11) acc[m][j] += c_gauss[k] * v;
                }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) s_h[m][r][c0 + j] = acc[m][j];
        }
    }
    __syncthreads();
    const float ca = p.coef[p.per_plane ? 2 * plane : 0], cb = p.coef[p.per_plane ? 2 * plane + 1 : 1];
    for (int i = tid; i < (LS_TILE / 4) * LS_TILE; i += LS_THREADS) {   // four adjacent rows per thread
        const int rg = i / LS_TILE, c = i - rg * LS_TILE, r0 = rg * 4;
        float acc[3][4];
#pragma unroll
        for (int m = 0; m < 3; m++) {
#pragma unroll
            for (int j = 0; j < 4; j++) acc[m][j] = 0.f;
#pragma unroll
            for (int kk = 0; kk < 14; kk++) {
                const float v = s_h[m][r0 + kk][c];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = kk - j;
                    if (k >= 0 && k < 11) acc[m][j] += c_gauss[k] * v;
                }
            }
        }
        const int gx = x0 + c;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int gy = y0 + r0 + j;
            if (gy >= p.H || gx >= p.W) continue;
            const size_t o = poff + (size_t)gy * p.W + gx;
            const float xv = __ldg(p.x + o), yv = __ldg(p.y + o);
            const float diff = xv - yv;
            const float sgn = diff > 0.0f ? 1.0f : (diff < 0.0f ? -1.0f : 0.0f); // torch.abs backward: sign
            p.dL_dx[o] = ca * (acc[0][j] + 2.0f * xv * acc[1][j] + yv * acc[2][j]) + cb * sgn;
        }
    }
}

} // namespace ws

using namespace ws;

extern "C" size_t ws_loss_bytes(int32_t planes, int32_t height, int32_t width)
{
    if (planes <= 0 || height <= 0 || width <= 0) return 0;
    const size_t n = (size_t)planes * height * width;
    const size_t tiles = (size_t)((width + LS_TILE - 1) / LS_TILE) * ((height + LS_TILE - 1) / LS_TILE);
    return 3 * (n * sizeof(float) + 256) + (size_t)planes * tiles * 2 * sizeof(float) + 1024;
}

namespace ws {
struct LossScratch {
    float *m0, *m1, *m2, *partials;
    static LossScratch from(char *p, size_t n, size_t ctas)
    {
        LossScratch s;
        s.m0 = carve<float>(p, n); s.m1 = carve<float>(p, n); s.m2 = carve<float>(p, n);
        s.partials = carve<float>(p, ctas * 2);
        return s;
    }
};
} // namespace ws

extern "C" int ws_l1_ssim_forward(int32_t planes, int32_t height, int32_t width, const float *img, const float *gt,
                                  float *sums, void *scratch, size_t scratch_bytes, ws_stream_t stream)
{
    if (planes <= 0 || height <= 0 || width <= 0 || !img || !gt || !sums || !scratch) {
        set_error("ws_l1_ssim_forward: bad arguments");
        return WS_E_INVALID;
    }
    if (scratch_bytes < ws_loss_bytes(planes, height, width)) {
        set_error("ws_l1_ssim_forward: scratch too small (%zu < %zu)", scratch_bytes, ws_loss_bytes(planes, height, width));
        return WS_E_SCRATCH;
    }
    if (planes > 65535) { set_error("ws_l1_ssim_forward: more than 65535 planes"); return WS_E_UNSUPPORTED; }
    int rc = upload_window();
    if (rc != WS_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    LossParams p;
    p.x = img; p.y = gt; p.planes = planes; p.H = height; p.W = width;
    p.tiles_x = (width + LS_TILE - 1) / LS_TILE; p.tiles_y = (height + LS_TILE - 1) / LS_TILE;
    const size_t ctas = (size_t)planes * p.tiles_x * p.tiles_y;
    LossScratch ls = LossScratch::from((char *)scratch, (size_t)planes * height * width, ctas);
    p.dm_dmu1 = ls.m0; p.dm_dexx = ls.m1; p.dm_dexy = ls.m2; p.partials = ls.partials;
    ssim_fwd<<<dim3(p.tiles_x, p.tiles_y, planes), LS_THREADS, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, false);
    loss_finish<<<planes, 256, 0, st>>>(ls.partials, p.tiles_x * p.tiles_y, sums);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}

extern "C" int ws_l1_ssim_backward(int32_t planes, int32_t height, int32_t width, const float *img, const float *gt,
                                   const float *coef, int32_t per_plane, const void *scratch, size_t scratch_bytes,
                                   float *dL_dimg, ws_stream_t stream)
{
    if (planes <= 0 || height <= 0 || width <= 0 || !img || !gt || !coef || !scratch || !dL_dimg) {
        set_error("ws_l1_ssim_backward: bad arguments");
        return WS_E_INVALID;
    }
    if (scratch_bytes < ws_loss_bytes(planes, height, width)) {
        set_error("ws_l1_ssim_backward: scratch too small");
        return WS_E_SCRATCH;
    }
    int rc = upload_window();
    if (rc != WS_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int tx = (width + LS_TILE - 1) / LS_TILE, ty = (height + LS_TILE - 1) / LS_TILE;
    LossScratch ls = LossScratch::from((char *)scratch, (size_t)planes * height * width, (size_t)planes * tx * ty);
    LossBwdParams p;
    p.x = img; p.y = gt; p.planes = planes; p.H = height; p.W = width;
    p.dm_dmu1 = ls.m0; p.dm_dexx = ls.m1; p.dm_dexy = ls.m2;
    p.coef = coef; p.per_plane = per_plane; p.dL_dx = dL_dimg;
    ssim_bwd<<<dim3(tx, ty, planes), LS_THREADS, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, false);
    return WS_OK;
}
