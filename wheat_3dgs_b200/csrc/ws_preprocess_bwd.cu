// ws_preprocess_bwd.cu -- backward of the per-Gaussian stages in ONE kernel (replaces
// computeCov2DCUDA, DGR backward.cu:144-274, and preprocessCUDA backward, :346-412, with SH
// backward :20-139 and cov3D backward :278-341).
//
// Input: the 48-byte accumulator row acc[g][12] the blend backward summed into
//   {dL_dcolor rgb, dL_ddepth | Sgx, Sgy, Sgxx, Sgxy | Sgyy, Sg, -, -}
// where g = G * dL_dalpha of a (pixel, Gaussian) pair, gx = g*dx, gxx = g*dx*dx, ... summed over the
// pixels.  With k = -0.5 * opacity:  dL_dconic = k * (Sgxx, Sgxy, Sgyy),  dL_dopacity = Sg,
// dL_dmean2D = k * (W * (A*Sgx + B*Sgy), H * (C*Sgy + B*Sgx))  (DGR backward.cu:583-598 with the
// per-Gaussian constants factored out of the pixel sums).
// Output: the eight gradient tensors of rasterize_gaussians_backward; EVERY element is written
// (zeros for culled Gaussians and for SH bands above the active degree), so the caller needs no
// zero-fill (the reference memsets ~304 MB per call at 1 M Gaussians,
// DGR rasterize_points.cu:153-162).  cov3D is recomputed from scale/rotation instead of being
// stored by the forward (saves 24 B/Gaussian written + read).
//
// The SH coefficients of the CTA's 256 Gaussians are staged in shared memory with coalesced
// 16-byte loads (as in the forward); each thread turns its own row in place into dL_dsh and the
// CTA writes the 256 x 3M block back with coalesced 16-byte stores.
#include "ws_common.cuh"

namespace ws {

#ifndef WS_PB_THREADS
#define WS_PB_THREADS 128
#endif
constexpr int PB_THREADS = WS_PB_THREADS; // Gaussians (threads) per CTA; the SH block of a CTA is PB_THREADS x 12 M bytes of shared
// memory.  128: the load / compute / store phases of twice as many resident CTAs overlap (0.148 -> 0.142 ms at config 2; 64: no gain)

#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f
__constant__ float b_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__constant__ float b_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

struct PreBwdParams {
    int P, D, M, W, H;
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *shs, *scales, *rotations, *cov3D_precomp;
    const float *view, *proj, *campos;
    const int32_t *radii;
    const uint8_t *clamped;
    const float *acc;
    const GRec *grec;
    float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales,
        *dL_drotations;
    int sh_vec4;
    const float *shs_rest;   // features_rest [P,M-1,3] (shs = features_dc) or NULL
    float *dL_dsh_rest;      // its gradient
    int raw;                 // scales/rotations/opacities are raw parameters: chain through exp/normalize/sigmoid
};

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3(float x, float y, float z) { F3 r = {x, y, z}; return r; }
__device__ __forceinline__ float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// forward cov3D (plain arithmetic is enough here: gradient tier)
__device__ __forceinline__ void cov3d_plain(const float *s, const float *q, float *c)
{
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    // M[c][r] = s_r * R[c][r], R column-major as in DGR forward.cu:135-139
    const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
    const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
    const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
    const float M00 = s[0] * R00, M01 = s[1] * R01, M02 = s[2] * R02;
    const float M10 = s[0] * R10, M11 = s[1] * R11, M12 = s[2] * R12;
    const float M20 = s[0] * R20, M21 = s[1] * R21, M22 = s[2] * R22;
    c[0] = M00 * M00 + M01 * M01 + M02 * M02;
    c[1] = M00 * M10 + M01 * M11 + M02 * M12;
    c[2] = M00 * M20 + M01 * M21 + M02 * M22;
    c[3] = M10 * M10 + M11 * M11 + M12 * M12;
    c[4] = M10 * M20 + M11 * M21 + M12 * M22;
    c[5] = M20 * M20 + M21 * M21 + M22 * M22;
}

// DGR backward.cu:278-341
__device__ __forceinline__ void cov3d_backward(const float *s, const float *q, const float *dc, float *dscale,
                                               float *drot)
{
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                     {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                     {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float dS[3][3] = {{dc[0], 0.5f * dc[1], 0.5f * dc[2]},
                      {0.5f * dc[1], dc[3], 0.5f * dc[4]},
                      {0.5f * dc[2], 0.5f * dc[4], dc[5]}};
    // dL_dM = 2 M dL_dSigma, M[c][rr] = s[rr] R[c][rr]; dMt[c][rr] = dM[rr][c]
    float dMt[3][3];
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int rr = 0; rr < 3; rr++)
            dMt[rr][c] = 2.0f * s[rr] * (R[0][rr] * dS[c][0] + R[1][rr] * dS[c][1] + R[2][rr] * dS[c][2]);
    // Rt[k][j] = R[j][k]
#pragma unroll
    for (int k = 0; k < 3; k++)
        dscale[k] = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int j = 0; j < 3; j++) dMt[k][j] *= s[k];
    drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) -
              4 * x * (dMt[2][2] + dMt[1][1]);
    drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) -
              4 * y * (dMt[2][2] + dMt[0][0]);
    drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) -
              4 * z * (dMt[1][1] + dMt[0][0]);
}

// A/B knob -DWS_PB_MINCTAS=4 (64 registers, 120 B of spills, 4 CTAs/SM instead of 3): 12 us faster on the raw-parameter
// path, 6 us slower on the standard one -> left off.
#ifdef WS_PB_MINCTAS
#define WS_PB_BOUNDS __launch_bounds__(PB_THREADS, WS_PB_MINCTAS)
#else
#define WS_PB_BOUNDS __launch_bounds__(PB_THREADS)
#endif
__global__ void WS_PB_BOUNDS preprocess_bwd(const PreBwdParams p)
{
    extern __shared__ float4 s_sh4[];
    __shared__ float s_view[16], s_proj[16];
    const int tid = threadIdx.x;
    const int base = blockIdx.x * PB_THREADS;
    const int idx = base + tid;
    if (tid < 16) s_view[tid] = p.view[tid];
    else if (tid < 32) s_proj[tid - 16] = p.proj[tid - 16];

    const int nvalid = min(PB_THREADS, p.P - base);
    const bool use_sh = (p.shs != nullptr && p.M > 0);
    const int row_f = 3 * p.M;
    const int srow = p.sh_vec4 ? (row_f / 4 + 1) * 4 : (row_f | 1); // row stride in floats
    float *s_sh = reinterpret_cast<float *>(s_sh4);
    if (use_sh) {
        if (p.sh_vec4) {
            const int row4 = row_f / 4, srow4 = row4 + 1;
            const float4 *src = reinterpret_cast<const float4 *>(p.shs) + (size_t)base * row4;
            const int n4 = nvalid * row4;
            for (int i = tid; i < n4; i += PB_THREADS) {
                int g = i / row4, c = i - g * row4;
                s_sh4[g * srow4 + c] = __ldg(src + i);
            }
        } else if (p.shs_rest) {
            const int rest_f = row_f - 3;
            const float *src = p.shs + (size_t)base * 3;
            for (int i = tid; i < nvalid * 3; i += PB_THREADS) {
                int g = i / 3, c = i - g * 3;
                s_sh[g * srow + c] = __ldg(src + i);
            }
            src = p.shs_rest + (size_t)base * rest_f;
            const int n = nvalid * rest_f;
            for (int i = tid; i < n; i += PB_THREADS) {
                int g = i / rest_f, c = i - g * rest_f;
                s_sh[g * srow + 3 + c] = __ldg(src + i);
            }
        } else {
            const float *src = p.shs + (size_t)base * row_f;
            const int n = nvalid * row_f;
            for (int i = tid; i < n; i += PB_THREADS) {
                int g = i / row_f, c = i - g * row_f;
                s_sh[g * srow + c] = __ldg(src + i);
            }
        }
    }
    __syncthreads();

    if (idx < p.P) {
        const float *v = s_view, *proj = s_proj;
        float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
        float a[12];
        float *row = s_sh + tid * srow;
        const bool vis = p.radii[idx] > 0;
        if (vis) {
            const float4 *ap = reinterpret_cast<const float4 *>(p.acc + (size_t)idx * 12);
            const float4 a0 = ap[0], a1 = ap[1], a2 = ap[2];
            a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
            {
                const float4 ga = p.grec[idx].a, gb = p.grec[idx].b; // {x, y, A, B}, {C, opacity, depth, tau}
                const float k = -0.5f * gb.y;
                a[4] = k * (float)p.W * (ga.z * a1.x + ga.w * a1.y);
                a[5] = k * (float)p.H * (gb.x * a1.y + ga.w * a1.x);
                a[6] = k * a1.z; a[7] = k * a1.w; a[8] = k * a2.x;
                a[9] = a2.y; a[10] = 0.f; a[11] = 0.f;
            }
            const float mx = p.means3D[3 * (size_t)idx], my = p.means3D[3 * (size_t)idx + 1],
                        mz = p.means3D[3 * (size_t)idx + 2];
            float s[3] = {0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f}, c3[6];
            float qn = 1.0f, qn_raw = 1.0f; // raw parameters: max(|q|, eps) and |q|
            if (p.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c3[k] = p.cov3D_precomp[6 * (size_t)idx + k];
            } else {
                float4 qq = reinterpret_cast<const float4 *>(p.rotations)[idx];
                float s0 = p.scales[3 * (size_t)idx], s1 = p.scales[3 * (size_t)idx + 1], s2 = p.scales[3 * (size_t)idx + 2];
                if (p.raw) {
                    qn = raw_quat_norm(qq);
                    qn_raw = sqrtf(qq.x * qq.x + qq.y * qq.y + qq.z * qq.z + qq.w * qq.w);
                    raw_activate(qq, s0, s1, s2);
                }
                q[0] = qq.x; q[1] = qq.y; q[2] = qq.z; q[3] = qq.w;
                s[0] = p.scale_modifier * s0; s[1] = p.scale_modifier * s1; s[2] = p.scale_modifier * s2;
                cov3d_plain(s, q, c3);
            }
            // ---- computeCov2DCUDA, DGR backward.cu:144-274 ----
            {
                const float dcx = a[6], dcy = a[7], dcz = a[8];
                float tx = v[0] * mx + v[4] * my + v[8] * mz + v[12];
                float ty = v[1] * mx + v[5] * my + v[9] * mz + v[13];
                const float tz = v[2] * mx + v[6] * my + v[10] * mz + v[14];
                const float limx = 1.3f * p.tanfovx, limy = 1.3f * p.tanfovy;
                const float txtz = tx / tz, tytz = ty / tz;
                tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
                ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
                const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
                const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
                const float h_x = p.focal_x, h_y = p.focal_y;
                const float J00 = h_x / tz, J02 = -(h_x * tx) / (tz * tz);
                const float J11 = h_y / tz, J12 = -(h_y * ty) / (tz * tz);
                // T[0][r] = W[0][r] J00 + W[2][r] J02, T[1][r] = W[1][r] J11 + W[2][r] J12
                const float T00 = v[0] * J00 + v[2] * J02, T01 = v[4] * J00 + v[6] * J02, T02 = v[8] * J00 + v[10] * J02;
                const float T10 = v[1] * J11 + v[2] * J12, T11 = v[5] * J11 + v[6] * J12, T12 = v[9] * J11 + v[10] * J12;
                // VT0 = Vrk T0, VT1 = Vrk T1
                const float VT00 = c3[0] * T00 + c3[1] * T01 + c3[2] * T02;
                const float VT01 = c3[1] * T00 + c3[3] * T01 + c3[4] * T02;
                const float VT02 = c3[2] * T00 + c3[4] * T01 + c3[5] * T02;
                const float VT10 = c3[0] * T10 + c3[1] * T11 + c3[2] * T12;
                const float VT11 = c3[1] * T10 + c3[3] * T11 + c3[4] * T12;
                const float VT12 = c3[2] * T10 + c3[4] * T11 + c3[5] * T12;
                const float ca = T00 * VT00 + T01 * VT01 + T02 * VT02 + 0.3f;
                const float cb = T00 * VT10 + T01 * VT11 + T02 * VT12;
                const float cc = T10 * VT10 + T11 * VT11 + T12 * VT12 + 0.3f;
                const float denom = ca * cc - cb * cb;
                float dL_da = 0, dL_db = 0, dL_dc = 0;
                const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
                if (denom2inv != 0) {
                    dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
                    dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
                    dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
                    dcov[0] = (T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc);
                    dcov[3] = (T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc);
                    dcov[5] = (T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc);
                    dcov[1] = 2 * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + 2 * T10 * T11 * dL_dc;
                    dcov[2] = 2 * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + 2 * T10 * T12 * dL_dc;
                    dcov[4] = 2 * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + 2 * T11 * T12 * dL_dc;
                }
                const float dL_dT00 = 2 * VT00 * dL_da + VT10 * dL_db;
                const float dL_dT01 = 2 * VT01 * dL_da + VT11 * dL_db;
                const float dL_dT02 = 2 * VT02 * dL_da + VT12 * dL_db;
                const float dL_dT10 = 2 * VT10 * dL_dc + VT00 * dL_db;
                const float dL_dT11 = 2 * VT11 * dL_dc + VT01 * dL_db;
                const float dL_dT12 = 2 * VT12 * dL_dc + VT02 * dL_db;
                const float dL_dJ00 = v[0] * dL_dT00 + v[4] * dL_dT01 + v[8] * dL_dT02;
                const float dL_dJ02 = v[2] * dL_dT00 + v[6] * dL_dT01 + v[10] * dL_dT02;
                const float dL_dJ11 = v[1] * dL_dT10 + v[5] * dL_dT11 + v[9] * dL_dT12;
                const float dL_dJ12 = v[2] * dL_dT10 + v[6] * dL_dT11 + v[10] * dL_dT12;
                const float tzi = 1.f / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
                const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
                const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
                const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tx) * tz3 * dL_dJ02 +
                                     (2 * h_y * ty) * tz3 * dL_dJ12;
                // transformVec4x3Transpose (auxiliary.h:87-95); ASSIGNED (backward.cu:273)
                dmean[0] = v[0] * dL_dtx + v[1] * dL_dty + v[2] * dL_dtz;
                dmean[1] = v[4] * dL_dtx + v[5] * dL_dty + v[6] * dL_dtz;
                dmean[2] = v[8] * dL_dtx + v[9] * dL_dty + v[10] * dL_dtz;
            }
            // ---- preprocessCUDA backward, DGR backward.cu:372-403 ----
            {
                const float m_w = 1.0f / ((proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15]) + 0.0000001f);
                const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
                const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
                const float g2x = a[4], g2y = a[5];
                dmean[0] += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
                dmean[1] += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
                dmean[2] += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
                const float mul3 = v[2] * mx + v[6] * my + v[10] * mz + v[14];
                const float gd = a[3];
                dmean[0] += (v[2] - v[3] * mul3) * gd;
                dmean[1] += (v[6] - v[7] * mul3) * gd;
                dmean[2] += (v[10] - v[11] * mul3) * gd;
            }
            // ---- SH backward, DGR backward.cu:20-139; row becomes dL_dsh in place ----
            if (use_sh) {
                const float dox = mx - p.campos[0], doy = my - p.campos[1], doz = mz - p.campos[2];
                const float len = sqrtf(dox * dox + doy * doy + doz * doz);
                const float x = dox / len, y = doy / len, z = doz / len;
                const uint8_t cl = p.clamped[idx];
                const F3 dRGB = f3(a[0] * ((cl & 1) ? 0.f : 1.f), a[1] * ((cl & 2) ? 0.f : 1.f),
                                   a[2] * ((cl & 4) ? 0.f : 1.f));
                F3 gx = f3(0, 0, 0), gy = f3(0, 0, 0), gz = f3(0, 0, 0); // dRGBdx, dRGBdy, dRGBdz
#define SHR(k) f3(row[3 * (k)], row[3 * (k) + 1], row[3 * (k) + 2])
#define SHW(k, wgt)                                                                            \
    {                                                                                          \
        const float _w = (wgt);                                                                \
        row[3 * (k)] = _w * dRGB.x; row[3 * (k) + 1] = _w * dRGB.y; row[3 * (k) + 2] = _w * dRGB.z; \
    }
#define AXPY(dst, s, vv) { const float _s = (s); dst.x += _s * vv.x; dst.y += _s * vv.y; dst.z += _s * vv.z; }
                const int deg = p.D;
                SHW(0, SH_C0);
                int written = 1;
                if (deg > 0) {
                    const F3 s1 = SHR(1), s2 = SHR(2), s3 = SHR(3);
                    SHW(1, -SH_C1 * y); SHW(2, SH_C1 * z); SHW(3, -SH_C1 * x);
                    AXPY(gx, -SH_C1, s3); AXPY(gy, -SH_C1, s1); AXPY(gz, SH_C1, s2);
                    written = 4;
                    if (deg > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        const F3 s4 = SHR(4), s5 = SHR(5), s6 = SHR(6), s7 = SHR(7), s8 = SHR(8);
                        SHW(4, b_SH_C2[0] * xy); SHW(5, b_SH_C2[1] * yz);
                        SHW(6, b_SH_C2[2] * (2.f * zz - xx - yy));
                        SHW(7, b_SH_C2[3] * xz); SHW(8, b_SH_C2[4] * (xx - yy));
                        AXPY(gx, b_SH_C2[0] * y, s4); AXPY(gx, b_SH_C2[2] * 2.f * -x, s6);
                        AXPY(gx, b_SH_C2[3] * z, s7); AXPY(gx, b_SH_C2[4] * 2.f * x, s8);
                        AXPY(gy, b_SH_C2[0] * x, s4); AXPY(gy, b_SH_C2[1] * z, s5);
                        AXPY(gy, b_SH_C2[2] * 2.f * -y, s6); AXPY(gy, b_SH_C2[4] * 2.f * -y, s8);
                        AXPY(gz, b_SH_C2[1] * y, s5); AXPY(gz, b_SH_C2[2] * 2.f * 2.f * z, s6);
                        AXPY(gz, b_SH_C2[3] * x, s7);
                        written = 9;
                        if (deg > 2) {
                            const F3 s9 = SHR(9), s10 = SHR(10), s11 = SHR(11), s12 = SHR(12), s13 = SHR(13),
                                     s14 = SHR(14), s15 = SHR(15);
                            SHW(9, b_SH_C3[0] * y * (3.f * xx - yy)); SHW(10, b_SH_C3[1] * xy * z);
                            SHW(11, b_SH_C3[2] * y * (4.f * zz - xx - yy));
                            SHW(12, b_SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                            SHW(13, b_SH_C3[4] * x * (4.f * zz - xx - yy));
                            SHW(14, b_SH_C3[5] * z * (xx - yy)); SHW(15, b_SH_C3[6] * x * (xx - 3.f * yy));
                            AXPY(gx, b_SH_C3[0] * 3.f * 2.f * xy, s9); AXPY(gx, b_SH_C3[1] * yz, s10);
                            AXPY(gx, b_SH_C3[2] * -2.f * xy, s11); AXPY(gx, b_SH_C3[3] * -3.f * 2.f * xz, s12);
                            AXPY(gx, b_SH_C3[4] * (-3.f * xx + 4.f * zz - yy), s13);
                            AXPY(gx, b_SH_C3[5] * 2.f * xz, s14); AXPY(gx, b_SH_C3[6] * 3.f * (xx - yy), s15);
                            AXPY(gy, b_SH_C3[0] * 3.f * (xx - yy), s9); AXPY(gy, b_SH_C3[1] * xz, s10);
                            AXPY(gy, b_SH_C3[2] * (-3.f * yy + 4.f * zz - xx), s11);
                            AXPY(gy, b_SH_C3[3] * -3.f * 2.f * yz, s12); AXPY(gy, b_SH_C3[4] * -2.f * xy, s13);
                            AXPY(gy, b_SH_C3[5] * -2.f * yz, s14); AXPY(gy, b_SH_C3[6] * -3.f * 2.f * xy, s15);
                            AXPY(gz, b_SH_C3[1] * xy, s10); AXPY(gz, b_SH_C3[2] * 4.f * 2.f * yz, s11);
                            AXPY(gz, b_SH_C3[3] * 3.f * (2.f * zz - xx - yy), s12);
                            AXPY(gz, b_SH_C3[4] * 4.f * 2.f * xz, s13); AXPY(gz, b_SH_C3[5] * (xx - yy), s14);
                            written = 16;
                        }
                    }
                }
#undef SHR
#undef SHW
#undef AXPY
                for (int k = 3 * written; k < row_f; k++) row[k] = 0.f; // bands above the active degree
                const float ddx = dot(gx, dRGB), ddy = dot(gy, dRGB), ddz = dot(gz, dRGB);
                // dnormvdv, auxiliary.h:107-117
                const float sum2 = dox * dox + doy * doy + doz * doz;
                const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
                dmean[0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
                dmean[1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
                dmean[2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
            }
            if (!p.cov3D_precomp) cov3d_backward(s, q, dcov, dscale, drot);
            if (p.raw) {
                // exp: d/dx = exp(x) = the activated scale; sigmoid: o (1 - o); normalize: (g - q^ (q^ . g)) / |q|
                // (below the clamp the norm is the constant eps: g / eps, as torch's clamp_min backward)
                if (!p.cov3D_precomp) {
                    const float inv_mod = p.scale_modifier != 0.f ? 1.0f / p.scale_modifier : 0.f;
#pragma unroll
                    for (int k = 0; k < 3; k++) dscale[k] *= s[k] * inv_mod;
                    const float qg = qn_raw < 1e-12f ? 0.f : q[0] * drot[0] + q[1] * drot[1] + q[2] * drot[2] + q[3] * drot[3];
#pragma unroll
                    for (int k = 0; k < 4; k++) drot[k] = (drot[k] - q[k] * qg) / qn;
                }
                const float o = p.grec[idx].b.y;
                a[9] *= o * (1.0f - o);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 12; k++) a[k] = 0.f;
            if (use_sh)
                for (int k = 0; k < row_f; k++) row[k] = 0.f;
        }
        const size_t i3 = 3 * (size_t)idx;
        p.dL_dmeans3D[i3] = dmean[0]; p.dL_dmeans3D[i3 + 1] = dmean[1]; p.dL_dmeans3D[i3 + 2] = dmean[2];
        p.dL_dmeans2D[i3] = a[4]; p.dL_dmeans2D[i3 + 1] = a[5]; p.dL_dmeans2D[i3 + 2] = 0.f;
        p.dL_dcolors[i3] = a[0]; p.dL_dcolors[i3 + 1] = a[1]; p.dL_dcolors[i3 + 2] = a[2];
        p.dL_dopacity[idx] = a[9];
#pragma unroll
        for (int k = 0; k < 6; k++) p.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
        if (p.dL_dscales) {
            p.dL_dscales[i3] = dscale[0]; p.dL_dscales[i3 + 1] = dscale[1]; p.dL_dscales[i3 + 2] = dscale[2];
        }
        if (p.dL_drotations) { // scalar stores: the tensor may be a 4-byte aligned slice of a gradient bucket
            const size_t i4 = 4 * (size_t)idx;
            p.dL_drotations[i4] = drot[0]; p.dL_drotations[i4 + 1] = drot[1];
            p.dL_drotations[i4 + 2] = drot[2]; p.dL_drotations[i4 + 3] = drot[3];
        }
    }
    if (use_sh && p.dL_dsh) {
        __syncthreads();
        if (p.sh_vec4) {
            const int row4 = row_f / 4, srow4 = row4 + 1;
            float4 *dst = reinterpret_cast<float4 *>(p.dL_dsh) + (size_t)base * row4;
            const int n4 = nvalid * row4;
            for (int i = tid; i < n4; i += PB_THREADS) {
                int g = i / row4, c = i - g * row4;
                dst[i] = s_sh4[g * srow4 + c];
            }
        } else if (p.shs_rest) {
            const int rest_f = row_f - 3;
            float *dst = p.dL_dsh + (size_t)base * 3;
            for (int i = tid; i < nvalid * 3; i += PB_THREADS) {
                int g = i / 3, c = i - g * 3;
                dst[i] = s_sh[g * srow + c];
            }
            dst = p.dL_dsh_rest + (size_t)base * rest_f;
            const int n = nvalid * rest_f;
            for (int i = tid; i < n; i += PB_THREADS) {
                int g = i / rest_f, c = i - g * rest_f;
                dst[i] = s_sh[g * srow + 3 + c];
            }
        } else {
            float *dst = p.dL_dsh + (size_t)base * row_f;
            const int n = nvalid * row_f;
            for (int i = tid; i < n; i += PB_THREADS) {
                int g = i / row_f, c = i - g * row_f;
                dst[i] = s_sh[g * srow + c];
            }
        }
    }
}

int launch_preprocess_bwd(const ws_view *v, const ws_gaussians *g, const int32_t *radii, GeomState &gs,
                          const ws_grads *gr, cudaStream_t st)
{
    PreBwdParams p;
    p.P = g->P; p.D = v->sh_degree; p.M = g->M; p.W = v->image_width; p.H = v->image_height;
    p.tanfovx = v->tanfovx; p.tanfovy = v->tanfovy;
    p.focal_y = p.H / (2.0f * v->tanfovy);
    p.focal_x = p.W / (2.0f * v->tanfovx);
    p.scale_modifier = v->scale_modifier;
    p.means3D = g->means3D; p.shs = g->shs; p.scales = g->scales; p.rotations = g->rotations;
    p.cov3D_precomp = g->cov3D_precomp;
    p.view = v->viewmatrix; p.proj = v->projmatrix; p.campos = v->campos;
    p.radii = radii; p.clamped = gs.clamped; p.acc = gs.acc; p.grec = gs.grec;
    p.dL_dmeans2D = gr->dL_dmeans2D; p.dL_dcolors = gr->dL_dcolors; p.dL_dopacity = gr->dL_dopacity;
    p.dL_dmeans3D = gr->dL_dmeans3D; p.dL_dcov3D = gr->dL_dcov3D; p.dL_dsh = gr->dL_dsh;
    p.dL_dscales = gr->dL_dscales; p.dL_drotations = gr->dL_drotations;
    const int row_f = 3 * g->M;
    const bool use_sh = g->shs != nullptr && g->M > 0;
    p.shs_rest = use_sh ? g->shs_rest : nullptr; p.dL_dsh_rest = gr->dL_dsh_rest; p.raw = g->raw_params;
    if (p.shs_rest && gr->dL_dsh && !gr->dL_dsh_rest) {
        set_error("preprocess backward: shs_rest given without dL_dsh_rest");
        return WS_E_INVALID;
    }
    p.sh_vec4 = (use_sh && !p.shs_rest && row_f % 4 == 0 && ((uintptr_t)g->shs % 16) == 0 &&
                 (!gr->dL_dsh || ((uintptr_t)gr->dL_dsh % 16) == 0)) ? 1 : 0;
    size_t smem = 0;
    if (use_sh)
        smem = p.sh_vec4 ? (size_t)PB_THREADS * (row_f / 4 + 1) * 16 : (size_t)PB_THREADS * (row_f | 1) * 4;
    if (smem > 200 * 1024) {
        set_error("preprocess backward: %d SH coefficients need %zu bytes of shared memory", g->M, smem);
        return WS_E_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
        WS_CUDA_CHECK(cudaFuncSetAttribute(preprocess_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int nb = (g->P + PB_THREADS - 1) / PB_THREADS;
    preprocess_bwd<<<nb, PB_THREADS, smem, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
