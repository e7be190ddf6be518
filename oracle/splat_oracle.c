/*
 * splat_oracle.c -- CPU restatement of the Wheat3DGS Gaussian-splatting hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (wheat_3dgs_b200/, the
 * drop-in packages, libwheatsplat.so) may import, link or call this file.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * leg use it, and there only as the checker / the timed CPU baseline.
 *
 * What it restates (file:line are into /root/reference/submodules/):
 *   DGR = diff-gaussian-rasterization, FSR = flashsplat-rasterization,
 *   KNN = simple-knn.
 *
 * Parity pin: the reference has no tests or golden vectors of its own
 * (SURVEY.md section 4).  This oracle is pinned against outputs of the
 * reference CUDA rasterizer itself, built for sm_100 from the sources in place
 * (oracle/build_ref.py -> oracle/_ref/) and run on a B200; the resulting
 * fixtures live under tests/golden/ (see tests/golden/README.md).
 *
 * Arithmetic notes.  The bit-exact tier (depth bits in the tile keys, radii,
 * tile rectangles) depends on how nvcc contracts a*b+c*d+... into FMAs in the
 * reference build.  The contraction was read from the reference's sm_100 SASS
 * (nvcc 12.9, default -fmad=true) and is spelled out here with explicit
 * fmaf(); this file must be compiled with -ffp-contract=off so that gcc adds
 * none of its own.  1.0f/x, a/b and sqrtf are IEEE-correct in the reference
 * build (no -use_fast_math, DGR/setup.py:29), as they are here.  expf differs
 * from CUDA's by <= 2 ulp; it only reaches the tolerance tier (color, depth,
 * alpha, gradients).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define BLOCK_X 16 /* DGR/cuda_rasterizer/config.h:16 */
#define BLOCK_Y 16 /* DGR/cuda_rasterizer/config.h:17 */
#define BLOCK_SIZE (BLOCK_X * BLOCK_Y)

/* DGR/cuda_rasterizer/auxiliary.h:22-39 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

typedef struct {
    int P, D, M;        /* #Gaussians, active SH degree, #SH coefficients per Gaussian */
    int W, H;
    float tanfovx, tanfovy, scale_modifier;
    const float *bg;         /* [3] */
    const float *view;       /* [16] column-major use: m[0],m[4],m[8],m[12] (auxiliary.h:58-66) */
    const float *proj;       /* [16] */
    const float *campos;     /* [3] */
    const float *means3D;    /* [P,3] */
    const float *shs;        /* [P,M,3] or NULL */
    const float *colors_precomp; /* [P,3] or NULL */
    const float *opacities;  /* [P] */
    const float *scales;     /* [P,3] or NULL */
    const float *rotations;  /* [P,4] or NULL */
    const float *cov3D_precomp; /* [P,6] or NULL */
} so_scene;

/* a*b + c*d + e*f + g  as nvcc contracts it in the reference build:
 * fadd(fma(e,f, fma(a,b, fmul(c,d))), g)  (read from SASS of preprocessCUDA) */
static inline float dot3_nv(float a, float b, float c, float d, float e, float f)
{
    return fmaf(e, f, fmaf(a, b, c * d));
}

/* DGR/cuda_rasterizer/auxiliary.h:41-44 : evaluated in double, then rounded;
 * nvcc fuses (v+1)*S-1 into one DFMA. */
static inline float ndc2pix(float v, int S)
{
    double t = (double)v + 1.0;
    t = fma(t, (double)S, -1.0);
    return (float)(t * 0.5);
}

/* float -> int conversion with CUDA semantics (F2I.TRUNC: saturate, NaN -> 0) */
static inline int f2i_trunc(float f)
{
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int)f;
}

/* DGR/cuda_rasterizer/auxiliary.h:46-56 */
static inline void get_rect(float px, float py, int max_radius, int gx, int gy, uint32_t *rmin,
                            uint32_t *rmax)
{
    float r = (float)max_radius;
    int x0 = f2i_trunc((px - r) * 0.0625f);
    int y0 = f2i_trunc((py - r) * 0.0625f);
    int x1 = f2i_trunc((((px + r) + 16.0f) - 1.0f) * 0.0625f);
    int y1 = f2i_trunc((((py + r) + 16.0f) - 1.0f) * 0.0625f);
    if (x0 < 0) x0 = 0;
    if (y0 < 0) y0 = 0;
    if (x1 < 0) x1 = 0;
    if (y1 < 0) y1 = 0;
    rmin[0] = (uint32_t)x0 < (uint32_t)gx ? (uint32_t)x0 : (uint32_t)gx;
    rmin[1] = (uint32_t)y0 < (uint32_t)gy ? (uint32_t)y0 : (uint32_t)gy;
    rmax[0] = (uint32_t)x1 < (uint32_t)gx ? (uint32_t)x1 : (uint32_t)gx;
    rmax[1] = (uint32_t)y1 < (uint32_t)gy ? (uint32_t)y1 : (uint32_t)gy;
}

/* DGR/cuda_rasterizer/forward.cu:118-152 (quaternion NOT normalised, :127) */
static void compute_cov3d(const float *scale, float mod, const float *rot, float *cov3D)
{
    float sx = mod * scale[0], sy = mod * scale[1], sz = mod * scale[2];
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    float yy = y * y, zz = z * z;
    /* rotation entries, rounding order read from SASS */
    float s_yz = yy + zz;
    float r00 = 1.0f - (s_yz + s_yz);
    float s_xz = fmaf(x, x, zz);
    float r11 = 1.0f - (s_xz + s_xz);
    float s_xy = fmaf(x, x, yy);
    float r22 = 1.0f - (s_xy + s_xy);
    float rz = r * z, rx = r * x, xz = x * z;
    float a;
    a = fmaf(x, y, -rz); float xy_m_rz = a + a;   /* 2(xy - rz) */
    a = fmaf(x, y, rz);  float xy_p_rz = a + a;   /* 2(xy + rz) */
    a = fmaf(r, y, xz);  float xz_p_ry = a + a;   /* 2(xz + ry) */
    a = fmaf(-r, y, xz); float xz_m_ry = a + a;   /* 2(xz - ry) */
    a = fmaf(y, z, -rx); float yz_m_rx = a + a;   /* 2(yz - rx) */
    a = fmaf(y, z, rx);  float yz_p_rx = a + a;   /* 2(yz + rx) */
    /* M = S * R, glm column-major: M[c][r] = s_r * R[c][r] */
    float M00 = sx * r00, M01 = sy * xy_m_rz, M02 = sz * xz_p_ry;
    float M10 = sx * xy_p_rz, M11 = sy * r11, M12 = sz * yz_m_rx;
    float M20 = sx * xz_m_ry, M21 = sy * yz_p_rx, M22 = sz * r22;
    /* Sigma = M^T M : Sigma[c][r] = M[r][0]M[c][0] + M[r][1]M[c][1] + M[r][2]M[c][2] */
    cov3D[0] = dot3_nv(M00, M00, M01, M01, M02, M02);
    cov3D[1] = dot3_nv(M00, M10, M01, M11, M02, M12);
    cov3D[2] = dot3_nv(M00, M20, M01, M21, M02, M22);
    cov3D[3] = dot3_nv(M10, M10, M11, M11, M12, M12);
    cov3D[4] = dot3_nv(M10, M20, M11, M21, M12, M22);
    cov3D[5] = dot3_nv(M20, M20, M21, M21, M22, M22);
}

/* DGR/cuda_rasterizer/forward.cu:74-113.  Returns T (2x3 used part) too, for the backward. */
static void compute_cov2d(const float *mean, float focal_x, float focal_y, float tan_fovx,
                          float tan_fovy, const float *c, const float *v, float *cov /*[3]*/)
{
    float x = mean[0], y = mean[1], z = mean[2];
    float tx = dot3_nv(x, v[0], y, v[4], z, v[8]) + v[12];
    float ty = dot3_nv(x, v[1], y, v[5], z, v[9]) + v[13];
    float tz = dot3_nv(x, v[2], y, v[6], z, v[10]) + v[14];
    float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    float txtz = tx / tz, tytz = ty / tz;
    float cx = fminf(limx, fmaxf(-limx, txtz));
    float cy = fminf(limy, fmaxf(-limy, tytz));
    tx = cx * tz;
    ty = cy * tz;
    float tz2 = tz * tz;
    float J00 = focal_x / tz;
    float J02 = ((-tx) * focal_x) / tz2;
    float J11 = focal_y / tz;
    float J12 = ((-ty) * focal_y) / tz2;
    /* T = W * J, T[c][r]; W[k][r] = view[4r + k] */
    float T00 = fmaf(v[2], J02, v[0] * J00);
    float T01 = fmaf(v[6], J02, v[4] * J00);
    float T02 = fmaf(v[10], J02, v[8] * J00);
    float T10 = fmaf(v[2], J12, v[1] * J11);
    float T11 = fmaf(v[6], J12, v[5] * J11);
    float T12 = fmaf(v[10], J12, v[9] * J11);
    /* A = T^T Vrk^T : A[c][r] = T[r][0]Vrk[0][c] + T[r][1]Vrk[1][c] + T[r][2]Vrk[2][c] */
    float A00 = dot3_nv(T00, c[0], T01, c[1], T02, c[2]);
    float A10 = dot3_nv(T00, c[1], T01, c[3], T02, c[4]);
    float A20 = dot3_nv(T00, c[2], T01, c[4], T02, c[5]);
    float A01 = dot3_nv(T10, c[0], T11, c[1], T12, c[2]);
    float A11 = dot3_nv(T10, c[1], T11, c[3], T12, c[4]);
    float A21 = dot3_nv(T10, c[2], T11, c[4], T12, c[5]);
    /* cov[c][r] = A[0][r]T[c][0] + A[1][r]T[c][1] + A[2][r]T[c][2] */
    cov[0] = dot3_nv(T00, A00, T01, A10, T02, A20) + 0.3f;
    cov[1] = dot3_nv(T00, A01, T01, A11, T02, A21);
    cov[2] = dot3_nv(T10, A01, T11, A11, T12, A21) + 0.3f;
}

/* DGR/cuda_rasterizer/forward.cu:20-71 */
static void color_from_sh(int idx, int deg, int M, const float *means, const float *campos,
                          const float *shs, uint8_t *clamped, float *rgb)
{
    float dx = means[3 * idx] - campos[0], dy = means[3 * idx + 1] - campos[1],
          dz = means[3 * idx + 2] - campos[2];
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float x = dx / len, y = dy / len, z = dz / len;
    const float *sh = shs + (size_t)idx * M * 3;
    for (int ch = 0; ch < 3; ch++) {
#define S(k) sh[3 * (k) + ch]
        float res = SH_C0 * S(0);
        if (deg > 0) {
            res = res - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) +
                      SH_C2[2] * (2.0f * zz - xx - yy) * S(6) + SH_C2[3] * xz * S(7) +
                      SH_C2[4] * (xx - yy) * S(8);
                if (deg > 2) {
                    res = res + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                          SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                          SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                          SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) +
                          SH_C3[5] * z * (xx - yy) * S(14) + SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        res += 0.5f;
        clamped[3 * idx + ch] = (res < 0);
        rgb[3 * idx + ch] = res < 0.0f ? 0.0f : res;
    }
}

/* DGR/cuda_rasterizer/forward.cu:155-256 (preprocessCUDA) with in_frustum (auxiliary.h:139-164).
 * FSR additions (FSR/cuda_rasterizer/forward.cu:237-238,258): image_xys[2,P] and gs_depth[P]
 * when non-NULL (caller pre-fills them with -1.0 / -0.1, FSR/rasterize_points.cu:77-78). */
void so_preprocess(const so_scene *s, int *radii, float *means2D, float *depths, float *cov3Ds,
                   float *rgb, float *conic_opacity, uint8_t *clamped, uint32_t *tiles_touched,
                   float *image_xys, float *gs_depth)
{
    const int P = s->P;
    const float focal_y = s->H / (2.0f * s->tanfovy); /* rasterizer_impl.cu:224-225 */
    const float focal_x = s->W / (2.0f * s->tanfovx);
    const int gx = (s->W + BLOCK_X - 1) / BLOCK_X, gy = (s->H + BLOCK_Y - 1) / BLOCK_Y;
    const float *v = s->view, *pm = s->proj;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        float x = s->means3D[3 * idx], y = s->means3D[3 * idx + 1], z = s->means3D[3 * idx + 2];
        float pvz = dot3_nv(x, v[2], y, v[6], z, v[10]) + v[14];
        if (pvz <= 0.2f) continue; /* auxiliary.h:154 (NaN is NOT culled, as in the reference) */
        float hx = dot3_nv(x, pm[0], y, pm[4], z, pm[8]) + pm[12];
        float hy = dot3_nv(x, pm[1], y, pm[5], z, pm[9]) + pm[13];
        float hw = dot3_nv(x, pm[3], y, pm[7], z, pm[11]) + pm[15];
        float p_w = 1.0f / (hw + 0.0000001f);
        float projx = hx * p_w, projy = hy * p_w;
        const float *cov3D;
        if (s->cov3D_precomp) {
            cov3D = s->cov3D_precomp + 6 * (size_t)idx;
        } else {
            compute_cov3d(s->scales + 3 * (size_t)idx, s->scale_modifier,
                          s->rotations + 4 * (size_t)idx, cov3Ds + 6 * (size_t)idx);
            cov3D = cov3Ds + 6 * (size_t)idx;
        }
        float cov[3];
        compute_cov2d(s->means3D + 3 * (size_t)idx, focal_x, focal_y, s->tanfovx, s->tanfovy, cov3D,
                      v, cov);
        float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
        if (det == 0.0f) continue;
        float det_inv = 1.0f / det;
        float conic[3] = {cov[2] * det_inv, cov[1] * (-det_inv), cov[0] * det_inv};
        float mid = 0.5f * (cov[0] + cov[2]);
        float sq = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
        float lambda1 = mid + sq, lambda2 = mid - sq;
        float my_radius = ceilf(3.0f * sqrtf(fmaxf(lambda1, lambda2)));
        float pix = ndc2pix(projx, s->W), piy = ndc2pix(projy, s->H);
        uint32_t rmin[2], rmax[2];
        get_rect(pix, piy, f2i_trunc(my_radius), gx, gy, rmin, rmax);
        if (image_xys) { /* written before the zero-tile cull, FSR forward.cu:237-238 */
            image_xys[idx] = pix;
            image_xys[P + idx] = piy;
        }
        if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
        if (s->colors_precomp == NULL)
            color_from_sh(idx, s->D, s->M, s->means3D, s->campos, s->shs, clamped, rgb);
        depths[idx] = pvz;
        radii[idx] = f2i_trunc(my_radius);
        means2D[2 * idx] = pix;
        means2D[2 * idx + 1] = piy;
        conic_opacity[4 * idx + 0] = conic[0];
        conic_opacity[4 * idx + 1] = conic[1];
        conic_opacity[4 * idx + 2] = conic[2];
        conic_opacity[4 * idx + 3] = s->opacities[idx];
        if (gs_depth) gs_depth[idx] = pvz;
        tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
    }
}

/* DGR/cuda_rasterizer/rasterizer_impl.cu:54-66 (checkFrustum / markVisible) */
void so_mark_visible(int P, const float *means3D, const float *v, uint8_t *present)
{
    for (int idx = 0; idx < P; idx++) {
        float x = means3D[3 * idx], y = means3D[3 * idx + 1], z = means3D[3 * idx + 2];
        float pvz = dot3_nv(x, v[2], y, v[6], z, v[10]) + v[14];
        present[idx] = !(pvz <= 0.2f);
    }
}

/* cub::DeviceScan::InclusiveSum, rasterizer_impl.cu:278.  Returns num_rendered (:282). */
uint32_t so_inclusive_scan(int P, const uint32_t *tiles_touched, uint32_t *point_offsets)
{
    uint32_t acc = 0;
    for (int i = 0; i < P; i++) {
        acc += tiles_touched[i];
        point_offsets[i] = acc;
    }
    return acc;
}

/* DGR/cuda_rasterizer/rasterizer_impl.cu:70-111 (duplicateWithKeys) */
void so_duplicate_with_keys(int P, int W, int H, const float *means2D, const float *depths,
                            const uint32_t *offsets, const int *radii, uint64_t *keys,
                            uint32_t *values)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < P; idx++) {
        if (radii[idx] > 0) {
            uint32_t off = (idx == 0) ? 0 : offsets[idx - 1];
            uint32_t rmin[2], rmax[2];
            get_rect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gx, gy, rmin, rmax);
            uint32_t dbits;
            memcpy(&dbits, &depths[idx], 4);
            for (uint32_t y = rmin[1]; y < rmax[1]; y++)
                for (uint32_t x = rmin[0]; x < rmax[0]; x++) {
                    uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
                    key <<= 32;
                    key |= dbits;
                    keys[off] = key;
                    values[off] = (uint32_t)idx;
                    off++;
                }
        }
    }
}

/* rasterizer_impl.cu:35-50 */
uint32_t so_get_higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* cub::DeviceRadixSort::SortPairs(keys, vals, n, 0, end_bit), rasterizer_impl.cu:304-309:
 * a stable LSD radix sort on bits [0, end_bit).  Bits >= end_bit are ignored by the
 * comparison but carried along with the key. */
void so_sort_pairs(uint64_t n, const uint64_t *keys_in, uint64_t *keys_out, const uint32_t *vals_in,
                   uint32_t *vals_out, int end_bit)
{
    uint64_t *ka = (uint64_t *)malloc(n * 8 + 8), *kb = (uint64_t *)malloc(n * 8 + 8);
    uint32_t *va = (uint32_t *)malloc(n * 4 + 4), *vb = (uint32_t *)malloc(n * 4 + 4);
    memcpy(ka, keys_in, n * 8);
    memcpy(va, vals_in, n * 4);
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << bits) - 1u;
        uint64_t hist[257];
        memset(hist, 0, sizeof(hist));
        for (uint64_t i = 0; i < n; i++) hist[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; d++) hist[d + 1] += hist[d];
        for (uint64_t i = 0; i < n; i++) {
            uint64_t dst = hist[(ka[i] >> shift) & mask]++;
            kb[dst] = ka[i];
            vb[dst] = va[i];
        }
        uint64_t *tk = ka; ka = kb; kb = tk;
        uint32_t *tv = va; va = vb; vb = tv;
    }
    memcpy(keys_out, ka, n * 8);
    memcpy(vals_out, va, n * 4);
    free(ka); free(kb); free(va); free(vb);
}

/* rasterizer_impl.cu:311 (memset) + :116-138 (identifyTileRanges).  ranges is uint2[T]. */
void so_identify_tile_ranges(uint64_t L, const uint64_t *keys, int T, uint32_t *ranges)
{
    memset(ranges, 0, (size_t)T * 8);
    for (uint64_t idx = 0; idx < L; idx++) {
        uint32_t currtile = (uint32_t)(keys[idx] >> 32);
        if (idx == 0)
            ranges[2 * currtile] = 0;
        else {
            uint32_t prevtile = (uint32_t)(keys[idx - 1] >> 32);
            if (currtile != prevtile) {
                ranges[2 * prevtile + 1] = (uint32_t)idx;
                ranges[2 * currtile] = (uint32_t)idx;
            }
        }
        if (idx == L - 1) ranges[2 * currtile + 1] = (uint32_t)L;
    }
}

/* DGR/cuda_rasterizer/forward.cu:261-381 (renderCUDA) and the FSR variant
 * FSR/cuda_rasterizer/forward.cu:269-418: gt_mask/unique_label/used_count/contrib_num are the
 * flashsplat additions and may be NULL (then this is exactly the DGR kernel).
 * used_count is accumulated in double (used_count64, [num_obj+1, P]) so that the oracle's
 * value is independent of the reference's non-deterministic atomic order. */
void so_render_forward(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                       const float *means2D, const float *features, const float *depths,
                       const float *conic_opacity, const float *bg, float *out_color,
                       float *out_depth, float *out_alpha, uint32_t *n_contrib,
                       const float *gt_mask, const int *unique_label, double *used_count64, int P,
                       int *out_contrib_num)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
            for (int py = ty * BLOCK_Y; py < ty * BLOCK_Y + BLOCK_Y && py < H; py++)
                for (int px = tx * BLOCK_X; px < tx * BLOCK_X + BLOCK_X && px < W; px++) {
                    const uint32_t pix_id = (uint32_t)W * py + px;
                    const float pixfx = (float)px, pixfy = (float)py;
                    int obj_id = 0;
                    if (gt_mask) obj_id = f2i_trunc(gt_mask[pix_id]); /* FSR forward.cu:306-311 */
                    float T = 1.0f, C[3] = {0, 0, 0}, weight = 0, D = 0;
                    uint32_t contributor = 0, last_contributor = 0;
                    for (uint32_t k = r0; k < r1; k++) {
                        contributor++;
                        const uint32_t g = point_list[k];
                        if (unique_label) { /* FSR forward.cu:357-362 */
                            int uid = unique_label[g];
                            if (uid >= 0 && obj_id != uid) continue;
                        }
                        float dx = means2D[2 * g] - pixfx, dy = means2D[2 * g + 1] - pixfy;
                        const float *co = conic_opacity + 4 * (size_t)g;
                        /* forward.cu:336, contraction as in the reference SASS */
                        float power = fmaf(fmaf(dx, dx * co[0], dy * (dy * co[2])), -0.5f,
                                           -(dy * (dx * co[1])));
                        if (power > 0.0f) continue;
                        float alpha = fminf(0.99f, co[3] * expf(power));
                        if (alpha < 1.0f / 255.0f) continue;
                        float test_T = T * (1 - alpha);
                        if (test_T < 0.0001f) break; /* done = true */
                        for (int ch = 0; ch < 3; ch++)
                            C[ch] = fmaf(T, alpha * features[3 * (size_t)g + ch], C[ch]);
                        weight = fmaf(T, alpha, weight);
                        D = fmaf(T, alpha * depths[g], D);
                        if (used_count64) { /* FSR forward.cu:398 */
#pragma omp atomic
                            used_count64[(size_t)obj_id * P + g] += (double)(alpha * T);
                        }
                        T = test_T;
                        last_contributor = contributor;
                    }
                    n_contrib[pix_id] = last_contributor;
                    for (int ch = 0; ch < 3; ch++)
                        out_color[(size_t)ch * H * W + pix_id] = fmaf(bg[ch], T, C[ch]);
                    out_alpha[pix_id] = weight;
                    out_depth[pix_id] = D;
                    if (out_contrib_num) out_contrib_num[pix_id] = (int)last_contributor;
                }
        }
}

/* DGR/cuda_rasterizer/backward.cu:415-601 (renderCUDA backward).  Per-pair values are
 * computed in float exactly as the reference does; the per-Gaussian sums (the reference's
 * atomicAdd targets) are accumulated in double so the oracle does not depend on atomic order.
 * Outputs (double, zero-initialised by the caller): dL_dmean2D[P,2], dL_dconic[P,3] (x,y,w),
 * dL_dopacity[P], dL_dcolors[P,3], dL_ddepths[P].
 * FSR variant (FSR backward.cu:455-462,529-536): gt_mask / unique_label gate, may be NULL. */
void so_render_backward(int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                        const float *bg, const float *means2D, const float *conic_opacity,
                        const float *colors, const float *depths, const float *alphas,
                        const uint32_t *n_contrib, const float *dL_dpixels,
                        const float *dL_dpixel_depths, const float *dL_dalphas, double *dL_dmean2D,
                        double *dL_dconic, double *dL_dopacity, double *dL_dcolors,
                        double *dL_ddepths, const float *gt_mask, const int *unique_label)
{
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
            for (int py = ty * BLOCK_Y; py < ty * BLOCK_Y + BLOCK_Y && py < H; py++)
                for (int px = tx * BLOCK_X; px < tx * BLOCK_X + BLOCK_X && px < W; px++) {
                    const uint32_t pix_id = (uint32_t)W * py + px;
                    const float pixfx = (float)px, pixfy = (float)py;
                    int obj_id = 0;
                    if (gt_mask) obj_id = f2i_trunc(gt_mask[pix_id]);
                    const float T_final = 1 - alphas[pix_id]; /* backward.cu:463 */
                    float T = T_final;
                    const uint32_t last_contributor = n_contrib[pix_id];
                    float accum_rec[3] = {0, 0, 0}, dL_dpixel[3], accum_depth_rec = 0,
                          accum_alpha_rec = 0;
                    for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[(size_t)i * H * W + pix_id];
                    const float dL_dpixel_depth = dL_dpixel_depths[pix_id];
                    const float dL_dalpha = dL_dalphas[pix_id];
                    float last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0;
                    /* contributor index c (1-based position in the tile list) runs from
                     * last_contributor down to 1 (backward.cu:516-518) */
                    for (uint32_t c = last_contributor; c >= 1; c--) {
                        const uint32_t g = point_list[r0 + c - 1];
                        (void)r1;
                        if (unique_label) {
                            int uid = unique_label[g];
                            if (uid >= 0 && obj_id != uid) continue;
                        }
                        const float dx = means2D[2 * g] - pixfx, dy = means2D[2 * g + 1] - pixfy;
                        const float *co = conic_opacity + 4 * (size_t)g;
                        const float power = fmaf(fmaf(dx, dx * co[0], dy * (dy * co[2])), -0.5f,
                                                 -(dy * (dx * co[1])));
                        if (power > 0.0f) continue;
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, co[3] * G);
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.f - alpha);
                        const float dchannel_dcolor = alpha * T;
                        float dL_dopa = 0.0f;
                        for (int ch = 0; ch < 3; ch++) {
                            const float cc = colors[3 * (size_t)g + ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = cc;
                            dL_dopa += (cc - accum_rec[ch]) * dL_dpixel[ch];
#pragma omp atomic
                            dL_dcolors[3 * (size_t)g + ch] += (double)(dchannel_dcolor * dL_dpixel[ch]);
                        }
                        const float c_d = depths[g];
                        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
                        last_depth = c_d;
                        dL_dopa += (c_d - accum_depth_rec) * dL_dpixel_depth;
#pragma omp atomic
                        dL_ddepths[g] += (double)(dchannel_dcolor * dL_dpixel_depth);
                        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
                        dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
                        dL_dopa *= T;
                        last_alpha = alpha;
                        float bg_dot_dpixel = 0;
                        for (int i = 0; i < 3; i++) bg_dot_dpixel += bg[i] * dL_dpixel[i];
                        dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                        const float dL_dG = co[3] * dL_dopa;
                        const float gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                        const float dG_ddely = -gdy * co[2] - gdx * co[1];
#pragma omp atomic
                        dL_dmean2D[2 * (size_t)g] += (double)(dL_dG * dG_ddelx * ddelx_dx);
#pragma omp atomic
                        dL_dmean2D[2 * (size_t)g + 1] += (double)(dL_dG * dG_ddely * ddely_dy);
#pragma omp atomic
                        dL_dconic[3 * (size_t)g] += (double)(-0.5f * gdx * dx * dL_dG);
#pragma omp atomic
                        dL_dconic[3 * (size_t)g + 1] += (double)(-0.5f * gdx * dy * dL_dG);
#pragma omp atomic
                        dL_dconic[3 * (size_t)g + 2] += (double)(-0.5f * gdy * dy * dL_dG);
#pragma omp atomic
                        dL_dopacity[g] += (double)(G * dL_dopa);
                    }
                }
        }
}

/* DGR/cuda_rasterizer/backward.cu:144-274 (computeCov2DCUDA).  dL_dconic is [P,3] = (x,y,w). */
static void cov2d_backward_one(int idx, const float *mean, const float *cov3D, float h_x, float h_y,
                               float tan_fovx, float tan_fovy, const float *vm,
                               const float *dL_dconic3, float *dL_dmean, float *dL_dcov)
{
    float dcx = dL_dconic3[0], dcy = dL_dconic3[1], dcz = dL_dconic3[2];
    float tx = vm[0] * mean[0] + vm[4] * mean[1] + vm[8] * mean[2] + vm[12];
    float ty = vm[1] * mean[0] + vm[5] * mean[1] + vm[9] * mean[2] + vm[13];
    float tz = vm[2] * mean[0] + vm[6] * mean[1] + vm[10] * mean[2] + vm[14];
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    /* glm column-major: J[c][r], W[c][r], T = W*J */
    float J[3][3] = {{h_x / tz, 0, -(h_x * tx) / (tz * tz)}, {0, h_y / tz, -(h_y * ty) / (tz * tz)}, {0, 0, 0}};
    float Wm[3][3] = {{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}};
    float Vrk[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
    float T[3][3], A[3][3], cov2D[3][3];
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++)
            T[c][r] = Wm[0][r] * J[c][0] + Wm[1][r] * J[c][1] + Wm[2][r] * J[c][2];
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++) /* A = T^T Vrk^T */
            A[c][r] = T[r][0] * Vrk[0][c] + T[r][1] * Vrk[1][c] + T[r][2] * Vrk[2][c];
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++)
            cov2D[c][r] = A[0][r] * T[c][0] + A[1][r] * T[c][1] + A[2][r] * T[c][2];
    float a = cov2D[0][0] += 0.3f;
    float b = cov2D[0][1];
    float c = cov2D[1][1] += 0.3f;
    float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
        dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
        dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
        dL_dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
        dL_dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
        dL_dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
        dL_dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
        dL_dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
        dL_dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
    } else {
        for (int i = 0; i < 6; i++) dL_dcov[i] = 0;
    }
    float dL_dT00 = 2 * (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_da +
                    (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_db;
    float dL_dT01 = 2 * (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_da +
                    (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_db;
    float dL_dT02 = 2 * (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_da +
                    (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_db;
    float dL_dT10 = 2 * (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc +
                    (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_db;
    float dL_dT11 = 2 * (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc +
                    (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_db;
    float dL_dT12 = 2 * (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc +
                    (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_db;
    float dL_dJ00 = Wm[0][0] * dL_dT00 + Wm[0][1] * dL_dT01 + Wm[0][2] * dL_dT02;
    float dL_dJ02 = Wm[2][0] * dL_dT00 + Wm[2][1] * dL_dT01 + Wm[2][2] * dL_dT02;
    float dL_dJ11 = Wm[1][0] * dL_dT10 + Wm[1][1] * dL_dT11 + Wm[1][2] * dL_dT12;
    float dL_dJ12 = Wm[2][0] * dL_dT10 + Wm[2][1] * dL_dT11 + Wm[2][2] * dL_dT12;
    float tzi = 1.f / tz, tz2 = tzi * tzi, tz3 = tz2 * tzi;
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tx) * tz3 * dL_dJ02 +
                   (2 * h_y * ty) * tz3 * dL_dJ12;
    /* transformVec4x3Transpose, auxiliary.h:87-95; ASSIGNS (backward.cu:273) */
    dL_dmean[0] = vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz;
    dL_dmean[1] = vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz;
    dL_dmean[2] = vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz;
    (void)idx;
}

/* DGR backward.cu:20-139 (SH backward) */
static void sh_backward_one(int idx, int deg, int M, const float *means, const float *campos,
                            const float *shs, const uint8_t *clamped, const float *dL_dcolor,
                            float *dL_dmeans, float *dL_dshs)
{
    float dox = means[3 * idx] - campos[0], doy = means[3 * idx + 1] - campos[1],
          doz = means[3 * idx + 2] - campos[2];
    float len = sqrtf(dox * dox + doy * doy + doz * doz);
    float x = dox / len, y = doy / len, z = doz / len;
    const float *sh = shs + (size_t)idx * M * 3;
    float *dL_dsh = dL_dshs + (size_t)idx * M * 3;
    float dRGB[3];
    for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[3 * idx + ch] * (clamped[3 * idx + ch] ? 0.f : 1.f);
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
#define SH(k, ch) sh[3 * (k) + (ch)]
#define DSH(k, w) for (int ch = 0; ch < 3; ch++) dL_dsh[3 * (k) + ch] = (w) * dRGB[ch]
    DSH(0, SH_C0);
    if (deg > 0) {
        DSH(1, -SH_C1 * y);
        DSH(2, SH_C1 * z);
        DSH(3, -SH_C1 * x);
        for (int ch = 0; ch < 3; ch++) {
            dRGBdx[ch] = -SH_C1 * SH(3, ch);
            dRGBdy[ch] = -SH_C1 * SH(1, ch);
            dRGBdz[ch] = SH_C1 * SH(2, ch);
        }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            DSH(4, SH_C2[0] * xy);
            DSH(5, SH_C2[1] * yz);
            DSH(6, SH_C2[2] * (2.f * zz - xx - yy));
            DSH(7, SH_C2[3] * xz);
            DSH(8, SH_C2[4] * (xx - yy));
            for (int ch = 0; ch < 3; ch++) {
                dRGBdx[ch] += SH_C2[0] * y * SH(4, ch) + SH_C2[2] * 2.f * -x * SH(6, ch) + SH_C2[3] * z * SH(7, ch) + SH_C2[4] * 2.f * x * SH(8, ch);
                dRGBdy[ch] += SH_C2[0] * x * SH(4, ch) + SH_C2[1] * z * SH(5, ch) + SH_C2[2] * 2.f * -y * SH(6, ch) + SH_C2[4] * 2.f * -y * SH(8, ch);
                dRGBdz[ch] += SH_C2[1] * y * SH(5, ch) + SH_C2[2] * 2.f * 2.f * z * SH(6, ch) + SH_C2[3] * x * SH(7, ch);
            }
            if (deg > 2) {
                DSH(9, SH_C3[0] * y * (3.f * xx - yy));
                DSH(10, SH_C3[1] * xy * z);
                DSH(11, SH_C3[2] * y * (4.f * zz - xx - yy));
                DSH(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                DSH(13, SH_C3[4] * x * (4.f * zz - xx - yy));
                DSH(14, SH_C3[5] * z * (xx - yy));
                DSH(15, SH_C3[6] * x * (xx - 3.f * yy));
                for (int ch = 0; ch < 3; ch++) {
                    dRGBdx[ch] += (SH_C3[0] * SH(9, ch) * 3.f * 2.f * xy + SH_C3[1] * SH(10, ch) * yz +
                                   SH_C3[2] * SH(11, ch) * -2.f * xy + SH_C3[3] * SH(12, ch) * -3.f * 2.f * xz +
                                   SH_C3[4] * SH(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                                   SH_C3[5] * SH(14, ch) * 2.f * xz + SH_C3[6] * SH(15, ch) * 3.f * (xx - yy));
                    dRGBdy[ch] += (SH_C3[0] * SH(9, ch) * 3.f * (xx - yy) + SH_C3[1] * SH(10, ch) * xz +
                                   SH_C3[2] * SH(11, ch) * (-3.f * yy + 4.f * zz - xx) +
                                   SH_C3[3] * SH(12, ch) * -3.f * 2.f * yz + SH_C3[4] * SH(13, ch) * -2.f * xy +
                                   SH_C3[5] * SH(14, ch) * -2.f * yz + SH_C3[6] * SH(15, ch) * -3.f * 2.f * xy);
                    dRGBdz[ch] += (SH_C3[1] * SH(10, ch) * xy + SH_C3[2] * SH(11, ch) * 4.f * 2.f * yz +
                                   SH_C3[3] * SH(12, ch) * 3.f * (2.f * zz - xx - yy) +
                                   SH_C3[4] * SH(13, ch) * 4.f * 2.f * xz + SH_C3[5] * SH(14, ch) * (xx - yy));
                }
            }
        }
    }
#undef SH
#undef DSH
    float ddx = dRGBdx[0] * dRGB[0] + dRGBdx[1] * dRGB[1] + dRGBdx[2] * dRGB[2];
    float ddy = dRGBdy[0] * dRGB[0] + dRGBdy[1] * dRGB[1] + dRGBdy[2] * dRGB[2];
    float ddz = dRGBdz[0] * dRGB[0] + dRGBdz[1] * dRGB[1] + dRGBdz[2] * dRGB[2];
    /* dnormvdv, auxiliary.h:107-117 */
    float sum2 = dox * dox + doy * doy + doz * doz;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dL_dmeans[3 * idx + 0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
    dL_dmeans[3 * idx + 1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
    dL_dmeans[3 * idx + 2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
}

/* DGR backward.cu:278-341 (computeCov3D backward) */
static void cov3d_backward_one(const float *scale, float mod, const float *rot, const float *dL_dcov3D,
                               float *dL_dscale, float *dL_drot)
{
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    /* R[c][r] column-major */
    float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                     {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                     {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    float M[3][3];
    for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++) M[c][rr] = s[rr] * R[c][rr];
    float dS[3][3] = {{dL_dcov3D[0], 0.5f * dL_dcov3D[1], 0.5f * dL_dcov3D[2]},
                      {0.5f * dL_dcov3D[1], dL_dcov3D[3], 0.5f * dL_dcov3D[4]},
                      {0.5f * dL_dcov3D[2], 0.5f * dL_dcov3D[4], dL_dcov3D[5]}};
    /* dL_dM = 2.0f * M * dL_dSigma : (2M)[c][r] then product */
    float M2[3][3], dM[3][3];
    for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++) M2[c][rr] = 2.0f * M[c][rr];
    for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++)
            dM[c][rr] = M2[0][rr] * dS[c][0] + M2[1][rr] * dS[c][1] + M2[2][rr] * dS[c][2];
    float Rt[3][3], dMt[3][3];
    for (int c = 0; c < 3; c++)
        for (int rr = 0; rr < 3; rr++) {
            Rt[c][rr] = R[rr][c];
            dMt[c][rr] = dM[rr][c];
        }
    for (int k = 0; k < 3; k++)
        dL_dscale[k] = Rt[k][0] * dMt[k][0] + Rt[k][1] * dMt[k][1] + Rt[k][2] * dMt[k][2];
    for (int k = 0; k < 3; k++)
        for (int j = 0; j < 3; j++) dMt[k][j] *= s[k];
    dL_drot[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
    dL_drot[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
    dL_drot[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
    dL_drot[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

/* BACKWARD::preprocess, DGR backward.cu:604-672: computeCov2DCUDA (:144-274) then
 * preprocessCUDA (:346-412).  Inputs are the float per-Gaussian sums from the render backward.
 * All outputs must be zero-initialised by the caller (RasterizeGaussiansBackwardCUDA,
 * DGR/rasterize_points.cu:153-162). dL_dconic here is [P,3] (x,y,w). dL_dmean2D is [P,3]
 * (z unused, stays 0). */
void so_preprocess_backward(const so_scene *s, const int *radii, const float *cov3Ds,
                            const uint8_t *clamped, const float *dL_dmean2D, const float *dL_dconic,
                            float *dL_dmean3D, float *dL_dcolor, const float *dL_ddepth,
                            float *dL_dcov3D, float *dL_dsh, float *dL_dscale, float *dL_drot)
{
    const float focal_y = s->H / (2.0f * s->tanfovy);
    const float focal_x = s->W / (2.0f * s->tanfovx);
    const float *view = s->view, *proj = s->proj;
#pragma omp parallel for schedule(static)
    for (int idx = 0; idx < s->P; idx++) {
        if (!(radii[idx] > 0)) continue;
        const float *m = s->means3D + 3 * (size_t)idx;
        const float *cov3D = s->cov3D_precomp ? s->cov3D_precomp + 6 * (size_t)idx : cov3Ds + 6 * (size_t)idx;
        cov2d_backward_one(idx, m, cov3D, focal_x, focal_y, s->tanfovx, s->tanfovy, view,
                           dL_dconic + 3 * (size_t)idx, dL_dmean3D + 3 * (size_t)idx,
                           dL_dcov3D + 6 * (size_t)idx);
        /* backward.cu:372-389 */
        float m_w = 1.0f / ((proj[3] * m[0] + proj[7] * m[1] + proj[11] * m[2] + proj[15]) + 0.0000001f);
        float mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
        float mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
        float g2x = dL_dmean2D[3 * (size_t)idx], g2y = dL_dmean2D[3 * (size_t)idx + 1];
        dL_dmean3D[3 * idx + 0] += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
        dL_dmean3D[3 * idx + 1] += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
        dL_dmean3D[3 * idx + 2] += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
        /* backward.cu:391-403 (the fork's depth gradient) */
        float mul3 = view[2] * m[0] + view[6] * m[1] + view[10] * m[2] + view[14];
        dL_dmean3D[3 * idx + 0] += (view[2] - view[3] * mul3) * dL_ddepth[idx];
        dL_dmean3D[3 * idx + 1] += (view[6] - view[7] * mul3) * dL_ddepth[idx];
        dL_dmean3D[3 * idx + 2] += (view[10] - view[11] * mul3) * dL_ddepth[idx];
        if (s->shs)
            sh_backward_one(idx, s->D, s->M, s->means3D, s->campos, s->shs, clamped, dL_dcolor,
                            dL_dmean3D, dL_dsh);
        if (s->scales)
            cov3d_backward_one(s->scales + 3 * (size_t)idx, s->scale_modifier,
                               s->rotations + 4 * (size_t)idx, dL_dcov3D + 6 * (size_t)idx,
                               dL_dscale + 3 * (size_t)idx, dL_drot + 4 * (size_t)idx);
    }
}

/* multi_instance_opt, run_3d_seg.py:54-72 (= utils/wheatgs_helper.py:68-86), gamma = slackness.
 * all_contrib f32[num_rows, P] -> labels u8[num_rows, P].
 * F.normalize(dim=0): v / max(||v||_2, 1e-12); argmax ties -> index 0. */
void so_multi_instance_opt(int num_rows, int P, const float *all_contrib, float gamma, uint8_t *labels)
{
#pragma omp parallel for schedule(static)
    for (int j = 0; j < P; j++) {
        /* torch .sum(dim=0) over a handful of rows: sequential float adds in row order */
        float sum = 0.f;
        for (int i = 0; i < num_rows; i++) sum += all_contrib[(size_t)i * P + j];
        for (int i = 0; i < num_rows; i++) {
            float a = all_contrib[(size_t)i * P + j];
            float rest = sum - a;
            float nrm = sqrtf(rest * rest + a * a);
            if (nrm < 1e-12f) nrm = 1e-12f;
            float n0 = rest / nrm + gamma, n1 = a / nrm;
            labels[(size_t)i * P + j] = n1 > n0 ? 1 : 0;
        }
    }
}

/* ---- simple-knn: KNN/simple_knn.cu ---- */
static uint32_t prep_morton(uint32_t x) /* :37-44 */
{
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}
static inline uint32_t f2u_trunc(float f) /* F2I.U32.TRUNC: saturate, NaN -> 0 */
{
    if (f != f || f <= 0.0f) return 0;
    if (f >= 4294967296.0f) return UINT32_MAX;
    return (uint32_t)f;
}
typedef struct { float minn[3], maxx[3]; } so_minmax;

static float dist_box_point(const so_minmax *box, const float *p) /* :120-130 */
{
    float diff[3] = {0, 0, 0};
    for (int k = 0; k < 3; k++)
        if (p[k] < box->minn[k] || p[k] > box->maxx[k])
            diff[k] = fminf(fabsf(p[k] - box->minn[k]), fabsf(p[k] - box->maxx[k]));
    /* nvcc contraction (KNN SASS): fma(z,z, fma(x,x, y*y)) */
    return fmaf(diff[2], diff[2], fmaf(diff[0], diff[0], diff[1] * diff[1]));
}
static void update_kbest3(const float *ref, const float *point, float *knn) /* :132-146 */
{
    float d[3] = {point[0] - ref[0], point[1] - ref[1], point[2] - ref[2]};
    /* nvcc contraction of d.x*d.x + d.y*d.y + d.z*d.z */
    float dist = fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1]));
    for (int j = 0; j < 3; j++)
        if (knn[j] > dist) { float t = knn[j]; knn[j] = dist; dist = t; }
}

/* SimpleKNN::knn, KNN/simple_knn.cu:186-222; morton codes / sorted indices can be returned
 * for stage-wise parity (may be NULL). */
void so_knn_dist2(int P, const float *points, float *mean_dists, uint32_t *morton_out,
                  uint32_t *indices_sorted_out)
{
    if (P <= 0) return;
    float minn[3] = {0, 0, 0}, maxx[3] = {0, 0, 0}; /* init = {0,0,0}, :192 */
    for (int i = 0; i < P; i++)
        for (int k = 0; k < 3; k++) {
            minn[k] = fminf(minn[k], points[3 * i + k]);
            maxx[k] = fmaxf(maxx[k], points[3 * i + k]);
        }
    uint32_t *morton = (uint32_t *)malloc((size_t)P * 4), *idx = (uint32_t *)malloc((size_t)P * 4);
    uint64_t *k64 = (uint64_t *)malloc((size_t)P * 8), *k64o = (uint64_t *)malloc((size_t)P * 8);
    uint32_t *idxo = (uint32_t *)malloc((size_t)P * 4);
    for (int i = 0; i < P; i++) { /* coord2Morton :46-62 */
        uint32_t c[3];
        for (int k = 0; k < 3; k++)
            c[k] = prep_morton(f2u_trunc(((points[3 * i + k] - minn[k]) / (maxx[k] - minn[k])) * 1023.0f));
        morton[i] = c[0] | (c[1] << 1) | (c[2] << 2);
        idx[i] = (uint32_t)i;
        k64[i] = morton[i];
    }
    so_sort_pairs((uint64_t)P, k64, k64o, idx, idxo, 32); /* :211-214, stable */
    if (morton_out) memcpy(morton_out, morton, (size_t)P * 4);
    if (indices_sorted_out) memcpy(indices_sorted_out, idxo, (size_t)P * 4);
    const int BOX = 1024;
    int nb = (P + BOX - 1) / BOX;
    so_minmax *boxes = (so_minmax *)malloc((size_t)nb * sizeof(so_minmax));
    for (int b = 0; b < nb; b++) { /* boxMinMax :79-118 */
        so_minmax me = {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
        for (int i = b * BOX; i < P && i < (b + 1) * BOX; i++)
            for (int k = 0; k < 3; k++) {
                me.minn[k] = fminf(me.minn[k], points[3 * idxo[i] + k]);
                me.maxx[k] = fmaxf(me.maxx[k], points[3 * idxo[i] + k]);
            }
        boxes[b] = me;
    }
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < P; i++) { /* boxMeanDist :148-184 */
        const float *point = points + 3 * (size_t)idxo[i];
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        int lo = i - 3 < 0 ? 0 : i - 3, hi = i + 3 > P - 1 ? P - 1 : i + 3;
        for (int j = lo; j <= hi; j++) {
            if (j == i) continue;
            update_kbest3(point, points + 3 * (size_t)idxo[j], best);
        }
        float reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
        for (int b = 0; b < nb; b++) {
            float dist = dist_box_point(&boxes[b], point);
            if (dist > reject || dist > best[2]) continue;
            int end = (b + 1) * BOX < P ? (b + 1) * BOX : P;
            for (int j = b * BOX; j < end; j++) {
                if (j == i) continue;
                update_kbest3(point, points + 3 * (size_t)idxo[j], best);
            }
        }
        mean_dists[idxo[i]] = (best[0] + best[1] + best[2]) / 3.0f;
    }
    free(morton); free(idx); free(k64); free(k64o); free(idxo); free(boxes);
}

int so_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void so_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
