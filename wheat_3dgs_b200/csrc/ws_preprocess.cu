// ws_preprocess.cu -- per-Gaussian forward stages:
//   preprocess_fwd   (replaces preprocessCUDA, DGR forward.cu:155-256; FSR forward.cu:155-264)
//   scan_block_sums  (replaces cub::DeviceScan::InclusiveSum, DGR rasterizer_impl.cu:278)
//   emit_keys        (replaces duplicateWithKeys, DGR rasterizer_impl.cu:70-111)
//   mark_visible     (replaces checkFrustum, DGR rasterizer_impl.cu:54-66)
//
// Bit-exact tier.  depth (key low word), radius and the tile rectangle must equal the
// reference's bit for bit, so every float op on that chain is an explicit _rn intrinsic in
// the order nvcc 12.9 contracts the reference source for sm_100 (read from its SASS):
//   a*b + c*d + e*f (+ g)  ->  fadd(fma(e,f, fma(a,b, fmul(c,d))), g)
//   1/x, a/b, sqrt         ->  IEEE (__frcp_rn, __fdiv_rn, __fsqrt_rn)
//   ndc2Pix                ->  double: DFMA((v+1), S, -1) * 0.5, rounded to float
// The SH -> RGB evaluation is in the tolerance tier and written plainly.
//
// Memory behaviour.  One thread per Gaussian, WS_PRE_THREADS (128) per CTA.  The SH block of a CTA
// (128 x 3M floats, contiguous in the [P,M,3] input) is brought in with fully coalesced
// 16-byte loads into padded shared memory (row stride 3M/4+1 float4 -> conflict-free
// LDS.128), instead of each thread striding through its own 192 bytes.  Outputs are three
// SoA float4 arrays (coalesced 512 B per warp store).  The CTA also reduces tiles_touched so
// the prefix sum only has to scan P/128 block sums.
#include "ws_common.cuh"

namespace ws {

__constant__ float c_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__constant__ float c_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                 0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};
#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f

struct PreParams {
    int P, D, M, W, H, gx, gy;
    float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
    const float *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *view, *proj, *campos;
    int32_t *radii;
    GRec *grec;
    uint32_t *tiles_touched, *block_sums;
    uint32_t *dkeys, *dvals; // depth bits (culled: 0xffffffff) and Gaussian id, input of the depth sort
    uint8_t *clamped;
    float *image_xys, *gs_depth; // flashsplat extras or NULL
    int sh_vec4;                 // 1: 3M % 4 == 0 and shs 16-byte aligned -> float4 staging
    const float *shs_rest;       // features_rest [P,M-1,3] (shs is then features_dc [P,1,3]) or NULL
    int raw;                     // opacities/scales/rotations are GaussianModel's raw parameters
};

// DGR forward.cu:118-152; op order from the reference SASS (see oracle/splat_oracle.c).
__device__ __forceinline__ void compute_cov3d(float sx, float sy, float sz, float r, float x, float y,
                                              float z, float *c)
{
    float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    float s_yz = __fadd_rn(yy, zz);
    float r00 = __fadd_rn(1.0f, -__fadd_rn(s_yz, s_yz));
    float s_xz = __fmaf_rn(x, x, zz);
    float r11 = __fadd_rn(1.0f, -__fadd_rn(s_xz, s_xz));
    float s_xy = __fmaf_rn(x, x, yy);
    float r22 = __fadd_rn(1.0f, -__fadd_rn(s_xy, s_xy));
    float rz = __fmul_rn(r, z), rx = __fmul_rn(r, x), xz = __fmul_rn(x, z);
    float a;
    a = __fmaf_rn(x, y, -rz); float xy_m_rz = __fadd_rn(a, a);
    a = __fmaf_rn(x, y, rz);  float xy_p_rz = __fadd_rn(a, a);
    a = __fmaf_rn(r, y, xz);  float xz_p_ry = __fadd_rn(a, a);
    a = __fmaf_rn(-r, y, xz); float xz_m_ry = __fadd_rn(a, a);
    a = __fmaf_rn(y, z, -rx); float yz_m_rx = __fadd_rn(a, a);
    a = __fmaf_rn(y, z, rx);  float yz_p_rx = __fadd_rn(a, a);
    float M00 = __fmul_rn(sx, r00), M01 = __fmul_rn(sy, xy_m_rz), M02 = __fmul_rn(sz, xz_p_ry);
    float M10 = __fmul_rn(sx, xy_p_rz), M11 = __fmul_rn(sy, r11), M12 = __fmul_rn(sz, yz_m_rx);
    float M20 = __fmul_rn(sx, xz_m_ry), M21 = __fmul_rn(sy, yz_p_rx), M22 = __fmul_rn(sz, r22);
    c[0] = dot3_nv(M00, M00, M01, M01, M02, M02);
    c[1] = dot3_nv(M00, M10, M01, M11, M02, M12);
    c[2] = dot3_nv(M00, M20, M01, M21, M02, M22);
    c[3] = dot3_nv(M10, M10, M11, M11, M12, M12);
    c[4] = dot3_nv(M10, M20, M11, M21, M12, M22);
    c[5] = dot3_nv(M20, M20, M21, M21, M22, M22);
}

// DGR forward.cu:74-113
__device__ __forceinline__ void compute_cov2d(float x, float y, float z, float focal_x, float focal_y,
                                              float tan_fovx, float tan_fovy, const float *c,
                                              const float *v, float &cxx, float &cxy, float &cyy)
{
    float tx = __fadd_rn(dot3_nv(x, v[0], y, v[4], z, v[8]), v[12]);
    float ty = __fadd_rn(dot3_nv(x, v[1], y, v[5], z, v[9]), v[13]);
    float tz = __fadd_rn(dot3_nv(x, v[2], y, v[6], z, v[10]), v[14]);
    float limx = __fmul_rn(1.3f, tan_fovx), limy = __fmul_rn(1.3f, tan_fovy);
    float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    float cx = fminf(limx, fmaxf(-limx, txtz));
    float cy = fminf(limy, fmaxf(-limy, tytz));
    tx = __fmul_rn(cx, tz);
    ty = __fmul_rn(cy, tz);
    float tz2 = __fmul_rn(tz, tz);
    float J00 = __fdiv_rn(focal_x, tz);
    float J02 = __fdiv_rn(__fmul_rn(-tx, focal_x), tz2);
    float J11 = __fdiv_rn(focal_y, tz);
    float J12 = __fdiv_rn(__fmul_rn(-ty, focal_y), tz2);
    float T00 = __fmaf_rn(v[2], J02, __fmul_rn(v[0], J00));
    float T01 = __fmaf_rn(v[6], J02, __fmul_rn(v[4], J00));
    float T02 = __fmaf_rn(v[10], J02, __fmul_rn(v[8], J00));
    float T10 = __fmaf_rn(v[2], J12, __fmul_rn(v[1], J11));
    float T11 = __fmaf_rn(v[6], J12, __fmul_rn(v[5], J11));
    float T12 = __fmaf_rn(v[10], J12, __fmul_rn(v[9], J11));
    float A00 = dot3_nv(T00, c[0], T01, c[1], T02, c[2]);
    float A10 = dot3_nv(T00, c[1], T01, c[3], T02, c[4]);
    float A20 = dot3_nv(T00, c[2], T01, c[4], T02, c[5]);
    float A01 = dot3_nv(T10, c[0], T11, c[1], T12, c[2]);
    float A11 = dot3_nv(T10, c[1], T11, c[3], T12, c[4]);
    float A21 = dot3_nv(T10, c[2], T11, c[4], T12, c[5]);
    cxx = __fadd_rn(dot3_nv(T00, A00, T01, A10, T02, A20), 0.3f);
    cxy = dot3_nv(T00, A01, T01, A11, T02, A21);
    cyy = __fadd_rn(dot3_nv(T10, A01, T11, A11, T12, A21), 0.3f);
}

__device__ __forceinline__ float ndc2pix(float v, int S)
{
    double t = __dadd_rn((double)v, 1.0);
    t = __fma_rn(t, (double)S, -1.0);
    return __double2float_rn(__dmul_rn(t, 0.5));
}

// DGR forward.cu:20-71 for one colour channel; sh points at this Gaussian's coefficient 0 of
// that channel, coefficients are `stride` floats apart.
__device__ __forceinline__ float sh_channel(int deg, const float *sh, int stride, float x, float y, float z)
{
    float res = SH_C0 * sh[0];
    if (deg > 0) {
        res = res - SH_C1 * y * sh[stride] + SH_C1 * z * sh[2 * stride] - SH_C1 * x * sh[3 * stride];
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            res = res + c_SH_C2[0] * xy * sh[4 * stride] + c_SH_C2[1] * yz * sh[5 * stride] +
                  c_SH_C2[2] * (2.0f * zz - xx - yy) * sh[6 * stride] + c_SH_C2[3] * xz * sh[7 * stride] +
                  c_SH_C2[4] * (xx - yy) * sh[8 * stride];
            if (deg > 2) {
                res = res + c_SH_C3[0] * y * (3.0f * xx - yy) * sh[9 * stride] +
                      c_SH_C3[1] * xy * z * sh[10 * stride] +
                      c_SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11 * stride] +
                      c_SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12 * stride] +
                      c_SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13 * stride] +
                      c_SH_C3[5] * z * (xx - yy) * sh[14 * stride] +
                      c_SH_C3[6] * x * (xx - 3.0f * yy) * sh[15 * stride];
            }
        }
    }
    return res + 0.5f;
}

__global__ void __launch_bounds__(WS_PRE_THREADS) preprocess_fwd(const PreParams p)
{
    extern __shared__ float4 s_sh4[]; // [256][row4] float4, row4 = 3M/4 + 1 (vec4) or scalar rows
    __shared__ float s_view[16], s_proj[16];
    __shared__ uint32_t s_warp_sum[WS_PRE_THREADS / 32];

    const int tid = threadIdx.x;
    const int base = blockIdx.x * WS_PRE_THREADS;
    const int idx = base + tid;
    if (tid < 16) s_view[tid] = p.view[tid];
    else if (tid < 32) s_proj[tid - 16] = p.proj[tid - 16];

    // ---- stage the CTA's SH block (coalesced) ----
    const int nvalid = min(WS_PRE_THREADS, p.P - base);
    const bool use_sh = (p.colors_precomp == nullptr);
    const int row_f = 3 * p.M; // floats per Gaussian
    if (use_sh) {
        if (p.sh_vec4) {
            const int row4 = row_f / 4, srow4 = row4 + 1;
            const float4 *src = reinterpret_cast<const float4 *>(p.shs) + (size_t)base * row4;
            const int n4 = nvalid * row4;
            for (int i = tid; i < n4; i += WS_PRE_THREADS) {
                int g = i / row4, c = i - g * row4;
                s_sh4[g * srow4 + c] = __ldg(src + i);
            }
        } else if (p.shs_rest) {
            // features_dc and features_rest staged side by side into the rows torch.cat would have built
            float *s_sh = reinterpret_cast<float *>(s_sh4);
            const int srow = row_f | 1, rest_f = row_f - 3;
            const float *src = p.shs + (size_t)base * 3;
            for (int i = tid; i < nvalid * 3; i += WS_PRE_THREADS) {
                int g = i / 3, c = i - g * 3;
                s_sh[g * srow + c] = __ldg(src + i);
            }
            src = p.shs_rest + (size_t)base * rest_f;
            const int n = nvalid * rest_f;
            for (int i = tid; i < n; i += WS_PRE_THREADS) {
                int g = i / rest_f, c = i - g * rest_f;
                s_sh[g * srow + 3 + c] = __ldg(src + i);
            }
        } else {
            float *s_sh = reinterpret_cast<float *>(s_sh4);
            const int srow = row_f | 1; // odd stride -> conflict-free
            const float *src = p.shs + (size_t)base * row_f;
            const int n = nvalid * row_f;
            for (int i = tid; i < n; i += WS_PRE_THREADS) {
                int g = i / row_f, c = i - g * row_f;
                s_sh[g * srow + c] = __ldg(src + i);
            }
        }
    }
    __syncthreads();

    uint32_t my_tiles = 0, my_dkey = 0xffffffffu;
    if (idx < p.P) {
        int my_rad = 0;
        const float *v = s_view, *pm = s_proj;
        const float x = p.means3D[3 * (size_t)idx], y = p.means3D[3 * (size_t)idx + 1],
                    z = p.means3D[3 * (size_t)idx + 2];
        // in_frustum, DGR auxiliary.h:139-164
        const float pvz = __fadd_rn(dot3_nv(x, v[2], y, v[6], z, v[10]), v[14]);
        if (!(pvz <= 0.2f)) { // NaN is not culled, as in the reference (FSETP.GTU)
            float hx = __fadd_rn(dot3_nv(x, pm[0], y, pm[4], z, pm[8]), pm[12]);
            float hy = __fadd_rn(dot3_nv(x, pm[1], y, pm[5], z, pm[9]), pm[13]);
            float hw = __fadd_rn(dot3_nv(x, pm[3], y, pm[7], z, pm[11]), pm[15]);
            float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
            float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
            float c3[6];
            if (p.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c3[k] = p.cov3D_precomp[6 * (size_t)idx + k];
            } else {
                float4 q = reinterpret_cast<const float4 *>(p.rotations)[idx];
                float s0 = p.scales[3 * (size_t)idx], s1 = p.scales[3 * (size_t)idx + 1],
                      s2 = p.scales[3 * (size_t)idx + 2];
                if (p.raw) raw_activate(q, s0, s1, s2);
                compute_cov3d(__fmul_rn(s0, p.scale_modifier), __fmul_rn(s1, p.scale_modifier),
                              __fmul_rn(s2, p.scale_modifier), q.x, q.y, q.z, q.w, c3);
            }
            float cxx, cxy, cyy;
            compute_cov2d(x, y, z, p.focal_x, p.focal_y, p.tanfovx, p.tanfovy, c3, v, cxx, cxy, cyy);
            const float det = __fmaf_rn(cxx, cyy, -__fmul_rn(cxy, cxy));
            if (det != 0.0f) {
                const float det_inv = __frcp_rn(det);
                const float conx = __fmul_rn(cyy, det_inv), cony = __fmul_rn(cxy, -det_inv),
                            conz = __fmul_rn(cxx, det_inv);
                const float mid = __fmul_rn(0.5f, __fadd_rn(cxx, cyy));
                const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
                const float lambda1 = __fadd_rn(mid, sq), lambda2 = __fadd_rn(mid, -sq);
                const int radius = __float2int_ru(__fmul_rn(3.0f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
                const float pix = ndc2pix(projx, p.W), piy = ndc2pix(projy, p.H);
                uint2 rmin, rmax;
                get_rect(pix, piy, radius, p.gx, p.gy, rmin, rmax);
                if (p.image_xys) { // FSR forward.cu:237-238 (before the zero-tile cull)
                    p.image_xys[idx] = pix;
                    p.image_xys[(size_t)p.P + idx] = piy;
                }
                const uint32_t tiles = (rmax.x - rmin.x) * (rmax.y - rmin.y);
                if (tiles != 0) {
                    float r, g, b;
                    uint8_t cl = 0;
                    if (use_sh) {
                        // DGR forward.cu:20-71
                        float dx = x - p.campos[0], dy = y - p.campos[1], dz = z - p.campos[2];
                        float len = sqrtf(dx * dx + dy * dy + dz * dz);
                        dx = dx / len; dy = dy / len; dz = dz / len;
                        const float *sh = p.sh_vec4
                                              ? reinterpret_cast<const float *>(s_sh4 + tid * (row_f / 4 + 1))
                                              : reinterpret_cast<const float *>(s_sh4) + tid * (row_f | 1);
                        r = sh_channel(p.D, sh + 0, 3, dx, dy, dz);
                        g = sh_channel(p.D, sh + 1, 3, dx, dy, dz);
                        b = sh_channel(p.D, sh + 2, 3, dx, dy, dz);
                        cl = (r < 0 ? 1 : 0) | (g < 0 ? 2 : 0) | (b < 0 ? 4 : 0);
                        r = fmaxf(r, 0.0f); g = fmaxf(g, 0.0f); b = fmaxf(b, 0.0f);
                    } else {
                        r = p.colors_precomp[3 * (size_t)idx];
                        g = p.colors_precomp[3 * (size_t)idx + 1];
                        b = p.colors_precomp[3 * (size_t)idx + 2];
                    }
                    const float opac = p.raw ? raw_sigmoid(p.opacities[idx]) : p.opacities[idx];
                    GRec rec;
                    rec.a = make_float4(pix, piy, conx, cony);
                    rec.b = make_float4(conz, opac, pvz, logf(255.0f * opac));
                    rec.c = make_float4(r, g, b, 0.0f);
                    rec.d = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                    p.grec[idx] = rec;
                    p.clamped[idx] = cl;
                    if (p.gs_depth) p.gs_depth[idx] = pvz; // FSR forward.cu:258
                    my_rad = radius;
                    my_tiles = tiles;
                    my_dkey = __float_as_uint(pvz);
                }
            }
        }
        p.radii[idx] = my_rad;
        p.tiles_touched[idx] = my_tiles;
        p.dkeys[idx] = my_dkey;
        p.dvals[idx] = (uint32_t)idx;
    }
    // block sum of tiles_touched
    uint32_t s = my_tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((tid & 31) == 0) s_warp_sum[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < WS_PRE_THREADS / 32; w++) t += s_warp_sum[w];
        p.block_sums[blockIdx.x] = t;
    }
}

int launch_preprocess(const ws_view *v, const ws_gaussians *g, int32_t *radii, GeomState &gs,
                      const ws_flash *fl, cudaStream_t st)
{
    PreParams p;
    p.P = g->P; p.D = v->sh_degree; p.M = g->M; p.W = v->image_width; p.H = v->image_height;
    p.gx = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X; p.gy = (p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    p.tanfovx = v->tanfovx; p.tanfovy = v->tanfovy;
    p.focal_y = p.H / (2.0f * v->tanfovy); // DGR rasterizer_impl.cu:224-225
    p.focal_x = p.W / (2.0f * v->tanfovx);
    p.scale_modifier = v->scale_modifier;
    p.means3D = g->means3D; p.shs = g->shs; p.colors_precomp = g->colors_precomp;
    p.opacities = g->opacities; p.scales = g->scales; p.rotations = g->rotations;
    p.cov3D_precomp = g->cov3D_precomp;
    p.shs_rest = g->shs ? g->shs_rest : nullptr; p.raw = g->raw_params;
    p.view = v->viewmatrix; p.proj = v->projmatrix; p.campos = v->campos;
    p.radii = radii; p.grec = gs.grec;
    p.tiles_touched = gs.tiles_touched; p.block_sums = gs.block_sums; p.clamped = gs.clamped;
    p.dkeys = gs.dkeys_a; p.dvals = gs.dvals_a;
    p.image_xys = fl ? fl->image_xys : nullptr;
    p.gs_depth = fl ? fl->gs_depth : nullptr;
    const int row_f = 3 * g->M;
    p.sh_vec4 = (g->shs && !p.shs_rest && row_f % 4 == 0 && ((uintptr_t)g->shs % 16) == 0) ? 1 : 0;
    size_t smem = 0;
    if (!g->colors_precomp)
        smem = p.sh_vec4 ? (size_t)WS_PRE_THREADS * (row_f / 4 + 1) * 16 : (size_t)WS_PRE_THREADS * (row_f | 1) * 4;
    if (smem > 200 * 1024) {
        set_error("preprocess: %d SH coefficients per Gaussian need %zu bytes of shared memory", g->M, smem);
        return WS_E_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
        WS_CUDA_CHECK(cudaFuncSetAttribute(preprocess_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int nb = (g->P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    preprocess_fwd<<<nb, WS_PRE_THREADS, smem, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

// ---- prefix sum over block sums: one CTA, carries across chunks of 1024 ----
__global__ void __launch_bounds__(1024) scan_block_sums(int nb, const uint32_t *__restrict__ sums,
                                                        uint32_t *__restrict__ offsets,
                                                        uint32_t *__restrict__ total)
{
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    pdl_wait();
    pdl_trigger();
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int c0 = 0; c0 < nb; c0 += 1024) {
        const int i = c0 + tid;
        const uint32_t v = i < nb ? sums[i] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) s_w[w] = inc;
        __syncthreads();
        if (w == 0) {
            uint32_t x = s_w[lane], xi = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t n = __shfl_up_sync(0xffffffffu, xi, o);
                if (lane >= o) xi += n;
            }
            s_w[lane] = xi - x; // exclusive warp offsets
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t excl = carry + s_w[w] + inc - v;
        if (i < nb) offsets[i] = excl;
        __syncthreads();
        if (tid == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (tid == 0) total[0] = s_carry;
}

// num_rendered for the host: one 8-byte store {seq, total} into a pinned, device-mapped mailbox the host spins on
// (ws_forward_prepare).  Runs on a side stream behind the scan, so the store's trip over PCIe holds nothing up.
__global__ void post_total(const uint32_t *__restrict__ total, unsigned long long *mailbox, uint32_t seq)
{
    *(volatile unsigned long long *)mailbox = ((unsigned long long)seq << 32) | (unsigned long long)total[0];
}

int launch_post_total(const uint32_t *total, unsigned long long *mailbox, uint32_t seq, cudaStream_t side)
{
    post_total<<<1, 1, 0, side>>>(total, mailbox, seq);
    WS_LAUNCH_CHECK(side, false);
    return WS_OK;
}

int launch_scan_u32(int nb, const uint32_t *sums, uint32_t *offsets, uint32_t *total, cudaStream_t st, bool debug)
{
    scan_block_sums<<<1, 1024, 0, st>>>(nb, sums, offsets, total);
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

int launch_scan_blocks(int P, GeomState &gs, cudaStream_t st, bool debug)
{
    const int nb = (P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    scan_block_sums<<<1, 1024, 0, st>>>(nb, gs.block_sums, gs.block_offsets, gs.total);
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

// ---- key emission: warp-flattened, coalesced ----
// Every warp owns 32 consecutive Gaussians.  Their tile counts are scanned in the warp and
// the warp's output run [base, base+total) is written 32 consecutive slots at a time; the
// owning Gaussian of a slot is found by a 5-step binary search over the inclusive scan held
// in registers.  Order of emission = ascending Gaussian id, then row-major (y, x) tiles,
// exactly the reference's (rasterizer_impl.cu:96-108), so a stable sort gives the same list.
__global__ void __launch_bounds__(WS_PRE_THREADS) emit_keys(int P, int gx, int gy,
                                                            const int32_t *__restrict__ radii,
                                                            const GRec *__restrict__ grec,
                                                            const uint32_t *__restrict__ tiles_touched,
                                                            const uint32_t *__restrict__ block_offsets,
                                                            uint64_t *__restrict__ keys,
                                                            uint32_t *__restrict__ vals)
{
    __shared__ uint32_t s_w[WS_PRE_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int idx = blockIdx.x * WS_PRE_THREADS + tid;
    uint32_t tt = 0, rminx = 0, rminy = 0, width = 1, dbits = 0;
    float4 ga = make_float4(0.f, 0.f, 0.f, 0.f), gb = ga;
    if (idx < P) {
        const int rad = radii[idx];
        if (rad > 0) {
            tt = tiles_touched[idx];
            ga = grec[idx].a; gb = grec[idx].b;
            uint2 rmin, rmax;
            get_rect(ga.x, ga.y, rad, gx, gy, rmin, rmax);
            rminx = rmin.x; rminy = rmin.y; width = rmax.x - rmin.x;
            dbits = __float_as_uint(gb.z);
        }
    }
    uint32_t inc = tt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    const uint32_t warp_total = __shfl_sync(0xffffffffu, inc, 31);
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    uint32_t warp_base = block_offsets[blockIdx.x];
    for (int k = 0; k < w; k++) warp_base += s_w[k];
    const uint32_t excl = inc - tt;
    const uint32_t first_idx = blockIdx.x * WS_PRE_THREADS + w * 32;
    const QuadTest qt = quad_test_of(ga, gb);
    for (uint32_t k0 = 0; k0 < warp_total; k0 += 32) {
        const uint32_t k = k0 + lane;
        int j = 0; // number of lanes whose inclusive sum <= k  == owner lane
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            uint32_t vv = __shfl_sync(0xffffffffu, inc, min(j + s - 1, 31));
            if (vv <= k) j += s;
        }
        j = min(j, 31);
        const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, j);
        const uint32_t o_minx = __shfl_sync(0xffffffffu, rminx, j);
        const uint32_t o_miny = __shfl_sync(0xffffffffu, rminy, j);
        const uint32_t o_w = __shfl_sync(0xffffffffu, width, j);
        const uint32_t o_d = __shfl_sync(0xffffffffu, dbits, j);
        const QuadTest oq = shfl_quad_test(qt, j);
        if (k < warp_total) {
            const uint32_t local = k - o_excl;
            const uint32_t ty = local / o_w, tx = local - ty * o_w;
            uint64_t key = (uint64_t)((o_miny + ty) * (uint32_t)gx + (o_minx + tx));
            key = (key << 32) | o_d;
            keys[(size_t)warp_base + k] = key;
            const uint32_t qm = quad_mask(oq, (float)((o_minx + tx) * WS_BLOCK_X), (float)((o_miny + ty) * WS_BLOCK_Y));
            vals[(size_t)warp_base + k] = (first_idx + j) | (qm << WS_ID_BITS);
        }
    }
}

int launch_emit_keys(int P, int W, int H, const int32_t *radii, GeomState &gs, uint64_t *keys,
                     uint32_t *vals, cudaStream_t st, bool debug)
{
    const int gx = (W + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    const int nb = (P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    emit_keys<<<nb, WS_PRE_THREADS, 0, st>>>(P, gx, gy, radii, gs.grec, gs.tiles_touched,
                                             gs.block_offsets, keys, vals);
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

// ---- depth-first binning (default) ----
// The reference sorts R = sum(tiles_touched) 64-bit tile|depth keys (6 byte-passes at 46 bits).
// The same list comes out of two smaller sorts: (1) the P Gaussians by their 32-bit depth bits
// (stable, so equal depths keep ascending id), (2) instances emitted in that order, stably sorted
// by their tile id alone (14 bits at config 2: 2 passes over 8-byte pairs instead of 6 over
// 12-byte pairs).  A stable sort by tile of a depth-ordered list is the stable sort by
// (tile, depth) of the id-ordered list: point_list is bit-identical (tests compare it, and the
// reconstructed tile|depth keys, with the reference's).
int launch_depth_sort(int P, GeomState &gs, cudaStream_t st, bool debug)
{
    uint32_t *kf, *vf;
    return launch_sort_pairs32((size_t)P, gs.dkeys_a, gs.dkeys_b, gs.dvals_a, gs.dvals_b, 32, gs.dsort, st, debug,
                               &kf, &vf); // 4 passes: the result is back in the "a" buffers
}
const uint32_t *depth_order(const GeomState &gs) { return gs.dvals_a; }

// tiles_touched summed over blocks of WS_PRE_THREADS Gaussians taken in depth order
// Also clears up to two word arrays for the stages behind it (the tile sort's histograms / status words and the tile
// ranges): two memset launches less between the host learning R and the blend kernel.
__global__ void __launch_bounds__(WS_PRE_THREADS) ordered_block_sums(int P, const uint32_t *__restrict__ order,
                                                                     const uint32_t *__restrict__ tiles_touched,
                                                                     uint32_t *__restrict__ block_sums,
                                                                     uint32_t *__restrict__ zero_a, size_t na,
                                                                     uint32_t *__restrict__ zero_b, size_t nb_words)
{
    __shared__ uint32_t s_w[WS_PRE_THREADS / 32];
    const int i = blockIdx.x * WS_PRE_THREADS + threadIdx.x;
    pdl_wait();
    pdl_trigger();
    {
        const size_t stride = (size_t)gridDim.x * WS_PRE_THREADS;
        for (size_t j = (size_t)i; j < na; j += stride) zero_a[j] = 0u;
        for (size_t j = (size_t)i; j < nb_words; j += stride) zero_b[j] = 0u;
    }
    uint32_t s = i < P ? tiles_touched[order[i]] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < WS_PRE_THREADS / 32; w++) t += s_w[w];
        block_sums[blockIdx.x] = t;
    }
}

// emit_keys over the depth-ordered Gaussians: key = tile id only, value = Gaussian id
__global__ void __launch_bounds__(WS_PRE_THREADS) emit_tiles_ordered(int P, int gx, int gy,
                                                                     const uint32_t *__restrict__ order,
                                                                     const int32_t *__restrict__ radii,
                                                                     const GRec *__restrict__ grec,
                                                                     const uint32_t *__restrict__ tiles_touched,
                                                                     const uint32_t *__restrict__ block_offsets,
                                                                     uint32_t *__restrict__ tkeys,
                                                                     uint32_t *__restrict__ vals,
                                                                     uint32_t *__restrict__ hist, int npass, int wbits)
{
    __shared__ uint32_t s_w[WS_PRE_THREADS / 32];
    // digit counts of the keys this CTA emits, all passes of the tile sort (pass p = key bits [8p, 8p+8)): the sort
    // needs them before its first pass and would otherwise read every key once more to get them
    __shared__ uint32_t s_hist[4 * 256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (hist)
        for (int j = tid; j < npass * 256; j += WS_PRE_THREADS) s_hist[j] = 0;
    pdl_wait();
    pdl_trigger();
    const int i = blockIdx.x * WS_PRE_THREADS + tid;
    uint32_t tt = 0, rminx = 0, rminy = 0, width = 1, gid = 0;
    float4 ga = make_float4(0.f, 0.f, 0.f, 0.f), gb = ga;
    if (i < P) {
        gid = order[i];
        const int rad = radii[gid];
        if (rad > 0) {
            tt = tiles_touched[gid];
            ga = grec[gid].a; gb = grec[gid].b;
            uint2 rmin, rmax;
            get_rect(ga.x, ga.y, rad, gx, gy, rmin, rmax);
            rminx = rmin.x; rminy = rmin.y; width = rmax.x - rmin.x;
        }
    }
    uint32_t inc = tt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    const uint32_t warp_total = __shfl_sync(0xffffffffu, inc, 31);
    if (lane == 31) s_w[w] = inc;
    __syncthreads();
    uint32_t warp_base = block_offsets[blockIdx.x];
    for (int k = 0; k < w; k++) warp_base += s_w[k];
    const uint32_t excl = inc - tt;
    const QuadTest qt = quad_test_of(ga, gb);
    for (uint32_t k0 = 0; k0 < warp_total; k0 += 32) {
        const uint32_t k = k0 + lane;
        int j = 0; // number of lanes whose inclusive sum <= k  == owner lane
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {
            uint32_t vv = __shfl_sync(0xffffffffu, inc, min(j + s - 1, 31));
            if (vv <= k) j += s;
        }
        j = min(j, 31);
        const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, j);
        const uint32_t o_minx = __shfl_sync(0xffffffffu, rminx, j);
        const uint32_t o_miny = __shfl_sync(0xffffffffu, rminy, j);
        const uint32_t o_w = __shfl_sync(0xffffffffu, width, j);
        const uint32_t o_id = __shfl_sync(0xffffffffu, gid, j);
        const QuadTest oq = shfl_quad_test(qt, j);
        uint32_t key = 0xffffffffu;
        if (k < warp_total) {
            const uint32_t local = k - o_excl;
            const uint32_t ty = local / o_w, tx = local - ty * o_w;
            key = (o_miny + ty) * (uint32_t)gx + (o_minx + tx);
            tkeys[(size_t)warp_base + k] = key;
            // which 8x8 quadrants of this tile the Gaussian can reach: the blend kernels' sub-tile cull, paid
            // once per instance here instead of once per (warp, record) in every blend pass
            const uint32_t qm = quad_mask(oq, (float)((o_minx + tx) * WS_BLOCK_X), (float)((o_miny + ty) * WS_BLOCK_Y));
            vals[(size_t)warp_base + k] = o_id | (qm << WS_ID_BITS);
        }
        if (hist) {
            // pass 0: neighbouring lanes hold neighbouring tiles, different bins.  Higher passes: runs of equal
            // digits (a Gaussian's row of tiles), one add per run by its first lane.
            const uint32_t dmask = (1u << wbits) - 1u;
            if (k < warp_total) atomicAdd(&s_hist[key & dmask], 1u);
            for (int p = 1; p < npass; p++) {
                const uint32_t d = key >> (wbits * p);   // 0xffffff.. for idle lanes: never equal to a live digit
                const uint32_t up = __shfl_up_sync(0xffffffffu, d, 1);
                const bool head = lane == 0 || d != up;
                const uint32_t heads = __ballot_sync(0xffffffffu, head);
                if (head && k < warp_total) {
                    const uint32_t after = heads & ~((2u << lane) - 1u); // heads strictly above this lane
                    const uint32_t next = after ? (uint32_t)__ffs(after) - 1u : 32u;
                    atomicAdd(&s_hist[p * 256 + (d & dmask)], next - (uint32_t)lane);
                }
            }
        }
    }
    if (hist) {
        __syncthreads();
        for (int j = tid; j < npass * 256; j += WS_PRE_THREADS) {
            const uint32_t c = s_hist[j];
            if (c) atomicAdd(hist + j, c);
        }
    }
}

int launch_emit_tiles_ordered(int P, int W, int H, const int32_t *radii, GeomState &gs, uint32_t *tkeys,
                              uint32_t *vals, cudaStream_t st, bool debug, uint32_t *hist, int npass, int wbits,
                              uint32_t *zero_a, size_t zero_a_words, uint32_t *zero_b, size_t zero_b_words)
{
    if (npass > 4) hist = nullptr; // (more than 2^32 tiles: not reachable; the sort then counts for itself)
    const int gx = (W + WS_BLOCK_X - 1) / WS_BLOCK_X, gy = (H + WS_BLOCK_Y - 1) / WS_BLOCK_Y;
    const int nb = (P + WS_PRE_THREADS - 1) / WS_PRE_THREADS;
    const uint32_t *order = depth_order(gs);
    WS_CUDA_CHECK(launch_pdl(ordered_block_sums, dim3(nb), dim3(WS_PRE_THREADS), 0, st, P, order, gs.tiles_touched,
                             gs.block_sums, zero_a, zero_a_words, zero_b, zero_b_words));
    WS_LAUNCH_CHECK(st, debug);
    WS_CUDA_CHECK(launch_pdl(scan_block_sums, dim3(1), dim3(1024), 0, st, nb, gs.block_sums, gs.block_offsets,
                             gs.total + 1));
    WS_LAUNCH_CHECK(st, debug);
    WS_CUDA_CHECK(launch_pdl(emit_tiles_ordered, dim3(nb), dim3(WS_PRE_THREADS), 0, st, P, gx, gy, order, radii, gs.grec,
                             gs.tiles_touched, gs.block_offsets, tkeys, vals, hist, npass, wbits));
    WS_LAUNCH_CHECK(st, debug);
    return WS_OK;
}

// ---- markVisible ----
__global__ void mark_visible_kernel(int P, const float *__restrict__ means3D, const float *__restrict__ view,
                                    uint8_t *__restrict__ present)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float x = means3D[3 * (size_t)idx], y = means3D[3 * (size_t)idx + 1], z = means3D[3 * (size_t)idx + 2];
    const float pvz = __fadd_rn(dot3_nv(x, view[2], y, view[6], z, view[10]), view[14]);
    present[idx] = (pvz <= 0.2f) ? 0 : 1;
}

} // namespace ws

extern "C" int ws_mark_visible(int32_t P, const float *means3D, const float *viewmatrix,
                               const float *projmatrix, uint8_t *present, ws_stream_t stream)
{
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) {
        ws::set_error("ws_mark_visible: bad arguments");
        return WS_E_INVALID;
    }
    if (P == 0) return WS_OK;
    ws::mark_visible_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, means3D, viewmatrix, present);
    WS_LAUNCH_CHECK((cudaStream_t)stream, false);
    return WS_OK;
}
