// ws_render_bwd.cu -- backward of the alpha blending (replaces renderCUDA backward,
// DGR backward.cu:415-601; flashsplat gate FSR backward.cu:455-462,529-536).
//
// Same tile / warp / lane mapping, record staging (TMA bulk copies, mbarrier double buffer) and
// sub-tile culling as the forward (ws_render_fwd.cu), walking the tile's records back to front
// from the largest last-contributor position in the tile (records behind it are never
// fetched; the reference walks them with `contributor >= last_contributor -> continue`).
//
// Gradient accumulation.  The reference issues 10 global float atomics per contributing
// (pixel, Gaussian) pair (backward.cu:555,563,589-598).  Here the 32 pixels of a warp look at
// the same record at the same time, so the 10 partial sums are first reduced across the warp
// with transposing shuffle butterflies (an 8-slot tree, 9 shuffles, and a 2-slot tree, 5
// shuffles: 44 instructions instead of the 62 of one 16-slot tree) and then 10 lanes issue ONE
// red.global.add.f32 each into a 48-byte per-Gaussian accumulator row acc[g][12] (2 sectors).
// (A shared-memory transpose was tried instead of the shuffles: fewer instructions, but the
// SM-wide shared-memory pipe made it 5 % slower.)  Float summation order differs from the
// reference's (non-deterministic) atomic order: gradients are in the 1e-3 relative tier.
//
// What is summed.  With g = G * dL_dalpha of a pair, every geometric gradient of the Gaussian is
// a per-Gaussian constant times a pixel moment of g:  dL_dopacity = S g,
// dL_dconic = -0.5 o (S g dx^2, S g dx dy, S g dy^2),  dL_dmean2D = -0.5 o (W (A S g dx + B S g dy),
// H (C S g dy + B S g dx))  (backward.cu:583-598 with the constants factored out of the sums), so
// the kernel accumulates {dL_dcolor rgb, dL_ddepth, S g dx, S g dy, S g dx^2, S g dx dy, S g dy^2, S g}
// and preprocess_bwd applies the constants once per Gaussian.  The colour/depth/alpha
// recurrences (backward.cu:538-575) are written as x += last_alpha * (y - x).
#include "ws_common.cuh"

namespace ws {

constexpr int RB_BATCH = 256;

struct RenderBwdParams {
    const uint2 *ranges;
    const float4 *q0, *q1, *q2;
    int W, H;
    const float *bg;
    const float *out_alpha;
    const uint32_t *n_contrib;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float *acc; // [P][12]
    const float *gt_mask;
    const int32_t *unique_label;
};

template <bool FLASH>
__global__ void __launch_bounds__(256) render_bwd(const RenderBwdParams p)
{
    __shared__ __align__(128) float4 s_q0[2][RB_BATCH];
    __shared__ __align__(128) float4 s_q1[2][RB_BATCH];
    __shared__ __align__(128) float4 s_q2[2][RB_BATCH];
    __shared__ __align__(8) uint64_t s_full[2];
    __shared__ uint32_t s_wmax[8];

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
    const int wx0 = blockIdx.x * WS_BLOCK_X + (w & 1) * 8;
    const int wy0 = blockIdx.y * WS_BLOCK_Y + (w >> 1) * 4;
    const int px = wx0 + (lane & 7), py = wy0 + (lane >> 3);
    const bool inside = px < p.W && py < p.H;
    const uint32_t pix_id = (uint32_t)p.W * py + px;
    const float pixfx = (float)px, pixfy = (float)py;
    const float fx0 = (float)wx0, fy0 = (float)wy0;
    const size_t HW = (size_t)p.H * p.W;

    const uint2 range = p.ranges[tile];
    const uint32_t last = inside ? p.n_contrib[pix_id] : 0u; // 1-based position of last blended record
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0) s_wmax[w] = wmax;
    if (tid == 0) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t bmax = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) bmax = max(bmax, s_wmax[k]);
    const int n = (int)min(bmax, range.y - range.x); // records [0, n) of the tile matter
    const int nb = (n + RB_BATCH - 1) / RB_BATCH;

    auto issue = [&](int b, int s) {
        const int cnt = min(RB_BATCH, n - b * RB_BATCH);
        const uint32_t bytes = (uint32_t)cnt * 16u;
        const size_t off = (size_t)range.x + (size_t)b * RB_BATCH;
        mbar_expect_tx(&s_full[s], 3u * bytes);
        tma_load_1d(&s_q0[s][0], p.q0 + off, bytes, &s_full[s]);
        tma_load_1d(&s_q1[s][0], p.q1 + off, bytes, &s_full[s]);
        tma_load_1d(&s_q2[s][0], p.q2 + off, bytes, &s_full[s]);
    };

    // per-pixel state, DGR backward.cu:461-496
    const float T_final = inside ? (1.0f - p.out_alpha[pix_id]) : 0.0f;
    float T = T_final;
    float dLr = 0.f, dLg = 0.f, dLb = 0.f, dLd = 0.f, dLa = 0.f;
    if (inside) {
        dLr = p.dL_dcolor[pix_id];
        dLg = p.dL_dcolor[HW + pix_id];
        dLb = p.dL_dcolor[2 * HW + pix_id];
        dLd = p.dL_ddepth[pix_id];
        dLa = p.dL_dalpha[pix_id];
    }
    float bgdot = 0.0f;
    bgdot += __ldg(p.bg + 0) * dLr;
    bgdot += __ldg(p.bg + 1) * dLg;
    bgdot += __ldg(p.bg + 2) * dLb;
    const float nTf_bg = -T_final * bgdot; // background term of backward.cu:577-580, times 1/(1-alpha)
    int obj_id = 0;
    if (FLASH) {
        if (inside && p.gt_mask) obj_id = __float2int_rz(p.gt_mask[pix_id]);
    }
    // running "what lies behind" sums, backward.cu:538-575: x <- last_alpha*last_x + (1-last_alpha)*x
    float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, last_depth = 0.f;
    float ac0 = 0.f, ac1 = 0.f, ac2 = 0.f, acd = 0.f, aca = 0.f;

    if (nb > 0 && tid == 0) issue(nb - 1, 0);
    for (int it = 0; it < nb; it++) {
        const int b = nb - 1 - it;
        const int s = it & 1;
        const int cnt = min(RB_BATCH, n - b * RB_BATCH);
        if (it + 1 < nb && tid == 0) issue(b - 1, (it + 1) & 1);
        mbar_wait(&s_full[s], (uint32_t)((it >> 1) & 1));

        for (int c0 = ((cnt - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
            const int r = c0 + lane;
            bool hit = false;
            if (r < cnt && (uint32_t)(b * RB_BATCH + r) < wmax)
                hit = rect_hit(s_q0[s][r], s_q1[s][r], fx0, fy0, 7.0f, 3.0f);
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int j = 31 - __clz(m);
                m &= ~(1u << j);
                const int rr = c0 + j;
                const uint32_t pos0 = (uint32_t)(b * RB_BATCH + rr);
                const float4 a = s_q0[s][rr];
                const float4 q = s_q1[s][rr];
                const float4 c = s_q2[s][rr];
                const uint32_t gid = __float_as_uint(c.w);
                float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = make_float4(0.f, 0.f, 0.f, 0.f);
                float2 vc = make_float2(0.f, 0.f);
                bool act = inside && pos0 < last; // backward.cu:516-518
                if (FLASH) {
                    if (p.unique_label != nullptr) {
                        const int uid = __ldg(p.unique_label + gid);
                        if (uid >= 0 && obj_id != uid) act = false;
                    }
                }
                bool contributed = false;
                if (act) {
                    const float dx = __fadd_rn(a.x, -pixfx), dy = __fadd_rn(a.y, -pixfy);
                    const float t_zz = __fmul_rn(dy, __fmul_rn(dy, q.x));
                    const float t_xx = __fmul_rn(dx, a.z);
                    const float t_xy = __fmul_rn(dy, __fmul_rn(dx, a.w));
                    const float power = __fmaf_rn(__fmaf_rn(dx, t_xx, t_zz), -0.5f, -t_xy);
                    if (!(power > 0.0f)) {
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, __fmul_rn(q.y, G));
                        if (!(alpha < 1.0f / 255.0f)) {
                            contributed = true;
                            const float inv = rcp_approx(1.0f - alpha);
                            T = T * inv; // backward.cu:534
                            const float wgt = alpha * T;
                            ac0 += last_alpha * (lc0 - ac0); lc0 = c.x;
                            ac1 += last_alpha * (lc1 - ac1); lc1 = c.y;
                            ac2 += last_alpha * (lc2 - ac2); lc2 = c.z;
                            acd += last_alpha * (last_depth - acd); last_depth = q.z;
                            aca += last_alpha * (1.0f - aca);
                            last_alpha = alpha;
                            float dL_dopa = (c.x - ac0) * dLr;
                            dL_dopa += (c.y - ac1) * dLg;
                            dL_dopa += (c.z - ac2) * dLb;
                            dL_dopa += (q.z - acd) * dLd;
                            dL_dopa += (1.0f - aca) * dLa;
                            dL_dopa = dL_dopa * T + nTf_bg * inv;
                            va = make_float4(wgt * dLr, wgt * dLg, wgt * dLb, wgt * dLd);
                            const float g = G * dL_dopa;
                            const float gx = g * dx, gy = g * dy;
                            vb = make_float4(gx, gy, gx * dx, gx * dy);
                            vc = make_float2(gy * dy, g);
                        }
                    }
                }
                if (__ballot_sync(0xffffffffu, contributed) == 0) continue;
                {
                    // transposing butterfly over 8 slots (lane l ends with slot l>>2 summed over the warp)
                    float v[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
                    float e0 = vc.x, e1 = vc.y; // and over the last 2 slots
                    {
                        const bool hi = (lane & 16) != 0;
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const float send = hi ? v[i] : v[i + 4];
                            const float keep = hi ? v[i + 4] : v[i];
                            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                        }
                        const float send = hi ? e0 : e1;
                        const float keep = hi ? e1 : e0;
                        e0 = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                    }
                    {
                        const bool hi = (lane & 8) != 0;
#pragma unroll
                        for (int i = 0; i < 2; i++) {
                            const float send = hi ? v[i] : v[i + 2];
                            const float keep = hi ? v[i + 2] : v[i];
                            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                        }
                        e0 += __shfl_xor_sync(0xffffffffu, e0, 8);
                    }
                    {
                        const bool hi = (lane & 4) != 0;
                        const float send = hi ? v[0] : v[1];
                        const float keep = hi ? v[1] : v[0];
                        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                        e0 += __shfl_xor_sync(0xffffffffu, e0, 4);
                    }
                    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
                    e0 += __shfl_xor_sync(0xffffffffu, e0, 2);
                    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
                    e0 += __shfl_xor_sync(0xffffffffu, e0, 1);
                    // lane l, l % 4 == 0: slot ((l>>4)&1)*4 + ((l>>3)&1)*2 + ((l>>2)&1) of v; lanes 1 / 17: slots 8 / 9
                    const int slot8 = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
                    if ((lane & 3) == 0) atomicAdd(p.acc + (size_t)gid * 12 + slot8, v[0]);
                    else if ((lane & 15) == 1) atomicAdd(p.acc + (size_t)gid * 12 + 8 + (lane >> 4), e0);
                }
            }
        }
        __syncthreads(); // everyone is done with stage s before it is refilled
    }
}

int launch_render_bwd(const ws_view *v, int P, GeomState &gs, BinningState &bs, ImageState &is,
                      const float *out_alpha, const float *dL_dcolor, const float *dL_ddepth,
                      const float *dL_dalpha, const ws_flash *fl, cudaStream_t st)
{
    WS_CUDA_CHECK(cudaMemsetAsync(gs.acc, 0, (size_t)P * 12 * sizeof(float), st));
    RenderBwdParams p;
    p.ranges = is.ranges; p.q0 = bs.q0; p.q1 = bs.q1; p.q2 = bs.q2;
    p.W = v->image_width; p.H = v->image_height;
    p.bg = v->bg; p.out_alpha = out_alpha; p.n_contrib = is.n_contrib;
    p.dL_dcolor = dL_dcolor; p.dL_ddepth = dL_ddepth; p.dL_dalpha = dL_dalpha;
    p.acc = gs.acc;
    p.gt_mask = fl ? fl->gt_mask : nullptr;
    p.unique_label = fl ? fl->unique_label : nullptr;
    dim3 grid((p.W + WS_BLOCK_X - 1) / WS_BLOCK_X, (p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (fl && fl->unique_label)
        render_bwd<true><<<grid, 256, 0, st>>>(p);
    else
        render_bwd<false><<<grid, 256, 0, st>>>(p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
