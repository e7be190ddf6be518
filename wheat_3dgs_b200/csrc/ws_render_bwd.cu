// ws_render_bwd.cu -- backward of the alpha blending (replaces renderCUDA backward,
// DGR backward.cu:415-601; flashsplat gate FSR backward.cu:455-462,529-536).
//
// Same warp / lane mapping (warp = 8x8 pixels, two pixels per lane, packed
// f32x2 arithmetic), record staging (quadrant-mask compaction of the tile's value list, TMA gather4 copies of
// the surviving Gaussians' rows, mbarrier double buffer) and sub-tile culling
// as the forward (ws_render_fwd.cu), walking the tile's records back to front from the largest
// last-contributor position in the tile (records behind it are never fetched; the reference
// walks them with `contributor >= last_contributor -> continue`).  A CTA is one half tile (two quadrants): an
// instance is fetched only if its mask touches one of them.  alpha is recomputed with the
// forward's exact operations, so the set of contributing pairs is the forward's.
//
// Per-pixel recurrence.  The reference keeps, per pixel, the colour / depth / alpha accumulated
// behind the current record (accum_rec, last_color, last_alpha ..., backward.cu:538-575) and
// forms dL_dalpha = T * sum_ch (c_ch - accum_rec_ch) dL_ch + ... .  With
//   dot_i = c_i . dL_dC + depth_i dL_dD + dL_dA   and   R_i = sum_{j behind i} alpha_j T_j dot_j
// the same quantity is  dL_dalpha_i = T_i dot_i - (R_i + T_final bg . dL_dC) / (1 - alpha_i):
// one scalar R per pixel instead of ten, accumulated back to front from small to large terms (no
// cancellation).  T_i is recovered as in the reference, T_i = T_{i+1} / (1 - alpha_i).
//
// Gradient accumulation.  The reference issues 10 global float atomics per contributing
// (pixel, Gaussian) pair (backward.cu:555,563,589-598).  Here the 64 pixels of a warp look at
// the same record at the same time: a lane adds its two pixels, the 10 partial sums are reduced
// across the warp with transposing shuffle butterflies (an 8-slot tree, 9 shuffles, and a 2-slot
// tree, 5 shuffles) and 10 lanes issue ONE red.global.add.f32 each into a 48-byte per-Gaussian
// accumulator row acc[g][12] (2 sectors).  (A shared-memory transpose was tried instead of the
// shuffles: fewer instructions, but the SM-wide shared-memory pipe made it 5 % slower.)  Float
// summation order differs from the reference's (non-deterministic) atomic order: gradients are
// in the 1e-3 relative tier.
//
// What is summed.  With g = G * dL_dalpha of a pair, every geometric gradient of the Gaussian is
// a per-Gaussian constant times a pixel moment of g:  dL_dopacity = S g,
// dL_dconic = -0.5 o (S g dx^2, S g dx dy, S g dy^2),  dL_dmean2D = -0.5 o (W (A S g dx + B S g dy),
// H (C S g dy + B S g dx))  (backward.cu:583-598 with the constants factored out of the sums), so
// the kernel accumulates {dL_dcolor rgb, dL_ddepth, S g dx, S g dy, S g dx^2, S g dx dy, S g dy^2, S g}
// and preprocess_bwd applies the constants once per Gaussian.
#include "ws_stage.cuh"

namespace ws {

#ifndef WS_RB_WARPS
#define WS_RB_WARPS 2
#endif
#ifndef WS_RB_NSTAGE
#define WS_RB_NSTAGE 3
#endif
#ifndef WS_RB_MINB
#define WS_RB_MINB 1
#endif
constexpr int RB_WARPS = WS_RB_WARPS;   // 8x8 pixel blocks (warps) per CTA: 2 = a 16x8 half tile, 4 = the tile
constexpr int RB_THREADS = 32 * RB_WARPS;
constexpr int RB_NSTAGE = WS_RB_NSTAGE;
using BwdRing = StageRing<RB_WARPS, RB_NSTAGE>;

struct RenderBwdParams {
    const uint2 *ranges;
    const uint32_t *vals; // sorted values: Gaussian id | quadrant mask << 28
    int W, H, tiles_x;
    const float *bg;
    const float *out_alpha;
    const uint32_t *n_contrib;
    const float *dL_dcolor, *dL_ddepth, *dL_dalpha;
    float *acc; // [P][12]
    const float *gt_mask;
    const int32_t *unique_label;
};

template <bool FLASH>
__global__ void __launch_bounds__(RB_THREADS, WS_RB_MINB) render_bwd(const __grid_constant__ TensorMap tmap, const RenderBwdParams p)
{
    __shared__ __align__(128) BwdRing sm;

    __shared__ uint32_t s_wmax[RB_WARPS];
    const int lane = threadIdx.x & 31;
    const uint32_t wl = threadIdx.x >> 5;     // warp within the CTA
    // the CTAs of a tile are adjacent in launch order
    const uint32_t tile = blockIdx.x / (4 / RB_WARPS);
    const uint32_t w = (blockIdx.x % (4 / RB_WARPS)) * RB_WARPS + wl; // quadrant of the tile
    const uint32_t cta_quads = ((1u << RB_WARPS) - 1u) << ((blockIdx.x % (4 / RB_WARPS)) * RB_WARPS);
    const int tile_y = (int)(tile / (uint32_t)p.tiles_x), tile_x = (int)(tile - (uint32_t)tile_y * p.tiles_x);
    const int wx0 = tile_x * WS_BLOCK_X + (int)(w & 1) * 8;
    const int wy0 = tile_y * WS_BLOCK_Y + (int)(w >> 1) * 8;
    const int px = wx0 + (lane & 7), py0 = wy0 + (lane >> 3), py1 = py0 + 4;
    const bool in0 = px < p.W && py0 < p.H, in1 = px < p.W && py1 < p.H;
    const uint32_t pix0 = (uint32_t)p.W * py0 + px, pix1 = pix0 + 4u * (uint32_t)p.W;
    const float pixfx = (float)px;
    const f2 npy = pk2(-(float)py0, -(float)py1);
    const size_t HW = (size_t)p.H * p.W;

    const uint2 range = p.ranges[tile];
    // 1-based position of the last blended record (0 outside the image: nothing is active)
    const uint32_t last0 = in0 ? p.n_contrib[pix0] : 0u, last1 = in1 ? p.n_contrib[pix1] : 0u;
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, max(last0, last1));
    if (lane == 0) s_wmax[wl] = wmax;
    sm.init();
    uint32_t bmax = 0;
#pragma unroll
    for (int k = 0; k < RB_WARPS; k++) bmax = max(bmax, s_wmax[k]);
    const int n = (int)min(bmax, range.y - range.x); // records [0, n) matter to this CTA
    const int nb = (n + BwdRing::BATCH - 1) / BwdRing::BATCH;
    const uint32_t *__restrict__ vals = p.vals + range.x;

    // per-pixel constants, DGR backward.cu:461-496 (lo = pixel 0, hi = pixel 1)
    const float Tf0 = in0 ? (1.0f - p.out_alpha[pix0]) : 0.0f, Tf1 = in1 ? (1.0f - p.out_alpha[pix1]) : 0.0f;
    float dl0[5] = {0.f, 0.f, 0.f, 0.f, 0.f}, dl1[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (in0) {
        dl0[0] = p.dL_dcolor[pix0]; dl0[1] = p.dL_dcolor[HW + pix0]; dl0[2] = p.dL_dcolor[2 * HW + pix0];
        dl0[3] = p.dL_ddepth[pix0]; dl0[4] = p.dL_dalpha[pix0];
    }
    if (in1) {
        dl1[0] = p.dL_dcolor[pix1]; dl1[1] = p.dL_dcolor[HW + pix1]; dl1[2] = p.dL_dcolor[2 * HW + pix1];
        dl1[3] = p.dL_ddepth[pix1]; dl1[4] = p.dL_dalpha[pix1];
    }
    const f2 dLr = pk2(dl0[0], dl1[0]), dLg = pk2(dl0[1], dl1[1]), dLb = pk2(dl0[2], dl1[2]),
             dLd = pk2(dl0[3], dl1[3]), dLa = pk2(dl0[4], dl1[4]);
    const float bgr = __ldg(p.bg + 0), bgg = __ldg(p.bg + 1), bgb = __ldg(p.bg + 2);
    // -T_final * (bg . dL_dpixel): the background term of backward.cu:577-580 before its 1/(1-alpha)
    const f2 ntfbg = pk2(-Tf0 * (bgr * dl0[0] + bgg * dl0[1] + bgb * dl0[2]),
                         -Tf1 * (bgr * dl1[0] + bgg * dl1[1] + bgb * dl1[2]));
    int obj0 = 0, obj1 = 0;
    if (FLASH) {
        if (in0 && p.gt_mask) obj0 = __float2int_rz(p.gt_mask[pix0]);
        if (in1 && p.gt_mask) obj1 = __float2int_rz(p.gt_mask[pix1]);
    }
    // loop-carried state as scalar halves (packed at the point of use, see ws_render_fwd.cu):
    // T = transmittance in front of the current record, R = sum over the records behind it of
    // alpha_j T_j (c_j . dL_dC + depth_j dL_dD + dL_dA)
    float Ta = Tf0, Tb = Tf1, Ra = 0.0f, Rb = 0.0f;

    // walking back to front: step `it` blends batch nb-1-it.  Ring prologue: steps 0 .. NSTAGE-2 requested,
    // the value of step NSTAGE-1 in flight
    uint32_t nv;
    {
        uint32_t pv[RB_NSTAGE - 1];
#pragma unroll
        for (int k = 0; k < RB_NSTAGE - 1; k++) pv[k] = BwdRing::load(vals, n, nb, nb - 1 - k, wl, lane);
        nv = BwdRing::load(vals, n, nb, nb - RB_NSTAGE, wl, lane);
#pragma unroll
        for (int k = 0; k < RB_NSTAGE - 1; k++)
            if (k < nb) sm.issue(&tmap, k, nb - 1 - k, pv[k], cta_quads, wl, lane);
    }
    for (int it = 0; it < nb; it++) {
        const int s = it % RB_NSTAGE;
        // slot (it-1) % NSTAGE was released by the barrier at the end of step it-1
        if (it + RB_NSTAGE - 1 < nb)
            sm.issue(&tmap, (it + RB_NSTAGE - 1) % RB_NSTAGE, nb - it - RB_NSTAGE, nv, cta_quads, wl, lane);
        nv = BwdRing::load(vals, n, nb, nb - 1 - it - RB_NSTAGE, wl, lane);
        sm.wait(it);

#pragma unroll 1
        for (int seg = RB_WARPS - 1; seg >= 0; seg--) {
          const int cnt = (int)sm.cnt[s][seg];
          const int sbase = seg * ST_SEG;
          {
            bool hit = false;
            if (lane < cnt) {
                const uint2 mt = sm.meta[s][sbase + lane];
                hit = ((mt.x >> (WS_ID_BITS + w)) & 1u) != 0u && mt.y < wmax;
            }
            uint32_t m = __ballot_sync(0xffffffffu, hit);
            while (m) {
                const int j = 31 - __clz(m);
                m &= ~(1u << j);
                const int rr = sbase + j;
                const uint2 meta = sm.meta[s][rr];
                const uint32_t pos0 = meta.y;
                const float4 a = sm.rec[s][rr].a;
                const float4 q = sm.rec[s][rr].b;
                const float4 c = sm.rec[s][rr].c;
                const uint32_t gid = meta.x & WS_ID_MASK;
                bool k0 = pos0 < last0, k1 = pos0 < last1; // backward.cu:516-518
                if (FLASH) {
                    if (p.unique_label != nullptr) {
                        const int uid = __ldg(p.unique_label + gid);
                        if (uid >= 0 && obj0 != uid) k0 = false;
                        if (uid >= 0 && obj1 != uid) k1 = false;
                    }
                }
                // alpha exactly as the forward computed it (same ops, same order)
                const float dx = __fadd_rn(a.x, -pixfx);
                const f2 dy = add2(bc2(a.y), npy);
                const f2 t_zz = mul2(dy, mul2(dy, bc2(q.x)));
                const float t_xx = __fmul_rn(dx, a.z);
                const f2 nt_xy = mul2(dy, bc2(__fmul_rn(dx, -a.w)));
                const f2 power = fma2(fma2(bc2(dx), bc2(t_xx), t_zz), bc2(-0.5f), nt_xy);
                const f2 G = expf2(power);
                const f2 og = mul2(bc2(q.y), G);
                const float al0 = fminf(0.99f, lo2(og)), al1 = fminf(0.99f, hi2(og));
                k0 = k0 && !(lo2(power) > 0.0f) && !(al0 < 1.0f / 255.0f);
                k1 = k1 && !(hi2(power) > 0.0f) && !(al1 < 1.0f / 255.0f);
                if (__ballot_sync(0xffffffffu, k0 || k1) == 0) continue;
                // a pixel that does not contribute runs along with alpha = 0 and G = 0: every
                // sum below gets an exact zero from it and its T and R stay as they are
                const f2 al = pk2(k0 ? al0 : 0.0f, k1 ? al1 : 0.0f);
                const f2 Gm = pk2(k0 ? lo2(G) : 0.0f, k1 ? hi2(G) : 0.0f);
                const f2 om = fma2(al, bc2(-1.0f), bc2(1.0f));
                const f2 inv = pk2(rcp_approx(lo2(om)), rcp_approx(hi2(om)));
                const f2 T = mul2(pk2(Ta, Tb), inv); // backward.cu:534
                const f2 R = pk2(Ra, Rb);
                const f2 dot = fma2(dLr, bc2(c.x), fma2(dLg, bc2(c.y), fma2(dLb, bc2(c.z), fma2(dLd, bc2(q.z), dLa))));
                // dL_dalpha = T (c.dL) - (R + T_final bg.dL) / (1 - alpha)   (backward.cu:538-580)
                const f2 dopa = fma2(T, dot, mul2(fma2(R, bc2(-1.0f), ntfbg), inv));
                const f2 wgt = mul2(al, T);
                const f2 Rn = fma2(wgt, dot, R);
                Ta = k0 ? lo2(T) : Ta; Tb = k1 ? hi2(T) : Tb;
                Ra = lo2(Rn); Rb = hi2(Rn);
                const f2 cr = mul2(wgt, dLr), cg = mul2(wgt, dLg), cb = mul2(wgt, dLb), cd = mul2(wgt, dLd);
                const f2 g = mul2(Gm, dopa);
                const f2 gx = mul2(g, bc2(dx)), gy = mul2(g, dy);
                const f2 gxx = mul2(gx, bc2(dx)), gxy = mul2(gx, dy), gyy = mul2(gy, dy);
                const float4 va = make_float4(lo2(cr) + hi2(cr), lo2(cg) + hi2(cg), lo2(cb) + hi2(cb), lo2(cd) + hi2(cd));
                const float4 vb = make_float4(lo2(gx) + hi2(gx), lo2(gy) + hi2(gy), lo2(gxx) + hi2(gxx), lo2(gxy) + hi2(gxy));
                const float2 vc = make_float2(lo2(gyy) + hi2(gyy), lo2(g) + hi2(g));
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
        }
        __syncthreads(); // stage s is free to be refilled
    }
}

int launch_render_bwd(const ws_view *v, int P, GeomState &gs, const uint32_t *vals, ImageState &is,
                      const float *out_alpha, const float *dL_dcolor, const float *dL_ddepth,
                      const float *dL_dalpha, const ws_flash *fl, cudaStream_t st)
{
    if ((uint32_t)P > WS_ID_MASK) {
        set_error("render: %d Gaussians exceed the 2^%u - 1 ids a sorted value can hold", P, WS_ID_BITS);
        return WS_E_UNSUPPORTED;
    }
    TensorMap tm;
    int rc = make_grec_tensor_map(&tm, gs.grec, P);
    if (rc != WS_OK) return rc;
    WS_CUDA_CHECK(cudaMemsetAsync(gs.acc, 0, (size_t)P * 12 * sizeof(float), st));
    RenderBwdParams p;
    p.ranges = is.ranges; p.vals = vals;
    p.W = v->image_width; p.H = v->image_height;
    p.bg = v->bg; p.out_alpha = out_alpha; p.n_contrib = is.n_contrib;
    p.dL_dcolor = dL_dcolor; p.dL_ddepth = dL_ddepth; p.dL_dalpha = dL_dalpha;
    p.acc = gs.acc;
    p.gt_mask = fl ? fl->gt_mask : nullptr;
    p.unique_label = fl ? fl->unique_label : nullptr;
    p.tiles_x = (p.W + WS_BLOCK_X - 1) / WS_BLOCK_X;
    const unsigned grid = (unsigned)(4 / RB_WARPS) * (unsigned)p.tiles_x * (unsigned)((p.H + WS_BLOCK_Y - 1) / WS_BLOCK_Y);
    if (fl && fl->unique_label)
        render_bwd<true><<<grid, RB_THREADS, 0, st>>>(tm, p);
    else
        render_bwd<false><<<grid, RB_THREADS, 0, st>>>(tm, p);
    WS_LAUNCH_CHECK(st, v->debug);
    return WS_OK;
}

} // namespace ws
